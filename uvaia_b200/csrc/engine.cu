// Host side of the C ABI (include/uvaia_b200.h): context, query preparation, database residency, round pipeline.
// Replaces the orchestration of src/nearest.c:233-338 (queue_t, batch loop, two OpenMP loops) around the kernels
// of encode.cu (K1), paircount.cu (K2) and replay.cu (K3).  No scoring happens on the host: without a CUDA device
// uvg_create() fails (there is no CPU fallback by design).
//
// Round pipeline.  References are scored in rounds, in file order.  Round j runs
//     stream A:  K2 pair counts of this rank's tile  --> stores every pair's counts into the score tile of the rank that
//                                                        OWNS the query (query tile qt belongs to rank qt % world; peer
//                                                        memory over NVLink), buffer j % 2
//     stream B:  [NCCL all-gather of the owners' early-abort bounds = the cross-rank barrier]  -->  K3 exact replay of
//                the owned queries over the whole round (all ranks' segments, in file order)
// K2 of round j+1 overlaps K3 of round j; K2 of round j+2 waits until every rank has replayed round j (buffer reuse).
#include "../../include/uvaia_b200.h"
#include "common.cuh"
#include "kernels.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include <dlfcn.h>
#include <unistd.h>

using namespace uvg;

namespace {
std::string g_create_error;
constexpr int NBUF = 2;
constexpr int FIN_DEVICE_MAX_K = 2048;            // heaps up to this many slots get their final sort on the device (k_finalise), larger ones on the host

// ---- NCCL through dlopen (the library loads without NCCL for single-GPU use) ------------------------------------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclInt8 = 0, ncclUint8 = 1, ncclInt32 = 2, ncclUint32 = 3, ncclInt64 = 4, ncclUint64 = 5 };
enum { ncclSum = 0, ncclProd = 1, ncclMax = 2, ncclMin = 3 };
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Bcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;      // optional
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool load(std::string& err)
    {
        if (handle) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so", nullptr};
        for (int i = 0; names[i] && !handle; ++i) handle = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
        if (!handle) { err = std::string("cannot load NCCL: ") + dlerror(); return false; }
        GetUniqueId = (decltype(GetUniqueId))dlsym(handle, "ncclGetUniqueId");
        CommInitRank = (decltype(CommInitRank))dlsym(handle, "ncclCommInitRank");
        AllGather = (decltype(AllGather))dlsym(handle, "ncclAllGather");
        AllReduce = (decltype(AllReduce))dlsym(handle, "ncclAllReduce");
        Bcast = (decltype(Bcast))dlsym(handle, "ncclBroadcast");
        CommDestroy = (decltype(CommDestroy))dlsym(handle, "ncclCommDestroy");
        GetErrorString = (decltype(GetErrorString))dlsym(handle, "ncclGetErrorString");
        if (!GetUniqueId || !CommInitRank || !AllGather || !AllReduce || !CommDestroy) { err = "NCCL symbols missing"; handle = nullptr; return false; }
        return true;
    }
};
NcclApi g_nccl;
}  // namespace

struct uvg_ctx {
    uvg_params prm{};
    std::string err;
    int K = 0, nchar = 0, trim = 0, W = 0, Wp = 0, n_chunks = 0;
    int rank = 0, world = 1;
    cudaStream_t st = nullptr, st_rp = nullptr, st_push = nullptr, st_copy = nullptr;
    cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_c = nullptr, ev_d = nullptr;
    cudaEvent_t ev_copied[2]{}, ev_stage_free[2]{};  // host->device staging ring of uvg_push_refs
    bool search_open = false;                         // rounds enqueued by uvg_search_async and not yet waited for
    size_t n_ev = 0;
    cudaEvent_t ev_pair[NBUF]{}, ev_replay[NBUF]{}, ev_ag[NBUF]{};
    // queries
    int n_query = 0, n_slots = 0, n_qtiles = 0, rows_cap = 0, qnib_stride = 0;
    int n_ptiles = 0;                                 // pivot form: thread-block tiles of PIVOT_SLOTS = 60 slots
    uvg_query_info qinfo{};
    int64_t qinfo_pivot_fix = -1;
    double set_queries_host_ms = 0, set_queries_device_ms = 0;                     // fix entries of the pivot form (-1: general kernel)
    size_t qcap[24]{};                                // bytes of the query-side device arrays, kept across query sets (pooled_alloc)
    QChunk* d_qdata = nullptr;
    uint32_t *d_qnib = nullptr, *d_qamb_off = nullptr, *d_qamb_ent = nullptr;
    // pivot form of the pair-count kernel (pairpivot.cu): chosen per query set by uvg_set_queries
    bool pivot_on = false;
    uint8_t* d_qstream = nullptr; uint32_t* d_qoff = nullptr;      // query stream of the pivot form: records of every (tile, chunk) + their offsets
    uint32_t *d_snp_off = nullptr, *d_snp_ent = nullptr;           // fix entries per slot
    uint4* d_pivot = nullptr;                         // [Wp/4][2]: b0, b1 planes of the pivot sequence
    uint8_t* d_wslot = nullptr;                       // [n_qtiles][16 warps][4]: slot of the tile each warp of k_pairpivot counts
    uint8_t* d_db2 = nullptr; int64_t db2_cap = 0;    // derived database (base, mc, one-hot planes: 3 KB records) of the resident references
    int32_t* d_ref_cnt = nullptr;                     // [2][db2_cap]: per reference popcount of its base plane / of its mc plane
    Threshold* d_thr_local = nullptr;                 // [rows_cap] bounds of the owned rows (written by K3)
    Threshold* d_thr_all[NBUF]{};                     // [world * rows_cap] what K2 reads (double buffered)
    // results of the owned rows in ONE block: heaps [rows_cap][K+1] | heap_n [rows_cap] | max_inc [rows_cap] (own_bytes), so
    // that a sharded search ends with a single all-gather into d_all [world][own_bytes]
    uint8_t* d_own = nullptr; uint8_t* d_all = nullptr; size_t own_bytes = 0, own_off_n = 0, own_off_mi = 0;
    // final tables of the owned rows (k_finalise), the block that travels: uvg_item [rows_cap][K] | heap_n | max_inc (fin_bytes).
    // Heaps beyond FIN_DEVICE_MAX_K slots are sorted on the host instead (fin_device false: the raw block travels).
    uint8_t* d_fin = nullptr; size_t fin_bytes = 0, fin_off_n = 0, fin_off_mi = 0; bool fin_device = false;
    HeapItem* d_heaps = nullptr;                      // = d_own
    int32_t *d_heap_n = nullptr, *d_max_inc = nullptr;    // inside d_own
    int32_t *d_batch_id = nullptr, *d_max_inc_batch = nullptr;
    bool ever_reduced = true;                         // sharded: the ever-best flags of all ranks have been combined (lazily, uvg_get_ever_best)
    // database (local shard)
    int64_t db_cap = 0, db_n = 0, db_scored = 0, global_base = 0, global_total = -1;
    int64_t search_limit = -1;                        // sharded, pipelined hosts: only rounds that end at or below this ordinal are searched yet
    uint4* d_db = nullptr;
    int32_t* d_non_n = nullptr;
    uint32_t* d_ramb_off = nullptr; uint32_t* d_ramb_ent = nullptr; int64_t amb_cap = 0, amb_n = 0;
    uint8_t* d_ever = nullptr; int64_t ever_cap = 0;
    // dense score tiles of a round: NBUF x [rows_cap + 2][world * seg_cap] x (4 counts)
    int64_t seg_cap = 0;                              // references per rank and round the tiles can hold
    uint4* d_tiles = nullptr; size_t tile_elems = 0;  // one buffer = tile_elems uint4
    size_t tiles_alloc = 0;                           // uint4 elements allocated (single GPU: kept across query sets)
    uint4* peer_tiles[MAX_RANKS]{};                   // base of d_tiles on every rank (peer-mapped)
    bool peer_ipc[MAX_RANKS]{};
    // sharded literal replay: packed planes and ambiguity lists of every rank (peer-mapped anew by every uvg_search: they move when they grow)
    const uint4* peer_db[MAX_RANKS]{}; const uint32_t* peer_ramb_off[MAX_RANKS]{}; const uint32_t* peer_ramb_ent[MAX_RANKS]{};
    bool peer_db_ipc[MAX_RANKS][3]{};
    bool peer_db_live = false;                        // other ranks hold mappings of d_db / d_ramb_*: growing them defers the free
    std::vector<void*> deferred_free;                       // mapped through a CUDA IPC handle (another process): closed with cudaIpcCloseMemHandle
    bool peers_ready = false;
    unsigned long long* d_counters = nullptr;
    unsigned long long* h_pinned = nullptr;           // small pinned read-back area
    void* d_ipc = nullptr;                            // staging for the handle exchange
    // push staging
    uint8_t* d_stage_text[2]{}; size_t stage_text_bytes = 0;
    int32_t *d_stage_nn = nullptr, *d_stage_namb = nullptr; uint8_t* d_stage_acc = nullptr; int64_t* d_stage_dst = nullptr; int64_t stage_rows = 0;
    int64_t* d_nacc = nullptr;
    // literal emulation of the reference's --pool batches (hazard H1, replay.cu): batch-wide limit of the consensus stage
    bool literal = false;                             // single GPU and not uvg_params.closed_form
    int32_t* d_bstate = nullptr;                      // [0..2] accumulators of the replay launches (ring), [4..5] limit of the batch in progress
    int64_t replay_seq = 0;                           // replay launches since the heaps were emptied
    std::vector<int64_t> file_starts{0};              // global ordinals at which a reference file starts (a batch ends with its file)
    uint32_t cons_amb_begin = 0, cons_amb_n = 0;      // ambiguity codes of the consensus slot inside d_qamb_ent
    int64_t* d_bounds = nullptr; int64_t* h_bounds = nullptr; size_t bounds_cap = 0, bounds_used = 0;   // batch boundaries of the rounds enqueued since the heaps were emptied
    uint8_t* d_snap_own = nullptr; int32_t* d_snap_batch = nullptr;         // speculation of a literal round: state at its start
    uint8_t* d_ever_round = nullptr; unsigned long long* d_spec = nullptr;  //     admissions and counters (admitted, flags) of the speculative pass
    // round schedule
    int64_t rounds_done = 0, round_local_start = 0, round_global_start = 0;
    int32_t next_round = 0;
    // stats
    uvg_stats stats{};
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pair_events;
    // nccl
    ncclComm_t comm = nullptr;
    // host cache of the heaps (filled lazily by uvg_get_heap, invalidated by every search / reset)
    int64_t work_seen = 0, work_last = 0;
    int64_t pivot_halves = 0;                         // pivot form: (slot, 64-site half) items a thread block of every tile counts, summed over the tiles
    int64_t work_host = 0;                            // executed work the host derives from the masks (pairpivot.cu does not count in the kernel)
    bool heaps_cached = false;
    std::vector<HeapItem> h_heaps; std::vector<int32_t> h_heap_n, h_max_inc;
    uint8_t* h_all = nullptr; size_t h_all_bytes = 0;      // pinned landing area of the result blocks
};

#define UVG_FAIL(ctx, code, ...) do { char b_[512]; snprintf(b_, sizeof b_, __VA_ARGS__); (ctx)->err = b_; return (code); } while (0)
#define UVG_CUDA(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    char b_[512]; snprintf(b_, sizeof b_, "%s:%d CUDA error: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); (ctx)->err = b_; \
    return e_ == cudaErrorMemoryAllocation ? UVG_ENOMEM : UVG_ECUDA; } } while (0)
#define UVG_NCCL(ctx, call) do { ncclResult_t r_ = (call); if (r_ != 0) { \
    char b_[512]; snprintf(b_, sizeof b_, "%s:%d NCCL error: %s", __FILE__, __LINE__, g_nccl.GetErrorString ? g_nccl.GetErrorString(r_) : "?"); \
    (ctx)->err = b_; return UVG_ENCCL; } } while (0)

static int32_t round_size_first(const uvg_params& p) { int32_t v = p.first_round > 0 ? p.first_round : 1024; return (v + GROUP - 1) / GROUP * GROUP; }
static int32_t round_size_max(const uvg_params& p)
{
    int32_t v = p.max_round > 0 ? p.max_round : 65536;
    v = (v + GROUP - 1) / GROUP * GROUP;
    return std::max(v, round_size_first(p));
}
static int32_t round_size_next(int32_t cur, int32_t mx) { int64_t n = (int64_t)cur * 4; return (int32_t)std::min<int64_t>(n, mx); }
// The rounds taper off towards the end of the database: the replay of a round runs under the pair counting of the next one,
// except the last replay, which nothing hides -- so once less than two nominal rounds remain, every round takes half of what
// is left (per rank, rounded up to a block's 64 references, not below ROUND_TAPER_MIN) and the last one is small.  In
// sharded mode this also shares the tail evenly between the ranks.  total < 0 = unknown (streaming hosts): nominal size.
constexpr int64_t ROUND_TAPER_MIN = 4096;
static int64_t round_size_at(int64_t S, int64_t T, int64_t total, int world)
{
    if (total < 0 || world < 1) return T;
    const int64_t rem = std::max<int64_t>(0, total - S), per = GROUP * RPL;
    const int64_t share = (rem + world - 1) / world;                          // what every rank still has to score
    if (share >= 2 * T) return T;
    const int64_t up = [&](int64_t v) { return std::max<int64_t>(per, (v + per - 1) / per * per); }(share <= ROUND_TAPER_MIN ? share : std::max<int64_t>(ROUND_TAPER_MIN, (share + 1) / 2));
    return std::min<int64_t>(T, up);
}

extern "C" int uvg_abi_version(void) { return UVG_ABI_VERSION; }

extern "C" const char* uvg_last_error(const uvg_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

extern "C" int uvg_create(const uvg_params* p, uvg_ctx** out)
{
    if (!p || !out) { g_create_error = "null argument"; return UVG_EINVAL; }
    *out = nullptr;
    if (p->abi_version != UVG_ABI_VERSION) { g_create_error = "ABI version mismatch"; return UVG_EINVAL; }
    if (p->nchar < 1 || p->nbest < 1) { g_create_error = "nchar and nbest must be positive"; return UVG_EINVAL; }
    if (p->nchar > (int32_t)AMB_SITE_MASK) { g_create_error = "alignments longer than 2^28 columns are not supported"; return UVG_EINVAL; }
    if (p->world > MAX_RANKS) { g_create_error = "at most 8 ranks (one box)"; return UVG_EINVAL; }
    if (p->world > 1 && (p->rank < 0 || p->rank >= p->world)) { g_create_error = "rank out of range"; return UVG_EINVAL; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev < 1) {
        g_create_error = std::string("no usable CUDA device (") + cudaGetErrorString(e) + "): uvaia_b200 has no CPU fallback";
        return UVG_ENODEV;
    }
    if (p->device < 0 || p->device >= ndev) { g_create_error = "device ordinal out of range"; return UVG_EINVAL; }
    uvg_ctx* c = new uvg_ctx();
    c->prm = *p;
    c->K = std::max(2, p->nbest);                                           // min_heap.c:58
    c->nchar = p->nchar;
    int trim = p->trim < 0 ? 0 : p->trim;                                    // fastaseq.c:708-712
    if (trim > p->nchar / 2.1) trim = (int)(p->nchar / 2.1);
    c->trim = trim;
    c->W = (p->nchar + 31) / 32;
    c->n_chunks = (c->W + WC - 1) / WC;
    c->Wp = c->n_chunks * WC;
    c->rank = p->world > 1 ? p->rank : 0;
    c->world = p->world > 1 ? p->world : 1;
    c->next_round = round_size_first(*p);
    if (c->prm.pool < 1) c->prm.pool = 1;
    c->literal = c->world == 1 && p->closed_form == 0;      // sharded: decided per query set (uvg_set_queries)
#define CK_(call) do { cudaError_t e2_ = (call); if (e2_ != cudaSuccess) { g_create_error = std::string("CUDA: ") + cudaGetErrorString(e2_); delete c; return UVG_ECUDA; } } while (0)
    CK_(cudaSetDevice(p->device));
    cudaDeviceProp prop;
    CK_(cudaGetDeviceProperties(&prop, p->device));
    if (prop.major < 10) { g_create_error = "device is not sm_100 or newer; this library ships sm_100a code only"; delete c; return UVG_ENODEV; }
    int lo = 0, hi = 0;
    CK_(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CK_(cudaStreamCreateWithPriority(&c->st, cudaStreamNonBlocking, lo));
    CK_(cudaStreamCreateWithPriority(&c->st_rp, cudaStreamNonBlocking, hi));       // the short replay / collective work goes first
    CK_(cudaStreamCreateWithPriority(&c->st_push, cudaStreamNonBlocking, hi));     // encoder: short memory-bound kernels
    CK_(cudaStreamCreateWithFlags(&c->st_copy, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
        CK_(cudaEventCreateWithFlags(&c->ev_copied[b], cudaEventDisableTiming));
        CK_(cudaEventCreateWithFlags(&c->ev_stage_free[b], cudaEventDisableTiming));
    }
    CK_(cudaEventCreate(&c->ev_a)); CK_(cudaEventCreate(&c->ev_b)); CK_(cudaEventCreate(&c->ev_c)); CK_(cudaEventCreate(&c->ev_d));
    for (int b = 0; b < NBUF; ++b) {
        CK_(cudaEventCreateWithFlags(&c->ev_pair[b], cudaEventDisableTiming));
        CK_(cudaEventCreateWithFlags(&c->ev_replay[b], cudaEventDisableTiming));
        CK_(cudaEventCreateWithFlags(&c->ev_ag[b], cudaEventDisableTiming));
    }
    CK_(upload_luts());
    CK_(paircount_setup());
    CK_(pairpivot_setup());
    CK_(replay_setup(c->K));
    CK_(cudaMalloc(&c->d_bstate, 32)); CK_(cudaMalloc(&c->d_spec, 16));
    CK_(cudaMalloc(&c->d_counters, 64)); CK_(cudaMalloc(&c->d_nacc, 64)); CK_(cudaMalloc(&c->d_ipc, 256 * MAX_RANKS + 256));
    CK_(cudaMemsetAsync(c->d_counters, 0, 64, c->st)); CK_(cudaStreamSynchronize(c->st));
    CK_(cudaMallocHost(&c->h_pinned, 2048));
#undef CK_
    *out = c;
    return UVG_OK;
}

static void close_peer_db(uvg_ctx* c);
static void close_peers(uvg_ctx* c)
{
    close_peer_db(c);
    for (int r = 0; r < MAX_RANKS; ++r) {
        if (c->peer_tiles[r] && r != c->rank && c->world > 1 && c->peer_ipc[r]) cudaIpcCloseMemHandle(c->peer_tiles[r]);
        c->peer_tiles[r] = nullptr; c->peer_ipc[r] = false;
    }
    c->peers_ready = false;
}

// query-side device arrays and the score tiles (peer-mapped in sharded mode) are kept for the next query set (same sizes
// in a serving loop: no cudaFree / cudaMalloc / IPC unmapping per uvg_set_queries, whose cost varies from microseconds to
// hundreds of milliseconds); pooled_alloc grows them when needed.  uvg_destroy releases everything.
template <typename T>
static cudaError_t pooled_alloc(uvg_ctx* c, int slot, T** p, size_t bytes)
{
    bytes = std::max<size_t>(bytes, 16);
    if (*p && c->qcap[slot] >= bytes) return cudaSuccess;
    cudaFree(*p); *p = nullptr; c->qcap[slot] = 0;
    const cudaError_t e = cudaMalloc(p, bytes);
    if (e == cudaSuccess) c->qcap[slot] = bytes;
    return e;
}

static void free_queries(uvg_ctx* c, bool final = false)
{
    c->pivot_on = false;
    if (!final) { c->seg_cap = 0; c->tile_elems = 0; return; }
    close_peers(c);
    for (size_t& b : c->qcap) b = 0;
    cudaFree(c->d_qdata); cudaFree(c->d_qnib); cudaFree(c->d_qamb_off); cudaFree(c->d_qamb_ent);
    cudaFree(c->d_qstream); cudaFree(c->d_qoff); cudaFree(c->d_snp_off); cudaFree(c->d_snp_ent); cudaFree(c->d_pivot); cudaFree(c->d_wslot); c->d_wslot = nullptr;
    c->d_qstream = nullptr; c->d_qoff = c->d_snp_off = c->d_snp_ent = nullptr; c->d_pivot = nullptr;
    cudaFree(c->d_thr_local); for (int b = 0; b < NBUF; ++b) { cudaFree(c->d_thr_all[b]); c->d_thr_all[b] = nullptr; }
    cudaFree(c->d_own); cudaFree(c->d_all); cudaFree(c->d_fin); c->d_own = c->d_all = c->d_fin = nullptr;
    cudaFree(c->d_snap_own); cudaFree(c->d_snap_batch); cudaFree(c->d_ever_round); c->d_snap_own = nullptr; c->d_snap_batch = nullptr; c->d_ever_round = nullptr;
    cudaFree(c->d_batch_id); cudaFree(c->d_max_inc_batch);
    cudaFree(c->d_tiles); c->d_tiles = nullptr; c->tiles_alloc = 0;
    c->d_qdata = nullptr; c->d_qnib = c->d_qamb_off = c->d_qamb_ent = nullptr; c->d_thr_local = nullptr;
    c->d_heaps = nullptr;
    c->d_heap_n = c->d_max_inc = c->d_batch_id = c->d_max_inc_batch = nullptr;
    c->seg_cap = 0; c->tile_elems = 0;
}

extern "C" void uvg_destroy(uvg_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->prm.device);
    if (c->st) cudaStreamSynchronize(c->st);
    if (c->st_rp) cudaStreamSynchronize(c->st_rp);
    if (c->st_push) cudaStreamSynchronize(c->st_push);
    if (c->st_copy) cudaStreamSynchronize(c->st_copy);
    free_queries(c, true);
    for (void* p : c->deferred_free) cudaFree(p);
    c->deferred_free.clear();
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    cudaFree(c->d_db); cudaFree(c->d_db2); cudaFree(c->d_ref_cnt); cudaFree(c->d_non_n); cudaFree(c->d_ramb_off); cudaFree(c->d_ramb_ent); cudaFree(c->d_ever);
    cudaFree(c->d_counters); cudaFree(c->d_nacc); cudaFree(c->d_ipc); cudaFree(c->d_bstate); cudaFree(c->d_spec);
    cudaFree(c->d_bounds); if (c->h_bounds) cudaFreeHost(c->h_bounds);
    cudaFree(c->d_stage_text[0]); cudaFree(c->d_stage_text[1]); cudaFree(c->d_stage_nn); cudaFree(c->d_stage_namb); cudaFree(c->d_stage_acc); cudaFree(c->d_stage_dst);
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    if (c->h_all) cudaFreeHost(c->h_all);
    for (auto& pr : c->pair_events) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
    if (c->ev_a) { cudaEventDestroy(c->ev_a); cudaEventDestroy(c->ev_b); cudaEventDestroy(c->ev_c); cudaEventDestroy(c->ev_d); }
    for (int b = 0; b < NBUF; ++b) {
        if (c->ev_pair[b]) cudaEventDestroy(c->ev_pair[b]);
        if (c->ev_replay[b]) cudaEventDestroy(c->ev_replay[b]);
        if (c->ev_ag[b]) cudaEventDestroy(c->ev_ag[b]);
    }
    for (int b = 0; b < 2; ++b) {
        if (c->ev_copied[b]) cudaEventDestroy(c->ev_copied[b]);
        if (c->ev_stage_free[b]) cudaEventDestroy(c->ev_stage_free[b]);
    }
    if (c->st) cudaStreamDestroy(c->st);
    if (c->st_rp) cudaStreamDestroy(c->st_rp);
    if (c->st_push) cudaStreamDestroy(c->st_push);
    if (c->st_copy) cudaStreamDestroy(c->st_copy);
    delete c;
}

static int sync_all(uvg_ctx* c)
{
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st_rp));
    UVG_CUDA(c, cudaStreamSynchronize(c->st_push));
    UVG_CUDA(c, cudaStreamSynchronize(c->st_copy));
    return UVG_OK;
}

// derived database of the pivot form: sized like the packed database, allocated when a query set first asks for it
static int ensure_db2(uvg_ctx* c)
{
    if (!c->pivot_on || c->db2_cap >= c->db_cap) return UVG_OK;
    int rc = sync_all(c);
    if (rc) return rc;
    const size_t per_pair = (size_t)c->n_chunks * CHUNK2_BYTES;             // bytes per record set of PRPL groups (capacities are multiples of it)
    const size_t pairs_old = (size_t)(c->db2_cap / (GROUP * PRPL)), pairs_new = (size_t)(c->db_cap / (GROUP * PRPL));
    uint8_t* n = nullptr;
    UVG_CUDA(c, cudaMalloc(&n, pairs_new * per_pair));
    // stream-ordered on c->st and waited for: the context's streams are non-blocking, a memset on the NULL stream is NOT ordered
    // before the derive kernel that follows (r02: a late memset zeroed freshly derived planes, one test in 357, timing dependent)
    if (c->d_db2 && c->db2_cap) UVG_CUDA(c, cudaMemcpyAsync(n, c->d_db2, pairs_old * per_pair, cudaMemcpyDeviceToDevice, c->st));
    UVG_CUDA(c, cudaMemsetAsync(n + pairs_old * per_pair, 0, (pairs_new - pairs_old) * per_pair, c->st));
    // the per-reference plane totals are re-derived with the planes: new array, the resident range is filled by the caller
    // (uvg_set_queries) or was filled by earlier pushes and is copied
    int32_t* cn = nullptr;
    UVG_CUDA(c, cudaMalloc(&cn, (size_t)c->db_cap * 2 * sizeof(int32_t)));
    UVG_CUDA(c, cudaMemsetAsync(cn, 0, (size_t)c->db_cap * 2 * sizeof(int32_t), c->st));
    if (c->d_ref_cnt && c->db2_cap) {
        UVG_CUDA(c, cudaMemcpyAsync(cn, c->d_ref_cnt, (size_t)c->db2_cap * 4, cudaMemcpyDeviceToDevice, c->st));
        UVG_CUDA(c, cudaMemcpyAsync(cn + c->db_cap, c->d_ref_cnt + c->db2_cap, (size_t)c->db2_cap * 4, cudaMemcpyDeviceToDevice, c->st));
    }
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    cudaFree(c->d_db2); cudaFree(c->d_ref_cnt);
    c->d_db2 = n; c->d_ref_cnt = cn; c->db2_cap = c->db_cap;
    return UVG_OK;
}

// host threads for the one-off query preparation (the reference does this serially, fastaseq.c:732-777)
static int host_threads(int64_t useful)
{
    const unsigned hw = std::thread::hardware_concurrency();
    return (int)std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(hw ? hw : 1, 32), useful));
}
template <typename F>
static void parallel_ranges(int n, int T, F&& fn)                  // fn(thread, begin, end) over T contiguous ranges of [0, n)
{
    if (T <= 1) { fn(0, 0, n); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < T; ++t) th.emplace_back([&, t] { fn(t, (int)((int64_t)n * t / T), (int)((int64_t)n * (t + 1) / T)); });
    for (auto& x : th) x.join();
}

// =====================================================================================================================
// queries
// =====================================================================================================================
static int reset_heaps(uvg_ctx* c)
{
    const int rows = c->rows_cap;
    if (c->n_query == 0) return UVG_OK;
    int rc = sync_all(c);
    if (rc) return rc;
    std::vector<int32_t> init(rows, c->nchar);                              // max_incompatible = nchar (nearest.c:375,387)
    UVG_CUDA(c, cudaMemcpyAsync(c->d_max_inc, init.data(), rows * 4, cudaMemcpyHostToDevice, c->st));
    UVG_CUDA(c, cudaMemcpyAsync(c->d_max_inc_batch, init.data(), rows * 4, cudaMemcpyHostToDevice, c->st));
    UVG_CUDA(c, cudaMemsetAsync(c->d_batch_id, 0xff, rows * 4, c->st));
    UVG_CUDA(c, cudaMemsetAsync(c->d_heap_n, 0, rows * 4, c->st));
    UVG_CUDA(c, cudaMemsetAsync(c->d_thr_local, 0, (size_t)rows * sizeof(Threshold), c->st));
    for (int b = 0; b < NBUF; ++b) UVG_CUDA(c, cudaMemsetAsync(c->d_thr_all[b], 0, (size_t)c->world * rows * sizeof(Threshold), c->st));
    UVG_CUDA(c, cudaMemsetAsync(c->d_heaps, 0, (size_t)rows * (c->K + 1) * sizeof(HeapItem), c->st));
    UVG_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 64, c->st));
    {   // every heap is empty: max_incompatible = nchar for all queries, so the first batch's limit is nchar
        const int32_t b0[8] = {0, 0, c->nchar, 0, c->nchar, c->nchar, 0, 0};
        UVG_CUDA(c, cudaMemcpyAsync(c->d_bstate, b0, sizeof b0, cudaMemcpyHostToDevice, c->st));
        c->replay_seq = 0;
        c->bounds_used = 0;                                                 // everything is idle here (sync_all above)
    }
    if (c->d_ever && c->ever_cap) UVG_CUDA(c, cudaMemsetAsync(c->d_ever, 0, c->ever_cap, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));                              // `init` goes out of scope
    c->db_scored = 0; c->heaps_cached = false; c->work_seen = 0; c->work_host = 0; c->search_open = false; c->n_ev = 0;
    c->rounds_done = 0; c->round_local_start = 0; c->round_global_start = 0;
    c->next_round = round_size_first(c->prm);
    uvg_stats z{};
    z.encode_ms = c->stats.encode_ms;
    c->stats = z;
    return UVG_OK;
}

extern "C" int uvg_set_queries(uvg_ctx* c, const uint8_t* text, int32_t n_query, uvg_query_info* info)
{
    if (!c || !text || n_query < 1) { if (c) c->err = "uvg_set_queries: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    int rc = sync_all(c);
    if (rc) return rc;
    const auto t_begin = std::chrono::steady_clock::now();
    free_queries(c);
    const int L = c->nchar, trim = c->trim, Q = n_query;
    const bool acgt = c->prm.acgt_mode != 0;
    uint8_t lut[3][256];
    build_luts(lut);
    const uint8_t* code_lut = lut[acgt ? 1 : 0];
    uint8_t up[256];
    for (int i = 0; i < 256; ++i) up[i] = (uint8_t)((i >= 'a' && i <= 'z') ? i - 32 : i);     // uppercase_string, nexus_common.c:105-111
    uint8_t code_raw[256];                                                  // 4-bit set of a raw (any case) byte
    for (int i = 0; i < 256; ++i) code_raw[i] = code_lut[up[i]];

    // ---- site partition (create_query_indices, src/fastaseq.c:732-777) ----
    // usable byte: --acgt: one of ACGT (is_site_acgt, utils.c:285-289); default: valid text (is_site_valid, utils.c:291-295)
    // Host threads sweep disjoint query ranges (row-major: one pass over the query text); per site and range: the first
    // usable byte, whether any / all queries are usable, whether the usable ones agree, and -- for the pivot form of the
    // pair-count kernel -- how many queries hold each plain base.  The ranges are merged in query order.
    bool pivot_ok = !acgt && (int64_t)c->Wp * 32 <= 65535 && !c->prm.early_abort && c->prm.pair_kernel != UVG_PAIR_GENERAL;
    std::vector<uint8_t> cons(L, 'N');
    std::vector<uint8_t> site_class(L, 0);                                  // 1 = idx, 2 = idx_m, 3 = idx_c
    std::vector<uint8_t> pv(pivot_ok ? L : 0, 1);                           // pivot base per site (one-hot code)
    int n_idx = 0, n_idx_m = 0, n_idx_c = 0;
    // Sharded contexts hold the same queries: once the communicator exists, rank 0 alone prepares them on the host and
    // the arrays reach the other ranks by broadcast over NVLink (8 processes preparing the same planes on the host cores
    // of one box cost 0.15 s per query set).  The other ranks only size their buffers.
    const bool do_prep = !(c->world > 1 && c->comm && g_nccl.Bcast && c->rank != 0);
    const bool bcast = c->world > 1 && c->comm && g_nccl.Bcast;
    if (do_prep) {
        const int T = host_threads(std::min<int64_t>(Q, std::max<int64_t>(1, ((int64_t)256 << 20) / ((int64_t)L * 20))));
        std::vector<uint8_t> first((size_t)T * L, 0), any((size_t)T * L, 0), all((size_t)T * L, 1), same((size_t)T * L, 1);
        std::vector<uint32_t> cnt(pivot_ok ? (size_t)T * L * 4 : 0, 0);                // [thread][base][site]
        // per raw byte: usable (0/1) and, for the pivot counts, the one-hot code of a usable plain base (else 0)
        uint8_t usable_raw[256], onehot_raw[256];
        for (int b = 0; b < 256; ++b) {
            const uint8_t ch = up[b];
            usable_raw[b] = acgt ? (ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T') : (lut[2][ch] != 0);
            const uint32_t code = code_lut[ch];
            onehot_raw[b] = (uint8_t)((usable_raw[b] && code && !(code & (code - 1))) ? code : 0);
        }
        parallel_ranges(Q, T, [&](int t, int qa, int qb) {
            uint8_t *f = &first[(size_t)t * L], *an = &any[(size_t)t * L], *al = &all[(size_t)t * L], *sm = &same[(size_t)t * L];
            uint32_t* cn = pivot_ok ? &cnt[(size_t)t * L * 4] : nullptr;
            for (int q = qa; q < qb; ++q) {
                const uint8_t* row = text + (size_t)q * L;
                for (int i = trim; i < L - trim; ++i) {                                 // branch-free: sequential updates of four byte arrays
                    const uint8_t b = row[i], ch = up[b], u = usable_raw[b], had = an[i], f0 = f[i];
                    al[i] &= u;
                    sm[i] &= (uint8_t)((uint8_t)!(u & had) | (uint8_t)(ch == f0));
                    f[i] = (u & (uint8_t)!had) ? ch : f0;
                    an[i] = had | u;
                }
                if (cn)
                    for (int i = trim; i < L - trim; ++i) {
                        const uint32_t oh = onehot_raw[row[i]];
                        cn[i] += oh & 1u; cn[(size_t)L + i] += (oh >> 1) & 1u; cn[(size_t)2 * L + i] += (oh >> 2) & 1u; cn[(size_t)3 * L + i] += oh >> 3;
                    }
            }
        });
        for (int i = trim; i < L - trim; ++i) {
            uint8_t f0 = 0; bool an = false, al = true, sm = true;
            for (int t = 0; t < T; ++t) {
                const size_t k = (size_t)t * L + i;
                if (!all[k]) al = false;
                if (!any[k]) continue;
                if (!same[k]) sm = false;
                if (!an) { an = true; f0 = first[k]; }
                else if (first[k] != f0) sm = false;
            }
            if (!an) continue;
            if (!sm) { site_class[i] = 1; ++n_idx; cons[i] = '#'; }
            else if (al) { site_class[i] = 3; ++n_idx_c; cons[i] = f0; }
            else { site_class[i] = 2; ++n_idx_m; cons[i] = f0; }
        }
        if (pivot_ok) {
            // the plain base most queries hold (ties: A before C before G before T; a site nobody holds a base at: A)
            for (int i = trim; i < L - trim; ++i) {
                uint32_t tot[4] = {0, 0, 0, 0};
                for (int t = 0; t < T; ++t) for (int k = 0; k < 4; ++k) tot[k] += cnt[((size_t)t * 4 + k) * L + i];
                int best = 0;
                for (int k = 1; k < 4; ++k) if (tot[k] > tot[best]) best = k;
                pv[i] = (uint8_t)(1u << best);
            }
        }
    }
    c->n_query = Q;
    c->n_slots = Q + 1;                                                     // + the consensus pseudo-query (nearest.c:428-433)
    c->n_qtiles = (c->n_slots + QTILE - 1) / QTILE;
    c->rows_cap = (c->n_qtiles + c->world - 1) / c->world * QTILE;
    // Literal emulation of the reference's --pool batches (hazard H1): on one GPU always (unless closed_form); sharded when all
    // queries fit one tile of 64 slots -- the searches in which the hazard actually fires (one or a few queries: every site is
    // a consensus site).  Then rank 0 owns every query, the batch-wide limit is the maximum over ITS rows and nothing has to
    // cross ranks but the packed planes of a truncated reference (peer-mapped, ensure_peer_db).  Larger sharded query sets
    // run the closed form and count the flags.
    c->literal = c->prm.closed_form == 0 && (c->world == 1 || c->n_qtiles == 1);
    c->qnib_stride = (L + 7) / 8;
    c->qinfo.n_query = Q; c->qinfo.nchar = L; c->qinfo.trim = trim;

    // ---- planes, nibble codes and ambiguity lists of every slot (zero outside the trimmed range); for the pivot form
    // also the (base, equals-pivot) planes and, per slot, the bases that differ from the pivot.  One pass, host threads
    // over disjoint slot ranges (every slot owns its words of the chunk arrays). ----
    const size_t n_chunk_total = (size_t)c->n_qtiles * c->n_chunks;
    std::vector<QChunk> qd(n_chunk_total);
    memset(qd.data(), 0, n_chunk_total * sizeof(QChunk));
    std::vector<uint32_t> qnib((size_t)c->n_slots * c->qnib_stride, 0);
    std::vector<std::vector<uint32_t>> amb(c->n_slots), snp(pivot_ok ? c->n_slots : 0);
    struct HostPlanes2 { uint4 pl[W4][QTILE][2]; };                          // host only: [unit][slot][plane 0 = plain base, 1 = pivot base], complemented below
    std::vector<HostPlanes2> qd2;
    if (pivot_ok) { qd2.resize(n_chunk_total); memset(qd2.data(), 0, n_chunk_total * sizeof(HostPlanes2)); }
    if (do_prep) parallel_ranges(c->n_slots, host_threads(c->n_slots), [&](int, int sa, int sb) {
        std::vector<uint32_t> plain_in_chunk(c->n_chunks);
        for (int q = sa; q < sb; ++q) {
            const int qt = q / QTILE, qi = q % QTILE;
            const bool is_cons = q == Q;
            const uint8_t* row = is_cons ? cons.data() : text + (size_t)q * L;
            std::fill(plain_in_chunk.begin(), plain_in_chunk.end(), 0u);
            // word by word: the bits of 32 sites are gathered in registers and stored once
            for (int w = trim >> 5; w <= (L - trim - 1) >> 5 && trim < L - trim; ++w) {
                const int s0 = std::max(w << 5, trim), s1 = std::min((w + 1) << 5, L - trim);
                uint32_t mb = 0, m1 = 0, m2 = 0, mseg = 0, me = 0, n_plain = 0;
                uint32_t nib[4] = {0, 0, 0, 0};
                for (int s = s0; s < s1; ++s) {
                    if (is_cons && site_class[s] != 3) continue;            // the consensus is scored on idx_c only
                    const uint32_t code = code_raw[row[s]];
                    const uint32_t j = (uint32_t)s & 31u, bit = 1u << j;
                    if (acgt && site_class[s] == 1) mseg |= bit;            // segregating-site mask (dist_unique, nearest.c:459-470)
                    nib[j >> 3] |= code << ((j & 7u) * 4u);
                    const uint32_t plain = (uint32_t)(code != 0u) & (uint32_t)((code & (code - 1u)) == 0u);
                    mb |= plain << j;
                    m1 |= (plain & (uint32_t)((code & 10u) != 0u)) << j;    // C or T
                    m2 |= (plain & (uint32_t)((code & 12u) != 0u)) << j;    // G or T
                    n_plain += plain;
                    if (code & (code - 1u)) { amb[q].push_back((uint32_t)s | (code << 28)); continue; }
                    if (pivot_ok && plain) {
                        // the planes treat EVERY plain base as the pivot's; a base that differs is settled by its fix entry
                        // (one reference that holds the pivot's base less, one that holds the query's base more), so a
                        // SNP alone does not make the kernel count its half from the planes
                        me |= bit;
                        if (code != pv[s]) snp[q].push_back((uint32_t)s | (code << 28));
                    }
                }
                const int k = w & 3;
                const size_t ci = (size_t)qt * c->n_chunks + w / WC;
                uint4* pl = qd[ci].pl[(w % WC) / 4][qi];
                (&pl[0].x)[k] = mb; (&pl[1].x)[k] = m1; (&pl[2].x)[k] = m2;
                if (acgt) (&pl[3].x)[k] = mseg;
                uint32_t* qn = &qnib[(size_t)q * c->qnib_stride];
                for (int i = 0; i < 4; ++i) if (w * 4 + i < c->qnib_stride) qn[w * 4 + i] = nib[i];
                plain_in_chunk[w / WC] += n_plain;
                if (pivot_ok) {
                    uint4* p2 = qd2[ci].pl[(w % WC) / 4][qi];
                    (&p2[0].x)[k] = mb; (&p2[1].x)[k] = me;
                }
            }
            uint32_t run = 0;
            for (int k = c->n_chunks - 1; k >= 0; --k) {
                qd[(size_t)qt * c->n_chunks + k].rem[qi] = run;
                run += plain_in_chunk[k];
            }
        }
    });
    std::vector<uint32_t> amb_off(c->n_slots + 1, 0), amb_ent;
    for (int q = 0; q < c->n_slots; ++q) {
        amb_off[q] = (uint32_t)amb_ent.size();
        amb_ent.insert(amb_ent.end(), amb[q].begin(), amb[q].end());
    }
    amb_off[c->n_slots] = (uint32_t)amb_ent.size();
    c->cons_amb_begin = amb_off[Q]; c->cons_amb_n = amb_off[Q + 1] - amb_off[Q];

    // ---- pivot form (pairpivot.cu): pivot planes, the query stream (per tile and chunk: what every warp does + the plane
    // words of the halves that are counted), fix entries per slot ----
    std::vector<uint32_t> snp_off, snp_ent, pivot_words, qoff;
    std::vector<uint8_t> wslot, qstream;                                    // wslot [tile][warp][j]: slot of the tile the warp counts
    int64_t n_fix_total = 0;
    if (pivot_ok && do_prep) {
        pivot_words.assign((size_t)(c->Wp / 4) * 8, 0u);                     // [unit][plane b0, b1][4 words]
        for (int s = 0; s < L; ++s) {
            const int w = s >> 5;
            uint32_t* u = &pivot_words[(size_t)(w >> 2) * 8];
            if (pv[s] & 10u) u[w & 3] |= 1u << (s & 31);
            if (pv[s] & 12u) u[4 + (w & 3)] |= 1u << (s & 31);
        }
        size_t n_snp = 0;
        for (int q = 0; q < c->n_slots; ++q) n_snp += snp[q].size();
        // the kernel works on the COMPLEMENTS (pairpivot.cu): plane 0 = sites where the slot holds no plain base (or that
        // are trimmed away / padding), plane 1 = sites where it does not hold the pivot base.  Padding slots stay all-zero
        // (nothing to do, never stored).  work[slot] = cost of the slot: 4 per half counted from the planes, 1 per all-N unit.
        std::vector<uint32_t> work((size_t)c->n_qtiles * QTILE, 0);
        parallel_ranges(c->n_qtiles, host_threads(c->n_qtiles), [&](int, int ta, int tb) {
            for (int qt = ta; qt < tb; ++qt)
                for (int ch = 0; ch < c->n_chunks; ++ch) {
                    HostPlanes2& blk = qd2[(size_t)qt * c->n_chunks + ch];
                    for (int u = 0; u < W4; ++u)
                        for (int qi = 0; qi < QTILE && qt * QTILE + qi < c->n_slots; ++qi) {
                            for (int pl = 0; pl < 2; ++pl) {
                                uint4& v = blk.pl[u][qi][pl];
                                v.x = ~v.x; v.y = ~v.y; v.z = ~v.z; v.w = ~v.w;
                            }
                            const uint4& e = blk.pl[u][qi][1];
                            const uint4& b = blk.pl[u][qi][0];
                            if ((b.x & b.y & b.z & b.w) == 0xffffffffu) work[(size_t)qt * QTILE + qi] += 1;               // table lookup
                            else work[(size_t)qt * QTILE + qi] += 4 * (((e.x | e.y) != 0u) + ((e.z | e.w) != 0u));       // halves counted from the planes
                        }
                }
        });
        // Pivot tiles: a thread block takes PIVOT_SLOTS = 60 consecutive slots; warp 0 only fills the staging ring (measured: a
        // warp that issues the TMA copies on top of its counting work sets the pace of the block and the others wait for it),
        // warps 1..15 count 4 slots each.  Which 4: the warps share one ring, so the block runs at the pace of the warp with the
        // most work -- longest-processing-time first: slots by decreasing work, each to the least loaded warp that has room.
        const int n_ptiles = (c->n_slots + PIVOT_SLOTS - 1) / PIVOT_SLOTS;
        c->n_ptiles = n_ptiles;
        wslot.assign((size_t)n_ptiles * PIVOT_WARPS * PIVOT_QW, 0xff);
        for (int pt = 0; pt < n_ptiles; ++pt) {
            int order[PIVOT_SLOTS], n_here = std::min(PIVOT_SLOTS, c->n_slots - pt * PIVOT_SLOTS);
            for (int i = 0; i < n_here; ++i) order[i] = i;
            const uint32_t* wk = &work[(size_t)pt * PIVOT_SLOTS];
            std::stable_sort(order, order + n_here, [&](int x, int y) { return wk[x] > wk[y]; });
            uint64_t load[PIVOT_WARPS] = {0}; int cnt[PIVOT_WARPS] = {0};
            for (int i = 0; i < n_here; ++i) {
                int best = -1;
                for (int w = 1; w < PIVOT_WARPS; ++w) if (cnt[w] < PIVOT_QW && (best < 0 || load[w] < load[best])) best = w;
                wslot[((size_t)pt * PIVOT_WARPS + best) * PIVOT_QW + cnt[best]++] = (uint8_t)order[i];
                load[best] += wk[order[i]];
            }
        }
        // the stream: per (pivot tile, chunk) the 16 warp heads, then the items of warp 1, warp 2, ... in the order the kernel
        // walks them (unit, slot of the warp, half).  Tiles are built in parallel, then concatenated.
        std::vector<std::vector<uint8_t>> tile_rec(n_ptiles);
        std::vector<std::vector<uint32_t>> tile_off(n_ptiles);
        std::vector<int64_t> tile_halves(n_ptiles, 0);
        parallel_ranges(n_ptiles, host_threads(n_ptiles), [&](int, int ta, int tb) {
            for (int pt = ta; pt < tb; ++pt) {
                std::vector<uint8_t>& rec = tile_rec[pt];
                std::vector<uint32_t>& off = tile_off[pt];
                off.reserve(c->n_chunks + 1);
                for (int ch = 0; ch < c->n_chunks; ++ch) {
                    off.push_back((uint32_t)rec.size());
                    const size_t h0 = rec.size();
                    rec.resize(h0 + QHEAD_BYTES, 0);
                    int n_items = 0;
                    for (int wp = 0; wp < PIVOT_WARPS; ++wp) {
                        QWarpHead hd{};
                        hd.item0 = (uint16_t)n_items;
                        for (int u = 0; u < W4; ++u)
                            for (int j = 0; j < PIVOT_QW; ++j) {
                                const int sl = wslot[((size_t)pt * PIVOT_WARPS + wp) * PIVOT_QW + j];
                                if (sl == 0xff) continue;
                                const int q = pt * PIVOT_SLOTS + sl;                         // global slot; its planes live in the 64-slot tile q / QTILE
                                const HostPlanes2& blk = qd2[(size_t)(q / QTILE) * c->n_chunks + ch];
                                const uint4& e = blk.pl[u][q % QTILE][1];
                                const uint4& b = blk.pl[u][q % QTILE][0];
                                if ((b.x & b.y & b.z & b.w) == 0xffffffffu) { hd.fulln |= (uint8_t)(1u << (u * 4 + j)); continue; }
                                const uint32_t bw[4] = {b.x, b.y, b.z, b.w}, ew[4] = {e.x, e.y, e.z, e.w};
                                for (int h = 0; h < 2; ++h) {
                                    if (!(ew[2 * h] | ew[2 * h + 1])) continue;
                                    hd.half |= (uint16_t)(1u << (u * 8 + j * 2 + h));
                                    const uint32_t it[2] = {bw[2 * h], bw[2 * h + 1]};                  // ne == nq: one plane (pairpivot.cu)
                                    const uint8_t* ib = reinterpret_cast<const uint8_t*>(it);
                                    rec.insert(rec.end(), ib, ib + QITEM_BYTES);
                                    ++n_items; ++hd.n_items;
                                }
                            }
                        memcpy(&rec[h0 + (size_t)wp * sizeof(QWarpHead)], &hd, sizeof hd);
                    }
                    if (rec.size() % 16) rec.resize(rec.size() + 16 - rec.size() % 16, 0);      // bulk copies move multiples of 16 bytes
                    tile_halves[pt] += n_items;
                }
                off.push_back((uint32_t)rec.size());
            }
        });
        int64_t halves_total = 0;
        size_t total = 0;
        for (int qt = 0; qt < n_ptiles; ++qt) { total += tile_rec[qt].size(); halves_total += tile_halves[qt]; }
        if (total > 0xfffffff0ull) pivot_ok = false;                        // offsets are 32 bits (never reached: 17 tiles ~ 1.5 MB)
        qstream.reserve(total);
        qoff.reserve((size_t)n_ptiles * (c->n_chunks + 1));
        for (int qt = 0; qt < n_ptiles && pivot_ok; ++qt) {
            const uint32_t base = (uint32_t)qstream.size();
            for (uint32_t o : tile_off[qt]) qoff.push_back(base + o);
            qstream.insert(qstream.end(), tile_rec[qt].begin(), tile_rec[qt].end());
        }
        c->pivot_halves = halves_total;
        // fix entries per slot: chunk << 16 | pivot's base << 10 | unit << 9 | word << 7 | query's base (A,C,G,T) << 5 | site in word
        snp_off.assign(c->n_slots + 1, 0);
        for (int q = 0; q < c->n_slots; ++q) {
            snp_off[q] = (uint32_t)snp_ent.size();
            for (uint32_t ent : snp[q]) {
                const uint32_t site = ent & AMB_SITE_MASK, code = ent >> 28, w = site >> 5;
                const uint32_t a = code == 1 ? 0 : code == 2 ? 1 : code == 4 ? 2 : 3;
                const uint32_t pc = pv[site], pa = pc == 1 ? 0 : pc == 2 ? 1 : pc == 4 ? 2 : 3;
                snp_ent.push_back((w / WC) << 16 | pa << 10 | ((w % WC) / 4) << 9 | (w & 3u) << 7 | a << 5 | (site & 31u));
            }
        }
        snp_off[c->n_slots] = (uint32_t)snp_ent.size();
        n_fix_total = (int64_t)n_snp;
        // a fix entry costs two scattered loads per thread block of every reference pair: divergent query sets (many plain
        // bases != pivot) stay on the general kernel
        if (c->prm.pair_kernel != UVG_PAIR_PIVOT && n_snp > (size_t)c->n_slots * (size_t)std::max(4, c->W / 8)) pivot_ok = false;
    }

    const auto t_host = std::chrono::steady_clock::now();
    if (bcast) {
        // what the other ranks cannot know: the form chosen, the sizes of the site sets and of the two variable lists
        int64_t hdr[8] = {pivot_ok ? 1 : 0, n_idx, n_idx_m, n_idx_c, (int64_t)amb_ent.size(), (int64_t)snp_ent.size() | ((int64_t)qstream.size() << 32),
                          n_fix_total | (c->pivot_halves << 32), 0x55564731};
        int64_t* d_hdr = nullptr;
        UVG_CUDA(c, cudaMalloc(&d_hdr, sizeof hdr));
        UVG_CUDA(c, cudaMemcpyAsync(d_hdr, hdr, sizeof hdr, cudaMemcpyHostToDevice, c->st));
        UVG_NCCL(c, g_nccl.Bcast(d_hdr, d_hdr, 8, ncclInt64, 0, c->comm, c->st));
        UVG_CUDA(c, cudaMemcpyAsync(hdr, d_hdr, sizeof hdr, cudaMemcpyDeviceToHost, c->st));
        UVG_CUDA(c, cudaStreamSynchronize(c->st));
        cudaFree(d_hdr);
        if (hdr[7] != 0x55564731) UVG_FAIL(c, UVG_ESTATE, "uvg_set_queries: the ranks disagree about the query broadcast");
        if (!do_prep) {
            pivot_ok = hdr[0] != 0; n_idx = (int)hdr[1]; n_idx_m = (int)hdr[2]; n_idx_c = (int)hdr[3]; n_fix_total = hdr[6] & 0xffffffffLL; c->pivot_halves = hdr[6] >> 32;
            amb_ent.assign((size_t)hdr[4], 0u);
            if (pivot_ok) {
                snp_ent.assign((size_t)(hdr[5] & 0xffffffffLL), 0u);
                snp_off.assign(c->n_slots + 1, 0u);
                qstream.assign((size_t)(hdr[5] >> 32), 0);
                c->n_ptiles = (c->n_slots + PIVOT_SLOTS - 1) / PIVOT_SLOTS;
                qoff.assign((size_t)c->n_ptiles * (c->n_chunks + 1), 0u);
                pivot_words.assign((size_t)(c->Wp / 4) * 8, 0u);
                wslot.assign((size_t)((c->n_slots + PIVOT_SLOTS - 1) / PIVOT_SLOTS) * PIVOT_WARPS * PIVOT_QW, 0);
            }
        }
    }
    c->qinfo.n_idx = n_idx; c->qinfo.n_idx_m = n_idx_m; c->qinfo.n_idx_c = n_idx_c;
    if (info) *info = c->qinfo;
    // ---- device allocations ----
    const int rows = c->rows_cap;
    UVG_CUDA(c, pooled_alloc(c, 0, &c->d_qdata, n_chunk_total * sizeof(QChunk)));
    UVG_CUDA(c, pooled_alloc(c, 1, &c->d_qnib, qnib.size() * 4));
    UVG_CUDA(c, pooled_alloc(c, 2, &c->d_qamb_off, amb_off.size() * 4));
    UVG_CUDA(c, pooled_alloc(c, 3, &c->d_qamb_ent, amb_ent.size() * 4));
    UVG_CUDA(c, pooled_alloc(c, 4, &c->d_thr_local, (size_t)rows * sizeof(Threshold)));
    for (int b = 0; b < NBUF; ++b) UVG_CUDA(c, pooled_alloc(c, 5 + b, &c->d_thr_all[b], (size_t)c->world * rows * sizeof(Threshold)));
    c->own_off_n = (size_t)rows * (c->K + 1) * sizeof(HeapItem);
    c->own_off_mi = c->own_off_n + (size_t)rows * 4;
    c->own_bytes = (c->own_off_mi + (size_t)rows * 4 + 15) / 16 * 16;
    UVG_CUDA(c, pooled_alloc(c, 7, &c->d_own, c->own_bytes));
    c->d_heaps = reinterpret_cast<HeapItem*>(c->d_own);
    c->d_heap_n = reinterpret_cast<int32_t*>(c->d_own + c->own_off_n); c->d_max_inc = reinterpret_cast<int32_t*>(c->d_own + c->own_off_mi);
    UVG_CUDA(c, pooled_alloc(c, 10, &c->d_batch_id, (size_t)rows * 4)); UVG_CUDA(c, pooled_alloc(c, 11, &c->d_max_inc_batch, (size_t)rows * 4));
    c->fin_device = c->K <= FIN_DEVICE_MAX_K;
    c->fin_off_n = (size_t)rows * c->K * sizeof(uvg_item);
    c->fin_off_mi = c->fin_off_n + (size_t)rows * 4;
    c->fin_bytes = (c->fin_off_mi + (size_t)rows * 4 + 15) / 16 * 16;
    if (c->fin_device) UVG_CUDA(c, pooled_alloc(c, 22, &c->d_fin, c->fin_bytes));
    if (c->world > 1) UVG_CUDA(c, pooled_alloc(c, 12, &c->d_all, (size_t)c->world * (c->fin_device ? c->fin_bytes : c->own_bytes)));
    if (c->literal) {
        UVG_CUDA(c, pooled_alloc(c, 13, &c->d_snap_own, c->own_bytes));
        UVG_CUDA(c, pooled_alloc(c, 14, &c->d_snap_batch, (size_t)rows * 8));
    }
    // dense score tiles: rounds are capped so that one buffer stays within ~3 GiB
    {
        int64_t cap = round_size_max(c->prm);
        const int64_t per_ref = (int64_t)(rows + 2) * c->world * 16;        // bytes of one column position over all segments
        const int64_t fit = ((int64_t)3 << 30) / per_ref / (GROUP * RPL) * (GROUP * RPL);
        // Single GPU: rounds shrink so that one buffer stays within ~3 GiB (nobody else depends on the round sizes).  Sharded:
        // the round schedule IS the row distribution the hosts deal with uvg_shard_of (first_round, max_round): it must not
        // depend on the query set, so the tiles take what max_round asks for (or the allocation fails loudly below).
        if (c->world == 1) cap = std::max<int64_t>(GROUP * RPL, std::min(cap, fit));
        c->seg_cap = cap;
        if (c->literal) UVG_CUDA(c, pooled_alloc(c, 21, &c->d_ever_round, (size_t)cap * c->world));      // admissions of one round (all segments)
        c->tile_elems = (size_t)(rows + 2) * c->world * cap;
        if (NBUF * c->tile_elems > c->tiles_alloc) {                       // the multi-GiB buffer is kept when the next query set fits
            if (c->world > 1) close_peers(c);                                // (every rank takes this branch alike: the size depends on shared parameters only);
                                                                             // the new buffer is mapped by the peers at the next search
            cudaFree(c->d_tiles); c->d_tiles = nullptr; c->tiles_alloc = 0;
            const cudaError_t e_t = cudaMalloc(&c->d_tiles, NBUF * c->tile_elems * sizeof(uint4));
            if (e_t != cudaSuccess) {
                cudaGetLastError();
                UVG_FAIL(c, UVG_ENOMEM, "uvg_set_queries: the score tiles of %d queries need %.1f GiB per rank at max_round %lld; create the contexts (and deal "
                         "the rows, uvg_shard_of) with a smaller max_round", Q, (double)(NBUF * c->tile_elems * sizeof(uint4)) / (1 << 30), (long long)cap);
            }
            c->tiles_alloc = NBUF * c->tile_elems;
        }
        if (c->world == 1) { c->peer_tiles[0] = c->d_tiles; c->peers_ready = true; }
    }
    UVG_CUDA(c, cudaMemcpyAsync(c->d_qdata, qd.data(), n_chunk_total * sizeof(QChunk), cudaMemcpyHostToDevice, c->st));
    UVG_CUDA(c, cudaMemcpyAsync(c->d_qnib, qnib.data(), qnib.size() * 4, cudaMemcpyHostToDevice, c->st));
    UVG_CUDA(c, cudaMemcpyAsync(c->d_qamb_off, amb_off.data(), amb_off.size() * 4, cudaMemcpyHostToDevice, c->st));
    if (!amb_ent.empty()) UVG_CUDA(c, cudaMemcpyAsync(c->d_qamb_ent, amb_ent.data(), amb_ent.size() * 4, cudaMemcpyHostToDevice, c->st));
    auto share = [&](void* d, size_t bytes) -> ncclResult_t {             // rank 0's array to everybody (stream order: after its upload)
        return bytes ? g_nccl.Bcast(d, d, bytes, ncclUint8, 0, c->comm, c->st) : 0;
    };
    if (bcast) {
        UVG_NCCL(c, share(c->d_qdata, n_chunk_total * sizeof(QChunk)));
        UVG_NCCL(c, share(c->d_qnib, qnib.size() * 4));
        UVG_NCCL(c, share(c->d_qamb_off, amb_off.size() * 4));
        UVG_NCCL(c, share(c->d_qamb_ent, amb_ent.size() * 4));
    }
    if (pivot_ok) {
        UVG_CUDA(c, pooled_alloc(c, 15, &c->d_qstream, qstream.size()));
        UVG_CUDA(c, pooled_alloc(c, 16, &c->d_snp_off, snp_off.size() * 4));
        UVG_CUDA(c, pooled_alloc(c, 17, &c->d_snp_ent, snp_ent.size() * 4));
        UVG_CUDA(c, pooled_alloc(c, 18, &c->d_pivot, pivot_words.size() * 4));
        UVG_CUDA(c, pooled_alloc(c, 19, &c->d_wslot, wslot.size()));
        UVG_CUDA(c, pooled_alloc(c, 20, &c->d_qoff, qoff.size() * 4));
        UVG_CUDA(c, cudaMemcpyAsync(c->d_wslot, wslot.data(), wslot.size(), cudaMemcpyHostToDevice, c->st));
        UVG_CUDA(c, cudaMemcpyAsync(c->d_qstream, qstream.data(), qstream.size(), cudaMemcpyHostToDevice, c->st));
        UVG_CUDA(c, cudaMemcpyAsync(c->d_qoff, qoff.data(), qoff.size() * 4, cudaMemcpyHostToDevice, c->st));
        UVG_CUDA(c, cudaMemcpyAsync(c->d_snp_off, snp_off.data(), snp_off.size() * 4, cudaMemcpyHostToDevice, c->st));
        if (!snp_ent.empty()) UVG_CUDA(c, cudaMemcpyAsync(c->d_snp_ent, snp_ent.data(), snp_ent.size() * 4, cudaMemcpyHostToDevice, c->st));
        UVG_CUDA(c, cudaMemcpyAsync(c->d_pivot, pivot_words.data(), pivot_words.size() * 4, cudaMemcpyHostToDevice, c->st));
        if (bcast) {
            UVG_NCCL(c, share(c->d_wslot, wslot.size()));
            UVG_NCCL(c, share(c->d_qstream, qstream.size()));
            UVG_NCCL(c, share(c->d_qoff, qoff.size() * 4));
            UVG_NCCL(c, share(c->d_snp_off, snp_off.size() * 4));
            UVG_NCCL(c, share(c->d_snp_ent, snp_ent.size() * 4));
            UVG_NCCL(c, share(c->d_pivot, pivot_words.size() * 4));
        }
        c->pivot_on = true;
        // references that are already resident get their derived planes now (one pass over the packed database)
        int rc2 = ensure_db2(c);
        if (rc2) return rc2;
        if (c->db2_cap) UVG_CUDA(c, cudaMemsetAsync(c->d_ref_cnt, 0, (size_t)c->db2_cap * 2 * sizeof(int32_t), c->st));
        UVG_CUDA(c, launch_derive_pivot(c->d_db, c->d_db2, c->d_pivot, c->d_ref_cnt, c->d_ref_cnt + c->db2_cap, 0, c->db_n, c->Wp, c->st));
        if (c->db_n) c->stats.other_kernel_launches += 1;
    }
    c->qinfo_pivot_fix = pivot_ok ? n_fix_total : -1;
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    rc = reset_heaps(c);
    const auto t_end = std::chrono::steady_clock::now();
    c->set_queries_host_ms = std::chrono::duration<double, std::milli>(t_host - t_begin).count();
    c->set_queries_device_ms = std::chrono::duration<double, std::milli>(t_end - t_host).count();
    return rc;
}

// ---- -k / --keep_resolved: exclude_redundant_query_sequences (src/fastaseq.c:797-841) ----
// The reference walks the pairs (i < j) of the ordered query set, skips a pair when either member is already out, and
// calls a pair a candidate when no site of idx holds two usable, different bytes (fastaseq.c:562-583 with maxdist 1).  For
// a candidate it compares which of the two has usable sites the other lacks, over idx and over idx_m (fastaseq.c:598-641),
// and drops one of them.  Whether a pair is a candidate and which side is more resolved do not depend on who is still in,
// so host threads classify blocks of rows in parallel (early exit at the first differing site) and the sequential walk
// replays only the candidates in (i, j) order.
extern "C" int uvg_prune_queries(const uint8_t* text, int32_t n_query, int32_t nchar, int32_t trim_arg, int32_t acgt_mode,
                                 int32_t keep_more_resolved, uint8_t* keep, int32_t* n_kept)
{
    if (!text || !keep || n_query < 1 || nchar < 1) return UVG_EINVAL;
    int trim = trim_arg < 0 ? 0 : trim_arg;                                  // fastaseq.c:708-712, as uvg_create
    if (trim > nchar / 2.1) trim = (int)(nchar / 2.1);
    const int L = nchar, Q = n_query;
    const bool acgt = acgt_mode != 0;
    uint8_t lut[3][256], up[256], usable[256];
    build_luts(lut);
    for (int i = 0; i < 256; ++i) up[i] = (uint8_t)((i >= 'a' && i <= 'z') ? i - 32 : i);
    for (int i = 0; i < 256; ++i) usable[i] = acgt ? (i == 'A' || i == 'C' || i == 'G' || i == 'T') : (lut[2][i] != 0);

    // site classes of create_query_indices (fastaseq.c:732-777): 1 = idx (segregating), 2 = idx_m (agree, somebody lacks the site)
    std::vector<uint32_t> idx, idx_m;
    {
        const int T = host_threads(std::min<int64_t>(Q, std::max<int64_t>(1, ((int64_t)256 << 20) / ((int64_t)L * 4))));
        std::vector<uint8_t> first((size_t)T * L, 0), all((size_t)T * L, 1), same((size_t)T * L, 1);
        parallel_ranges(Q, T, [&](int t, int qa, int qb) {
            uint8_t *f = &first[(size_t)t * L], *al = &all[(size_t)t * L], *sm = &same[(size_t)t * L];
            for (int q = qa; q < qb; ++q) {
                const uint8_t* row = text + (size_t)q * L;
                for (int i = trim; i < L - trim; ++i) {
                    const uint8_t ch = up[row[i]];
                    if (!usable[ch]) { al[i] = 0; continue; }
                    if (!f[i]) f[i] = ch;
                    else if (ch != f[i]) sm[i] = 0;
                }
            }
        });
        for (int i = trim; i < L - trim; ++i) {
            uint8_t f0 = 0; bool al = true, sm = true;
            for (int t = 0; t < T; ++t) {
                const size_t k = (size_t)t * L + i;
                if (!all[k]) al = false;
                if (!first[k]) continue;
                if (!same[k]) sm = false;
                if (!f0) f0 = first[k];
                else if (first[k] != f0) sm = false;
            }
            if (!f0) continue;
            if (!sm) idx.push_back((uint32_t)i);
            else if (!al) idx_m.push_back((uint32_t)i);
        }
    }
    // per query: the bytes at idx (0 where not usable), and bitmaps of the usable sites of idx and of idx_m
    const size_t n1 = idx.size(), n2 = idx_m.size(), n1p = (n1 + 31) / 32 * 32, w1 = n1p / 64 + 1, w2 = n2 / 64 + 1;
    std::vector<uint8_t> col((size_t)Q * n1p, 0);
    std::vector<uint64_t> bm1((size_t)Q * w1, 0), bm2((size_t)Q * w2, 0);
    const int T = host_threads(Q);
    parallel_ranges(Q, T, [&](int, int qa, int qb) {
        for (int q = qa; q < qb; ++q) {
            const uint8_t* row = text + (size_t)q * L;
            for (size_t k = 0; k < n1; ++k) {
                const uint8_t ch = up[row[idx[k]]];
                if (usable[ch]) { col[(size_t)q * n1p + k] = ch; bm1[(size_t)q * w1 + k / 64] |= 1ull << (k % 64); }
            }
            for (size_t k = 0; k < n2; ++k)
                if (usable[up[row[idx_m[k]]]]) bm2[(size_t)q * w2 + k / 64] |= 1ull << (k % 64);
        }
    });
    auto resolved = [](const uint64_t* a, const uint64_t* b, size_t w) {      // left_is_resolved_right: -1 / 0 / +1, 0xff if both
        bool left = false, right = false;
        for (size_t k = 0; k < w; ++k) { left |= (a[k] & ~b[k]) != 0; right |= (~a[k] & b[k]) != 0; }
        return left && right ? 0xff : left ? -1 : right ? 1 : 0;
    };
    std::vector<uint8_t> valid(Q, 1);
    const int BLOCK = 256;
    std::vector<std::vector<uint32_t>> cand(BLOCK);
    for (int i0 = 0; i0 < Q - 1; i0 += BLOCK) {
        const int i1 = std::min(Q - 1, i0 + BLOCK);
        const std::vector<uint8_t> snap = valid;                  // rows / columns already out never come back
        parallel_ranges(i1 - i0, std::min(T, i1 - i0), [&](int, int a, int b) {
            for (int i = i0 + a; i < i0 + b; ++i) {
                cand[i - i0].clear();
                if (!snap[i]) continue;
                const uint8_t* ri = &col[(size_t)i * n1p];
                for (int j = i + 1; j < Q; ++j) {
                    if (!snap[j]) continue;
                    const uint8_t* rj = &col[(size_t)j * n1p];
                    bool differ = false;
                    for (size_t k = 0; k < n1p && !differ; k += 32) {
                        uint8_t acc = 0;
                        for (int u = 0; u < 32; ++u) acc |= (uint8_t)((ri[k + u] != 0) & (rj[k + u] != 0) & (ri[k + u] != rj[k + u]));
                        differ = acc != 0;
                    }
                    if (!differ) cand[i - i0].push_back((uint32_t)j);
                }
            }
        });
        for (int i = i0; i < i1; ++i) for (uint32_t j : cand[i - i0]) {
            if (!valid[i] || !valid[j]) continue;
            int red1 = resolved(&bm1[(size_t)i * w1], &bm1[(size_t)j * w1], w1);
            if (red1 > 1) continue;                                // both hold sites the other lacks: not equivalent
            const int red2 = resolved(&bm2[(size_t)i * w2], &bm2[(size_t)j * w2], w2);
            if (red2 > 1) continue;
            if (!red1 && !red2) valid[j] = 0;                      // identical, gaps and Ns included
            red1 += red2;
            if (!red1) continue;                                   // opposite signs: complementary information
            if (keep_more_resolved) { if (red1 > 0) valid[i] = 0; else valid[j] = 0; }
            else                    { if (red1 > 0) valid[j] = 0; else valid[i] = 0; }
        }
    }
    int32_t n = 0;
    for (int q = 0; q < Q; ++q) { keep[q] = valid[q]; n += valid[q]; }
    if (n_kept) *n_kept = n;
    return UVG_OK;
}

extern "C" int uvg_reset_search(uvg_ctx* c)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (c->n_query == 0) UVG_FAIL(c, UVG_ESTATE, "uvg_reset_search: no queries set");
    return reset_heaps(c);
}

// =====================================================================================================================
// database residency (K1)
// =====================================================================================================================
template <typename T>
static cudaError_t grow_array(T** ptr, size_t old_count, size_t new_count, bool zero_new, cudaStream_t st, std::vector<void*>* defer = nullptr)
{
    T* n = nullptr;
    cudaError_t e = cudaMalloc(&n, new_count * sizeof(T));
    if (e != cudaSuccess) return e;
    if (*ptr && old_count) {
        e = cudaMemcpyAsync(n, *ptr, old_count * sizeof(T), cudaMemcpyDeviceToDevice, st);
        if (e != cudaSuccess) return e;
    }
    if (zero_new && new_count > old_count) {
        e = cudaMemsetAsync(n + old_count, 0, (new_count - old_count) * sizeof(T), st);
        if (e != cudaSuccess) return e;
    }
    e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return e;
    if (defer && *ptr) defer->push_back(*ptr);      // another rank may still read the old array (ensure_peer_db releases it)
    else cudaFree(*ptr);
    *ptr = n;
    return cudaSuccess;
}

// Growing a device array moves it: nothing may be in flight (a host that reserves db_capacity never pays for this).
static int ensure_db_capacity(uvg_ctx* c, int64_t need_refs)
{
    if (need_refs <= c->db_cap) return UVG_OK;
    int rc = sync_all(c);
    if (rc) return rc;
    int64_t cap = std::max<int64_t>(need_refs, c->db_cap * 2);
    cap = std::max<int64_t>(cap, c->prm.db_capacity);
    cap = std::max<int64_t>(cap, 4096);
    cap = (cap + GROUP * PRPL - 1) / (GROUP * PRPL) * (GROUP * PRPL);     // whole records of the derived database (PRPL >= RPL groups)
    const size_t per_group = (size_t)(c->Wp / 4) * NPLANE * GROUP;            // uint4 per group
    const size_t og = (size_t)(c->db_cap / GROUP), ng = (size_t)(cap / GROUP);
    std::vector<void*>* defer = c->peer_db_live ? &c->deferred_free : nullptr;
    UVG_CUDA(c, grow_array(&c->d_db, og * per_group, ng * per_group, true, c->st_push, defer));
    UVG_CUDA(c, grow_array(&c->d_non_n, (size_t)c->db_cap, (size_t)cap, true, c->st_push));
    UVG_CUDA(c, grow_array(&c->d_ramb_off, c->db_cap ? (size_t)c->db_cap + 1 : 0, (size_t)cap + 1, true, c->st_push, defer));
    c->db_cap = cap;
    return ensure_db2(c);
}

static int ensure_amb_capacity(uvg_ctx* c, int64_t need)
{
    if (need <= c->amb_cap) return UVG_OK;
    if (need > 0xfffffff0LL) UVG_FAIL(c, UVG_ENOMEM, "more than 2^32 ambiguity codes in the resident references");
    int rc = sync_all(c);
    if (rc) return rc;
    int64_t cap = std::max<int64_t>(need, c->amb_cap * 2);
    cap = std::max<int64_t>(cap, std::max<int64_t>(1 << 16, c->prm.db_capacity * 4));   // ~1.7 codes per SARS-CoV-2 genome
    UVG_CUDA(c, grow_array(&c->d_ramb_ent, (size_t)c->amb_cap, (size_t)cap, false, c->st_push, c->peer_db_live ? &c->deferred_free : nullptr));
    c->amb_cap = cap;
    return UVG_OK;
}

static int ensure_ever(uvg_ctx* c, int64_t n_global)
{
    if (n_global <= c->ever_cap) return UVG_OK;
    int64_t cap = std::max<int64_t>(n_global, c->ever_cap * 2);
    cap = std::max<int64_t>(cap, std::max<int64_t>(4096, c->prm.db_capacity * (int64_t)c->world));
    int rc = sync_all(c);
    if (rc) return rc;
    UVG_CUDA(c, grow_array(&c->d_ever, (size_t)c->ever_cap, (size_t)cap, true, c->st));
    c->ever_cap = cap;
    return UVG_OK;
}

static int ensure_stage_rows(uvg_ctx* c, int64_t rows)
{
    if (rows <= c->stage_rows) return UVG_OK;
    UVG_CUDA(c, cudaStreamSynchronize(c->st_push));
    cudaFree(c->d_stage_nn); cudaFree(c->d_stage_namb); cudaFree(c->d_stage_acc); cudaFree(c->d_stage_dst);
    c->d_stage_nn = c->d_stage_namb = nullptr; c->d_stage_acc = nullptr; c->d_stage_dst = nullptr; c->stage_rows = 0;
    UVG_CUDA(c, cudaMalloc(&c->d_stage_nn, rows * 4));
    UVG_CUDA(c, cudaMalloc(&c->d_stage_namb, rows * 4));
    UVG_CUDA(c, cudaMalloc(&c->d_stage_acc, rows));
    UVG_CUDA(c, cudaMalloc(&c->d_stage_dst, rows * 8));
    c->stage_rows = rows;
    return UVG_OK;
}

// encode n rows already on the device (stream st_push); appends the accepted ones.  Search rounds that are in flight
// only read rows below db_n, so the encoder may run under them.
static int push_device_rows(uvg_ctx* c, const uint8_t* d_text, int64_t n, int32_t min_non_n, int32_t* non_n_out,
                            uint8_t* accepted_out, int64_t* n_acc_out)
{
    int rc = ensure_stage_rows(c, n);
    if (rc) return rc;
    rc = ensure_db_capacity(c, c->db_n + n);
    if (rc) return rc;
    cudaStream_t sp = c->st_push;
    UVG_CUDA(c, launch_count_non_n(d_text, n, c->nchar, min_non_n, c->prm.acgt_mode, c->d_stage_nn, c->d_stage_acc, c->d_stage_namb, sp));
    UVG_CUDA(c, launch_scan_accept(c->d_stage_acc, c->d_stage_namb, n, c->db_n, (uint32_t)c->amb_n, c->d_stage_dst, c->d_ramb_off, c->d_nacc, sp));
    int64_t* h_n = reinterpret_cast<int64_t*>(c->h_pinned);
    UVG_CUDA(c, cudaMemcpyAsync(h_n, c->d_nacc, 16, cudaMemcpyDeviceToHost, sp));
    if (non_n_out) UVG_CUDA(c, cudaMemcpyAsync(non_n_out, c->d_stage_nn, n * 4, cudaMemcpyDeviceToHost, sp));
    if (accepted_out) UVG_CUDA(c, cudaMemcpyAsync(accepted_out, c->d_stage_acc, n, cudaMemcpyDeviceToHost, sp));
    UVG_CUDA(c, cudaStreamSynchronize(sp));
    const int64_t n_acc = h_n[0], n_amb = h_n[1];
    rc = ensure_amb_capacity(c, c->amb_n + n_amb);
    if (rc) return rc;
    UVG_CUDA(c, launch_encode(d_text, n, c->nchar, c->Wp, c->prm.acgt_mode, c->d_stage_dst, c->d_stage_nn, c->d_db, c->d_ramb_off,
                              c->d_ramb_ent, c->d_non_n, sp));
    c->stats.other_kernel_launches += 3;
    if (c->pivot_on && n_acc > 0) {
        // rows above db_n still hold the totals of dropped / cleared references: zero the range first
        UVG_CUDA(c, cudaMemsetAsync(c->d_ref_cnt + c->db_n, 0, (size_t)n_acc * 4, sp));
        UVG_CUDA(c, cudaMemsetAsync(c->d_ref_cnt + c->db2_cap + c->db_n, 0, (size_t)n_acc * 4, sp));
        UVG_CUDA(c, launch_derive_pivot(c->d_db, c->d_db2, c->d_pivot, c->d_ref_cnt, c->d_ref_cnt + c->db2_cap, c->db_n, c->db_n + n_acc, c->Wp, sp));
        c->stats.other_kernel_launches += 1;
    }
    c->db_n += n_acc;
    c->amb_n += n_amb;
    if (n_acc_out) *n_acc_out += n_acc;
    return UVG_OK;
}

extern "C" int uvg_push_refs_device(uvg_ctx* c, const uint8_t* d_text, int64_t n, int32_t min_non_n, int32_t* non_n_out,
                                    uint8_t* accepted_out, int64_t* n_accepted)
{
    if (!c || (!d_text && n > 0) || n < 0) { if (c) c->err = "uvg_push_refs_device: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (n_accepted) *n_accepted = 0;
    if (n == 0) return UVG_OK;
    UVG_CUDA(c, cudaEventRecord(c->ev_c, c->st_push));
    const int64_t step = 1 << 16;
    for (int64_t i = 0; i < n; i += step) {
        const int64_t m = std::min(step, n - i);
        int rc = push_device_rows(c, d_text + i * (int64_t)c->nchar, m, min_non_n, non_n_out ? non_n_out + i : nullptr,
                                  accepted_out ? accepted_out + i : nullptr, n_accepted);
        if (rc) return rc;
    }
    UVG_CUDA(c, cudaEventRecord(c->ev_d, c->st_push));
    UVG_CUDA(c, cudaEventSynchronize(c->ev_d));            // the caller may reuse d_text
    float ms = 0; cudaEventElapsedTime(&ms, c->ev_c, c->ev_d);
    c->stats.encode_ms = ms;
    return UVG_OK;
}

// Host text -> device: a two-buffer staging ring.  The copy of piece i+1 (copy engine, stream st_copy) runs under the
// encoder kernels of piece i (stream st_push) and under whatever search rounds are in flight.
extern "C" int uvg_push_refs(uvg_ctx* c, const uint8_t* text, int64_t n, int32_t min_non_n, int32_t* non_n_out,
                             uint8_t* accepted_out, int64_t* n_accepted)
{
    if (!c || (!text && n > 0) || n < 0) { if (c) c->err = "uvg_push_refs: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (n_accepted) *n_accepted = 0;
    if (n == 0) return UVG_OK;
    int64_t rows_per_piece = std::max<int64_t>(1, std::min<int64_t>(n, (128ll << 20) / c->nchar));
    rows_per_piece = std::min<int64_t>(rows_per_piece, 1 << 16);
    const size_t need = (size_t)rows_per_piece * c->nchar;
    if (need > c->stage_text_bytes) {
        UVG_CUDA(c, cudaStreamSynchronize(c->st_push)); UVG_CUDA(c, cudaStreamSynchronize(c->st_copy));
        for (int b = 0; b < 2; ++b) { cudaFree(c->d_stage_text[b]); c->d_stage_text[b] = nullptr; }
        c->stage_text_bytes = 0;
        for (int b = 0; b < 2; ++b) UVG_CUDA(c, cudaMalloc(&c->d_stage_text[b], need));
        c->stage_text_bytes = need;
    }
    const int64_t n_pieces = (n + rows_per_piece - 1) / rows_per_piece;
    auto issue_copy = [&](int64_t k) -> cudaError_t {
        const int b = (int)(k & 1);
        const int64_t i = k * rows_per_piece, m = std::min(rows_per_piece, n - i);
        cudaError_t e = cudaStreamWaitEvent(c->st_copy, c->ev_stage_free[b], 0);          // encoder of piece k-2 is done with it
        if (e == cudaSuccess) e = cudaMemcpyAsync(c->d_stage_text[b], text + i * (int64_t)c->nchar, (size_t)m * c->nchar, cudaMemcpyHostToDevice, c->st_copy);
        if (e == cudaSuccess) e = cudaEventRecord(c->ev_copied[b], c->st_copy);
        return e;
    };
    UVG_CUDA(c, cudaEventRecord(c->ev_c, c->st_push));
    UVG_CUDA(c, issue_copy(0));
    for (int64_t k = 0; k < n_pieces; ++k) {
        const int b = (int)(k & 1);
        const int64_t i = k * rows_per_piece, m = std::min(rows_per_piece, n - i);
        if (k + 1 < n_pieces) UVG_CUDA(c, issue_copy(k + 1));
        UVG_CUDA(c, cudaStreamWaitEvent(c->st_push, c->ev_copied[b], 0));
        int rc = push_device_rows(c, c->d_stage_text[b], m, min_non_n, non_n_out ? non_n_out + i : nullptr,
                                  accepted_out ? accepted_out + i : nullptr, n_accepted);
        if (rc) return rc;
        UVG_CUDA(c, cudaEventRecord(c->ev_stage_free[b], c->st_push));
    }
    UVG_CUDA(c, cudaEventRecord(c->ev_d, c->st_push));
    UVG_CUDA(c, cudaEventSynchronize(c->ev_d));            // rows are resident when the call returns
    float ms = 0; cudaEventElapsedTime(&ms, c->ev_c, c->ev_d);
    c->stats.encode_ms = ms;
    return UVG_OK;
}

extern "C" int uvg_drop_refs(uvg_ctx* c)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (c->world > 1) UVG_FAIL(c, UVG_ESTATE, "uvg_drop_refs is not available in sharded mode");
    if (c->search_open) UVG_FAIL(c, UVG_ESTATE, "uvg_drop_refs: a search is in flight (call uvg_wait first)");
    if (c->db_scored != c->db_n) UVG_FAIL(c, UVG_ESTATE, "uvg_drop_refs: %lld resident references have not been searched yet",
                                         (long long)(c->db_n - c->db_scored));
    c->global_base += c->db_n;
    c->db_n = 0; c->db_scored = 0; c->round_local_start = 0; c->amb_n = 0;
    return UVG_OK;
}

extern "C" int uvg_clear_refs(uvg_ctx* c)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    int rc = sync_all(c);
    if (rc) return rc;
    c->search_open = false; c->n_ev = 0;
    c->global_base = 0; c->db_n = 0; c->db_scored = 0; c->round_local_start = 0; c->global_total = -1; c->amb_n = 0; c->search_limit = -1;
    c->file_starts.assign(1, 0);
    return c->n_query ? reset_heaps(c) : UVG_OK;
}

extern "C" int64_t uvg_db_count(const uvg_ctx* c) { return c ? c->db_n : 0; }

// The reference fills its batches per file: a batch ends at end-of-file and the next file starts a new one
// (nearest.c:245-286), so where the files end is part of what --pool means.
extern "C" int uvg_mark_file_end(uvg_ctx* c, int64_t n_accepted_so_far)
{
    if (!c) return UVG_EINVAL;
    const int64_t at = n_accepted_so_far >= 0 ? n_accepted_so_far : c->global_base + c->db_n;
    if (at < c->file_starts.back()) UVG_FAIL(c, UVG_EINVAL, "uvg_mark_file_end: %lld is before the previous file end", (long long)at);
    if (c->world == 1 && at < c->global_base + c->db_scored) UVG_FAIL(c, UVG_ESTATE, "uvg_mark_file_end: references beyond %lld were searched already", (long long)at);
    if (at > c->file_starts.back()) c->file_starts.push_back(at);
    return UVG_OK;
}


// =====================================================================================================================
// packed reference cache: the resident packed references leave / enter the device as they are, so a
// host that searched a FASTA file once can keep the 3 bits per site the encoder made of it (planes + ambiguity lists +
// non_n) and skip reading, parsing and encoding the text next time: 0.375 bytes per site cross PCIe instead of 1.
// Blob: header | planes of whole groups of 32 references | non_n[n] | ambiguity offsets [n+1] (relative) | entries.
// =====================================================================================================================
namespace {
struct PackedHeader {
    char magic[8];
    int32_t version, nchar, Wp, acgt_mode;
    int64_t n_refs, n_amb;
    uint64_t planes_bytes, reserved[2];
};
static_assert(sizeof(PackedHeader) == 64, "packed header is 64 bytes");
const char PACKED_MAGIC[8] = {'U', 'V', 'G', 'P', 'A', 'C', 'K', '1'};
inline size_t pad16(size_t b) { return (b + 15) / 16 * 16; }
struct PackedLayout { size_t planes, non_n, off, ent, total; };
inline PackedLayout packed_layout(int Wp, int64_t n, int64_t n_amb)
{
    PackedLayout l;
    l.planes = (size_t)((n + GROUP - 1) / GROUP) * (size_t)(Wp / 4) * NPLANE * GROUP * sizeof(uint4);
    l.non_n = pad16((size_t)n * 4); l.off = pad16((size_t)(n + 1) * 4); l.ent = pad16((size_t)n_amb * 4);
    l.total = sizeof(PackedHeader) + l.planes + l.non_n + l.off + l.ent;
    return l;
}
}  // namespace

static int packed_range(uvg_ctx* c, int64_t first, int64_t n, uint32_t* amb_begin, uint32_t* amb_end, const char* who)
{
    if (first < 0 || n < 1 || first + n > c->db_n) UVG_FAIL(c, UVG_EINVAL, "%s: references [%lld, %lld) are not resident (%lld are)", who,
                                                            (long long)first, (long long)(first + n), (long long)c->db_n);
    if (first % GROUP) UVG_FAIL(c, UVG_EINVAL, "%s: the first reference must be a multiple of %d", who, GROUP);
    UVG_CUDA(c, cudaStreamSynchronize(c->st_push));
    UVG_CUDA(c, cudaMemcpy(amb_begin, c->d_ramb_off + first, 4, cudaMemcpyDeviceToHost));
    UVG_CUDA(c, cudaMemcpy(amb_end, c->d_ramb_off + first + n, 4, cudaMemcpyDeviceToHost));
    return UVG_OK;
}

extern "C" int uvg_packed_bytes(uvg_ctx* c, int64_t first, int64_t n, size_t* bytes)
{
    if (!c || !bytes) { if (c) c->err = "uvg_packed_bytes: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    uint32_t a = 0, b = 0;
    int rc = packed_range(c, first, n, &a, &b, "uvg_packed_bytes");
    if (rc) return rc;
    *bytes = packed_layout(c->Wp, n, (int64_t)b - a).total;
    return UVG_OK;
}

extern "C" int uvg_export_packed(uvg_ctx* c, int64_t first, int64_t n, void* blob, size_t capacity)
{
    if (!c || !blob) { if (c) c->err = "uvg_export_packed: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    uint32_t a = 0, b = 0;
    int rc = packed_range(c, first, n, &a, &b, "uvg_export_packed");
    if (rc) return rc;
    const int64_t n_amb = (int64_t)b - a;
    const PackedLayout l = packed_layout(c->Wp, n, n_amb);
    if (capacity < l.total) UVG_FAIL(c, UVG_EINVAL, "uvg_export_packed: the blob needs %zu bytes, %zu given", l.total, capacity);
    uint8_t* out = static_cast<uint8_t*>(blob);
    PackedHeader h{};
    memcpy(h.magic, PACKED_MAGIC, 8);
    h.version = 1; h.nchar = c->nchar; h.Wp = c->Wp; h.acgt_mode = c->prm.acgt_mode != 0; h.n_refs = n; h.n_amb = n_amb; h.planes_bytes = l.planes;
    memcpy(out, &h, sizeof h);
    uint8_t* planes = out + sizeof h;
    uint8_t* non_n = planes + l.planes;
    uint32_t* off = reinterpret_cast<uint32_t*>(non_n + l.non_n);
    uint8_t* ent = reinterpret_cast<uint8_t*>(off) + l.off;
    const size_t per_group = (size_t)(c->Wp / 4) * NPLANE * GROUP;                    // uint4 per group
    cudaStream_t sp = c->st_push;
    UVG_CUDA(c, cudaMemcpyAsync(planes, c->d_db + (size_t)(first / GROUP) * per_group, l.planes, cudaMemcpyDeviceToHost, sp));
    memset(non_n + (size_t)n * 4, 0, l.non_n - (size_t)n * 4);
    UVG_CUDA(c, cudaMemcpyAsync(non_n, c->d_non_n + first, (size_t)n * 4, cudaMemcpyDeviceToHost, sp));
    memset(reinterpret_cast<uint8_t*>(off) + (size_t)(n + 1) * 4, 0, l.off - (size_t)(n + 1) * 4);
    UVG_CUDA(c, cudaMemcpyAsync(off, c->d_ramb_off + first, (size_t)(n + 1) * 4, cudaMemcpyDeviceToHost, sp));
    if (l.ent) memset(ent + (size_t)n_amb * 4, 0, l.ent - (size_t)n_amb * 4);
    if (n_amb) UVG_CUDA(c, cudaMemcpyAsync(ent, c->d_ramb_ent + a, (size_t)n_amb * 4, cudaMemcpyDeviceToHost, sp));
    UVG_CUDA(c, cudaStreamSynchronize(sp));
    for (int64_t i = 0; i <= n; ++i) off[i] -= a;
    if (n % GROUP) {                                                                  // lanes above the range: whatever was there before
        uint4* g = reinterpret_cast<uint4*>(planes) + (size_t)(n / GROUP) * per_group;
        for (size_t u = 0; u < per_group; u += GROUP)
            for (int lane = (int)(n % GROUP); lane < GROUP; ++lane) g[u + lane] = make_uint4(0, 0, 0, 0);
    }
    return UVG_OK;
}

extern "C" int uvg_import_packed(uvg_ctx* c, const void* blob, size_t bytes, int64_t* n_refs)
{
    if (!c || !blob || bytes < sizeof(PackedHeader)) { if (c) c->err = "uvg_import_packed: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    PackedHeader h;
    memcpy(&h, blob, sizeof h);
    if (memcmp(h.magic, PACKED_MAGIC, 8) || h.version != 1) UVG_FAIL(c, UVG_EINVAL, "uvg_import_packed: not a packed reference blob (magic / version)");
    if (h.nchar != c->nchar || h.Wp != c->Wp || h.acgt_mode != (c->prm.acgt_mode != 0))
        UVG_FAIL(c, UVG_EINVAL, "uvg_import_packed: the blob was made for nchar %d, acgt_mode %d; this context has nchar %d, acgt_mode %d",
                 h.nchar, h.acgt_mode, c->nchar, (int)(c->prm.acgt_mode != 0));
    if (h.n_refs < 1 || h.n_amb < 0) UVG_FAIL(c, UVG_EINVAL, "uvg_import_packed: corrupt header");
    const int64_t n = h.n_refs, n_amb = h.n_amb;
    const PackedLayout l = packed_layout(c->Wp, n, n_amb);
    if (l.planes != h.planes_bytes || bytes < l.total) UVG_FAIL(c, UVG_EINVAL, "uvg_import_packed: truncated blob (%zu bytes, %zu expected)", bytes, l.total);
    if (c->db_n % GROUP) UVG_FAIL(c, UVG_ESTATE, "uvg_import_packed: %lld references are resident; packed blobs append at multiples of %d",
                                  (long long)c->db_n, GROUP);
    const uint8_t* planes = static_cast<const uint8_t*>(blob) + sizeof h;
    const uint8_t* non_n = planes + l.planes;
    const uint32_t* off = reinterpret_cast<const uint32_t*>(non_n + l.non_n);
    const uint8_t* ent = reinterpret_cast<const uint8_t*>(off) + l.off;
    if (off[0] != 0 || off[n] != (uint32_t)n_amb) UVG_FAIL(c, UVG_EINVAL, "uvg_import_packed: corrupt ambiguity offsets");
    int rc = ensure_db_capacity(c, c->db_n + n);
    if (rc) return rc;
    rc = ensure_amb_capacity(c, c->amb_n + n_amb);
    if (rc) return rc;
    cudaStream_t sp = c->st_push;
    UVG_CUDA(c, cudaEventRecord(c->ev_c, sp));
    const size_t per_group = (size_t)(c->Wp / 4) * NPLANE * GROUP;
    UVG_CUDA(c, cudaMemcpyAsync(c->d_db + (size_t)(c->db_n / GROUP) * per_group, planes, l.planes, cudaMemcpyHostToDevice, sp));
    UVG_CUDA(c, cudaMemcpyAsync(c->d_non_n + c->db_n, non_n, (size_t)n * 4, cudaMemcpyHostToDevice, sp));
    std::vector<uint32_t> abs_off((size_t)n + 1);
    for (int64_t i = 0; i <= n; ++i) {
        if (i && off[i] < off[i - 1]) UVG_FAIL(c, UVG_EINVAL, "uvg_import_packed: corrupt ambiguity offsets");
        abs_off[i] = off[i] + (uint32_t)c->amb_n;
    }
    UVG_CUDA(c, cudaMemcpyAsync(c->d_ramb_off + c->db_n, abs_off.data(), (size_t)(n + 1) * 4, cudaMemcpyHostToDevice, sp));
    if (n_amb) UVG_CUDA(c, cudaMemcpyAsync(c->d_ramb_ent + c->amb_n, ent, (size_t)n_amb * 4, cudaMemcpyHostToDevice, sp));
    if (c->pivot_on) {
        UVG_CUDA(c, cudaMemsetAsync(c->d_ref_cnt + c->db_n, 0, (size_t)n * 4, sp));
        UVG_CUDA(c, cudaMemsetAsync(c->d_ref_cnt + c->db2_cap + c->db_n, 0, (size_t)n * 4, sp));
        UVG_CUDA(c, launch_derive_pivot(c->d_db, c->d_db2, c->d_pivot, c->d_ref_cnt, c->d_ref_cnt + c->db2_cap, c->db_n, c->db_n + n, c->Wp, sp));
        c->stats.other_kernel_launches += 1;
    }
    UVG_CUDA(c, cudaEventRecord(c->ev_d, sp));
    UVG_CUDA(c, cudaEventSynchronize(c->ev_d));                                       // the caller may reuse the blob
    float ms = 0; cudaEventElapsedTime(&ms, c->ev_c, c->ev_d);
    c->stats.encode_ms = ms;
    c->db_n += n;
    c->amb_n += n_amb;
    if (n_refs) *n_refs = n;
    return UVG_OK;
}

// =====================================================================================================================
// search: rounds of K2 (pair counts -> score tiles of the owners) and K3 (exact replay)
// =====================================================================================================================
static int event_pair(uvg_ctx* c, size_t idx, cudaEvent_t* a, cudaEvent_t* b)
{
    while (c->pair_events.size() <= idx) {
        cudaEvent_t x, y;
        UVG_CUDA(c, cudaEventCreate(&x)); UVG_CUDA(c, cudaEventCreate(&y));
        c->pair_events.emplace_back(x, y);
    }
    *a = c->pair_events[idx].first; *b = c->pair_events[idx].second;
    return UVG_OK;
}

// sharded mode: every rank maps the score tiles of all ranks (CUDA IPC; the handles travel through a tiny NCCL all-gather)
static int ensure_peers(uvg_ctx* c)
{
    if (c->peers_ready) return UVG_OK;
    // what every rank publishes: the IPC handle of its score tiles (ranks in other processes map it), and process id, device
    // and plain pointer (ranks of the SAME process -- one host thread and one context per GPU -- cannot open their own
    // process's IPC handles: they enable peer access and use the pointer)
    struct PeerInfo { cudaIpcMemHandle_t handle; int64_t pid; uint64_t ptr; int32_t device; int32_t pad[3]; };
    static_assert(sizeof(cudaIpcMemHandle_t) == 64 && sizeof(PeerInfo) == 96, "peer record");
    PeerInfo mine{};
    UVG_CUDA(c, cudaIpcGetMemHandle(&mine.handle, c->d_tiles));
    mine.pid = (int64_t)getpid(); mine.ptr = (uint64_t)(uintptr_t)c->d_tiles; mine.device = c->prm.device;
    uint8_t* stage = static_cast<uint8_t*>(c->d_ipc);
    UVG_CUDA(c, cudaMemcpyAsync(stage, &mine, sizeof mine, cudaMemcpyHostToDevice, c->st));
    UVG_NCCL(c, g_nccl.AllGather(stage, stage + sizeof mine, sizeof mine, ncclUint8, c->comm, c->st));
    std::vector<PeerInfo> all(c->world);
    UVG_CUDA(c, cudaMemcpyAsync(all.data(), stage + sizeof mine, sizeof(PeerInfo) * (size_t)c->world, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) { c->peer_tiles[r] = c->d_tiles; c->peer_ipc[r] = false; continue; }
        if (all[r].pid == mine.pid) {
            const cudaError_t e = cudaDeviceEnablePeerAccess(all[r].device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) UVG_CUDA(c, e);
            cudaGetLastError();
            c->peer_tiles[r] = reinterpret_cast<uint4*>((uintptr_t)all[r].ptr);
            c->peer_ipc[r] = false;
            continue;
        }
        void* p = nullptr;
        UVG_CUDA(c, cudaIpcOpenMemHandle(&p, all[r].handle, cudaIpcMemLazyEnablePeerAccess));
        c->peer_tiles[r] = static_cast<uint4*>(p);
        c->peer_ipc[r] = true;
    }
    c->peers_ready = true;
    return UVG_OK;
}

// sharded literal replay: every rank maps the packed database of every rank (planes, ambiguity offsets and entries)
static void close_peer_db(uvg_ctx* c)
{
    for (int r = 0; r < MAX_RANKS; ++r) {
        const void* ptr[3] = {c->peer_db[r], c->peer_ramb_off[r], c->peer_ramb_ent[r]};
        for (int k = 0; k < 3; ++k) if (ptr[k] && c->peer_db_ipc[r][k]) cudaIpcCloseMemHandle(const_cast<void*>(ptr[k]));
        c->peer_db[r] = nullptr; c->peer_ramb_off[r] = c->peer_ramb_ent[r] = nullptr;
        for (int k = 0; k < 3; ++k) c->peer_db_ipc[r][k] = false;
    }
}
static int ensure_peer_db(uvg_ctx* c)
{
    struct DbInfo { cudaIpcMemHandle_t handle[3]; uint64_t ptr[3]; int64_t pid; int32_t device; int32_t pad; };
    static_assert(sizeof(DbInfo) == 232, "peer database record");
    int rc = sync_all(c);                                                    // nothing of an earlier search reads the old mappings
    if (rc) return rc;
    close_peer_db(c);
    DbInfo mine{};
    void* own[3] = {c->d_db, c->d_ramb_off, c->d_ramb_ent};
    for (int k = 0; k < 3; ++k) {
        mine.ptr[k] = (uint64_t)(uintptr_t)own[k];
        if (own[k]) UVG_CUDA(c, cudaIpcGetMemHandle(&mine.handle[k], own[k]));
    }
    mine.pid = (int64_t)getpid(); mine.device = c->prm.device;
    uint8_t* stage = static_cast<uint8_t*>(c->d_ipc);
    UVG_CUDA(c, cudaMemcpyAsync(stage, &mine, sizeof mine, cudaMemcpyHostToDevice, c->st));
    UVG_NCCL(c, g_nccl.AllGather(stage, stage + sizeof mine, sizeof mine, ncclUint8, c->comm, c->st));
    std::vector<DbInfo> all(c->world);
    UVG_CUDA(c, cudaMemcpyAsync(all.data(), stage + sizeof mine, sizeof(DbInfo) * (size_t)c->world, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    // every rank has passed its sync_all: nobody reads an array that was replaced since the last exchange
    for (void* p : c->deferred_free) cudaFree(p);
    c->deferred_free.clear();
    c->peer_db_live = true;
    for (int r = 0; r < c->world; ++r) {
        void* got[3] = {nullptr, nullptr, nullptr};
        for (int k = 0; k < 3; ++k) {
            if (!all[r].ptr[k]) continue;                                    // that rank holds no references (or no ambiguity entries)
            if (r == c->rank) got[k] = own[k];
            else if (all[r].pid == mine.pid) {
                const cudaError_t e = cudaDeviceEnablePeerAccess(all[r].device, 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) UVG_CUDA(c, e);
                cudaGetLastError();
                got[k] = reinterpret_cast<void*>((uintptr_t)all[r].ptr[k]);
            } else {
                UVG_CUDA(c, cudaIpcOpenMemHandle(&got[k], all[r].handle[k], cudaIpcMemLazyEnablePeerAccess));
                c->peer_db_ipc[r][k] = true;
            }
        }
        c->peer_db[r] = static_cast<const uint4*>(got[0]);
        c->peer_ramb_off[r] = static_cast<const uint32_t*>(got[1]); c->peer_ramb_ent[r] = static_cast<const uint32_t*>(got[2]);
    }
    return UVG_OK;
}

static void fill_pair_params(uvg_ctx* c, PairParams& pp, int64_t lb, int64_t le, int buf, int64_t col0)
{
    pp.db = c->d_db; pp.qdata = c->d_qdata; pp.thr = c->d_thr_all[buf];
    pp.qnib = c->d_qnib; pp.qamb_off = c->d_qamb_off; pp.qamb_ent = c->d_qamb_ent;
    pp.rb_cnt = c->d_ref_cnt; pp.mc_cnt = c->d_ref_cnt ? c->d_ref_cnt + c->db2_cap : nullptr;
    pp.wslot = c->d_wslot;
    pp.db2 = c->d_db2; pp.qstream = c->d_qstream; pp.qoff = c->d_qoff; pp.snp_off = c->d_snp_off; pp.snp_ent = c->d_snp_ent;
    pp.ramb_off = c->d_ramb_off; pp.ramb_ent = c->d_ramb_ent; pp.non_n = c->d_non_n;
    pp.qnib_stride = c->qnib_stride;
    pp.Wp = c->Wp; pp.n_chunks = c->n_chunks; pp.n_qtiles = c->n_qtiles; pp.n_slots = c->n_slots; pp.n_ptiles = c->n_ptiles;
    pp.local_begin = lb; pp.local_end = le;
    for (int r = 0; r < c->world; ++r) pp.out[r] = c->peer_tiles[r] + (size_t)buf * c->tile_elems;
    pp.row_stride = (int64_t)c->world * c->seg_cap; pp.col0 = col0;
    pp.world = c->world; pp.rows_cap = c->rows_cap;
    pp.early_abort = c->prm.early_abort; pp.work = c->d_counters + 2;
}

// one scoring round: this rank's tile [lstart, lstart+cnt) -> score tiles of the owners, then replay of the owned rows
// over the segments of all ranks.  Asynchronous on the two streams: there is no host synchronisation inside a search.
static int run_round(uvg_ctx* c, int64_t lstart, int64_t cnt, int64_t T, int64_t round_first, int64_t round_count, size_t* n_ev)
{
    const int b = (int)(c->rounds_done % NBUF);
    // buffer b was last read by the replay of round j-2: single GPU -> our own replay; sharded -> every rank's replay,
    // which the all-gather of round j-1 (issued after it on each rank's stream B) has already waited for
    UVG_CUDA(c, cudaStreamWaitEvent(c->st, c->world > 1 ? c->ev_ag[(b + 1) % NBUF] : c->ev_replay[b], 0));
    // diagnostic (UVG_DEBUG_NO_OVERLAP=1): the pair counting of round j also waits for the replay of round j-1, so that its
    // kernel times are those of an undisturbed GPU (how much the replay running beside it costs; profiles/)
    static const bool no_overlap = getenv("UVG_DEBUG_NO_OVERLAP") && atoi(getenv("UVG_DEBUG_NO_OVERLAP")) != 0;
    if (no_overlap && c->rounds_done > 0) UVG_CUDA(c, cudaStreamWaitEvent(c->st, c->ev_replay[(b + 1) % NBUF], 0));
    cudaEvent_t ea, eb;
    int rc = event_pair(c, *n_ev, &ea, &eb);
    if (rc) return rc;
    UVG_CUDA(c, cudaEventRecord(ea, c->st));
    if (cnt > 0) {
        PairParams pp{};
        fill_pair_params(c, pp, lstart, lstart + cnt, b, (int64_t)c->rank * c->seg_cap);
        UVG_CUDA(c, c->pivot_on ? launch_pairpivot(pp, c->st) : launch_paircount(pp, c->prm.acgt_mode, c->st));
        c->stats.pair_kernel_launches += 1;
        if (c->pivot_on) {   // executed (pair, 4-word unit) iterations per lane of this launch: blocks per tile x halves / 2 x references per lane
            const int64_t per = GROUP * PRPL, pairs = (lstart + cnt + per - 1) / per - lstart / per;
            c->work_host += pairs * c->pivot_halves * PRPL / 2;
        }
    }
    UVG_CUDA(c, cudaEventRecord(eb, c->st));
    ++*n_ev;
    UVG_CUDA(c, cudaEventRecord(c->ev_pair[b], c->st));

    UVG_CUDA(c, cudaStreamWaitEvent(c->st_rp, c->ev_pair[b], 0));
    if (c->world > 1) {
        // bounds of the owned rows (state after round j-1) to everybody; doubles as the barrier "all tiles of round j are
        // written".  K2 of round j+1 reads buffer (j+1) % 2.
        UVG_NCCL(c, g_nccl.AllGather(c->d_thr_local, c->d_thr_all[(b + 1) % NBUF], (size_t)c->rows_cap * sizeof(Threshold), ncclUint8,
                                     c->comm, c->st_rp));
        UVG_CUDA(c, cudaEventRecord(c->ev_ag[b], c->st_rp));
    }
    ReplayParams rp{};
    rp.tile = c->d_tiles + (size_t)b * c->tile_elems; rp.row_stride = (int64_t)c->world * c->seg_cap;
    rp.nseg = c->world; rp.seg_stride = (int32_t)c->seg_cap;
    for (int g = 0; g < c->world; ++g) {
        const int64_t gb = round_first + (int64_t)g * T;
        rp.seg_global[g] = gb;
        rp.seg_cnt[g] = (int32_t)std::max<int64_t>(0, std::min<int64_t>(T, round_first + round_count - gb));
    }
    rp.rows_cap = c->rows_cap;
    rp.heaps = c->d_heaps; rp.heap_n = c->d_heap_n; rp.max_inc = c->d_max_inc;
    rp.batch_id = c->d_batch_id; rp.max_inc_batch = c->d_max_inc_batch;
    rp.thr = c->world > 1 ? c->d_thr_local : c->d_thr_all[b];              // single GPU: straight into what K2 of round j+2 reads
    rp.ever_best = c->d_ever; rp.counters = c->d_counters;
    rp.n_query = c->n_query; rp.world = c->world; rp.rank = c->rank;
    rp.K = c->K; rp.acgt_mode = c->prm.acgt_mode; rp.pool = c->prm.pool;
    cudaEvent_t ra, rb;
    rc = event_pair(c, *n_ev, &ra, &rb);
    if (rc) return rc;
    UVG_CUDA(c, cudaEventRecord(ra, c->st_rp));
    rp.ord_begin = round_first; rp.ord_end = round_first + round_count;
    // the reference's batches that touch this round: --pool accepted references, restarting with every file
    // (nearest.c:245-286).  pieces = the round cut at the batch boundaries; bounds = first ordinals of those batches
    struct Piece { int64_t o, e; bool start; };
    std::vector<Piece> pieces;
    std::vector<int64_t> bounds;
    {
        const int64_t pool = c->prm.pool, end = round_first + round_count;
        size_t k = std::upper_bound(c->file_starts.begin(), c->file_starts.end(), round_first) - c->file_starts.begin() - 1;
        for (int64_t o = round_first; o < end;) {
            while (k + 1 < c->file_starts.size() && c->file_starts[k + 1] <= o) ++k;
            const int64_t fs = c->file_starts[k];
            int64_t nb = fs + ((o - fs) / pool + 1) * pool;
            if (k + 1 < c->file_starts.size()) nb = std::min(nb, c->file_starts[k + 1]);
            pieces.push_back({o, std::min(end, nb), (o - fs) % pool == 0});
            bounds.push_back(fs + (o - fs) / pool * pool);                    // first ordinal of the batch that holds o
            if (nb >= end) bounds.push_back(nb);
            o = std::min(end, nb);
        }
    }
    if (!bounds.empty()) {
        // device copy of the bounds: appended to a buffer that lives until the heaps are emptied (rounds in flight read it)
        const size_t need = c->bounds_used + bounds.size();
        if (need > c->bounds_cap) {
            int rc2 = sync_all(c);
            if (rc2) return rc2;
            const size_t cap = std::max<size_t>(need * 2, 1 << 16);
            int64_t* nd = nullptr; int64_t* nh = nullptr;
            UVG_CUDA(c, cudaMalloc(&nd, cap * 8));
            UVG_CUDA(c, cudaMallocHost(&nh, cap * 8));
            cudaFree(c->d_bounds); if (c->h_bounds) cudaFreeHost(c->h_bounds);
            c->d_bounds = nd; c->h_bounds = nh; c->bounds_cap = cap; c->bounds_used = 0;
        }
        memcpy(c->h_bounds + c->bounds_used, bounds.data(), bounds.size() * 8);
        UVG_CUDA(c, cudaMemcpyAsync(c->d_bounds + c->bounds_used, c->h_bounds + c->bounds_used, bounds.size() * 8, cudaMemcpyHostToDevice, c->st_rp));
        rp.bounds = c->d_bounds + c->bounds_used; rp.n_bounds = (int32_t)bounds.size();
        c->bounds_used += bounds.size();
    }
    if (!c->literal) {
        UVG_CUDA(c, launch_replay(rp, c->st_rp));
        c->stats.other_kernel_launches += 1;
    } else if (round_count > 0) {
        // Literal emulation of the reference's batches (hazard H1), speculatively.  The reference's result can differ from the
        // pool-independent closed form only for pairs the closed-form pass FLAGS (a superset, replay.cu); rounds without a flag
        // -- all of them for ordinary query sets -- are settled by ONE closed-form launch.  So: (1) keep the state of the
        // round's start, (2) closed-form pass over the whole round with its admissions marked in a scratch array, (3) commit if it
        // raised no flag, else (4) put the state back and replay the round literally, one launch per piece of a batch, the
        // batch-wide limit travelling through d_bstate.  Everything is decided on the device: the launches of (4) return at once
        // when the round is settled, no host synchronisation.
        const size_t rows = (size_t)c->rows_cap;
        UVG_CUDA(c, cudaMemcpyAsync(c->d_snap_own, c->d_own, c->own_bytes, cudaMemcpyDeviceToDevice, c->st_rp));
        UVG_CUDA(c, cudaMemcpyAsync(c->d_snap_batch, c->d_batch_id, rows * 4, cudaMemcpyDeviceToDevice, c->st_rp));
        UVG_CUDA(c, cudaMemcpyAsync(c->d_snap_batch + rows, c->d_max_inc_batch, rows * 4, cudaMemcpyDeviceToDevice, c->st_rp));
        UVG_CUDA(c, cudaMemsetAsync(c->d_spec, 0, 16, c->st_rp));
        UVG_CUDA(c, cudaMemsetAsync(c->d_ever_round, 0, (size_t)round_count, c->st_rp));
        ReplayParams sp = rp;
        sp.counters = c->d_spec; sp.ever_round = c->d_ever_round; sp.ord_round0 = round_first;
        UVG_CUDA(c, launch_replay(sp, c->st_rp));
        UVG_CUDA(c, launch_spec_commit(c->d_spec, c->d_ever_round, round_count, round_first, c->d_ever, c->d_counters, c->st_rp));
        const int64_t i0 = c->replay_seq;
        SpecRestore sr{};
        sr.spec = c->d_spec; sr.snap_own = reinterpret_cast<const uint4*>(c->d_snap_own); sr.own = reinterpret_cast<uint4*>(c->d_own);
        sr.own_vec = (int64_t)(c->own_bytes / 16);
        sr.snap_batch_id = c->d_snap_batch; sr.batch_id = c->d_batch_id; sr.snap_mib = c->d_snap_batch + rows; sr.mib = c->d_max_inc_batch; sr.rows = (int32_t)rows;
        sr.snap_max_inc = reinterpret_cast<const int32_t*>(c->d_snap_own + c->own_off_mi);
        sr.n_query = c->n_query;
        sr.open_batch = pieces[0].start ? -1 : (int32_t)bounds[0];
        sr.b_prev = c->d_bstate + (i0 + 2) % 3; sr.b_carry = c->d_bstate + 4 + (i0 + 1) % 2; sr.b_acc = c->d_bstate + i0 % 3;
        UVG_CUDA(c, launch_spec_restore(sr, c->st_rp));
        rp.literal = 1; rp.spec_flags = c->d_spec + 1;
        rp.seg_local0 = lstart;
        for (int g = 0; g < c->world; ++g) {
            rp.seg_db[g] = c->world > 1 ? c->peer_db[g] : c->d_db;
            rp.seg_ramb_off[g] = c->world > 1 ? c->peer_ramb_off[g] : c->d_ramb_off;
            rp.seg_ramb_ent[g] = c->world > 1 ? c->peer_ramb_ent[g] : c->d_ramb_ent;
        }
        rp.qcons = c->d_qdata + (size_t)(c->n_query / QTILE) * c->n_chunks; rp.cons_qi = c->n_query % QTILE;
        rp.cons_nib = c->d_qnib + (size_t)c->n_query * c->qnib_stride;
        rp.camb = c->d_qamb_ent + c->cons_amb_begin; rp.n_camb = (int32_t)c->cons_amb_n;
        rp.Wp = c->Wp;
        for (const Piece& pc : pieces) {
            const int64_t i = c->replay_seq++;
            rp.ord_begin = pc.o; rp.ord_end = pc.e; rp.batch_start = pc.start;
            rp.b_prev = c->d_bstate + (i + 2) % 3; rp.b_acc = c->d_bstate + i % 3; rp.b_zero = c->d_bstate + (i + 1) % 3;
            rp.b_carry_in = c->d_bstate + 4 + (i + 1) % 2; rp.b_carry_out = c->d_bstate + 4 + i % 2;
            UVG_CUDA(c, launch_replay(rp, c->st_rp));
        }
        c->stats.other_kernel_launches += 3 + (int64_t)pieces.size();
    }
    UVG_CUDA(c, cudaEventRecord(rb, c->st_rp));
    ++*n_ev;
    UVG_CUDA(c, cudaEventRecord(c->ev_replay[b], c->st_rp));
    return UVG_OK;
}

// enqueue the rounds that cover every resident reference not scored yet; returns without waiting
static int search_enqueue(uvg_ctx* c, bool taper)
{
    if (c->n_query == 0) UVG_FAIL(c, UVG_ESTATE, "uvg_search: no queries set");
    if (c->world > 1 && !c->comm) UVG_FAIL(c, UVG_ESTATE, "uvg_search: sharded context without uvg_nccl_init");
    if (c->world > 1 && c->global_total < 0) UVG_FAIL(c, UVG_ESTATE, "uvg_search: sharded context without uvg_set_global_count");
    int rc;
    if (c->world > 1) { rc = ensure_peers(c); if (rc) return rc; }
    if (c->world > 1 && c->literal) { rc = ensure_peer_db(c); if (rc) return rc; }       // collective: the arrays move when they grow
    const int32_t mx = (int32_t)std::min<int64_t>(round_size_max(c->prm), c->seg_cap);
    if (c->next_round > mx) c->next_round = mx;
    c->heaps_cached = false;
    // every array the rounds touch is sized before the first launch (growing one synchronises everything first)
    {
        const int64_t upto = c->world == 1 ? c->global_base + c->db_n : c->global_total;
        if (upto > 0x7fffffffLL) UVG_FAIL(c, UVG_EINVAL, "more than 2^31 references are not supported");
        rc = ensure_ever(c, upto);
        if (rc) return rc;
    }
    if (!c->search_open) {
        c->search_open = true; c->n_ev = 0;
        c->stats.pair_kernel_ms = 0; c->stats.replay_ms = 0;
        UVG_CUDA(c, cudaEventRecord(c->ev_a, c->st));
    }
    for (;;) {
        // single GPU: uvg_search knows that the resident rows are all there is and tapers the rounds towards their end;
        // uvg_search_async (more rows follow) keeps the nominal sizes
        const int64_t T = c->world == 1 ? (taper ? round_size_at(c->db_scored, c->next_round, c->db_n, 1) : (int64_t)c->next_round) : round_size_at(c->round_global_start, c->next_round, c->prm.nominal_tail ? -1 : c->global_total, c->world);
        int64_t lstart, cnt, rfirst, rcount;
        if (c->world == 1) {
            cnt = std::min<int64_t>(T, c->db_n - c->db_scored);
            if (cnt <= 0) break;
            lstart = c->db_scored; rfirst = c->global_base + lstart; rcount = cnt;
        } else {
            const int64_t S = c->round_global_start;
            if (S >= c->global_total) break;
            if (c->search_limit >= 0 && std::min(S + (int64_t)c->world * T, c->global_total) > c->search_limit) break;   // its rows are still on their way
            const int64_t gb = S + (int64_t)c->rank * T, ge = std::min(gb + T, c->global_total);
            cnt = std::max<int64_t>(0, ge - gb);
            lstart = c->round_local_start; rfirst = S; rcount = std::min<int64_t>((int64_t)c->world * T, c->global_total - S);
            if (lstart + cnt > c->db_n) UVG_FAIL(c, UVG_ESTATE, "rank %d holds %lld references but round %lld needs local rows up to %lld",
                                                 c->rank, (long long)c->db_n, (long long)c->rounds_done, (long long)(lstart + cnt));
        }
        rc = run_round(c, lstart, cnt, T, rfirst, rcount, &c->n_ev);
        if (rc) return rc;
        if (c->world > 1 && c->round_global_start + (int64_t)c->world * T >= c->global_total && lstart + cnt != c->db_n)
            UVG_FAIL(c, UVG_ESTATE, "rank %d holds %lld references but the round schedule (first_round %d, max_round %d, %lld in all) gives it %lld: "
                     "the rows were dealt with another schedule (uvg_shard_of)", c->rank, (long long)c->db_n, round_size_first(c->prm), round_size_max(c->prm),
                     (long long)c->global_total, (long long)(lstart + cnt));
        if (c->world == 1) { c->db_scored += cnt; c->stats.n_refs += cnt; }
        else { c->round_global_start += (int64_t)c->world * T; c->round_local_start += T; c->db_scored = std::min(c->db_n, c->round_local_start); c->stats.n_refs += rcount; }
        ++c->rounds_done; ++c->stats.n_rounds;
        c->next_round = round_size_next(c->next_round, mx);
    }
    return UVG_OK;
}

extern "C" int uvg_search_async(uvg_ctx* c)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    return search_enqueue(c, false);
}

extern "C" int uvg_wait(uvg_ctx* c)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (!c->search_open) return sync_all(c);
    // the search ends when the last replay has finished: join stream B into stream A
    UVG_CUDA(c, cudaEventRecord(c->ev_d, c->st_rp));
    UVG_CUDA(c, cudaStreamWaitEvent(c->st, c->ev_d, 0));
    if (c->fin_device && c->n_query > 0) {
        // the final tables are made where the heaps are: slot swap + stable descending sort per row, sizes and limits behind them
        UVG_CUDA(c, launch_finalise(c->d_heaps, c->d_heap_n, c->K, c->rows_cap, c->d_fin, c->st));
        UVG_CUDA(c, cudaMemcpyAsync(c->d_fin + c->fin_off_n, c->d_heap_n, (size_t)c->rows_cap * 4, cudaMemcpyDeviceToDevice, c->st));
        UVG_CUDA(c, cudaMemcpyAsync(c->d_fin + c->fin_off_mi, c->d_max_inc, (size_t)c->rows_cap * 4, cudaMemcpyDeviceToDevice, c->st));
        c->stats.other_kernel_launches += 1;
    }
    if (c->world > 1) {
        // the tables of all queries to every rank: ONE all-gather over NVLink (heaps, sizes and limits are one block).  The
        // ever-best flags are combined lazily by uvg_get_ever_best (nobody needs them inside the search).
        UVG_NCCL(c, g_nccl.AllGather(c->fin_device ? c->d_fin : c->d_own, c->d_all, c->fin_device ? c->fin_bytes : c->own_bytes, ncclUint8, c->comm, c->st));
        c->ever_reduced = false;
    }
    UVG_CUDA(c, cudaEventRecord(c->ev_b, c->st));
    UVG_CUDA(c, cudaMemcpyAsync(c->h_pinned + 20, c->d_counters + 2, 16, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaEventSynchronize(c->ev_b));
    int rc = sync_all(c);
    if (rc) return rc;
    c->search_open = false;
    // the final all-gather came after every rank's last replay: nobody reads an array that grew (and moved) during the search
    for (void* p : c->deferred_free) cudaFree(p);
    c->deferred_free.clear();
    {   // executed (pair, 4-word unit) iterations per lane: whole units + halves / 2 (pairpivot.cu counts halves of two words)
        const int64_t total = (int64_t)c->h_pinned[20] + c->work_host;
        c->work_last = total - c->work_seen;
        c->work_seen = total;
    }
    float ms = 0; cudaEventElapsedTime(&ms, c->ev_a, c->ev_b);
    c->stats.search_ms = ms;
    for (size_t i = 0; i < c->n_ev; ++i) {        // events alternate: pair-count launch, replay launch
        float t = 0; cudaEventElapsedTime(&t, c->pair_events[i].first, c->pair_events[i].second);
        if (i % 2 == 0) c->stats.pair_kernel_ms += t; else c->stats.replay_ms += t;
    }
    c->n_ev = 0;
    return UVG_OK;
}

extern "C" int uvg_search(uvg_ctx* c)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    int rc = search_enqueue(c, true);
    if (rc) return rc;
    return uvg_wait(c);
}

// =====================================================================================================================
// results
// =====================================================================================================================
static_assert(sizeof(uvg_item) == 32, "k_finalise writes uvg_item rows");
static int compare_items_desc(const void* pa, const void* pb)
{   // compare_q_item_score (src/min_heap.c:41-47): first non-zero b[i]-a[i]
    const HeapItem* a = static_cast<const HeapItem*>(pa);
    const HeapItem* b = static_cast<const HeapItem*>(pb);
    for (int i = 0; i < 6; ++i) {
        int r = b->s[i] - a->s[i];
        if (r) return r;
    }
    return 0;
}

static int cache_heaps(uvg_ctx* c)
{
    if (c->search_open) { int rc = uvg_wait(c); if (rc) return rc; }
    if (c->heaps_cached) return UVG_OK;
    const size_t rows_r = (size_t)c->rows_cap, rows = (size_t)c->world * rows_r, stride = c->K + 1;
    c->h_heap_n.resize(rows); c->h_max_inc.resize(rows);
    // one device -> host copy of the result blocks of all ranks (pinned); sizes and limits are sliced out on the host
    const size_t blk_bytes = c->fin_device ? c->fin_bytes : c->own_bytes, off_n = c->fin_device ? c->fin_off_n : c->own_off_n,
                 off_mi = c->fin_device ? c->fin_off_mi : c->own_off_mi;
    const size_t total = (size_t)c->world * blk_bytes;
    if (total > c->h_all_bytes) {
        if (c->h_all) cudaFreeHost(c->h_all);
        c->h_all = nullptr; c->h_all_bytes = 0;
        UVG_CUDA(c, cudaMallocHost(&c->h_all, total));
        c->h_all_bytes = total;
    }
    const uint8_t* src = c->world > 1 ? c->d_all : (c->fin_device ? c->d_fin : c->d_own);
    UVG_CUDA(c, cudaMemcpyAsync(c->h_all, src, total, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    if (!c->fin_device) c->h_heaps.resize(rows * stride);
    for (int r = 0; r < c->world; ++r) {
        const uint8_t* blk = c->h_all + (size_t)r * blk_bytes;
        if (!c->fin_device) memcpy(c->h_heaps.data() + (size_t)r * rows_r * stride, blk, rows_r * stride * sizeof(HeapItem));
        memcpy(c->h_heap_n.data() + (size_t)r * rows_r, blk + off_n, rows_r * 4);
        memcpy(c->h_max_inc.data() + (size_t)r * rows_r, blk + off_mi, rows_r * 4);
    }
    c->heaps_cached = true;
    return UVG_OK;
}

static size_t row_of_query(const uvg_ctx* c, int iq)
{
    const int qt = iq / QTILE;
    return (size_t)(qt % c->world) * c->rows_cap + (size_t)(qt / c->world) * QTILE + iq % QTILE;
}

// heap_finalise_heap_qsort (src/min_heap.c:149-153) of one cached heap: slot 0 <-> slot n, then libc qsort over [0, n)
static int finalise_row(const uvg_ctx* c, int iq, uvg_item* out)
{
    const size_t row = row_of_query(c, iq);
    const int n = c->h_heap_n[row];
    if (n == 0) return 0;
    if (c->fin_device) {                                                      // sorted on the device (k_finalise): rows of rank r start at block r
        const size_t r = row / c->rows_cap, lrow = row % c->rows_cap;
        memcpy(out, c->h_all + r * c->fin_bytes + lrow * c->K * sizeof(uvg_item), (size_t)n * sizeof(uvg_item));
        return n;
    }
    const HeapItem* a = c->h_heaps.data() + row * (c->K + 1);
    std::vector<HeapItem> arr(n);
    arr[0] = a[n];
    for (int i = 1; i < n; ++i) arr[i] = a[i];
    qsort(arr.data(), n, sizeof(HeapItem), compare_items_desc);
    for (int i = 0; i < n; ++i) {
        for (int k = 0; k < 6; ++k) out[i].score[k] = arr[i].s[k];
        out[i].ref = arr[i].r;
    }
    return n;
}

extern "C" int uvg_get_heap(uvg_ctx* c, int32_t iq, uvg_item* out, int32_t* n_out)
{
    if (!c || !out || !n_out) { if (c) c->err = "uvg_get_heap: bad arguments"; return UVG_EINVAL; }
    if (iq < 0 || iq >= c->n_query) UVG_FAIL(c, UVG_EINVAL, "uvg_get_heap: query %d out of range", iq);
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    int rc = cache_heaps(c);
    if (rc) return rc;
    *n_out = finalise_row(c, iq, out);
    return UVG_OK;
}

// every query's table in one call: out holds n_query x max(2, nbest) items (row iq at out + iq * max(2, nbest)), n_out[iq] rows
extern "C" int uvg_get_all_heaps(uvg_ctx* c, uvg_item* out, int32_t* n_out)
{
    if (!c || !out || !n_out) { if (c) c->err = "uvg_get_all_heaps: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    int rc = cache_heaps(c);
    if (rc) return rc;
    if (c->fin_device) {                                                      // whole query tiles are contiguous on both sides
        for (int qt = 0; qt < c->n_qtiles; ++qt) {
            const int q0 = qt * QTILE, nq = std::min(QTILE, c->n_query - q0);
            if (nq <= 0) break;
            const size_t row = row_of_query(c, q0), r = row / c->rows_cap, lrow = row % c->rows_cap;
            memcpy(out + (size_t)q0 * c->K, c->h_all + r * c->fin_bytes + lrow * c->K * sizeof(uvg_item), (size_t)nq * c->K * sizeof(uvg_item));
            for (int i = 0; i < nq; ++i) n_out[q0 + i] = c->h_heap_n[row + i];
        }
        return UVG_OK;
    }
    parallel_ranges(c->n_query, host_threads(c->n_query / 64), [&](int, int qa, int qb) {
        for (int iq = qa; iq < qb; ++iq) n_out[iq] = finalise_row(c, iq, out + (size_t)iq * c->K);
    });
    return UVG_OK;
}

extern "C" int uvg_get_ever_best(uvg_ctx* c, int64_t first, int64_t n, uint8_t* flags)
{
    if (!c || !flags || first < 0 || n < 0) { if (c) c->err = "uvg_get_ever_best: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (c->search_open) { int rc = uvg_wait(c); if (rc) return rc; }
    if (c->world > 1 && !c->ever_reduced) {
        // sharded: every rank marked the references its own queries admitted; combine (collective: every rank calls this)
        if (c->global_total > 0) UVG_NCCL(c, g_nccl.AllReduce(c->d_ever, c->d_ever, (size_t)c->global_total, ncclUint8, ncclMax, c->comm, c->st));
        c->ever_reduced = true;
    }
    memset(flags, 0, (size_t)n);
    const int64_t m = std::min(n, c->ever_cap - first);
    if (m > 0) {
        UVG_CUDA(c, cudaMemcpyAsync(flags, c->d_ever + first, (size_t)m, cudaMemcpyDeviceToHost, c->st));
        UVG_CUDA(c, cudaStreamSynchronize(c->st));
    }
    return UVG_OK;
}

extern "C" int uvg_get_stats(uvg_ctx* c, uvg_stats* out)
{
    if (!c || !out) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (c->search_open) { int rc = uvg_wait(c); if (rc) return rc; }
    UVG_CUDA(c, cudaMemcpyAsync(c->h_pinned + 16, c->d_counters, 24, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    c->stats.n_admitted = (int64_t)c->h_pinned[16]; c->stats.n_h1_flags = (int64_t)c->h_pinned[17];
    c->stats.warp_chunks_executed = c->work_last / 32;          // the counter is in (pair, 4-word unit) per lane; 16 pairs x 2 units = one (warp, chunk) of the general kernel
    c->stats.pivot_fix_entries = c->pivot_on ? c->qinfo_pivot_fix : -1;
    c->stats.set_queries_host_ms = c->set_queries_host_ms; c->stats.set_queries_device_ms = c->set_queries_device_ms;
    if (c->n_query) {
        int rc = cache_heaps(c);
        if (rc) return rc;
        int64_t mm = 0;
        for (int q = 0; q < c->n_query; ++q) mm = std::max<int64_t>(mm, c->h_max_inc[row_of_query(c, q)] - 1);
        c->stats.max_mismatch = mm;
    }
    *out = c->stats;
    return UVG_OK;
}

// =====================================================================================================================
// sharding helpers + NCCL
// =====================================================================================================================
extern "C" int uvg_shard_of(int64_t ordinal, int64_t n_total, int32_t world, int32_t first_round, int32_t max_round, int32_t* rank, int64_t* local)
{
    if (ordinal < 0 || world < 1 || (n_total >= 0 && ordinal >= n_total)) return UVG_EINVAL;
    uvg_params p{}; p.first_round = first_round; p.max_round = max_round;
    int64_t T = round_size_first(p), S = 0, Ls = 0;
    const int32_t mx = round_size_max(p);
    for (;;) {
        const int64_t Te = round_size_at(S, T, n_total, world);
        if (ordinal < S + (int64_t)world * Te) {
            if (rank) *rank = (int32_t)((ordinal - S) / Te);
            if (local) *local = Ls + (ordinal - S) % Te;
            return UVG_OK;
        }
        S += (int64_t)world * Te; Ls += Te;
        T = round_size_next((int32_t)T, mx);
    }
}

extern "C" int uvg_nccl_unique_id(void* out128)
{
    if (!out128) return UVG_EINVAL;
    if (!g_nccl.load(g_create_error)) return UVG_ENCCL;
    ncclUniqueId id;
    if (g_nccl.GetUniqueId(&id) != 0) { g_create_error = "ncclGetUniqueId failed"; return UVG_ENCCL; }
    memcpy(out128, &id, 128);
    return UVG_OK;
}

extern "C" int uvg_nccl_init(uvg_ctx* c, const void* id128)
{
    if (!c || !id128) return UVG_EINVAL;
    if (c->world < 2) UVG_FAIL(c, UVG_ESTATE, "uvg_nccl_init on a single-GPU context");
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (!g_nccl.load(c->err)) return UVG_ENCCL;
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    UVG_NCCL(c, g_nccl.CommInitRank(&c->comm, c->world, id, c->rank));
    return UVG_OK;
}

extern "C" int uvg_set_global_count(uvg_ctx* c, int64_t n_total)
{
    if (!c || n_total < 0) return UVG_EINVAL;
    c->global_total = n_total;
    return UVG_OK;
}

extern "C" int uvg_set_search_limit(uvg_ctx* c, int64_t n_resident)
{
    if (!c) return UVG_EINVAL;
    c->search_limit = n_resident < 0 ? -1 : n_resident;
    return UVG_OK;
}

// =====================================================================================================================
// bench / test utilities
// =====================================================================================================================
extern "C" int uvg_synth_rows_device(uvg_ctx* c, const uint8_t* clades, int32_t n_clades, uint64_t seed, int64_t first, int64_t n,
                                     double n_frac_mean, double ambiguous_frac, uint8_t* d_text)
{
    if (!c || !clades || n_clades < 1 || n < 0 || !d_text) { if (c) c->err = "uvg_synth_rows_device: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    uint8_t* d_clades = nullptr;
    const size_t bytes = (size_t)n_clades * c->nchar;
    UVG_CUDA(c, cudaMalloc(&d_clades, bytes));
    UVG_CUDA(c, cudaMemcpyAsync(d_clades, clades, bytes, cudaMemcpyHostToDevice, c->st));
    cudaError_t e = launch_synth_rows(d_clades, n_clades, c->nchar, seed, first, n, n_frac_mean, ambiguous_frac, d_text, c->st);
    cudaError_t e2 = cudaStreamSynchronize(c->st);
    cudaFree(d_clades);
    UVG_CUDA(c, e); UVG_CUDA(c, e2);
    return UVG_OK;
}

extern "C" void* uvg_device_alloc(uvg_ctx* c, size_t bytes)
{
    if (!c) return nullptr;
    void* p = nullptr;
    if (cudaSetDevice(c->prm.device) != cudaSuccess) return nullptr;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) { c->err = std::string("cudaMalloc: ") + cudaGetErrorString(e); return nullptr; }
    return p;
}
extern "C" void uvg_device_free(uvg_ctx* c, void* p) { if (c) { cudaSetDevice(c->prm.device); cudaFree(p); } }
extern "C" int uvg_memcpy_d2h(uvg_ctx* c, void* dst, const void* d_src, size_t bytes)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    UVG_CUDA(c, cudaMemcpyAsync(dst, d_src, bytes, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    return UVG_OK;
}
extern "C" int uvg_memcpy_h2d(uvg_ctx* c, void* d_dst, const void* src, size_t bytes)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    UVG_CUDA(c, cudaMemcpyAsync(d_dst, src, bytes, cudaMemcpyHostToDevice, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    return UVG_OK;
}
extern "C" void* uvg_host_alloc_pinned(size_t bytes) { void* p = nullptr; return cudaMallocHost(&p, bytes) == cudaSuccess ? p : nullptr; }
extern "C" void uvg_host_free_pinned(void* p) { if (p) cudaFreeHost(p); }

// scores of resident references [first, first+n) against every slot into tile buffer 0 (test hooks only; single GPU)
static int debug_score(uvg_ctx* c, int64_t first, int64_t n, const char* who)
{
    if (c->world > 1) UVG_FAIL(c, UVG_ESTATE, "%s: not available in sharded mode", who);
    if (c->n_query == 0 || first + n > c->db_n || n > c->seg_cap) UVG_FAIL(c, UVG_ESTATE, "%s: range not resident, too long, or no queries", who);
    int rc = sync_all(c);
    if (rc) return rc;
    PairParams pp{};
    fill_pair_params(c, pp, first, first + n, 0, 0);
    pp.early_abort = 0; pp.work = nullptr;
    UVG_CUDA(c, c->pivot_on ? launch_pairpivot(pp, c->st) : launch_paircount(pp, c->prm.acgt_mode, c->st));
    return UVG_OK;
}

extern "C" int uvg_debug_pair_counts_rows(uvg_ctx* c, int32_t q_first, int32_t q_count, int64_t first, int64_t n, int32_t* out)
{
    if (!c || !out || first < 0 || n < 1 || q_first < 0 || q_count < 1) { if (c) c->err = "uvg_debug_pair_counts_rows: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (q_first + q_count > c->n_query) UVG_FAIL(c, UVG_ESTATE, "uvg_debug_pair_counts_rows: no such queries");
    int rc = debug_score(c, first, n, "uvg_debug_pair_counts_rows");
    if (rc) return rc;
    UVG_CUDA(c, cudaMemcpy2DAsync(out, (size_t)n * 16, c->d_tiles + (size_t)q_first * c->seg_cap, (size_t)c->seg_cap * 16,
                                  (size_t)n * 16, (size_t)q_count, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    return UVG_OK;
}

extern "C" int uvg_debug_pair_counts(uvg_ctx* c, int64_t first, int64_t n, int32_t* out)
{
    if (!c) return UVG_EINVAL;
    return uvg_debug_pair_counts_rows(c, 0, c->n_query, first, n, out);
}

extern "C" int uvg_debug_ref_meta(uvg_ctx* c, int64_t first, int64_t n, int32_t* non_n, int32_t* c0, int32_t* ci)
{
    if (!c || first < 0 || n < 1) { if (c) c->err = "uvg_debug_ref_meta: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    int rc = debug_score(c, first, n, "uvg_debug_ref_meta");
    if (rc) return rc;
    std::vector<uint4> cons((size_t)n);
    UVG_CUDA(c, cudaMemcpyAsync(cons.data(), c->d_tiles + (size_t)c->rows_cap * c->seg_cap, (size_t)n * 16, cudaMemcpyDeviceToHost, c->st));
    if (non_n) UVG_CUDA(c, cudaMemcpyAsync(non_n, c->d_non_n + first, (size_t)n * 4, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    const bool acgt = c->prm.acgt_mode != 0;
    for (int64_t i = 0; i < n; ++i) {
        if (c0) c0[i] = (int32_t)cons[i].x;
        if (ci) ci[i] = acgt ? (int32_t)(cons[i].y - cons[i].x) : (int32_t)(cons[i].w - cons[i].x);
    }
    return UVG_OK;
}

extern "C" int uvg_debug_unpack_ref(uvg_ctx* c, int64_t local, uint8_t* codes_out)
{
    if (!c || !codes_out || local < 0 || local >= c->db_n) { if (c) c->err = "uvg_debug_unpack_ref: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    uint8_t* d = nullptr;
    UVG_CUDA(c, cudaMalloc(&d, c->nchar));
    cudaError_t e = launch_unpack_ref(c->d_db, c->d_ramb_off, c->d_ramb_ent, local, c->nchar, c->Wp, d, c->st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(codes_out, d, c->nchar, cudaMemcpyDeviceToHost, c->st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->st);
    cudaFree(d);
    UVG_CUDA(c, e);
    return UVG_OK;
}
