// Host side of the C ABI (include/uvaia_b200.h): context, query preparation, database residency, round driver.
// Replaces the orchestration of src/nearest.c:233-338 (queue_t, batch loop, two OpenMP loops) around the kernels
// of encode.cu (K1), paircount.cu (K2) and replay.cu (K3).  No scoring happens on the host: without a CUDA device
// uvg_create() fails (there is no CPU fallback by design).
#include "../../include/uvaia_b200.h"
#include "common.cuh"
#include "kernels.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <dlfcn.h>

using namespace uvg;

namespace {
std::string g_create_error;

// ---- NCCL through dlopen (the library loads without NCCL for single-GPU use) ------------------------------------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { ncclInt8 = 0, ncclUint8 = 1, ncclInt32 = 2, ncclUint32 = 3, ncclInt64 = 4, ncclUint64 = 5 };
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
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
        CommDestroy = (decltype(CommDestroy))dlsym(handle, "ncclCommDestroy");
        GetErrorString = (decltype(GetErrorString))dlsym(handle, "ncclGetErrorString");
        if (!GetUniqueId || !CommInitRank || !AllGather || !CommDestroy) { err = "NCCL symbols missing"; return false; }
        return true;
    }
};
NcclApi g_nccl;
}  // namespace

struct uvg_ctx {
    uvg_params prm{};
    std::string err;
    int K = 0, nchar = 0, trim = 0, W = 0, Wp = 0, n_chunks = 0, ambw = 0;
    int rank = 0, world = 1;
    cudaStream_t st = nullptr;
    cudaEvent_t ev_a = nullptr, ev_b = nullptr, ev_c = nullptr, ev_d = nullptr;
    // queries
    int n_query = 0, n_qtiles = 0;
    uvg_query_info qinfo{};
    QChunk* d_qdata = nullptr;
    uint4* d_cons_acgt = nullptr; uint2* d_cons_vo = nullptr;
    Threshold* d_thr = nullptr;
    HeapItem* d_heaps = nullptr;
    int32_t *d_heap_n = nullptr, *d_max_inc = nullptr, *d_batch_id = nullptr, *d_max_inc_batch = nullptr;
    // database (local shard)
    int64_t db_cap = 0, db_n = 0, db_scored = 0, global_base = 0, global_total = -1;
    uint4* d_db = nullptr; uint32_t* d_refamb = nullptr;
    int32_t *d_non_n = nullptr, *d_c0 = nullptr, *d_ci = nullptr;
    uint8_t* d_ever = nullptr; int64_t ever_cap = 0;
    // dense score tile of one round: [n_query][score_cap] x (4 counts)
    int64_t score_cap = 0;
    uint4* d_scores = nullptr;
    unsigned long long* d_counters = nullptr;
    unsigned long long* h_pinned = nullptr;          // small pinned read-back area
    // push staging
    uint8_t* d_stage_text = nullptr; size_t stage_text_bytes = 0;
    int32_t* d_stage_nn = nullptr; uint8_t* d_stage_acc = nullptr; int64_t* d_stage_dst = nullptr; int64_t stage_rows = 0;
    int64_t* d_nacc = nullptr;
    // round schedule
    int64_t rounds_done = 0, round_local_start = 0, round_global_start = 0;
    int32_t next_round = 0;
    // stats
    uvg_stats stats{};
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pair_events;
    // nccl
    ncclComm_t comm = nullptr;
    int32_t* dbg_counts = nullptr;
    // host cache of the heaps (filled lazily by uvg_get_heap, invalidated by every search / reset)
    int64_t work_base = 0;
    bool heaps_cached = false;
    std::vector<HeapItem> h_heaps; std::vector<int32_t> h_heap_n, h_max_inc;
};

#define UVG_FAIL(ctx, code, ...) do { char b_[512]; snprintf(b_, sizeof b_, __VA_ARGS__); (ctx)->err = b_; return (code); } while (0)
#define UVG_CUDA(ctx, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { \
    char b_[512]; snprintf(b_, sizeof b_, "%s:%d CUDA error: %s", __FILE__, __LINE__, cudaGetErrorString(e_)); (ctx)->err = b_; \
    return e_ == cudaErrorMemoryAllocation ? UVG_ENOMEM : UVG_ECUDA; } } while (0)

static int32_t round_size_first(const uvg_params& p) { int32_t v = p.first_round > 0 ? p.first_round : 1024; return (v + GROUP - 1) / GROUP * GROUP; }
static int32_t round_size_max(const uvg_params& p)
{
    int32_t v = p.max_round > 0 ? p.max_round : 65536;
    v = (v + GROUP - 1) / GROUP * GROUP;
    return std::max(v, round_size_first(p));
}
static int32_t round_size_next(int32_t cur, int32_t mx) { int64_t n = (int64_t)cur * 4; return (int32_t)std::min<int64_t>(n, mx); }

extern "C" int uvg_abi_version(void) { return UVG_ABI_VERSION; }

extern "C" const char* uvg_last_error(const uvg_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

extern "C" int uvg_create(const uvg_params* p, uvg_ctx** out)
{
    if (!p || !out) { g_create_error = "null argument"; return UVG_EINVAL; }
    *out = nullptr;
    if (p->abi_version != UVG_ABI_VERSION) { g_create_error = "ABI version mismatch"; return UVG_EINVAL; }
    if (p->nchar < 1 || p->nbest < 1) { g_create_error = "nchar and nbest must be positive"; return UVG_EINVAL; }
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
    c->ambw = (c->Wp + 31) / 32;
    c->rank = p->world > 1 ? p->rank : 0;
    c->world = p->world > 1 ? p->world : 1;
    c->next_round = round_size_first(*p);
    if (c->prm.pool < 1) c->prm.pool = 1;
#define CK_(call) do { cudaError_t e2_ = (call); if (e2_ != cudaSuccess) { g_create_error = std::string("CUDA: ") + cudaGetErrorString(e2_); delete c; return UVG_ECUDA; } } while (0)
    CK_(cudaSetDevice(p->device));
    cudaDeviceProp prop;
    CK_(cudaGetDeviceProperties(&prop, p->device));
    if (prop.major < 10) { g_create_error = "device is not sm_100 or newer; this library ships sm_100a code only"; delete c; return UVG_ENODEV; }
    CK_(cudaStreamCreateWithFlags(&c->st, cudaStreamNonBlocking));
    CK_(cudaEventCreate(&c->ev_a)); CK_(cudaEventCreate(&c->ev_b)); CK_(cudaEventCreate(&c->ev_c)); CK_(cudaEventCreate(&c->ev_d));
    CK_(upload_luts());
    CK_(paircount_setup());
    CK_(replay_setup(c->K));
    CK_(cudaMalloc(&c->d_counters, 64)); CK_(cudaMalloc(&c->d_nacc, 64));
    CK_(cudaMemset(c->d_counters, 0, 64));
    CK_(cudaMallocHost(&c->h_pinned, 1024));
#undef CK_
    *out = c;
    return UVG_OK;
}

static void free_queries(uvg_ctx* c)
{
    cudaFree(c->d_qdata); cudaFree(c->d_cons_acgt); cudaFree(c->d_cons_vo); cudaFree(c->d_thr); cudaFree(c->d_heaps);
    cudaFree(c->d_heap_n); cudaFree(c->d_max_inc); cudaFree(c->d_batch_id); cudaFree(c->d_max_inc_batch);
    cudaFree(c->d_scores);
    c->d_qdata = nullptr; c->d_cons_acgt = nullptr; c->d_cons_vo = nullptr; c->d_thr = nullptr; c->d_heaps = nullptr;
    c->d_heap_n = c->d_max_inc = c->d_batch_id = c->d_max_inc_batch = nullptr;
    c->d_scores = nullptr; c->score_cap = 0;
}

extern "C" void uvg_destroy(uvg_ctx* c)
{
    if (!c) return;
    cudaSetDevice(c->prm.device);
    if (c->st) cudaStreamSynchronize(c->st);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    free_queries(c);
    cudaFree(c->d_db); cudaFree(c->d_refamb); cudaFree(c->d_non_n); cudaFree(c->d_c0); cudaFree(c->d_ci); cudaFree(c->d_ever);
    cudaFree(c->d_counters); cudaFree(c->d_nacc);
    cudaFree(c->d_stage_text); cudaFree(c->d_stage_nn); cudaFree(c->d_stage_acc); cudaFree(c->d_stage_dst);
    cudaFree(c->dbg_counts);
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    for (auto& pr : c->pair_events) { cudaEventDestroy(pr.first); cudaEventDestroy(pr.second); }
    if (c->ev_a) { cudaEventDestroy(c->ev_a); cudaEventDestroy(c->ev_b); cudaEventDestroy(c->ev_c); cudaEventDestroy(c->ev_d); }
    if (c->st) cudaStreamDestroy(c->st);
    delete c;
}

// =====================================================================================================================
// queries
// =====================================================================================================================
static int reset_heaps(uvg_ctx* c)
{
    const int Q = c->n_query;
    if (Q == 0) return UVG_OK;
    std::vector<int32_t> init(Q, c->nchar);                                 // max_incompatible = nchar (nearest.c:375,387)
    UVG_CUDA(c, cudaMemcpyAsync(c->d_max_inc, init.data(), Q * 4, cudaMemcpyHostToDevice, c->st));
    UVG_CUDA(c, cudaMemcpyAsync(c->d_max_inc_batch, init.data(), Q * 4, cudaMemcpyHostToDevice, c->st));
    UVG_CUDA(c, cudaMemsetAsync(c->d_batch_id, 0xff, Q * 4, c->st));
    UVG_CUDA(c, cudaMemsetAsync(c->d_heap_n, 0, Q * 4, c->st));
    UVG_CUDA(c, cudaMemsetAsync(c->d_thr, 0, (size_t)c->n_qtiles * QTILE * sizeof(Threshold), c->st));
    UVG_CUDA(c, cudaMemsetAsync(c->d_heaps, 0, (size_t)Q * (c->K + 1) * sizeof(HeapItem), c->st));
    UVG_CUDA(c, cudaMemsetAsync(c->d_counters, 0, 64, c->st));
    if (c->d_ever && c->ever_cap) UVG_CUDA(c, cudaMemsetAsync(c->d_ever, 0, c->ever_cap, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));                              // `init` goes out of scope
    c->db_scored = 0; c->heaps_cached = false;
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
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    free_queries(c);
    const int L = c->nchar, trim = c->trim, Wp = c->Wp, Q = n_query;
    const bool acgt = c->prm.acgt_mode != 0;
    uint8_t lut[3][256];
    build_luts(lut);
    const uint8_t* code_lut = lut[acgt ? 1 : 0];
    uint8_t up[256];
    for (int i = 0; i < 256; ++i) up[i] = (uint8_t)((i >= 'a' && i <= 'z') ? i - 32 : i);     // uppercase_string, nexus_common.c:105-111

    // ---- site partition (create_query_indices, src/fastaseq.c:732-777) ----
    // usable byte: --acgt: one of ACGT (is_site_acgt, utils.c:285-289); default: valid text (is_site_valid, utils.c:291-295)
    std::vector<uint8_t> cons(L, 'N');
    std::vector<uint8_t> site_class(L, 0);                                  // 1 = idx, 2 = idx_m, 3 = idx_c
    int n_idx = 0, n_idx_m = 0, n_idx_c = 0;
    for (int i = trim; i < L - trim; ++i) {
        uint8_t first = 0; bool any = false, all = true, same = true;
        for (int q = 0; q < Q; ++q) {
            uint8_t ch = up[text[(size_t)q * L + i]];
            bool usable = acgt ? (ch == 'A' || ch == 'C' || ch == 'G' || ch == 'T') : (lut[2][ch] != 0);
            if (!usable) { all = false; continue; }
            if (!any) { any = true; first = ch; }
            else if (ch != first) same = false;
        }
        if (!any) continue;
        if (!same) { site_class[i] = 1; ++n_idx; cons[i] = '#'; }
        else if (all) { site_class[i] = 3; ++n_idx_c; cons[i] = first; }
        else { site_class[i] = 2; ++n_idx_m; cons[i] = first; }
    }
    c->n_query = Q;
    c->n_qtiles = (Q + QTILE - 1) / QTILE;
    c->qinfo.n_query = Q; c->qinfo.nchar = L; c->qinfo.trim = trim;
    c->qinfo.n_idx = n_idx; c->qinfo.n_idx_m = n_idx_m; c->qinfo.n_idx_c = n_idx_c;
    if (info) *info = c->qinfo;

    // ---- bit-planes of every query (zero outside the trimmed range) ----
    const size_t n_chunk_total = (size_t)c->n_qtiles * c->n_chunks;
    std::vector<QChunk> qd(n_chunk_total);
    memset(qd.data(), 0, n_chunk_total * sizeof(QChunk));
    std::vector<uint32_t> one_count((size_t)Q * c->n_chunks, 0);            // potential ACGT matches per (query, chunk)
    for (int q = 0; q < Q; ++q) {
        const int qt = q / QTILE, qi = q % QTILE, warp = qi / QPW;
        const uint8_t* row = text + (size_t)q * L;
        for (int w = 0; w < c->W; ++w) {
            uint32_t a = 0, cc = 0, g = 0, t = 0, im = 0;
            const int s0 = w * 32, s1 = std::min(L, s0 + 32);
            for (int s = std::max(s0, trim); s < std::min(s1, L - trim); ++s) {
                const uint8_t code = code_lut[up[row[s]]];
                const uint32_t bit = 1u << (s - s0);
                if (code & 1) a |= bit;
                if (code & 2) cc |= bit;
                if (code & 4) g |= bit;
                if (code & 8) t |= bit;
                if (site_class[s] == 1) im |= bit;
            }
            const uint32_t v = a | cc | g | t;
            const uint32_t multi = (a & cc) | (g & t) | ((a | cc) & (g | t));
            const uint32_t one = v & ~multi;
            QChunk& ch = qd[(size_t)qt * c->n_chunks + w / WC];
            ch.acgt[w % WC][qi] = make_uint4(a, cc, g, t);
            ch.vo[w % WC][qi] = make_uint2(v, acgt ? im : one);
            if (!acgt && (v & ~one)) ch.slow[w % WC] |= 1u << warp;
            one_count[(size_t)q * c->n_chunks + w / WC] += (uint32_t)__builtin_popcount(acgt ? v : one);
        }
        uint32_t run = 0;
        for (int k = c->n_chunks - 1; k >= 0; --k) {
            qd[(size_t)qt * c->n_chunks + k].rem[qi] = run;
            run += one_count[(size_t)q * c->n_chunks + k];
        }
    }
    // ---- consensus pseudo-query restricted to idx_c (queue_distance_to_consensus, nearest.c:428-433) ----
    std::vector<uint4> ca(Wp, make_uint4(0, 0, 0, 0));
    std::vector<uint2> cv(Wp, make_uint2(0, 0));
    for (int s = trim; s < L - trim; ++s) {
        if (site_class[s] != 3) continue;
        const uint8_t code = code_lut[cons[s]];
        const uint32_t bit = 1u << (s & 31);
        uint4& x = ca[s >> 5];
        if (code & 1) x.x |= bit;
        if (code & 2) x.y |= bit;
        if (code & 4) x.z |= bit;
        if (code & 8) x.w |= bit;
        if (code) cv[s >> 5].x |= bit;
        if (code && !(code & (code - 1))) cv[s >> 5].y |= bit;
    }

    // ---- device allocations ----
    const size_t Qp = (size_t)c->n_qtiles * QTILE;
    UVG_CUDA(c, cudaMalloc(&c->d_qdata, n_chunk_total * sizeof(QChunk)));
    UVG_CUDA(c, cudaMalloc(&c->d_cons_acgt, Wp * sizeof(uint4)));
    UVG_CUDA(c, cudaMalloc(&c->d_cons_vo, Wp * sizeof(uint2)));
    UVG_CUDA(c, cudaMalloc(&c->d_thr, Qp * sizeof(Threshold)));
    UVG_CUDA(c, cudaMalloc(&c->d_heaps, (size_t)Q * (c->K + 1) * sizeof(HeapItem)));
    UVG_CUDA(c, cudaMalloc(&c->d_heap_n, Q * 4)); UVG_CUDA(c, cudaMalloc(&c->d_max_inc, Q * 4));
    UVG_CUDA(c, cudaMalloc(&c->d_batch_id, Q * 4)); UVG_CUDA(c, cudaMalloc(&c->d_max_inc_batch, Q * 4));
    // dense score tile: rounds are capped so that n_query x round x 16 B stays within ~2 GiB
    {
        int64_t cap = round_size_max(c->prm);
        const int64_t fit = ((int64_t)2 << 30) / ((int64_t)Q * 16) / (GROUP * RPL) * (GROUP * RPL);
        cap = std::max<int64_t>(GROUP * RPL, std::min(cap, fit));
        c->score_cap = cap;
        UVG_CUDA(c, cudaMalloc(&c->d_scores, (size_t)Q * cap * sizeof(uint4)));
    }
    UVG_CUDA(c, cudaMemcpyAsync(c->d_qdata, qd.data(), n_chunk_total * sizeof(QChunk), cudaMemcpyHostToDevice, c->st));
    UVG_CUDA(c, cudaMemcpyAsync(c->d_cons_acgt, ca.data(), Wp * sizeof(uint4), cudaMemcpyHostToDevice, c->st));
    UVG_CUDA(c, cudaMemcpyAsync(c->d_cons_vo, cv.data(), Wp * sizeof(uint2), cudaMemcpyHostToDevice, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    return reset_heaps(c);
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
static cudaError_t grow_array(T** ptr, size_t old_count, size_t new_count, bool zero_new, cudaStream_t st)
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
    cudaFree(*ptr);
    *ptr = n;
    return cudaSuccess;
}

static int ensure_db_capacity(uvg_ctx* c, int64_t need_refs)
{
    if (need_refs <= c->db_cap) return UVG_OK;
    int64_t cap = std::max<int64_t>(need_refs, c->db_cap * 2);
    cap = std::max<int64_t>(cap, c->prm.db_capacity);
    cap = std::max<int64_t>(cap, 4096);
    cap = (cap + GROUP * RPL - 1) / (GROUP * RPL) * (GROUP * RPL);
    const size_t og = (size_t)(c->db_cap / GROUP), ng = (size_t)(cap / GROUP);
    UVG_CUDA(c, grow_array(&c->d_db, og * c->Wp * GROUP, ng * c->Wp * GROUP, false, c->st));
    UVG_CUDA(c, grow_array(&c->d_refamb, og * c->ambw, ng * c->ambw, true, c->st));
    UVG_CUDA(c, grow_array(&c->d_non_n, (size_t)c->db_cap, (size_t)cap, true, c->st));
    UVG_CUDA(c, grow_array(&c->d_c0, (size_t)c->db_cap, (size_t)cap, true, c->st));
    UVG_CUDA(c, grow_array(&c->d_ci, (size_t)c->db_cap, (size_t)cap, true, c->st));
    c->db_cap = cap;
    return UVG_OK;
}

static int ensure_ever(uvg_ctx* c, int64_t n_global)
{
    if (n_global <= c->ever_cap) return UVG_OK;
    int64_t cap = std::max<int64_t>(n_global, c->ever_cap * 2);
    cap = std::max<int64_t>(cap, 4096);
    UVG_CUDA(c, grow_array(&c->d_ever, (size_t)c->ever_cap, (size_t)cap, true, c->st));
    c->ever_cap = cap;
    return UVG_OK;
}

static int ensure_stage_rows(uvg_ctx* c, int64_t rows)
{
    if (rows <= c->stage_rows) return UVG_OK;
    cudaFree(c->d_stage_nn); cudaFree(c->d_stage_acc); cudaFree(c->d_stage_dst);
    c->d_stage_nn = nullptr; c->d_stage_acc = nullptr; c->d_stage_dst = nullptr; c->stage_rows = 0;
    UVG_CUDA(c, cudaMalloc(&c->d_stage_nn, rows * 4));
    UVG_CUDA(c, cudaMalloc(&c->d_stage_acc, rows));
    UVG_CUDA(c, cudaMalloc(&c->d_stage_dst, rows * 8));
    c->stage_rows = rows;
    return UVG_OK;
}

// encode n rows already on the device; appends the accepted ones
static int push_device_rows(uvg_ctx* c, const uint8_t* d_text, int64_t n, int32_t min_non_n, int32_t* non_n_out,
                            uint8_t* accepted_out, int64_t* n_acc_out)
{
    int rc = ensure_stage_rows(c, n);
    if (rc) return rc;
    rc = ensure_db_capacity(c, c->db_n + n);
    if (rc) return rc;
    UVG_CUDA(c, launch_count_non_n(d_text, n, c->nchar, min_non_n, c->d_stage_nn, c->d_stage_acc, c->st));
    UVG_CUDA(c, launch_scan_accept(c->d_stage_acc, n, c->db_n, c->d_stage_dst, c->d_nacc, c->st));
    UVG_CUDA(c, launch_encode(d_text, n, c->nchar, c->W, c->Wp, c->ambw, c->prm.acgt_mode, c->d_stage_dst, c->d_stage_nn,
                              c->d_db, c->d_refamb, c->d_non_n, c->st));
    c->stats.other_kernel_launches += 3;
    int64_t* h_n = reinterpret_cast<int64_t*>(c->h_pinned);
    UVG_CUDA(c, cudaMemcpyAsync(h_n, c->d_nacc, 8, cudaMemcpyDeviceToHost, c->st));
    if (non_n_out) UVG_CUDA(c, cudaMemcpyAsync(non_n_out, c->d_stage_nn, n * 4, cudaMemcpyDeviceToHost, c->st));
    if (accepted_out) UVG_CUDA(c, cudaMemcpyAsync(accepted_out, c->d_stage_acc, n, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    c->db_n += *h_n;
    if (n_acc_out) *n_acc_out += *h_n;
    return UVG_OK;
}

extern "C" int uvg_push_refs_device(uvg_ctx* c, const uint8_t* d_text, int64_t n, int32_t min_non_n, int32_t* non_n_out,
                                    uint8_t* accepted_out, int64_t* n_accepted)
{
    if (!c || (!d_text && n > 0) || n < 0) { if (c) c->err = "uvg_push_refs_device: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (n_accepted) *n_accepted = 0;
    if (n == 0) return UVG_OK;
    UVG_CUDA(c, cudaEventRecord(c->ev_c, c->st));
    const int64_t step = 1 << 16;
    for (int64_t i = 0; i < n; i += step) {
        const int64_t m = std::min(step, n - i);
        int rc = push_device_rows(c, d_text + i * (int64_t)c->nchar, m, min_non_n, non_n_out ? non_n_out + i : nullptr,
                                  accepted_out ? accepted_out + i : nullptr, n_accepted);
        if (rc) return rc;
    }
    UVG_CUDA(c, cudaEventRecord(c->ev_d, c->st));
    UVG_CUDA(c, cudaEventSynchronize(c->ev_d));
    float ms = 0; cudaEventElapsedTime(&ms, c->ev_c, c->ev_d);
    c->stats.encode_ms = ms;
    return UVG_OK;
}

extern "C" int uvg_push_refs(uvg_ctx* c, const uint8_t* text, int64_t n, int32_t min_non_n, int32_t* non_n_out,
                             uint8_t* accepted_out, int64_t* n_accepted)
{
    if (!c || (!text && n > 0) || n < 0) { if (c) c->err = "uvg_push_refs: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (n_accepted) *n_accepted = 0;
    if (n == 0) return UVG_OK;
    // stage through a device text buffer of <= 256 MiB per step
    const int64_t rows_per_step = std::max<int64_t>(1, std::min<int64_t>(n, (256ll << 20) / c->nchar));
    const size_t need = (size_t)rows_per_step * c->nchar;
    if (need > c->stage_text_bytes) {
        cudaFree(c->d_stage_text); c->d_stage_text = nullptr; c->stage_text_bytes = 0;
        UVG_CUDA(c, cudaMalloc(&c->d_stage_text, need));
        c->stage_text_bytes = need;
    }
    UVG_CUDA(c, cudaEventRecord(c->ev_c, c->st));
    for (int64_t i = 0; i < n; i += rows_per_step) {
        const int64_t m = std::min(rows_per_step, n - i);
        UVG_CUDA(c, cudaMemcpyAsync(c->d_stage_text, text + i * (int64_t)c->nchar, (size_t)m * c->nchar, cudaMemcpyHostToDevice, c->st));
        int rc = push_device_rows(c, c->d_stage_text, m, min_non_n, non_n_out ? non_n_out + i : nullptr,
                                  accepted_out ? accepted_out + i : nullptr, n_accepted);
        if (rc) return rc;
    }
    UVG_CUDA(c, cudaEventRecord(c->ev_d, c->st));
    UVG_CUDA(c, cudaEventSynchronize(c->ev_d));
    float ms = 0; cudaEventElapsedTime(&ms, c->ev_c, c->ev_d);
    c->stats.encode_ms = ms;
    return UVG_OK;
}

extern "C" int uvg_drop_refs(uvg_ctx* c)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (c->world > 1) UVG_FAIL(c, UVG_ESTATE, "uvg_drop_refs is not available in sharded mode");
    if (c->db_scored != c->db_n) UVG_FAIL(c, UVG_ESTATE, "uvg_drop_refs: %lld resident references have not been searched yet",
                                         (long long)(c->db_n - c->db_scored));
    if (c->d_refamb) UVG_CUDA(c, cudaMemsetAsync(c->d_refamb, 0, (size_t)(c->db_cap / GROUP) * c->ambw * 4, c->st));
    c->global_base += c->db_n;
    c->db_n = 0; c->db_scored = 0; c->round_local_start = 0;
    return UVG_OK;
}

extern "C" int uvg_clear_refs(uvg_ctx* c)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (c->d_refamb) UVG_CUDA(c, cudaMemsetAsync(c->d_refamb, 0, (size_t)(c->db_cap / GROUP) * c->ambw * 4, c->st));
    c->global_base = 0; c->db_n = 0; c->db_scored = 0; c->round_local_start = 0; c->global_total = -1;
    return c->n_query ? reset_heaps(c) : UVG_OK;
}

extern "C" int64_t uvg_db_count(const uvg_ctx* c) { return c ? c->db_n : 0; }

// =====================================================================================================================
// search: rounds of K2 (pair counts -> survivors) and K3 (exact replay)
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

static void fill_pair_params(uvg_ctx* c, PairParams& pp, int64_t lb, int64_t le)
{
    pp.db = c->d_db; pp.refamb = c->d_refamb; pp.qdata = c->d_qdata; pp.thr = c->d_thr;
    pp.Wp = c->Wp; pp.n_chunks = c->n_chunks; pp.ambw = c->ambw; pp.n_qtiles = c->n_qtiles; pp.n_query = c->n_query;
    pp.local_begin = lb; pp.local_end = le; pp.global_minus_local = 0;
    pp.scores = c->d_scores; pp.score_stride = c->score_cap; pp.score_col0 = 0;
    pp.early_abort = c->prm.early_abort; pp.work = c->d_counters + 2;
}

// one scoring round over the local tile [lstart, lstart+cnt): consensus stage, pair counts into the dense tile, replay.
// Everything is asynchronous on the context's stream: there is no host synchronisation inside a search.
static int run_round(uvg_ctx* c, int64_t lstart, int64_t cnt, int64_t T, int64_t tile_global, int64_t round_first,
                     int64_t round_count, size_t* n_pair_events)
{
    (void)T; (void)round_first; (void)round_count;
    if (c->world > 1) UVG_FAIL(c, UVG_ESTATE, "sharded search is not wired to the dense score tile yet");
    if (cnt <= 0) return UVG_OK;
    PairParams pp{};
    fill_pair_params(c, pp, lstart, lstart + cnt);
    UVG_CUDA(c, launch_consensus(c->d_db, c->d_cons_acgt, c->d_cons_vo, c->Wp, c->prm.acgt_mode, pp.local_begin, pp.local_end,
                                 c->d_c0, c->d_ci, c->st));
    cudaEvent_t ea, eb;
    int rc = event_pair(c, *n_pair_events, &ea, &eb);
    if (rc) return rc;
    UVG_CUDA(c, cudaEventRecord(ea, c->st));
    UVG_CUDA(c, launch_paircount(pp, c->prm.acgt_mode, c->st));
    UVG_CUDA(c, cudaEventRecord(eb, c->st));
    ++*n_pair_events;
    ReplayParams rp{};
    rp.scores = c->d_scores; rp.row_stride = c->score_cap;
    rp.nseg = 1; rp.seg_stride = (int32_t)cnt; rp.seg_cnt[0] = (int32_t)cnt; rp.seg_global[0] = tile_global;
    rp.non_n = c->d_non_n + lstart; rp.c0 = c->d_c0 + lstart; rp.ci = c->d_ci + lstart;
    rp.heaps = c->d_heaps; rp.heap_n = c->d_heap_n; rp.max_inc = c->d_max_inc;
    rp.batch_id = c->d_batch_id; rp.max_inc_batch = c->d_max_inc_batch; rp.thr = c->d_thr;
    rp.ever_best = c->d_ever; rp.counters = c->d_counters;
    rp.q_first = 0; rp.n_query = c->n_query; rp.K = c->K; rp.acgt_mode = c->prm.acgt_mode; rp.pool = c->prm.pool;
    cudaEvent_t ra, rb;
    rc = event_pair(c, *n_pair_events, &ra, &rb);
    if (rc) return rc;
    UVG_CUDA(c, cudaEventRecord(ra, c->st));
    UVG_CUDA(c, launch_replay(rp, c->st));
    UVG_CUDA(c, cudaEventRecord(rb, c->st));
    ++*n_pair_events;
    c->stats.pair_kernel_launches += 1; c->stats.other_kernel_launches += 2;
    return UVG_OK;
}

extern "C" int uvg_search(uvg_ctx* c)
{
    if (!c) return UVG_EINVAL;
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (c->n_query == 0) UVG_FAIL(c, UVG_ESTATE, "uvg_search: no queries set");
    if (c->world > 1 && !c->comm) UVG_FAIL(c, UVG_ESTATE, "uvg_search: sharded context without uvg_nccl_init");
    if (c->world > 1 && c->global_total < 0) UVG_FAIL(c, UVG_ESTATE, "uvg_search: sharded context without uvg_set_global_count");
    const int32_t mx = (int32_t)std::min<int64_t>(round_size_max(c->prm), c->score_cap);
    size_t n_ev = 0;
    c->heaps_cached = false;
    c->stats.pair_kernel_ms = 0; c->stats.replay_ms = 0;
    UVG_CUDA(c, cudaMemcpyAsync(c->h_pinned + 20, c->d_counters + 2, 8, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    c->work_base = (int64_t)c->h_pinned[20];
    UVG_CUDA(c, cudaEventRecord(c->ev_a, c->st));
    for (;;) {
        const int64_t T = c->next_round;
        int64_t lstart, cnt, tile_global, rfirst, rcount;
        if (c->world == 1) {
            cnt = std::min<int64_t>(T, c->db_n - c->db_scored);
            if (cnt <= 0) break;
            lstart = c->db_scored; tile_global = c->global_base + lstart; rfirst = tile_global; rcount = cnt;
        } else {
            const int64_t S = c->round_global_start;
            if (S >= c->global_total) break;
            const int64_t gb = S + (int64_t)c->rank * T, ge = std::min(gb + T, c->global_total);
            cnt = std::max<int64_t>(0, ge - gb);
            lstart = c->round_local_start; tile_global = gb; rfirst = S; rcount = std::min<int64_t>((int64_t)c->world * T, c->global_total - S);
            if (lstart + cnt > c->db_n) UVG_FAIL(c, UVG_ESTATE, "rank %d holds %lld references but round %lld needs local rows up to %lld",
                                                 c->rank, (long long)c->db_n, (long long)c->rounds_done, (long long)(lstart + cnt));
        }
        int rc = ensure_ever(c, rfirst + rcount);
        if (rc) return rc;
        if (rfirst + rcount > 0x7fffffffLL) UVG_FAIL(c, UVG_EINVAL, "more than 2^31 references are not supported");
        rc = run_round(c, lstart, cnt, T, tile_global, rfirst, rcount, &n_ev);
        if (rc) return rc;
        if (c->world == 1) { c->db_scored += cnt; c->stats.n_refs += cnt; }
        else { c->round_global_start += (int64_t)c->world * T; c->round_local_start += T; c->db_scored = std::min(c->db_n, c->round_local_start); c->stats.n_refs += rcount; }
        ++c->rounds_done; ++c->stats.n_rounds;
        c->next_round = round_size_next(c->next_round, mx);
    }
    UVG_CUDA(c, cudaEventRecord(c->ev_b, c->st));
    UVG_CUDA(c, cudaEventSynchronize(c->ev_b));
    float ms = 0; cudaEventElapsedTime(&ms, c->ev_a, c->ev_b);
    c->stats.search_ms = ms;
    for (size_t i = 0; i < n_ev; ++i) {           // events alternate: pair-count launch, replay launch
        float t = 0; cudaEventElapsedTime(&t, c->pair_events[i].first, c->pair_events[i].second);
        if (i % 2 == 0) c->stats.pair_kernel_ms += t; else c->stats.replay_ms += t;
    }
    return UVG_OK;
}

// =====================================================================================================================
// results
// =====================================================================================================================
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
    if (c->heaps_cached) return UVG_OK;
    const size_t Q = c->n_query, stride = c->K + 1;
    c->h_heaps.resize(Q * stride); c->h_heap_n.resize(Q); c->h_max_inc.resize(Q);
    UVG_CUDA(c, cudaMemcpyAsync(c->h_heaps.data(), c->d_heaps, Q * stride * sizeof(HeapItem), cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaMemcpyAsync(c->h_heap_n.data(), c->d_heap_n, Q * 4, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaMemcpyAsync(c->h_max_inc.data(), c->d_max_inc, Q * 4, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    c->heaps_cached = true;
    return UVG_OK;
}

extern "C" int uvg_get_heap(uvg_ctx* c, int32_t iq, uvg_item* out, int32_t* n_out)
{
    if (!c || !out || !n_out) { if (c) c->err = "uvg_get_heap: bad arguments"; return UVG_EINVAL; }
    if (iq < 0 || iq >= c->n_query) UVG_FAIL(c, UVG_EINVAL, "uvg_get_heap: query %d out of range", iq);
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    int rc = cache_heaps(c);
    if (rc) return rc;
    const int n = c->h_heap_n[iq];
    *n_out = n;
    if (n == 0) return UVG_OK;
    const HeapItem* a = c->h_heaps.data() + (size_t)iq * (c->K + 1);
    // heap_finalise_heap_qsort (src/min_heap.c:149-153): slot 0 <-> slot n, then libc qsort over [0, n)
    std::vector<HeapItem> arr(n);
    arr[0] = a[n];
    for (int i = 1; i < n; ++i) arr[i] = a[i];
    qsort(arr.data(), n, sizeof(HeapItem), compare_items_desc);
    for (int i = 0; i < n; ++i) {
        for (int k = 0; k < 6; ++k) out[i].score[k] = arr[i].s[k];
        out[i].ref = arr[i].r;
    }
    return UVG_OK;
}

extern "C" int uvg_get_ever_best(uvg_ctx* c, int64_t first, int64_t n, uint8_t* flags)
{
    if (!c || !flags || first < 0 || n < 0) { if (c) c->err = "uvg_get_ever_best: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
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
    UVG_CUDA(c, cudaMemcpyAsync(c->h_pinned + 16, c->d_counters, 24, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    c->stats.n_admitted = (int64_t)c->h_pinned[16]; c->stats.n_h1_flags = (int64_t)c->h_pinned[17];
    c->stats.warp_chunks_executed = (int64_t)c->h_pinned[18] - c->work_base;
    if (c->n_query) {
        int rc = cache_heaps(c);
        if (rc) return rc;
        int64_t mm = 0;
        for (int q = 0; q < c->n_query; ++q) mm = std::max<int64_t>(mm, c->h_max_inc[q] - 1);
        c->stats.max_mismatch = mm;
    }
    *out = c->stats;
    return UVG_OK;
}

// =====================================================================================================================
// sharding helpers + NCCL
// =====================================================================================================================
extern "C" int uvg_shard_of(int64_t ordinal, int32_t world, int32_t first_round, int32_t max_round, int32_t* rank, int64_t* local)
{
    if (ordinal < 0 || world < 1) return UVG_EINVAL;
    uvg_params p{}; p.first_round = first_round; p.max_round = max_round;
    int64_t T = round_size_first(p), S = 0, Ls = 0;
    const int32_t mx = round_size_max(p);
    for (;;) {
        if (ordinal < S + (int64_t)world * T) {
            if (rank) *rank = (int32_t)((ordinal - S) / T);
            if (local) *local = Ls + (ordinal - S) % T;
            return UVG_OK;
        }
        S += (int64_t)world * T; Ls += T;
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
    ncclResult_t r = g_nccl.CommInitRank(&c->comm, c->world, id, c->rank);
    if (r != 0) UVG_FAIL(c, UVG_ENCCL, "ncclCommInitRank failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    return UVG_OK;
}

extern "C" int uvg_set_global_count(uvg_ctx* c, int64_t n_total)
{
    if (!c || n_total < 0) return UVG_EINVAL;
    c->global_total = n_total;
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

extern "C" int uvg_debug_pair_counts(uvg_ctx* c, int64_t first, int64_t n, int32_t* out)
{
    if (!c || !out || first < 0 || n < 1) { if (c) c->err = "uvg_debug_pair_counts: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (c->n_query == 0 || first + n > c->db_n || n > c->score_cap) UVG_FAIL(c, UVG_ESTATE, "uvg_debug_pair_counts: range not resident, too long, or no queries");
    PairParams pp{};
    fill_pair_params(c, pp, first, first + n);
    pp.early_abort = 0; pp.work = nullptr;
    UVG_CUDA(c, launch_paircount(pp, c->prm.acgt_mode, c->st));
    UVG_CUDA(c, cudaMemcpy2DAsync(out, (size_t)n * 16, c->d_scores, (size_t)c->score_cap * 16, (size_t)n * 16, (size_t)c->n_query,
                                  cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    return UVG_OK;
}

extern "C" int uvg_debug_pair_counts_rows(uvg_ctx* c, int32_t q_first, int32_t q_count, int64_t first, int64_t n, int32_t* out)
{
    if (!c || !out || first < 0 || n < 1 || q_first < 0 || q_count < 1) { if (c) c->err = "uvg_debug_pair_counts_rows: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (c->n_query == 0 || first + n > c->db_n || n > c->score_cap || q_first + q_count > c->n_query)
        UVG_FAIL(c, UVG_ESTATE, "uvg_debug_pair_counts_rows: range not resident, too long, or no such queries");
    PairParams pp{};
    fill_pair_params(c, pp, first, first + n);
    pp.early_abort = 0; pp.work = nullptr;
    UVG_CUDA(c, launch_paircount(pp, c->prm.acgt_mode, c->st));
    UVG_CUDA(c, cudaMemcpy2DAsync(out, (size_t)n * 16, c->d_scores + (size_t)q_first * c->score_cap, (size_t)c->score_cap * 16,
                                  (size_t)n * 16, (size_t)q_count, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    return UVG_OK;
}

extern "C" int uvg_debug_ref_meta(uvg_ctx* c, int64_t first, int64_t n, int32_t* non_n, int32_t* c0, int32_t* ci)
{
    if (!c || first < 0 || n < 1) { if (c) c->err = "uvg_debug_ref_meta: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    if (c->n_query == 0 || first + n > c->db_n) UVG_FAIL(c, UVG_ESTATE, "uvg_debug_ref_meta: range not resident or no queries");
    UVG_CUDA(c, launch_consensus(c->d_db, c->d_cons_acgt, c->d_cons_vo, c->Wp, c->prm.acgt_mode, first, first + n, c->d_c0, c->d_ci, c->st));
    if (non_n) UVG_CUDA(c, cudaMemcpyAsync(non_n, c->d_non_n + first, (size_t)n * 4, cudaMemcpyDeviceToHost, c->st));
    if (c0) UVG_CUDA(c, cudaMemcpyAsync(c0, c->d_c0 + first, (size_t)n * 4, cudaMemcpyDeviceToHost, c->st));
    if (ci) UVG_CUDA(c, cudaMemcpyAsync(ci, c->d_ci + first, (size_t)n * 4, cudaMemcpyDeviceToHost, c->st));
    UVG_CUDA(c, cudaStreamSynchronize(c->st));
    return UVG_OK;
}

extern "C" int uvg_debug_unpack_ref(uvg_ctx* c, int64_t local, uint8_t* codes_out)
{
    if (!c || !codes_out || local < 0 || local >= c->db_n) { if (c) c->err = "uvg_debug_unpack_ref: bad arguments"; return UVG_EINVAL; }
    UVG_CUDA(c, cudaSetDevice(c->prm.device));
    uint8_t* d = nullptr;
    UVG_CUDA(c, cudaMalloc(&d, c->nchar));
    cudaError_t e = launch_unpack_ref(c->d_db, local, c->nchar, c->Wp, d, c->st);
    if (e == cudaSuccess) e = cudaMemcpyAsync(codes_out, d, c->nchar, cudaMemcpyDeviceToHost, c->st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->st);
    cudaFree(d);
    UVG_CUDA(c, e);
    return UVG_OK;
}
