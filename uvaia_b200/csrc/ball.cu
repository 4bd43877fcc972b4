// uvg_ball_*: the radius search of the reference program `uvaiaball` (src/ball.c:138-291) on the GPU -- which reference
// sequences are within `dist` differences of SOME query.  Replaces seq_ball_against_query_structure (src/fastaseq.c:660-696)
// with quick_pairwise_score_truncated_idx_indelcheck / quick_pairwise_score_acgt (fastaseq.c:562-583) inside the OpenMP loop
// ball.c:240-244, and the reference filter ball.c:201,216-222.
//
// The ball distance compares BYTES (upper-cased), not IUPAC sets: two usable bytes that differ count one, whatever they are
// (U differs from T, R differs from A; letters outside the IUPAC alphabet are usable text).  So this path works on the
// sequence text itself, not on the packed 4-bit planes of the n-best search.  With radius R = dist + 1 and the site sets of
// create_query_indices (fastaseq.c:732-777; idx_c / idx_m: the queries that hold a usable byte agree, idx: they differ):
//     c  = #{ i in idx_c + idx_m : ref_i usable, ref_i != consensus_i }     (both stages truncate at R; a reference with c >= R is out)
//     2c < R                      -> kept without looking at any query      (the loop at fastaseq.c:687 tests min_dist + c_dist with
//                                                                            min_dist still holding c: the reference's own rule)
//     else kept iff some query q has  #{ i in idx : ref_i, q_i usable, ref_i != q_i } < R - c
// K-B1 k_ball_cons: one block per reference row streams the row once (HBM-bound: 1 byte per site): valid_ref_sites for the
// filter, c against the consensus row.  K-B2 k_ball_query: only for the rows the first kernel could not decide -- the row's
// bytes at the idx sites go to shared memory, the threads take queries (idx bytes of the queries are stored transposed, four
// sites per word: coalesced across threads), byte-SIMD compares, early exit as soon as one query is inside.
#include "../../include/uvaia_b200.h"
#include "common.cuh"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace {

__constant__ uint8_t c_up[256];        // upper case
__constant__ uint8_t c_usable[256];    // of an UPPER-CASE byte: default mode valid text (not in NnXx-?Oo.), --acgt: A C G T
__constant__ uint8_t c_validtxt[256];  // of an upper-case byte: not in NnXx-?Oo. (quick_count_sequence_non_N counts these over the whole row)

constexpr int BALL_THREADS = 256;

// Scratch of uvg_ball_classify_device_noalloc (the entry point that allocates nothing): what k_ball_cons hands to k_ball_query
// -- per row the consensus-stage differences c, and valid_ref_sites when the caller does not want them.  File-scope device
// memory: every GPU that loads the library holds these 512 KiB until the process exits.  In HOST code the name below is only a
// stand-in for the symbol, not a device address: the address comes from cudaGetSymbolAddress (noalloc_scratch_of), per device.
struct BallScratch {
    int32_t c[UVG_BALL_NOALLOC_MAX_ROWS];
    int32_t non_n[UVG_BALL_NOALLOC_MAX_ROWS];
};
static __device__ __align__(256) BallScratch g_ball_scratch;

// keep codes: 0 = too ambiguous (filtered), 1 = outside the ball, 2 = inside, 3 = undecided after the consensus stages
__global__ void __launch_bounds__(BALL_THREADS) k_ball_cons(const uint8_t* __restrict__ text, int64_t n, int L, const uint8_t* __restrict__ cc,
                                                           int min_non_n, int R, int32_t* __restrict__ c_out, int32_t* __restrict__ nonn_out,
                                                           uint8_t* __restrict__ keep)
{
    __shared__ int s_red[2][BALL_THREADS / 32];
    const int64_t row = blockIdx.x;
    if (row >= n) return;
    const uint8_t* r = text + row * (int64_t)L;
    int nn = 0, c = 0;
    for (int i = threadIdx.x; i < L; i += BALL_THREADS) {
        const uint8_t b = c_up[r[i]], k = cc[i];
        nn += c_validtxt[b];
        c += (k != 0) & (c_usable[b] != 0) & (b != k);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, o); c += __shfl_xor_sync(0xffffffffu, c, o); }
    if ((threadIdx.x & 31) == 0) { s_red[0][threadIdx.x >> 5] = nn; s_red[1][threadIdx.x >> 5] = c; }
    __syncthreads();
    if (threadIdx.x == 0) {
        nn = c = 0;
        for (int k = 0; k < BALL_THREADS / 32; ++k) { nn += s_red[0][k]; c += s_red[1][k]; }
        nonn_out[row] = nn; c_out[row] = c;
        keep[row] = nn < min_non_n ? 0 : c >= R ? 1 : 2 * c < R ? 2 : 3;
    }
}

__global__ void __launch_bounds__(BALL_THREADS) k_ball_query(const uint8_t* __restrict__ text, int64_t n, int L, const uint32_t* __restrict__ idx,
                                                            int n_idx, int n_words, const uint32_t* __restrict__ qxT, int n_query, int q_stride, int R,
                                                            const int32_t* __restrict__ c_in, uint8_t* __restrict__ keep)
{
    extern __shared__ __align__(16) uint8_t s_rx[];            // the row's bytes at the idx sites (0 = not usable), 4 per word
    __shared__ int s_found;
    const int64_t row = blockIdx.x;
    if (row >= n || keep[row] != 3) return;
    const uint8_t* r = text + row * (int64_t)L;
    for (int k = threadIdx.x; k < n_words * 4; k += BALL_THREADS) {
        uint8_t b = 0;
        if (k < n_idx) { b = c_up[r[idx[k]]]; if (!c_usable[b]) b = 0; }
        s_rx[k] = b;
    }
    if (threadIdx.x == 0) s_found = 0;
    __syncthreads();
    const int lim = R - c_in[row];                             // some query with fewer than `lim` differences on idx
    const uint32_t* rx = reinterpret_cast<const uint32_t*>(s_rx);
    for (int q = threadIdx.x; q < n_query && !s_found; q += BALL_THREADS) {
        int d = 0;
        for (int w = 0; w < n_words; ++w) {
            const uint32_t a = qxT[(size_t)w * q_stride + q], b = rx[w];
            // bytes that differ where both are non-zero (usable)
            const uint32_t m = __vcmpne4(a, b) & __vcmpne4(a, 0u) & __vcmpne4(b, 0u);
            d += __popc(m) >> 3;
            if ((w & 15) == 15 && d >= lim) break;
        }
        if (d < lim) s_found = 1;
    }
    __syncthreads();
    if (threadIdx.x == 0) keep[row] = s_found ? 2 : 1;
}

}  // namespace

struct uvg_ball {
    uvg_ball_params prm{};
    std::string err;
    int nchar = 0, trim = 0, dist = 0;
    bool acgt = false;
    cudaStream_t st = nullptr, st_copy = nullptr;
    cudaEvent_t ev_copied[2]{}, ev_free[2]{};
    // queries
    int n_query = 0, n_idx = 0, n_words = 0, q_stride = 0;
    uvg_query_info qinfo{};
    uint8_t* d_cc = nullptr; uint32_t* d_idx = nullptr; uint32_t* d_qxT = nullptr;
    // staging
    uint8_t* d_text[2]{}; size_t text_bytes = 0;
    int32_t *d_c = nullptr, *d_nonn = nullptr; uint8_t* d_keep = nullptr; int64_t rows_cap = 0;
};

static std::string g_ball_error;
#define BALL_FAIL(b, code, ...) do { char m_[512]; snprintf(m_, sizeof m_, __VA_ARGS__); (b)->err = m_; return (code); } while (0)
#define BALL_CUDA(b, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { char m_[512]; snprintf(m_, sizeof m_, "%s:%d CUDA error: %s", __FILE__, __LINE__, \
    cudaGetErrorString(e_)); (b)->err = m_; return e_ == cudaErrorMemoryAllocation ? UVG_ENOMEM : UVG_ECUDA; } } while (0)

extern "C" const char* uvg_ball_last_error(const uvg_ball* b) { return b ? b->err.c_str() : g_ball_error.c_str(); }

extern "C" int uvg_ball_create(const uvg_ball_params* p, uvg_ball** out)
{
    if (!p || !out) { g_ball_error = "null argument"; return UVG_EINVAL; }
    *out = nullptr;
    if (p->abi_version != UVG_ABI_VERSION) { g_ball_error = "ABI version mismatch"; return UVG_EINVAL; }
    if (p->nchar < 1) { g_ball_error = "nchar must be positive"; return UVG_EINVAL; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev < 1) {
        g_ball_error = std::string("no usable CUDA device (") + cudaGetErrorString(e) + "): uvaia_b200 has no CPU fallback";
        return UVG_ENODEV;
    }
    if (p->device < 0 || p->device >= ndev) { g_ball_error = "device ordinal out of range"; return UVG_EINVAL; }
    uvg_ball* b = new uvg_ball();
    b->prm = *p;
    b->nchar = p->nchar;
    int trim = p->trim < 0 ? 0 : p->trim;                                    // fastaseq.c:708-712
    if (trim > p->nchar / 2.1) trim = (int)(p->nchar / 2.1);
    b->trim = trim;
    int dist = p->dist < 0 ? 0 : p->dist;                                    // fastaseq.c:713-715
    if (dist > (p->nchar - 2 * trim) / 10) dist = (int)((p->nchar - 2 * trim) / 10);
    b->dist = dist;
    b->acgt = p->acgt_mode != 0;
#define CK_(call) do { cudaError_t e2_ = (call); if (e2_ != cudaSuccess) { g_ball_error = std::string("CUDA: ") + cudaGetErrorString(e2_); delete b; return UVG_ECUDA; } } while (0)
    CK_(cudaSetDevice(p->device));
    cudaDeviceProp prop;
    CK_(cudaGetDeviceProperties(&prop, p->device));
    if (prop.major < 10) { g_ball_error = "device is not sm_100 or newer; this library ships sm_100a code only"; delete b; return UVG_ENODEV; }
    CK_(cudaStreamCreateWithFlags(&b->st, cudaStreamNonBlocking));
    CK_(cudaStreamCreateWithFlags(&b->st_copy, cudaStreamNonBlocking));
    for (int k = 0; k < 2; ++k) {
        CK_(cudaEventCreateWithFlags(&b->ev_copied[k], cudaEventDisableTiming));
        CK_(cudaEventCreateWithFlags(&b->ev_free[k], cudaEventDisableTiming));
    }
    uint8_t up[256], usable[256], validtxt[256];
    for (int i = 0; i < 256; ++i) {
        up[i] = (uint8_t)((i >= 'a' && i <= 'z') ? i - 32 : i);              // uppercase_string, nexus_common.c:105-111
        validtxt[i] = 1;
    }
    for (const char* q = "NnXx-?Oo."; *q; ++q) validtxt[(uint8_t)*q] = 0;    // is_indel, utils.c:263
    for (int i = 0; i < 256; ++i) usable[i] = b->acgt ? (i == 'A' || i == 'C' || i == 'G' || i == 'T') : validtxt[i];
    CK_(cudaMemcpyToSymbol(c_up, up, 256)); CK_(cudaMemcpyToSymbol(c_usable, usable, 256)); CK_(cudaMemcpyToSymbol(c_validtxt, validtxt, 256));
#undef CK_
    *out = b;
    return UVG_OK;
}

extern "C" void uvg_ball_destroy(uvg_ball* b)
{
    if (!b) return;
    cudaSetDevice(b->prm.device);
    if (b->st) cudaStreamSynchronize(b->st);
    if (b->st_copy) cudaStreamSynchronize(b->st_copy);
    cudaFree(b->d_cc); cudaFree(b->d_idx); cudaFree(b->d_qxT);
    cudaFree(b->d_text[0]); cudaFree(b->d_text[1]); cudaFree(b->d_c); cudaFree(b->d_nonn); cudaFree(b->d_keep);
    for (int k = 0; k < 2; ++k) { if (b->ev_copied[k]) cudaEventDestroy(b->ev_copied[k]); if (b->ev_free[k]) cudaEventDestroy(b->ev_free[k]); }
    if (b->st) cudaStreamDestroy(b->st);
    if (b->st_copy) cudaStreamDestroy(b->st_copy);
    delete b;
}

extern "C" int32_t uvg_ball_radius(const uvg_ball* b) { return b ? b->dist : -1; }

extern "C" int uvg_ball_set_queries(uvg_ball* b, const uint8_t* text, int32_t n_query, uvg_query_info* info)
{
    if (!b || !text || n_query < 1) { if (b) b->err = "uvg_ball_set_queries: bad arguments"; return UVG_EINVAL; }
    BALL_CUDA(b, cudaSetDevice(b->prm.device));
    BALL_CUDA(b, cudaStreamSynchronize(b->st));
    const int L = b->nchar, trim = b->trim, Q = n_query;
    uint8_t up[256], usable[256];
    for (int i = 0; i < 256; ++i) up[i] = (uint8_t)((i >= 'a' && i <= 'z') ? i - 32 : i);
    for (int i = 0; i < 256; ++i) {
        bool v = true;
        for (const char* q = "NnXx-?Oo."; *q; ++q) if (i == (uint8_t)*q) v = false;
        usable[i] = b->acgt ? (i == 'A' || i == 'C' || i == 'G' || i == 'T') : v;
    }
    // create_query_indices (fastaseq.c:732-777): per site of the trimmed range the first usable byte, whether all queries are
    // usable, whether the usable ones agree
    std::vector<uint8_t> cc(L, 0);
    std::vector<uint32_t> idx;
    int n_idx_m = 0, n_idx_c = 0;
    for (int i = trim; i < L - trim; ++i) {
        uint8_t first = 0; bool all = true, same = true;
        for (int q = 0; q < Q; ++q) {
            const uint8_t ch = up[text[(size_t)q * L + i]];
            if (!usable[ch]) { all = false; continue; }
            if (!first) first = ch;
            else if (ch != first) { same = false; break; }
        }
        if (!first) continue;
        if (!same) idx.push_back((uint32_t)i);
        else { cc[i] = first; if (all) ++n_idx_c; else ++n_idx_m; }
    }
    b->n_query = Q; b->n_idx = (int)idx.size();
    b->n_words = (b->n_idx + 3) / 4;
    b->q_stride = (Q + 31) / 32 * 32;
    if ((size_t)b->n_words * 4 > 200 * 1024) BALL_FAIL(b, UVG_EINVAL, "uvg_ball_set_queries: %d segregating sites exceed what a thread block stages (204,800)", b->n_idx);
    std::vector<uint32_t> qxT((size_t)std::max(1, b->n_words) * b->q_stride, 0u);
    for (int k = 0; k < b->n_idx; ++k)
        for (int q = 0; q < Q; ++q) {
            const uint8_t ch = up[text[(size_t)q * L + idx[k]]];
            if (usable[ch]) qxT[(size_t)(k >> 2) * b->q_stride + q] |= (uint32_t)ch << ((k & 3) * 8);
        }
    cudaFree(b->d_cc); cudaFree(b->d_idx); cudaFree(b->d_qxT);
    b->d_cc = nullptr; b->d_idx = nullptr; b->d_qxT = nullptr;
    BALL_CUDA(b, cudaMalloc(&b->d_cc, L));
    BALL_CUDA(b, cudaMalloc(&b->d_idx, std::max<size_t>(4, idx.size() * 4)));
    BALL_CUDA(b, cudaMalloc(&b->d_qxT, qxT.size() * 4));
    BALL_CUDA(b, cudaMemcpyAsync(b->d_cc, cc.data(), L, cudaMemcpyHostToDevice, b->st));
    if (!idx.empty()) BALL_CUDA(b, cudaMemcpyAsync(b->d_idx, idx.data(), idx.size() * 4, cudaMemcpyHostToDevice, b->st));
    BALL_CUDA(b, cudaMemcpyAsync(b->d_qxT, qxT.data(), qxT.size() * 4, cudaMemcpyHostToDevice, b->st));
    BALL_CUDA(b, cudaStreamSynchronize(b->st));
    if (b->n_words * 4 > 48 * 1024)
        BALL_CUDA(b, cudaFuncSetAttribute(k_ball_query, cudaFuncAttributeMaxDynamicSharedMemorySize, b->n_words * 4));
    b->qinfo.n_query = Q; b->qinfo.nchar = L; b->qinfo.trim = trim;
    b->qinfo.n_idx = b->n_idx; b->qinfo.n_idx_m = n_idx_m; b->qinfo.n_idx_c = n_idx_c;
    if (info) *info = b->qinfo;
    return UVG_OK;
}

extern "C" int uvg_ball_push_refs(uvg_ball* b, const uint8_t* text, int64_t n, int32_t min_non_n, uint8_t* keep_out, int32_t* non_n_out)
{
    if (!b || (!text && n > 0) || n < 0 || !keep_out) { if (b) b->err = "uvg_ball_push_refs: bad arguments"; return UVG_EINVAL; }
    if (b->n_query == 0) BALL_FAIL(b, UVG_ESTATE, "uvg_ball_push_refs: no queries set");
    BALL_CUDA(b, cudaSetDevice(b->prm.device));
    if (n == 0) return UVG_OK;
    const int L = b->nchar, R = b->dist + 1;
    int64_t rows = std::max<int64_t>(1, std::min<int64_t>(n, (128ll << 20) / L));
    rows = std::min<int64_t>(rows, 1 << 16);
    if ((size_t)rows * L > b->text_bytes || rows > b->rows_cap) {
        BALL_CUDA(b, cudaStreamSynchronize(b->st)); BALL_CUDA(b, cudaStreamSynchronize(b->st_copy));
        for (int k = 0; k < 2; ++k) { cudaFree(b->d_text[k]); b->d_text[k] = nullptr; }
        cudaFree(b->d_c); cudaFree(b->d_nonn); cudaFree(b->d_keep);
        b->d_c = b->d_nonn = nullptr; b->d_keep = nullptr; b->text_bytes = 0; b->rows_cap = 0;
        for (int k = 0; k < 2; ++k) BALL_CUDA(b, cudaMalloc(&b->d_text[k], (size_t)rows * L));
        BALL_CUDA(b, cudaMalloc(&b->d_c, rows * 4 * 2)); BALL_CUDA(b, cudaMalloc(&b->d_nonn, rows * 4 * 2)); BALL_CUDA(b, cudaMalloc(&b->d_keep, rows * 2));
        b->text_bytes = (size_t)rows * L; b->rows_cap = rows;
    }
    const int64_t pieces = (n + rows - 1) / rows;
    auto copy_in = [&](int64_t k) -> cudaError_t {
        const int s = (int)(k & 1);
        const int64_t i = k * rows, m = std::min(rows, n - i);
        cudaError_t e = cudaStreamWaitEvent(b->st_copy, b->ev_free[s], 0);
        if (e == cudaSuccess) e = cudaMemcpyAsync(b->d_text[s], text + i * (int64_t)L, (size_t)m * L, cudaMemcpyHostToDevice, b->st_copy);
        if (e == cudaSuccess) e = cudaEventRecord(b->ev_copied[s], b->st_copy);
        return e;
    };
    BALL_CUDA(b, copy_in(0));
    for (int64_t k = 0; k < pieces; ++k) {
        const int s = (int)(k & 1);
        const int64_t i = k * rows, m = std::min(rows, n - i);
        if (k + 1 < pieces) BALL_CUDA(b, copy_in(k + 1));                   // the next piece crosses PCIe under this piece's kernels
        BALL_CUDA(b, cudaStreamWaitEvent(b->st, b->ev_copied[s], 0));
        int32_t* d_c = b->d_c + s * rows; int32_t* d_nn = b->d_nonn + s * rows; uint8_t* d_keep = b->d_keep + s * rows;
        k_ball_cons<<<(unsigned)m, BALL_THREADS, 0, b->st>>>(b->d_text[s], m, L, b->d_cc, min_non_n, R, d_c, d_nn, d_keep);
        BALL_CUDA(b, cudaGetLastError());
        k_ball_query<<<(unsigned)m, BALL_THREADS, (size_t)std::max(1, b->n_words) * 4, b->st>>>(b->d_text[s], m, L, b->d_idx, b->n_idx, b->n_words, b->d_qxT,
                                                                                             b->n_query, b->q_stride, R, d_c, d_keep);
        BALL_CUDA(b, cudaGetLastError());
        BALL_CUDA(b, cudaEventRecord(b->ev_free[s], b->st));
        BALL_CUDA(b, cudaMemcpyAsync(keep_out + i, d_keep, (size_t)m, cudaMemcpyDeviceToHost, b->st));
        if (non_n_out) BALL_CUDA(b, cudaMemcpyAsync(non_n_out + i, d_nn, (size_t)m * 4, cudaMemcpyDeviceToHost, b->st));
    }
    BALL_CUDA(b, cudaStreamSynchronize(b->st));
    return UVG_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// uvg_ball_classify_device_noalloc: the two kernels of uvg_ball_push_refs for callers that must not allocate (code under CUDA
// graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Rows and results are the caller's DEVICE buffers; the scratch between the kernels is the file-scope
// array g_ball_scratch.  No cudaMalloc / cudaFree / allocating library routine is called here.
// ---------------------------------------------------------------------------------------------------------------------
namespace {
constexpr int BALL_MAX_DEVICES = 64;
std::mutex g_scratch_mutex;
BallScratch* g_scratch_ptr[BALL_MAX_DEVICES] = {};       // device address of g_ball_scratch on every device that asked for it

// once per device, with that device current: the device address of the symbol (the host-side name is NOT one)
cudaError_t noalloc_scratch_of(int device, BallScratch** out)
{
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    if (!g_scratch_ptr[device]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_ball_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[device] = static_cast<BallScratch*>(ptr);
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}
}  // namespace

extern "C" int uvg_ball_classify_device_noalloc(uvg_ball* b, const uint8_t* d_text, int64_t n, int32_t min_non_n, uint8_t* d_keep, int32_t* d_non_n)
{
    if (!b || (!d_text && n > 0) || n < 0 || (!d_keep && n > 0)) { if (b) b->err = "uvg_ball_classify_device_noalloc: bad arguments"; return UVG_EINVAL; }
    if (b->n_query == 0) BALL_FAIL(b, UVG_ESTATE, "uvg_ball_classify_device_noalloc: no queries set");
    if (n > (int64_t)UVG_BALL_NOALLOC_MAX_ROWS)
        BALL_FAIL(b, UVG_EINVAL, "uvg_ball_classify_device_noalloc: %lld rows exceed the %d rows the static scratch array holds "
                  "(UVG_BALL_NOALLOC_MAX_ROWS); split the call or use uvg_ball_push_refs", (long long)n, (int)UVG_BALL_NOALLOC_MAX_ROWS);
    if (b->prm.device >= BALL_MAX_DEVICES) BALL_FAIL(b, UVG_EINVAL, "uvg_ball_classify_device_noalloc: device ordinal %d beyond %d", b->prm.device, BALL_MAX_DEVICES);
    BALL_CUDA(b, cudaSetDevice(b->prm.device));
    if (n == 0) return UVG_OK;
    BallScratch* scratch = nullptr;
    BALL_CUDA(b, noalloc_scratch_of(b->prm.device, &scratch));
    const int L = b->nchar, R = b->dist + 1;
    // nothing in the scratch has to start at zero: k_ball_cons writes c / non_n of every row < n before k_ball_query reads them
    // (same stream).  It keeps the last call's contents; two calls that overlap on the GPU (other streams, other contexts of the
    // same device, other threads) would race on it -- every call runs on the context's stream and is waited for here.
    int32_t* d_c = scratch->c;
    int32_t* d_nn = d_non_n ? d_non_n : scratch->non_n;
    k_ball_cons<<<(unsigned)n, BALL_THREADS, 0, b->st>>>(d_text, n, L, b->d_cc, min_non_n, R, d_c, d_nn, d_keep);
    BALL_CUDA(b, cudaGetLastError());
    k_ball_query<<<(unsigned)n, BALL_THREADS, (size_t)std::max(1, b->n_words) * 4, b->st>>>(d_text, n, L, b->d_idx, b->n_idx, b->n_words, b->d_qxT,
                                                                                         b->n_query, b->q_stride, R, d_c, d_keep);
    BALL_CUDA(b, cudaGetLastError());
    BALL_CUDA(b, cudaStreamSynchronize(b->st));
    return UVG_OK;
}
