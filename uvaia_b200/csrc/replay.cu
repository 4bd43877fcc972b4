// K3: exact replay of the reference's per-query sequential admission over the survivors of one scoring round.
// Replaces heap_insert / heap_bubble_up / heap_bubble_down (src/min_heap.c:93-147) and the max_incompatible update
// of queue_update_min_heaps (src/nearest.c:488, 504-508 / 450, 472-475).  The reference admits a pair only if
//   D = valid - acgt  <  heap.max_incompatible      (a limit that is re-derived from the worst kept item after every
//                                                    insertion and can go UP as well as down, README.md:337-342)
// and then only if its 6-tuple is strictly better than the heap root, so the result depends on the order in which
// references arrive.  The pair-count kernel therefore only filters with a STALE root (safe: the root never gets
// worse) and this kernel re-applies the exact rule to the survivors of each query in file order.
// One thread block per query: the heap lives in shared memory, survivors are ordered by a direct-address window
// (at most one survivor per (query, reference)), and one thread performs the sequential admission.
#include "common.cuh"
#include "kernels.h"
#include <algorithm>

namespace uvg {

constexpr int RP_THREADS = 128;
constexpr int RP_WIN = 1024;                 // references per ordering window
constexpr int RP_MAXWIN_WORDS = 256;         // occupancy bitmap: up to 8192 windows = 8.4M references per round
constexpr int RP_HEAP_SMEM_ITEMS = 4096;     // heaps up to this many slots live in shared memory

__device__ __forceinline__ bool item_less(const HeapItem& x, const HeapItem& y)
{   // x is worse than y in the order of compare_q_item_score (src/min_heap.c:41-47)
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        if (x.s[i] != y.s[i]) return x.s[i] < y.s[i];
    }
    return false;
}
__device__ __forceinline__ void heap_down(HeapItem* a, int n, int p)
{   // heap_bubble_down, src/min_heap.c:119-134 (left child examined first, strict comparisons)
    for (;;) {
        int c = 2 * p, m = p;
        if (c <= n && item_less(a[c], a[m])) m = c;
        if (c + 1 <= n && item_less(a[c + 1], a[m])) m = c + 1;
        if (m == p) return;
        HeapItem t = a[p]; a[p] = a[m]; a[m] = t;
        p = m;
    }
}
__device__ __forceinline__ void heap_up(HeapItem* a, int i)
{   // heap_bubble_up, src/min_heap.c:136-147
    while (i > 1) {
        int par = i >> 1;
        if (!item_less(a[i], a[par])) return;
        HeapItem t = a[par]; a[par] = a[i]; a[i] = t;
        i = par;
    }
}

__global__ void __launch_bounds__(1024) k_bucket_scan(const int32_t* __restrict__ qcount, int32_t* __restrict__ qoffset,
                                                      int32_t* __restrict__ qcursor, int n_query)
{
    __shared__ int s_warp[32];
    __shared__ int s_run;
    if (threadIdx.x == 0) s_run = 0;
    __syncthreads();
    for (int base = 0; base < n_query; base += blockDim.x) {
        int i = base + threadIdx.x;
        int a = (i < n_query) ? qcount[i] : 0;
        int incl = a;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if ((int)(threadIdx.x & 31) >= o) incl += t; }
        if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = incl;
        __syncthreads();
        if (threadIdx.x < 32) {
            int v = s_warp[threadIdx.x], inc2 = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, inc2, o); if ((int)threadIdx.x >= o) inc2 += t; }
            s_warp[threadIdx.x] = inc2 - v;
        }
        __syncthreads();
        int excl = incl - a + s_warp[threadIdx.x >> 5];
        int run = s_run;
        if (i < n_query) { qoffset[i] = run + excl; qcursor[i] = 0; }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) s_run = run + excl + a;
        __syncthreads();
    }
    if (threadIdx.x == 0) qoffset[n_query] = s_run;
}

__global__ void __launch_bounds__(256) k_bucket_scatter(const Survivor* __restrict__ surv, const int32_t* __restrict__ surv_ci,
                                                        int64_t n, const int32_t* __restrict__ qoffset, int32_t* __restrict__ qcursor,
                                                        Survivor* __restrict__ out, int32_t* __restrict__ out_ci)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        Survivor s = surv[i];
        int pos = qoffset[s.q] + atomicAdd(qcursor + s.q, 1);
        out[pos] = s;
        out_ci[pos] = surv_ci[i];
    }
}

__global__ void __launch_bounds__(RP_THREADS) k_replay(const ReplayParams p, int heap_in_smem)
{
    extern __shared__ __align__(16) uint8_t smem[];
    Survivor* s_rec = reinterpret_cast<Survivor*>(smem);                       // RP_WIN
    int32_t* s_ci = reinterpret_cast<int32_t*>(smem + RP_WIN * sizeof(Survivor));
    int32_t* s_slot = s_ci + RP_WIN;                                           // RP_WIN
    uint32_t* s_occ = reinterpret_cast<uint32_t*>(s_slot + RP_WIN);            // RP_MAXWIN_WORDS
    int32_t* s_misc = reinterpret_cast<int32_t*>(s_occ + RP_MAXWIN_WORDS);     // 64 ints
    HeapItem* s_heap = reinterpret_cast<HeapItem*>(s_misc + 64);

    const int q = blockIdx.x;
    const int cnt = p.qcount[q];
    if (cnt == 0) return;
    const int off = p.qoffset[q];
    const int K = p.K;
    HeapItem* gheap = p.heaps + (size_t)q * (K + 1);
    HeapItem* a = heap_in_smem ? s_heap : gheap;
    int n = p.heap_n[q];
    if (heap_in_smem) for (int i = 1 + threadIdx.x; i <= n; i += blockDim.x) a[i] = gheap[i];
    const int nwin = (p.round_count + RP_WIN - 1) / RP_WIN;
    for (int i = threadIdx.x; i < RP_MAXWIN_WORDS; i += blockDim.x) s_occ[i] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
        int w = (p.bucketed[off + i].r - p.round_first) / RP_WIN;
        atomicOr(s_occ + (w >> 5), 1u << (w & 31));
    }
    // sequential state (thread 0)
    int max_inc = p.max_inc[q], batch_id = p.batch_id[q], max_inc_batch = p.max_inc_batch[q];
    int admitted = 0, h1 = 0;
    __syncthreads();
    for (int w = 0; w < nwin; ++w) {
        if (!((s_occ[w >> 5] >> (w & 31)) & 1u)) continue;                      // uniform across the block
        const int w0 = p.round_first + w * RP_WIN;
        for (int i = threadIdx.x; i < RP_WIN; i += blockDim.x) s_slot[i] = -1;
        __syncthreads();
        for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
            int d = p.bucketed[off + i].r - w0;
            if (d >= 0 && d < RP_WIN) s_slot[d] = i;
        }
        __syncthreads();
        // ordered compaction: thread t owns slots [t*8, t*8+8)
        constexpr int PER = RP_WIN / RP_THREADS;
        int mine = 0;
#pragma unroll
        for (int k = 0; k < PER; ++k) mine += (s_slot[threadIdx.x * PER + k] >= 0);
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(0xffffffffu, incl, o); if ((int)(threadIdx.x & 31) >= o) incl += t; }
        if ((threadIdx.x & 31) == 31) s_misc[threadIdx.x >> 5] = incl;
        __syncthreads();
        int before = incl - mine;
        for (int k = 0; k < (int)(threadIdx.x >> 5); ++k) before += s_misc[k];
        int total = 0;
        for (int k = 0; k < RP_THREADS / 32; ++k) total += s_misc[k];
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            int i = s_slot[threadIdx.x * PER + k];
            if (i >= 0) { s_rec[before] = p.bucketed[off + i]; s_ci[before] = p.bucketed_ci[off + i]; ++before; }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int t = 0; t < total; ++t) {
                const Survivor& sv = s_rec[t];
                const int b = sv.r / p.pool;
                if (b != batch_id) { batch_id = b; max_inc_batch = max_inc; }
                const int D = p.acgt_mode ? (sv.s[1] - sv.s[0]) : (sv.s[3] - sv.s[0]);
                if (D >= max_inc) continue;                                     // nearest.c:488-496 / 450-460 (closed form)
                HeapItem it;
#pragma unroll
                for (int k = 0; k < 6; ++k) it.s[k] = sv.s[k];
                it.r = sv.r; it.pad = 0;
                if (n == K) {                                                   // heap_insert, min_heap.c:95-106
                    if (!item_less(a[1], it)) continue;
                    a[1] = it;
                    heap_down(a, n, 1);
                } else {                                                        // min_heap.c:108-116
                    ++n;
                    a[n] = it;
                    heap_up(a, n);
                }
                ++admitted;
                p.ever_best[sv.r] = 1;                                          // is_best, nearest.c:505
                if (s_ci[t] >= max_inc_batch) ++h1;                             // hazard H1 (necessary condition)
                if (n == K) max_inc = p.acgt_mode ? (a[1].s[1] - a[1].s[0] + 1) : (a[1].s[3] - a[1].s[0] + 1);   // :475 / :508
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        s_misc[40] = n;
        p.heap_n[q] = n; p.max_inc[q] = max_inc; p.batch_id[q] = batch_id; p.max_inc_batch[q] = max_inc_batch;
        p.qcount[q] = 0;
        if (admitted) atomicAdd(p.counters + 0, (unsigned long long)admitted);
        if (h1) atomicAdd(p.counters + 1, (unsigned long long)h1);
    }
    __syncthreads();
    n = s_misc[40];
    if (n == K) {
        // threshold snapshot for the next rounds: heap root + admission staircase (see Threshold in common.cuh)
        int dm = 0;
        for (int i = 1 + threadIdx.x; i <= n; i += blockDim.x) {
            const int d = p.acgt_mode ? (a[i].s[1] - a[i].s[0]) : (a[i].s[3] - a[i].s[0]);
            dm = max(dm, d);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) dm = max(dm, __shfl_xor_sync(0xffffffffu, dm, o));
        if ((threadIdx.x & 31) == 0) s_misc[48 + (threadIdx.x >> 5)] = dm;
        __syncthreads();
        for (int k = 0; k < RP_THREADS / 32; ++k) dm = max(dm, s_misc[48 + k]);
        int sh = 0;
        while ((dm >> sh) >= NBAR) ++sh;
        if (threadIdx.x < NBAR) {
            const int lo = (int)threadIdx.x << sh;
            int mn = 0x7fffffff;
            for (int i = 1; i <= n; ++i) {
                const int d = p.acgt_mode ? (a[i].s[1] - a[i].s[0]) : (a[i].s[3] - a[i].s[0]);
                if (d >= lo) mn = min(mn, a[i].s[0]);
            }
            p.thr[q].bar[threadIdx.x] = mn;
        }
        if (threadIdx.x == 0) {
#pragma unroll
            for (int k = 0; k < 6; ++k) p.thr[q].s[k] = a[1].s[k];
            p.thr[q].dmax = dm + 1;
            p.thr[q].full = 1 + sh;
        }
    }
    if (heap_in_smem) for (int i = 1 + threadIdx.x; i <= n; i += blockDim.x) gheap[i] = a[i];
}

static size_t replay_smem_bytes(int K, int* in_smem)
{
    size_t fixed = RP_WIN * sizeof(Survivor) + RP_WIN * 4 * 2 + RP_MAXWIN_WORDS * 4 + 64 * 4;
    *in_smem = (K + 1) <= RP_HEAP_SMEM_ITEMS;
    return fixed + (*in_smem ? (size_t)(K + 1) * sizeof(HeapItem) : 0);
}

cudaError_t replay_setup(int K)
{
    int in_smem;
    size_t bytes = replay_smem_bytes(K, &in_smem);
    return cudaFuncSetAttribute(k_replay, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

__global__ void __launch_bounds__(256) k_bucket_count(const Survivor* __restrict__ surv, int64_t n, int32_t* __restrict__ qcount)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(qcount + surv[i].q, 1);
}

cudaError_t launch_bucket_count(const Survivor* seg, int64_t n, int32_t* qcount, cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    int64_t blocks = std::min<int64_t>((n + 255) / 256, 148 * 32);
    k_bucket_count<<<(unsigned)blocks, 256, 0, st>>>(seg, n, qcount);
    return cudaGetLastError();
}
cudaError_t launch_bucket_scan(const int32_t* qcount, int32_t* qoffset, int32_t* qcursor, int n_query, cudaStream_t st)
{
    k_bucket_scan<<<1, 1024, 0, st>>>(qcount, qoffset, qcursor, n_query);
    return cudaGetLastError();
}
cudaError_t launch_bucket_scatter(const Survivor* seg, const int32_t* seg_ci, int64_t n, const int32_t* qoffset, int32_t* qcursor,
                                  Survivor* out, int32_t* out_ci, cudaStream_t st)
{
    if (n <= 0) return cudaSuccess;
    int64_t blocks = std::min<int64_t>((n + 255) / 256, 148 * 32);
    k_bucket_scatter<<<(unsigned)blocks, 256, 0, st>>>(seg, seg_ci, n, qoffset, qcursor, out, out_ci);
    return cudaGetLastError();
}
cudaError_t launch_replay_only(const ReplayParams& p, cudaStream_t st)
{
    if (p.n_query <= 0) return cudaSuccess;
    if ((p.round_count + RP_WIN - 1) / RP_WIN > RP_MAXWIN_WORDS * 32) return cudaErrorInvalidValue;
    int in_smem;
    size_t bytes = replay_smem_bytes(p.K, &in_smem);
    k_replay<<<p.n_query, RP_THREADS, bytes, st>>>(p, in_smem);
    return cudaGetLastError();
}

}  // namespace uvg
