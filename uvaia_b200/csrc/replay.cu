// K3: exact replay of the reference's per-query sequential admission over the dense score tile of one round.
// Replaces heap_insert / heap_bubble_up / heap_bubble_down (src/min_heap.c:93-147) and the max_incompatible update
// of queue_update_min_heaps (src/nearest.c:488, 504-508 / 450, 472-475).  The reference admits a pair only if
//   D = valid - acgt  <  heap.max_incompatible      (a limit that is re-derived from the worst kept item after every
//                                                    insertion and can go UP as well as down, README.md:337-342)
// and then only if its 6-tuple is strictly better than the heap root, so the result depends on the order in which
// references arrive: a parallel top-k is not bit-exact.  Here the pair-count kernel stores the four counts of EVERY
// pair of the round (16 B per pair, coalesced; a few ms of HBM traffic per million references), and one thread block
// per query streams its row in file order: all threads pre-filter 1024 references at a time with a bound that is
// safe for the whole window (Threshold in common.cuh), then one thread applies the exact rule to the few candidates.
#include "common.cuh"
#include "kernels.h"
#include <algorithm>

namespace uvg {

constexpr int RP_THREADS = 128;
constexpr int RP_WIN = 1024;                 // references per window
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
__device__ __forceinline__ int item_mismatch(const HeapItem& it, int acgt_mode)
{
    return acgt_mode ? (it.s[1] - it.s[0]) : (it.s[3] - it.s[0]);
}

struct RpState { int n, max_inc, dmax, root_s0; };

__global__ void __launch_bounds__(RP_THREADS) k_replay(const ReplayParams p, int heap_in_smem)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint4* s_sc = reinterpret_cast<uint4*>(smem);                                  // RP_WIN raw counts
    uint32_t* s_mask = reinterpret_cast<uint32_t*>(smem + RP_WIN * sizeof(uint4)); // RP_WIN / 32 candidate bits
    int32_t* s_misc = reinterpret_cast<int32_t*>(s_mask + RP_WIN / 32);            // 32 ints
    HeapItem* s_heap = reinterpret_cast<HeapItem*>(s_misc + 32);
    RpState* st = reinterpret_cast<RpState*>(s_misc);

    const int q = p.q_first + blockIdx.x;                 // global query index
    const int K = p.K, acgt = p.acgt_mode;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    HeapItem* gheap = p.heaps + (size_t)q * (K + 1);
    HeapItem* a = heap_in_smem ? s_heap : gheap;
    const uint4* row = p.scores + (size_t)blockIdx.x * p.row_stride;
    {
        const int n0 = p.heap_n[q];
        if (heap_in_smem) for (int i = 1 + threadIdx.x; i <= n0; i += blockDim.x) a[i] = gheap[i];
        if (threadIdx.x == 0) {
            st->n = n0; st->max_inc = p.max_inc[q];
            const Threshold& t = p.thr[q];
            st->dmax = t.full ? t.dmax : 0x7fffffff;
            st->root_s0 = t.full ? t.s[0] : -0x7fffffff;
        }
    }
    int batch_id = p.batch_id[q], max_inc_batch = p.max_inc_batch[q];      // used by thread 0 only
    int admitted = 0, h1 = 0;
    __syncthreads();

    const int total = p.nseg * p.seg_stride;
    for (int w0 = 0; w0 < total; w0 += RP_WIN) {
        // ---- parallel part: stage the window and flag candidates with a bound that holds for the whole window ----
        const int dmax = st->dmax, root_s0 = st->root_s0;
        bool any_valid = false;
#pragma unroll
        for (int k = 0; k < RP_WIN / RP_THREADS; ++k) {
            const int i = k * RP_THREADS + threadIdx.x, pos = w0 + i;
            bool cand = false;
            if (pos < total) {
                const int g = pos / p.seg_stride, within = pos - g * p.seg_stride;
                if (within < p.seg_cnt[g]) {
                    const uint4 sc = row[pos];
                    s_sc[i] = sc;
                    const int D = acgt ? (int)(sc.y - sc.x) : (int)(sc.w - sc.x);
                    cand = (D >= 0) && (D < dmax) && ((int)sc.x >= root_s0);
                    any_valid = true;
                }
            }
            const uint32_t m = __ballot_sync(0xffffffffu, cand);
            if (lane == 0) s_mask[k * (RP_THREADS / 32) + warp] = m;
        }
        if (!__syncthreads_or(any_valid ? 1 : 0)) continue;      // padding between segments
        // ---- sequential part: exact admission in file order ----
        if (threadIdx.x == 0) {
            int n = st->n, max_inc = st->max_inc;
            bool changed = false;
            for (int mw = 0; mw < RP_WIN / 32; ++mw) {
                uint32_t m = s_mask[mw];
                while (m) {
                    const int b = __ffs(m) - 1;
                    m &= m - 1;
                    const int i = mw * 32 + b, pos = w0 + i;
                    const uint4 sc = s_sc[i];
                    const int D = acgt ? (int)(sc.y - sc.x) : (int)(sc.w - sc.x);
                    const int g = pos / p.seg_stride, within = pos - g * p.seg_stride;
                    const int r = (int)(p.seg_global[g] + within);
                    const int bt = r / p.pool;
                    if (bt != batch_id) { batch_id = bt; max_inc_batch = max_inc; }
                    if (D >= max_inc) continue;                                     // nearest.c:488-496 / 450-460 (closed form)
                    HeapItem it;
                    const int c0 = p.c0[pos], nn = p.non_n[pos];
                    if (!acgt) { it.s[0] = sc.x; it.s[1] = sc.y; it.s[2] = sc.z; it.s[3] = sc.w; it.s[4] = (int)sc.x - c0; it.s[5] = nn; }
                    else { it.s[0] = sc.x; it.s[1] = sc.y; it.s[2] = (int)sc.x - c0; it.s[3] = nn; it.s[4] = D - (int)sc.z; it.s[5] = sc.z; }
                    it.r = r; it.pad = 0;
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
                    changed = true;
                    p.ever_best[r] = 1;                                             // is_best, nearest.c:505
                    if (p.ci[pos] >= max_inc_batch) ++h1;                           // hazard H1 (necessary condition)
                    if (n == K) max_inc = item_mismatch(a[1], acgt) + 1;            // nearest.c:475 / :508
                }
            }
            st->n = n; st->max_inc = max_inc;
            if (changed && n == K) {
                int dm = 0;
                for (int i = 1; i <= n; ++i) dm = max(dm, item_mismatch(a[i], acgt));
                st->dmax = dm + 1; st->root_s0 = a[1].s[0];
            }
        }
        __syncthreads();
    }

    const int n = st->n;
    if (threadIdx.x == 0) {
        p.heap_n[q] = n; p.max_inc[q] = st->max_inc; p.batch_id[q] = batch_id; p.max_inc_batch[q] = max_inc_batch;
        if (admitted) atomicAdd(p.counters + 0, (unsigned long long)admitted);
        if (h1) atomicAdd(p.counters + 1, (unsigned long long)h1);
    }
    if (n == K) {
        // threshold snapshot for the next rounds: heap root + admission staircase (see Threshold in common.cuh)
        int dm = 0;
        for (int i = 1 + threadIdx.x; i <= n; i += blockDim.x) dm = max(dm, item_mismatch(a[i], acgt));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) dm = max(dm, __shfl_xor_sync(0xffffffffu, dm, o));
        if (lane == 0) s_misc[16 + warp] = dm;
        __syncthreads();
        for (int k = 0; k < RP_THREADS / 32; ++k) dm = max(dm, s_misc[16 + k]);
        int sh = 0;
        while ((dm >> sh) >= NBAR) ++sh;
        if (threadIdx.x < NBAR) {
            const int lo = (int)threadIdx.x << sh;
            int mn = 0x7fffffff;
            for (int i = 1; i <= n; ++i)
                if (item_mismatch(a[i], acgt) >= lo) mn = min(mn, a[i].s[0]);
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
    size_t fixed = RP_WIN * sizeof(uint4) + RP_WIN / 32 * 4 + 32 * 4;
    *in_smem = (K + 1) <= RP_HEAP_SMEM_ITEMS;
    return fixed + (*in_smem ? (size_t)(K + 1) * sizeof(HeapItem) : 0);
}

cudaError_t replay_setup(int K)
{
    int in_smem;
    size_t bytes = replay_smem_bytes(K, &in_smem);
    return cudaFuncSetAttribute(k_replay, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

cudaError_t launch_replay(const ReplayParams& p, cudaStream_t st)
{
    if (p.n_query <= 0) return cudaSuccess;
    int in_smem;
    size_t bytes = replay_smem_bytes(p.K, &in_smem);
    k_replay<<<p.n_query, RP_THREADS, bytes, st>>>(p, in_smem);
    return cudaGetLastError();
}

}  // namespace uvg
