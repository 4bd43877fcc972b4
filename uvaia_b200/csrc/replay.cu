// K3: exact replay of the reference's per-query sequential admission over the dense score tile of one round.
// Replaces heap_insert / heap_bubble_up / heap_bubble_down (src/min_heap.c:93-147) and the max_incompatible update
// of queue_update_min_heaps (src/nearest.c:488, 504-508 / 450, 472-475).  The reference admits a pair only if
//   D = valid - acgt  <  heap.max_incompatible      (a limit that is re-derived from the worst kept item after every
//                                                    insertion and can go UP as well as down, README.md:337-342)
// and then only if its 6-tuple is strictly better than the heap root, so the result depends on the order in which
// references arrive: a parallel top-k is not bit-exact.  Here the pair-count kernel stores the four counts of EVERY
// pair of the round (16 B per pair), and one thread block per query streams its row in file order.  The state that
// decides admission (heap size, max_incompatible, root) only changes when a pair IS admitted, which is rare
// (~K ln(R/K) times per query), so all threads test 1,024 references at a time against the CURRENT state -- an exact
// test, not a bound -- one thread admits the first survivor, and the rest of the window is re-tested from shared
// memory against the new state.
#include "common.cuh"
#include "kernels.h"
#include <algorithm>

namespace uvg {

constexpr int RP_THREADS = 128;
constexpr int RP_WIN = 1024;                 // references per window
constexpr int RP_PER = RP_WIN / RP_THREADS;
constexpr int RP_HEAP_SMEM_ITEMS = 1024;     // heaps up to this many slots live in shared memory

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

struct RpState {
    int n, max_inc;
    int root[4];          // first keys of the heap root (valid when n == K)
    int cur_batch, mib;   // batch (of --pool accepted references) of the last candidate and max_incompatible at its start
    int restart;          // first position to re-test after an admission, or -1
};

__global__ void __launch_bounds__(RP_THREADS) k_replay(const ReplayParams p, int heap_in_smem)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint4* s_sc = reinterpret_cast<uint4*>(smem);                                  // RP_WIN raw counts
    uint32_t* s_mask = reinterpret_cast<uint32_t*>(smem + RP_WIN * sizeof(uint4)); // candidate bits
    uint32_t* s_fmask = s_mask + RP_WIN / 32;                                      // hazard-H1 bits
    int32_t* s_misc = reinterpret_cast<int32_t*>(s_fmask + RP_WIN / 32);           // 32 ints
    HeapItem* s_heap = reinterpret_cast<HeapItem*>(s_misc + 32);
    RpState* st = reinterpret_cast<RpState*>(s_misc);

    const int row = blockIdx.x;
    const int q = ((row / QTILE) * p.world + p.rank) * QTILE + row % QTILE;        // query this row belongs to
    if (q >= p.n_query) return;
    const int K = p.K, acgt = p.acgt_mode;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    HeapItem* gheap = p.heaps + (size_t)row * (K + 1);
    HeapItem* a = heap_in_smem ? s_heap : gheap;
    const uint4* trow = p.tile + (size_t)row * p.row_stride;
    const uint4* tcons = p.tile + (size_t)p.rows_cap * p.row_stride;
    const uint4* tmeta = tcons + p.row_stride;
    {
        const int n0 = p.heap_n[row];
        if (heap_in_smem) for (int i = 1 + threadIdx.x; i <= n0; i += blockDim.x) a[i] = gheap[i];
        __syncthreads();
        if (threadIdx.x == 0) {
            st->n = n0; st->max_inc = p.max_inc[row];
            for (int k = 0; k < 4; ++k) st->root[k] = n0 == K ? a[1].s[k] : 0;
            st->cur_batch = p.batch_id[row]; st->mib = p.max_inc_batch[row];
            st->restart = -1;
        }
    }
    int admitted = 0, h1 = 0;                                                      // thread 0 only
    __syncthreads();

    for (int g = 0; g < p.nseg; ++g) {
        const int seg_n = p.seg_cnt[g];
        const int64_t seg_global = p.seg_global[g];
        const int64_t seg_col = (int64_t)g * p.seg_stride;
        for (int w0 = 0; w0 < seg_n; w0 += RP_WIN) {
            const int cntw = min(RP_WIN, seg_n - w0);
#pragma unroll
            for (int k = 0; k < RP_PER; ++k) {
                const int i = k * RP_THREADS + threadIdx.x;
                if (i < cntw) s_sc[i] = trow[seg_col + w0 + i];
            }
            int start = 0;
            for (;;) {
                __syncthreads();
                // ---- all threads: test the window against the current state ----
                const int n = st->n, max_inc = st->max_inc, mib = st->mib, cur_batch = st->cur_batch;
                const int r0 = st->root[0], r1 = st->root[1], r2 = st->root[2], r3 = st->root[3];
#pragma unroll
                for (int k = 0; k < RP_PER; ++k) {
                    const int i = k * RP_THREADS + threadIdx.x;
                    bool cand = false, flag = false;
                    if (i >= start && i < cntw) {
                        const uint4 sc = s_sc[i];
                        const int x = (int)sc.x, y = (int)sc.y, z = (int)sc.z, w = (int)sc.w;
                        const int D = acgt ? (y - x) : (w - x);
                        if (D < max_inc) {
                            // strictly better than the root is decided on all six keys by thread 0; here: not worse on the
                            // keys the pair record holds
                            if (n < K) cand = true;
                            else if (acgt) cand = x > r0 || (x == r0 && y >= r1);
                            else cand = x > r0 || (x == r0 && (y > r1 || (y == r1 && (z > r2 || (z == r2 && w >= r3)))));
                        } else if (mib < max_inc) {
                            // hazard H1: max_incompatible rose inside the current batch; the reference would
                            // have scored this pair with consensus counts truncated at the batch-start limit.  Necessary
                            // condition for its result to differ from the closed form:
                            const int64_t ord = seg_global + w0 + i;
                            if ((int)(ord / p.pool) == cur_batch) {
                                const uint4 cc = tcons[seg_col + w0 + i];
                                const int ci = acgt ? (int)(cc.y - cc.x) : (int)(cc.w - cc.x);
                                flag = ci >= mib && D - ci + mib < max_inc;
                            }
                        }
                    }
                    const uint32_t m = __ballot_sync(0xffffffffu, cand), f = __ballot_sync(0xffffffffu, flag);
                    if (lane == 0) { s_mask[k * (RP_THREADS / 32) + warp] = m; s_fmask[k * (RP_THREADS / 32) + warp] = f; }
                }
                __syncthreads();
                // ---- thread 0: admit in file order until the state changes ----
                if (threadIdx.x == 0) {
                    int nn = n, mi = max_inc, restart = -1, cb = cur_batch, mb = mib;
                    for (int mw = start >> 5; mw < (cntw + 31) >> 5 && restart < 0; ++mw) {
                        uint32_t m = s_mask[mw];
                        while (m) {
                            const int b = __ffs(m) - 1;
                            m &= m - 1;
                            const int i = mw * 32 + b;
                            if (i < start) continue;
                            const uint4 sc = s_sc[i];
                            const int64_t col = seg_col + w0 + i;
                            const int r = (int)(seg_global + w0 + i);
                            const int bt = r / p.pool;
                            if (bt != cb) { cb = bt; mb = mi; }
                            const uint4 cc = tcons[col];
                            const int c0 = (int)cc.x, nonn = (int)tmeta[col].x;
                            HeapItem it;
                            if (!acgt) { it.s[0] = sc.x; it.s[1] = sc.y; it.s[2] = sc.z; it.s[3] = sc.w; it.s[4] = (int)sc.x - c0; it.s[5] = nonn; }
                            else { it.s[0] = sc.x; it.s[1] = sc.y; it.s[2] = (int)sc.x - c0; it.s[3] = nonn; it.s[4] = (int)(sc.y - sc.x) - (int)sc.z; it.s[5] = sc.z; }
                            it.r = r; it.pad = 0;
                            if (nn == K) {                                                  // heap_insert, min_heap.c:95-106
                                if (!item_less(a[1], it)) continue;
                                a[1] = it;
                                heap_down(a, nn, 1);
                            } else {                                                        // min_heap.c:108-116
                                ++nn;
                                a[nn] = it;
                                heap_up(a, nn);
                            }
                            ++admitted;
                            p.ever_best[r] = 1;                                             // is_best, nearest.c:505
                            {   // the closed form may differ from the reference for THIS pair too (consensus stage truncated)
                                const int ci = acgt ? (int)(cc.y - cc.x) : (int)(cc.w - cc.x);
                                if (ci >= mb) ++h1;
                            }
                            if (nn == K) {                                                  // nearest.c:475 / :508
                                mi = item_mismatch(a[1], acgt) + 1;
                                for (int k = 0; k < 4; ++k) st->root[k] = a[1].s[k];
                                restart = i + 1;
                                break;
                            }
                        }
                    }
                    // hazard flags of the positions that were passed with the state they were tested against
                    const int upto = restart < 0 ? cntw : restart;
                    for (int mw = start >> 5; mw < (upto + 31) >> 5; ++mw) {
                        uint32_t f = s_fmask[mw];
                        if (mw == (start >> 5)) f &= 0xffffffffu << (start & 31);
                        if (mw == (upto >> 5) && (upto & 31)) f &= (1u << (upto & 31)) - 1u;
                        h1 += __popc(f);
                    }
                    st->n = nn; st->max_inc = mi; st->cur_batch = cb; st->mib = mb; st->restart = restart;
                }
                __syncthreads();
                const int restart = st->restart;
                if (restart < 0 || restart >= cntw) break;
                start = restart;
            }
            __syncthreads();
        }
    }

    const int n = st->n;
    if (threadIdx.x == 0) {
        p.heap_n[row] = n; p.max_inc[row] = st->max_inc; p.batch_id[row] = st->cur_batch; p.max_inc_batch[row] = st->mib;
        if (admitted) atomicAdd(p.counters + 0, (unsigned long long)admitted);
        if (h1) atomicAdd(p.counters + 1, (unsigned long long)h1);
    }
    if (n == K) {
        // bounds for the early abort of later rounds (see Threshold in common.cuh)
        int dm = 0;
        for (int i = 1 + threadIdx.x; i <= n; i += blockDim.x) dm = max(dm, item_mismatch(a[i], acgt));
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) dm = max(dm, __shfl_xor_sync(0xffffffffu, dm, o));
        if (lane == 0) s_misc[16 + warp] = dm;
        __syncthreads();
        if (threadIdx.x == 0) {
            for (int k = 0; k < RP_THREADS / 32; ++k) dm = max(dm, s_misc[16 + k]);
            Threshold t;
            t.full = 1; t.dmax = dm + 1; t.root_s0 = a[1].s[0]; t.pad = 0;
            p.thr[row] = t;
        }
    }
    if (heap_in_smem) for (int i = 1 + threadIdx.x; i <= n; i += blockDim.x) gheap[i] = a[i];
}

static size_t replay_smem_bytes(int K, int* in_smem)
{
    size_t fixed = RP_WIN * sizeof(uint4) + 2 * (RP_WIN / 32) * 4 + 32 * 4;
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
    if (p.rows_cap <= 0) return cudaSuccess;
    int in_smem;
    size_t bytes = replay_smem_bytes(p.K, &in_smem);
    k_replay<<<p.rows_cap, RP_THREADS, bytes, st>>>(p, in_smem);
    return cudaGetLastError();
}

}  // namespace uvg
