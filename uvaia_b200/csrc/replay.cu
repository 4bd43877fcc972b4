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
#include <cstdlib>
#include "common.cuh"
#include "kernels.h"
#include <algorithm>

namespace uvg {

// Block size NT (template): 128 threads x windows of 1,024 references when there are many rows (one GPU: a block per query
// fills the machine), 256 threads x windows of 2,048 when a rank owns few rows (8 GPUs: 128 rows) -- a row is streamed by ONE
// block, so its memory-level parallelism is what the replays at the end of a sharded search cost.  Both at 78 registers: the
// blocks live beside the pair-count kernel of the next round, whose two blocks per SM fill the register file, and every
// register a replay block holds keeps pair-count blocks off its SM (512-thread blocks at 114 registers took whole SMs:
// profiles/replay_interference_r02an.log).
constexpr int RP_PER = 8;                    // references per thread and window
constexpr int REPLAY_GRID_CAP = 0;           // most thread blocks of a replay launch (0 = one per row); see replay_row
constexpr int RP_HEAP_SMEM_ITEMS = 1024;     // heaps up to this many slots live in shared memory
#define RP_CONSTS(NT) constexpr int RP_THREADS = (NT); constexpr int RP_WIN = (NT) * RP_PER; (void)RP_WIN;

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
    int cb_begin, cb_end, bk;   // closed form: the reference's batch of the last candidate [cb_begin, cb_end) (global ordinals), its index in p.bounds
    int mib;                    //              max_incompatible of THIS query at the start of that batch
    int restart;          // first position to re-test after an admission, or -1
    int pend;             // literal mode: window position of a pair whose consensus stage the reference truncated, or -1
    int t[4];             // literal mode: the truncated consensus counts of that pair (cons_prefix)
};
static_assert(sizeof(RpState) <= 16 * 4, "RpState lives in the first 16 ints of s_misc");

// ---------------------------------------------------------------------------------------------------------------------
// Literal emulation of hazard H1 (nearest.c:289-294, 428-433, 485-501).  The reference scores the
// consensus stage of a whole batch (--pool accepted references, a batch also ends with its file) with the batch-wide
// limit B = max over the queries of max_incompatible AT THE START of the batch, through
// biomcmc_pairwise_score_matches_truncated_idx (alignment.c:855-884) / quick_pairwise_score_acgt_and_valid
// (fastaseq.c:585-596): the loop over idx_c stops right after the site at which (valid - acgt) reaches B.  For a
// reference whose full consensus mismatches ci are >= B the stored counts are those of that PREFIX of idx_c, and their
// mismatch count is exactly B.  A query whose own limit m rose above B inside the batch then accepts the pair if
// B + (D - ci) < m and scores it with the prefix counts.  cons_prefix recomputes the prefix counts of one such pair from
// the packed planes: block-wide, three passes (per-thread ranges -> the range that holds the B-th mismatch -> binary
// search of the site inside it -> counts up to that site).
// ---------------------------------------------------------------------------------------------------------------------
struct ConsSrc {            // what cons_prefix reads (all device pointers)
    const uint4* db; const QChunk* qcons;      // packed planes of the local references; first chunk of the consensus slot's tile
    int cons_qi;                               // slot of the consensus inside its tile
    const uint32_t* cons_nib;                  // 4-bit codes of the consensus (idx_c sites only)
    const uint32_t* camb; int n_camb;          // ambiguity codes of the consensus (site | code << 28)
    const uint32_t* ramb_off; const uint32_t* ramb_ent;
    int Wp, acgt;
};

template <int NT>
__device__ __forceinline__ int block_sum(int v, int* s_red)
{   // sum over the block; s_red holds RP_THREADS / 32 + 1 ints
    RP_CONSTS(NT)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    int t = 0;
#pragma unroll
    for (int k = 0; k < RP_THREADS / 32; ++k) t += s_red[k];
    return t;
}

// counts [acgt, text, partial, valid] of reference `lr` against the consensus over the sites <= p of the units [ua, ub)
// (plane part only: sites where both hold a plain base)
__device__ __forceinline__ void cons_units(const ConsSrc& c, int64_t lr, int ua, int ub, int64_t p, int& m_out, int& v_out)
{
    const int64_t group = lr / GROUP; const int lane = (int)(lr % GROUP);
    int m = 0, v = 0;
    for (int u = ua; u < ub; ++u) {
        if ((int64_t)u * 128 > p) break;
        const uint4 rb = c.db[db_unit_index(group, u, 0, lane, c.Wp)], r0 = c.db[db_unit_index(group, u, 1, lane, c.Wp)],
                    r1 = c.db[db_unit_index(group, u, 2, lane, c.Wp)];
        const uint4* q = c.qcons[u / W4].pl[u % W4][c.cons_qi];
        const uint4 qb = q[0], q0 = q[1], q1 = q[2];
        const uint32_t R[3][4] = {{rb.x, rb.y, rb.z, rb.w}, {r0.x, r0.y, r0.z, r0.w}, {r1.x, r1.y, r1.z, r1.w}};
        const uint32_t Qv[3][4] = {{qb.x, qb.y, qb.z, qb.w}, {q0.x, q0.y, q0.z, q0.w}, {q1.x, q1.y, q1.z, q1.w}};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int64_t s0 = ((int64_t)u * 4 + k) * 32;
            if (s0 > p) break;
            const uint32_t keep = (p - s0 >= 31) ? 0xffffffffu : ((2u << (uint32_t)(p - s0)) - 1u);
            const uint32_t V = R[0][k] & Qv[0][k] & keep;
            const uint32_t M = V & ~(R[1][k] ^ Qv[1][k]) & ~(R[2][k] ^ Qv[2][k]);
            v += __popc(V); m += __popc(M);
        }
    }
    m_out = m; v_out = v;
}

// sites outside the planes (an ambiguity code on either side), entries e = tid, tid + stride, ...: adds to [text, partial, valid]
// the events with lo <= site <= p.  Never an ACGT match: every such valid site is a mismatch of the truncation rule.
__device__ __forceinline__ void cons_events(const ConsSrc& c, int64_t lr, int64_t lo, int64_t p, int tid, int stride, int& t, int& pa, int& v)
{
    if (c.acgt) return;                                           // --acgt: only A, C, G, T are coded, no lists
    const uint32_t r0 = c.ramb_off[lr], r1 = c.ramb_off[lr + 1];
    for (uint32_t e = r0 + tid; e < r1; e += stride) {
        const uint32_t ent = c.ramb_ent[e], site = ent & AMB_SITE_MASK, rc = ent >> 28;
        if ((int64_t)site < lo || (int64_t)site > p) continue;
        const uint32_t qc = (c.cons_nib[site >> 3] >> ((site & 7u) * 4u)) & 15u;
        if (!qc) continue;
        ++v; pa += (qc & rc) != 0u; t += qc == rc;
    }
    const int64_t group = lr / GROUP; const int lane = (int)(lr % GROUP);
    const uint32_t* dbw = reinterpret_cast<const uint32_t*>(c.db);
    for (int e = tid; e < c.n_camb; e += stride) {
        const uint32_t ent = c.camb[e], site = ent & AMB_SITE_MASK, qc = ent >> 28;
        if ((int64_t)site < lo || (int64_t)site > p) continue;
        const uint32_t w = site >> 5, b = site & 31u;
        const size_t base = db_unit_index(group, (int)(w >> 2), 0, lane, c.Wp) * 4 + (w & 3u);
        if (!((dbw[base] >> b) & 1u)) continue;                   // the reference holds no plain base here (its own codes: loop above)
        const uint32_t x = (dbw[base + GROUP * 4] >> b) & 1u, y = (dbw[base + 2 * GROUP * 4] >> b) & 1u;
        ++v; pa += (qc >> (x + 2u * y)) & 1u;
    }
}

// block-wide; returns (through st->t) the counts the reference's truncated consensus stage holds for reference `lr` at limit B:
// default mode [acgt, text, partial, valid], --acgt [mismatches, valid, -, -]
template <int NT>
__device__ void cons_prefix(const ConsSrc& c, int64_t lr, int B, RpState* st, int* s_red, int* s_scan)
{
    RP_CONSTS(NT)
    const int tid = threadIdx.x;
    const int nU = c.Wp / 4, per = (nU + RP_THREADS - 1) / RP_THREADS;
    const int ua = min(nU, tid * per), ub = min(nU, ua + per);
    const int64_t BIG = (int64_t)c.Wp * 32;
    // pass 1: mismatches (valid - acgt) of every thread's range
    int m, v, et = 0, ep = 0, ev = 0;
    cons_units(c, lr, ua, ub, BIG, m, v);
    if (ub > ua) cons_events(c, lr, (int64_t)ua * 128, (int64_t)ub * 128 - 1, 0, 1, et, ep, ev);
    __syncthreads();
    s_scan[tid] = (v - m) + ev;
    __syncthreads();
    if (tid == 0) {
        int run = 0, win = -1;
        for (int k = 0; k < RP_THREADS; ++k) {
            const int x = s_scan[k];
            if (win < 0 && run + x >= B) { win = k; s_scan[RP_THREADS] = B - run; }
            run += x;
        }
        s_scan[RP_THREADS + 1] = win;
    }
    __syncthreads();
    const int win = s_scan[RP_THREADS + 1], need = s_scan[RP_THREADS];
    int64_t pstar = BIG;                                          // fewer than B mismatches in all: nothing was cut
    if (win >= 0) {
        // pass 2: smallest site p of the winner's range with `need` mismatches in [range start, p]
        const int wa = min(nU, win * per), wb = min(nU, wa + per);
        int64_t lo = (int64_t)wa * 128, hi = (int64_t)wb * 128 - 1;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            int mm = 0, vv = 0, a = 0, b2 = 0, e = 0;
            if (tid < wb - wa) cons_units(c, lr, wa + tid, wa + tid + 1, mid, mm, vv);
            cons_events(c, lr, (int64_t)wa * 128, mid, tid, RP_THREADS, a, b2, e);
            const int cnt = block_sum<NT>((vv - mm) + e, s_red);
            if (cnt >= need) hi = mid; else lo = mid + 1;
        }
        pstar = lo;
    }
    // pass 3: the four counts over the sites <= pstar
    cons_units(c, lr, ua, ub, pstar, m, v);
    et = ep = ev = 0;
    cons_events(c, lr, 0, pstar, tid, RP_THREADS, et, ep, ev);
    const int tm = block_sum<NT>(m, s_red), tv = block_sum<NT>(v, s_red), tt = block_sum<NT>(et, s_red), tp = block_sum<NT>(ep, s_red), te = block_sum<NT>(ev, s_red);
    if (tid == 0) {
        if (c.acgt) { st->t[0] = tv - tm; st->t[1] = tv; st->t[2] = 0; st->t[3] = 0; }
        else { st->t[0] = tm; st->t[1] = tm + tt; st->t[2] = tm + tp; st->t[3] = tv + te; }
    }
    __syncthreads();
}

// One row (= one owned query) of a round.  The thread blocks of k_replay are few and long-lived, and they sit beside the
// pair-count kernel of the next round, whose two blocks per SM fill the register file: every resident replay block keeps a
// pair-count block off its SM for as long as it lives.  So the replay runs with a capped grid (launch_replay) and each
// block takes several rows in turn.
template <int NT>
__device__ __forceinline__ void replay_row(const ReplayParams& p, const int heap_in_smem, const int row)
{
    RP_CONSTS(NT)
    extern __shared__ __align__(16) uint8_t smem[];
    uint4* s_sc = reinterpret_cast<uint4*>(smem);                                  // RP_WIN raw counts
    uint32_t* s_mask = reinterpret_cast<uint32_t*>(smem + RP_WIN * sizeof(uint4)); // candidate bits
    uint32_t* s_fmask = s_mask + RP_WIN / 32;                                      // hazard-H1 bits (closed form: flags; literal: truncated pairs)
    int32_t* s_misc = reinterpret_cast<int32_t*>(s_fmask + RP_WIN / 32);           // 64 ints: state | 16 per-warp maxima | 16 per-warp sums
    int32_t* s_scan = s_misc + 64;                                                 // RP_THREADS + 2 ints (cons_prefix)
    HeapItem* s_heap = reinterpret_cast<HeapItem*>(s_scan + RP_THREADS + 32);
    RpState* st = reinterpret_cast<RpState*>(s_misc);
    int* s_red = s_misc + 32;                                                      // RP_THREADS / 32 ints (block_sum)
    static_assert(NT / 32 <= 16, "per-warp scratch");

    const int q = ((row / QTILE) * p.world + p.rank) * QTILE + row % QTILE;        // query this row belongs to
    if (q >= p.n_query) return;
    // literal launches of a round that the speculative closed-form pass has already settled (no pair could have been
    // scored differently by the reference: 0 flags) have nothing to do
    if (p.spec_flags && *p.spec_flags == 0ull) return;
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
            st->mib = p.max_inc_batch[row];
            // the batch of the last candidate of an earlier launch: still open if it reaches into this launch
            const int b0 = p.batch_id[row];
            int k = 0, e = b0;
            if (p.n_bounds > 0 && b0 >= (int)p.bounds[0]) {
                while (k + 1 < p.n_bounds && (int)p.bounds[k + 1] <= b0) ++k;
                e = k + 1 < p.n_bounds ? (int)p.bounds[k + 1] : b0;
            }
            st->cb_begin = b0; st->cb_end = e; st->bk = k;
            // literal launches never straddle a batch boundary: the first launch of a batch notes where it starts and the
            // query's limit there (what the speculative pass of a later round compares with, k_spec_restore)
            if (p.literal && p.batch_start) { st->cb_begin = (int)p.ord_begin; st->mib = st->max_inc; }
            st->restart = -1; st->pend = -1;
        }
    }
    // literal mode: the batch-wide limit of the consensus stage (nearest.c:290-291).  A launch never straddles a batch
    // boundary; the first launch of a batch takes the maximum the previous launch accumulated, the others the carried value.
    const int lit = p.literal;
    const int B = lit ? (p.batch_start ? *p.b_prev : *p.b_carry_in) : 0;
    if (lit && row == 0 && threadIdx.x == 0) { *p.b_carry_out = B; *p.b_zero = 0; }
    ConsSrc csrc;
    if (lit) {
        csrc.db = nullptr; csrc.qcons = p.qcons; csrc.cons_qi = p.cons_qi; csrc.cons_nib = p.cons_nib; csrc.camb = p.camb; csrc.n_camb = p.n_camb;
        csrc.ramb_off = nullptr; csrc.ramb_ent = nullptr; csrc.Wp = p.Wp; csrc.acgt = acgt;
    }
    int admitted = 0, h1 = 0;                                                      // thread 0 only
    __syncthreads();

    for (int g = 0; g < p.nseg; ++g) {
        const int64_t seg_global = p.seg_global[g];
        const int64_t seg_col = (int64_t)g * p.seg_stride;
        // the launch covers the global ordinals [ord_begin, ord_end) of the round (the whole round unless a batch ends inside it)
        const int seg_lo = (int)max((int64_t)0, p.ord_begin - seg_global);
        const int seg_n = (int)min((int64_t)p.seg_cnt[g], p.ord_end - seg_global);
        // the windows of a row are a serial chain (the state of one feeds the next), so what a window costs is latency: its
        // records are fetched into registers one window ahead, and a window without a candidate -- almost all of them once the
        // heap is full -- takes ONE barrier: every thread tests the records it staged itself, and __syncthreads_or tells
        // whether anybody has to look closer
        uint4 ahead[RP_PER];
#pragma unroll
        for (int k = 0; k < RP_PER; ++k) {
            const int i = k * RP_THREADS + threadIdx.x;
            ahead[k] = seg_lo + i < seg_n ? trow[seg_col + seg_lo + i] : make_uint4(0u, 0u, 0u, 0u);
        }
        for (int w0 = seg_lo; w0 < seg_n; w0 += RP_WIN) {
            const int cntw = min(RP_WIN, seg_n - w0);
#pragma unroll
            for (int k = 0; k < RP_PER; ++k) s_sc[k * RP_THREADS + threadIdx.x] = ahead[k];
#pragma unroll
            for (int k = 0; k < RP_PER; ++k) {
                const int i = w0 + RP_WIN + k * RP_THREADS + threadIdx.x;
                if (i < seg_n) ahead[k] = trow[seg_col + i];
            }
            int start = 0;
            for (;;) {
                // no barrier here: a thread tests the records it wrote itself, and the state was published before the last
                // barrier of the previous pass
                // ---- all threads: test the window against the current state ----
                const int n = st->n, max_inc = st->max_inc, mib = st->mib, cb_begin = st->cb_begin, cb_end = st->cb_end;
                const int r0 = st->root[0], r1 = st->root[1], r2 = st->root[2], r3 = st->root[3];
                uint32_t any = 0;
#pragma unroll
                for (int k = 0; k < RP_PER; ++k) {
                    const int i = k * RP_THREADS + threadIdx.x;
                    bool cand = false, flag = false;
                    if (i >= start && i < cntw) {
                        const uint4 sc = s_sc[i];
                        const int x = (int)sc.x, y = (int)sc.y, z = (int)sc.z, w = (int)sc.w;
                        const int D = acgt ? (y - x) : (w - x);
                        if (lit && B < max_inc) {
                            // this query's limit rose above the batch-wide limit inside the batch: references whose consensus
                            // stage reached B were truncated there (their stored consensus mismatches ARE B)
                            const uint4 cc = tcons[seg_col + w0 + i];
                            const int ci = acgt ? (int)(cc.y - cc.x) : (int)(cc.w - cc.x);
                            if (ci >= B) {
                                // accepted by the three staged tests iff B + (D - ci) < max_inc (nearest.c:488-496); the truncated
                                // item can only be worse than the full one on the first key, so x < root.s0 cannot enter a full heap
                                flag = (D - ci + B < max_inc) && (n < K || x >= r0);
                            } else if (D < max_inc) {
                                if (n < K) cand = true;
                                else if (acgt) cand = x > r0 || (x == r0 && y >= r1);
                                else cand = x > r0 || (x == r0 && (y > r1 || (y == r1 && (z > r2 || (z == r2 && w >= r3)))));
                            }
                        } else if (D < max_inc) {
                            // strictly better than the root is decided on all six keys by thread 0; here: not worse on the
                            // keys the pair record holds
                            if (n < K) cand = true;
                            else if (acgt) cand = x > r0 || (x == r0 && y >= r1);
                            else cand = x > r0 || (x == r0 && (y > r1 || (y == r1 && (z > r2 || (z == r2 && w >= r3)))));
                        } else if (!lit && mib < max_inc) {
                            // hazard H1: max_incompatible rose inside the current batch; the reference would
                            // have scored this pair with consensus counts truncated at the batch-start limit.  Necessary
                            // condition for its result to differ from the closed form:
                            const int64_t ord = seg_global + w0 + i;
                            if (ord >= cb_begin && ord < cb_end) {
                                const uint4 cc = tcons[seg_col + w0 + i];
                                const int ci = acgt ? (int)(cc.y - cc.x) : (int)(cc.w - cc.x);
                                flag = ci >= mib && D - ci + mib < max_inc;
                            }
                        }
                    }
                    const uint32_t m = __ballot_sync(0xffffffffu, cand), f = __ballot_sync(0xffffffffu, flag);
                    if (lane == 0) { s_mask[k * (RP_THREADS / 32) + warp] = m; s_fmask[k * (RP_THREADS / 32) + warp] = f; }
                    any |= m | f;
                }
                if (!__syncthreads_or((int)(any != 0u))) break;                    // nothing to admit, nothing to flag
                // ---- thread 0: admit in file order until the state changes ----
                if (threadIdx.x == 0) {
                    int nn = n, mi = max_inc, restart = -1, cbb = cb_begin, cbe = cb_end, bk = st->bk, mb = mib, pend = -1;
                    // heap_insert (min_heap.c:93-117) + the max_incompatible update (nearest.c:475 / :508); true = the state changed
                    auto admit = [&](const HeapItem& it) -> bool {
                        if (nn == K) {
                            if (!item_less(a[1], it)) return false;
                            a[1] = it;
                            heap_down(a, nn, 1);
                        } else {
                            ++nn;
                            a[nn] = it;
                            heap_up(a, nn);
                        }
                        ++admitted;
                        if (p.ever_round) p.ever_round[it.r - p.ord_round0] = 1;           // speculative pass: committed by k_spec_commit
                        else p.ever_best[it.r] = 1;                                         // is_best, nearest.c:505
                        if (nn == K) {
                            mi = item_mismatch(a[1], acgt) + 1;
                            for (int k = 0; k < 4; ++k) st->root[k] = a[1].s[k];
                        }
                        return true;
                    };
                    for (int mw = start >> 5; mw < (cntw + 31) >> 5 && restart < 0; ++mw) {
                        const uint32_t hm = lit ? s_fmask[mw] : 0u;
                        uint32_t m = s_mask[mw] | hm;
                        while (m) {
                            const int b = __ffs(m) - 1;
                            m &= m - 1;
                            const int i = mw * 32 + b;
                            if (i < start) continue;
                            if ((hm >> b) & 1u) { pend = i; restart = i; break; }       // truncated in the reference: all threads recount the prefix
                            const uint4 sc = s_sc[i];
                            const int64_t col = seg_col + w0 + i;
                            const int r = (int)(seg_global + w0 + i);
                            if (!lit && (r < cbb || r >= cbe)) {
                                // a new batch of the reference program starts for this query: its limit at the start of that batch
                                while (bk + 1 < p.n_bounds && (int)p.bounds[bk + 1] <= r) ++bk;
                                cbb = (int)p.bounds[bk]; cbe = bk + 1 < p.n_bounds ? (int)p.bounds[bk + 1] : 0x7fffffff;
                                mb = mi;
                            }
                            const uint4 cc = tcons[col];
                            const int c0 = (int)cc.x, nonn = (int)tmeta[col].x;
                            HeapItem it;
                            if (!acgt) { it.s[0] = sc.x; it.s[1] = sc.y; it.s[2] = sc.z; it.s[3] = sc.w; it.s[4] = (int)sc.x - c0; it.s[5] = nonn; }
                            else { it.s[0] = sc.x; it.s[1] = sc.y; it.s[2] = (int)sc.x - c0; it.s[3] = nonn; it.s[4] = (int)(sc.y - sc.x) - (int)sc.z; it.s[5] = sc.z; }
                            it.r = r; it.pad = 0;
                            if (!admit(it)) continue;
                            if (!lit) {   // the closed form may differ from the reference for THIS pair too (consensus stage truncated)
                                const int ci = acgt ? (int)(cc.y - cc.x) : (int)(cc.w - cc.x);
                                if (ci >= mb) ++h1;
                            }
                            if (nn == K) { restart = i + 1; break; }
                        }
                    }
                    if (!lit) {
                        // hazard flags of the positions that were passed with the state they were tested against
                        const int upto = restart < 0 ? cntw : restart;
                        for (int mw = start >> 5; mw < (upto + 31) >> 5; ++mw) {
                            uint32_t f = s_fmask[mw];
                            if (mw == (start >> 5)) f &= 0xffffffffu << (start & 31);
                            if (mw == (upto >> 5) && (upto & 31)) f &= (1u << (upto & 31)) - 1u;
                            h1 += __popc(f);
                        }
                    }
                    st->n = nn; st->max_inc = mi; st->cb_begin = cbb; st->cb_end = cbe; st->bk = bk; st->mib = mb; st->restart = restart; st->pend = pend;
                }
                __syncthreads();
                const int pend = st->pend;
                if (pend >= 0) {
                    // literal mode: the pair at `pend` reaches the heap with the TRUNCATED consensus counts (block-wide recount)
                    const int64_t ord = seg_global + w0 + pend;
                    csrc.db = p.seg_db[g]; csrc.ramb_off = p.seg_ramb_off[g]; csrc.ramb_ent = p.seg_ramb_ent[g];     // the rank that holds the reference
                    cons_prefix<NT>(csrc, p.seg_local0 + w0 + pend, B, st, s_red, s_scan);
                    if (threadIdx.x == 0) {
                        const uint4 sc = s_sc[pend];
                        const int64_t col = seg_col + w0 + pend;
                        const uint4 cc = tcons[col];
                        const int nonn = (int)tmeta[col].x;
                        const int t0 = st->t[0], t1 = st->t[1], t2 = st->t[2], t3 = st->t[3];
                        HeapItem it;
                        if (!acgt) {
                            // score[i] = stage counts + truncated consensus counts; score[4] = the stage counts' ACGT matches (nearest.c:499-500)
                            it.s[0] = (int)sc.x - (int)cc.x + t0; it.s[1] = (int)sc.y - (int)cc.y + t1; it.s[2] = (int)sc.z - (int)cc.z + t2;
                            it.s[3] = (int)sc.w - (int)cc.w + t3; it.s[4] = (int)sc.x - (int)cc.x; it.s[5] = nonn;
                        } else {
                            // tile: sc = (matches, valid, mismatches on idx); cc = (consensus matches, consensus valid).  nearest.c:456-470
                            const int V = (int)sc.y - (int)cc.y + t1;                           // valid pairs with the prefix of idx_c
                            const int M = ((int)sc.y - (int)sc.x) - ((int)cc.y - (int)cc.x) + t0;   // mismatches: stages + B
                            it.s[0] = V - M; it.s[1] = V; it.s[2] = (V - M) - (t1 - t0); it.s[3] = nonn; it.s[4] = M - (int)sc.z; it.s[5] = sc.z;
                        }
                        it.r = (int)ord; it.pad = 0;
                        int nn = st->n, mi = st->max_inc;
                        ++h1;
                        if (nn == K) {
                            if (item_less(a[1], it)) {
                                a[1] = it; heap_down(a, nn, 1);
                                ++admitted; p.ever_best[it.r] = 1;
                                mi = item_mismatch(a[1], acgt) + 1;
                                for (int k = 0; k < 4; ++k) st->root[k] = a[1].s[k];
                            }
                        } else {
                            ++nn; a[nn] = it; heap_up(a, nn);
                            ++admitted; p.ever_best[it.r] = 1;
                            if (nn == K) { mi = item_mismatch(a[1], acgt) + 1; for (int k = 0; k < 4; ++k) st->root[k] = a[1].s[k]; }
                        }
                        st->n = nn; st->max_inc = mi; st->pend = -1; st->restart = pend + 1;
                    }
                    __syncthreads();
                }
                const int restart = st->restart;
                if (restart < 0 || restart >= cntw) break;
                start = restart;
            }
        }
    }

    const int n = st->n;
    if (threadIdx.x == 0) {
        p.heap_n[row] = n; p.max_inc[row] = st->max_inc; p.batch_id[row] = st->cb_begin; p.max_inc_batch[row] = st->mib;
        if (lit) atomicMax(p.b_acc, st->max_inc);                                  // batch-wide limit of the next batch = max over the queries
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

template <int NT>
__global__ void __launch_bounds__(NT, NT == 128 ? 6 : 3) k_replay(const ReplayParams p, int heap_in_smem)
{
    for (int row = blockIdx.x; row < p.rows_cap; row += gridDim.x) {
        replay_row<NT>(p, heap_in_smem, row);
        __syncthreads();                                                           // the next row reuses the shared memory
    }
}

// ---- speculation of a literal round (engine.cu run_round) ----
__global__ void k_spec_commit(const unsigned long long* __restrict__ spec, const uint8_t* __restrict__ ever_round, int64_t n, int64_t ord_round0,
                              uint8_t* __restrict__ ever_best, unsigned long long* __restrict__ counters)
{
    if (spec[1] != 0ull) return;                                                   // flagged: the literal launches redo the round
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        if (ever_round[i]) ever_best[ord_round0 + i] = 1;
    if (blockIdx.x == 0 && threadIdx.x == 0 && spec[0]) atomicAdd(counters, spec[0]);
}

__global__ void __launch_bounds__(256) k_spec_restore(const SpecRestore r)
{
    if (r.spec[1] == 0ull) return;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < r.own_vec; i += (int64_t)gridDim.x * blockDim.x) r.own[i] = r.snap_own[i];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < r.rows; i += gridDim.x * blockDim.x) { r.batch_id[i] = r.snap_batch_id[i]; r.mib[i] = r.snap_mib[i]; }
    if (blockIdx.x != 0) return;
    // batch-wide limits of the consensus stage as the reference had them at the start of the round: for a batch that starts
    // there the maximum of the queries' limits now; for a batch in progress the maximum of the limits at ITS start (a query
    // whose last candidate is older than that batch has not changed since)
    __shared__ int s_max[2][8];
    int now = 0, open = 0;
    for (int i = threadIdx.x; i < r.n_query; i += blockDim.x) {
        const int m = r.snap_max_inc[i];
        now = max(now, m);
        open = max(open, r.snap_batch_id[i] == r.open_batch ? r.snap_mib[i] : m);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { now = max(now, __shfl_xor_sync(0xffffffffu, now, o)); open = max(open, __shfl_xor_sync(0xffffffffu, open, o)); }
    if ((threadIdx.x & 31) == 0) { s_max[0][threadIdx.x >> 5] = now; s_max[1][threadIdx.x >> 5] = open; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int k = 0; k < 8; ++k) { now = max(now, s_max[0][k]); open = max(open, s_max[1][k]); }
        *r.b_prev = now; *r.b_carry = open; *r.b_acc = 0;
    }
}

cudaError_t launch_spec_commit(const unsigned long long* spec, const uint8_t* ever_round, int64_t n, int64_t ord_round0, uint8_t* ever_best,
                               unsigned long long* counters, cudaStream_t st)
{
    k_spec_commit<<<(unsigned)std::max<int64_t>(1, std::min<int64_t>(64, (n + 255) / 256)), 256, 0, st>>>(spec, ever_round, n, ord_round0, ever_best, counters);
    return cudaGetLastError();
}

cudaError_t launch_spec_restore(const SpecRestore& r, cudaStream_t st)
{
    k_spec_restore<<<64, 256, 0, st>>>(r);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------------
// Final tables on the device: heap_finalise_heap_qsort (src/min_heap.c:149-153) of every owned row -- slot 0 <-> slot n,
// then the descending sort of slots [0, n) by compare_q_item_score (:41-47).  glibc's qsort is a stable merge sort for
// arrays this small, so rows that compare equal keep their order in the heap array; a rank sort (position = number of rows
// that go before) gives exactly that order.  One block per row; rows leave as uvg_item (six scores + 64-bit ordinal).
// ---------------------------------------------------------------------------------------------------------------------
struct FinalItem { int32_t s[6]; long long ref; };
static_assert(sizeof(FinalItem) == 32, "uvg_item layout");

__global__ void __launch_bounds__(128) k_finalise(const HeapItem* __restrict__ heaps, const int32_t* __restrict__ heap_n, int K, int rows, FinalItem* __restrict__ out)
{
    extern __shared__ uint4 s_fin[];
    HeapItem* a = reinterpret_cast<HeapItem*>(s_fin);
    for (int row = blockIdx.x; row < rows; row += gridDim.x) {
        const int n = heap_n[row];
        const HeapItem* h = heaps + (size_t)row * (K + 1);
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x) a[i] = h[i == 0 ? n : i];
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x) {
            const HeapItem me = a[i];
            int pos = 0;
            for (int j = 0; j < n; ++j) {
                const HeapItem o = a[j];
                int c = 0;                                                   // > 0: o sorts before me
#pragma unroll
                for (int k = 5; k >= 0; --k) { const int d = o.s[k] - me.s[k]; c = d ? d : c; }
                pos += (c > 0) | ((c == 0) & (j < i));
            }
            FinalItem f;
#pragma unroll
            for (int k = 0; k < 6; ++k) f.s[k] = me.s[k];
            f.ref = me.r;
            out[(size_t)row * K + pos] = f;
        }
    }
}

cudaError_t launch_finalise(const HeapItem* heaps, const int32_t* heap_n, int K, int rows, void* out, cudaStream_t st)
{
    if (rows <= 0) return cudaSuccess;
    const size_t smem = (size_t)(K + 1) * sizeof(HeapItem);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k_finalise, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    k_finalise<<<(unsigned)std::min(rows, 148 * 8), 128, smem, st>>>(heaps, heap_n, K, rows, static_cast<FinalItem*>(out));
    return cudaGetLastError();
}

static size_t replay_smem_bytes(int K, int NT, int* in_smem)
{
    const size_t RP_WIN = (size_t)NT * RP_PER;
    size_t fixed = RP_WIN * sizeof(uint4) + 2 * (RP_WIN / 32) * 4 + 64 * 4 + ((size_t)NT + 32) * 4;
    *in_smem = (K + 1) <= RP_HEAP_SMEM_ITEMS;
    return fixed + (*in_smem ? (size_t)(K + 1) * sizeof(HeapItem) : 0);
}

cudaError_t replay_setup(int K)
{
    int in_smem;
    cudaError_t e = cudaFuncSetAttribute(k_replay<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)replay_smem_bytes(K, 128, &in_smem));
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k_replay<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)replay_smem_bytes(K, 256, &in_smem));
    return e;
}

cudaError_t launch_replay(const ReplayParams& p, cudaStream_t st)
{
    if (p.rows_cap <= 0) return cudaSuccess;
    // experiment knobs: UVG_REPLAY_GRID = most blocks of a launch (0: one per row), UVG_REPLAY_WIDE = 1: 256-thread blocks when rows are few
    static const int grid_cap = getenv("UVG_REPLAY_GRID") ? atoi(getenv("UVG_REPLAY_GRID")) : REPLAY_GRID_CAP;
    static const int wide = getenv("UVG_REPLAY_WIDE") ? atoi(getenv("UVG_REPLAY_WIDE")) : 1;
    const int grid = grid_cap > 0 ? std::min(p.rows_cap, grid_cap) : p.rows_cap;
    int in_smem;
    if (wide && p.rows_cap <= 320) {
        const size_t bytes = replay_smem_bytes(p.K, 256, &in_smem);
        k_replay<256><<<grid, 256, bytes, st>>>(p, in_smem);
    } else {
        const size_t bytes = replay_smem_bytes(p.K, 128, &in_smem);
        k_replay<128><<<grid, 128, bytes, st>>>(p, in_smem);
    }
    return cudaGetLastError();
}

}  // namespace uvg
