// K2: pair-count kernel for sm_100a.  Replaces the reference's scoring calls
//   biomcmc_pairwise_score_matches_truncated_idx   biomcmc alignment.c:855-884   (default mode, x3 per pair)
//   quick_pairwise_score_acgt_and_valid            src/fastaseq.c:585-596        (--acgt mode)
//   queue_distance_to_consensus                    src/nearest.c:428-433         (the consensus is one more query slot)
//   the count part of queue_update_min_heaps       src/nearest.c:466-470, 499-501
// by closed-form bit-plane arithmetic.  Per 32-site column word and (query, reference) pair, with the
// three planes of common.cuh (base = plain base present, b0, b1 = which base):
//   V = r.base & q.base                         both are plain bases                       -> valid  (3 LOP3 in all)
//   M = V & ~(r.b0 ^ q.b0) & ~(r.b1 ^ q.b1)     ... and the same base                      -> acgt = text = partial
// popcount(M) is not taken per word: the M words of a pair go through a carry-save adder tree (Harley-Seal, two LOP3
// per 3:2 compressor: xor3 = 0x96, majority = 0xE8) that emits one weight-4 word per four column words, so the XU pipe
// (POPC, 16 lanes/clk/SM) sees 1.25 POPC per pair-word instead of 2 and the ALU pipe 4.5 LOP3 instead of 5; the
// popcounts are summed by IMAD on the FMA pipe into packed 16-bit counters.
// Sites where either sequence holds an ambiguity code are outside the planes; their exact contribution to valid /
// partial / text (never to acgt: alignment.c:871-874 needs a plain base) is added in the epilogue from the sparse
// (site, code) lists of the reference and of the query.
// Lane = 2 references (one of each of the block's two groups), warp = 8 query slots whose planes are broadcast from
// shared memory, block = 64 slots x 64 references.  Both operands are staged by TMA bulk copies (cp.async.bulk +
// mbarrier) through a 6-deep ring, so the inner loop issues only LDS, LOP3 (ALU), POPC (XU) and IMAD (FMA).
// The epilogue stores the four counts of every pair straight into the score tile of the rank that owns the query
// (peer-mapped memory over NVLink when the database is sharded): the all-to-all of the sharded search is fused here.
#include "common.cuh"
#include "kernels.h"
#include "pair_epilogue.cuh"
#include "sm100_prims.cuh"

namespace uvg {

// One pipeline stage: a query chunk (64 slots x WC words) and the same WC words of the block's 64 references.
struct __align__(16) Stage {
    QChunk q;
    uint4  ref[RPL][W4][NPLANE][GROUP];
};
constexpr uint32_t STAGE_REF_BYTES = W4 * NPLANE * GROUP * sizeof(uint4);
constexpr size_t PAIR_SMEM = STAGES * sizeof(Stage) + 64 /*barriers*/ + QTILE * sizeof(Threshold);
constexpr int FLUSH_CHUNKS = 65535 / (WC * 32);        // packed 16-bit counters are flushed before they can overflow
constexpr int ABORT_EVERY = 4;                         // chunks between two early-abort checks

// MODE 0: default (IUPAC) scoring; MODE 1: --acgt.  LONG = alignments of 65,536 columns or more: the packed counters are
// flushed into 32-bit totals every FLUSH_CHUNKS chunks.
template <int MODE, bool LONG>
__global__ void __launch_bounds__(PAIR_THREADS, 2) k_paircount(const PairParams p, const uint32_t k4, const uint32_t k64k)
{
    extern __shared__ __align__(128) uint8_t smem[];
    Stage* stage = reinterpret_cast<Stage*>(smem);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * sizeof(Stage));
    Threshold* s_thr = reinterpret_cast<Threshold*>(smem + STAGES * sizeof(Stage) + 64);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int qt = blockIdx.x % p.n_qtiles;
    const int64_t group0 = (p.local_begin / (GROUP * RPL) + blockIdx.x / p.n_qtiles) * RPL;
    const QChunk* qsrc = p.qdata + (size_t)qt * p.n_chunks;
    const size_t group_stride = (size_t)(p.Wp / 4) * NPLANE * GROUP;         // uint4 elements between consecutive groups
    const uint4* rsrc = p.db + (size_t)group0 * group_stride;
    const int owner = qt % p.world, row0 = (qt / p.world) * QTILE;           // where this query tile's rows live
    bool in_range[RPL];
    int64_t local[RPL];
#pragma unroll
    for (int g = 0; g < RPL; ++g) {
        local[g] = (group0 + g) * GROUP + lane;
        in_range[g] = local[g] >= p.local_begin && local[g] < p.local_end;
    }

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(full + s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < QTILE) s_thr[threadIdx.x] = p.thr[(size_t)owner * p.rows_cap + row0 + threadIdx.x];
    __syncthreads();
    auto issue = [&](int c, int s) {
        mbar_expect_tx(full + s, (uint32_t)(sizeof(QChunk) + RPL * STAGE_REF_BYTES));
        bulk_g2s(&stage[s].q, qsrc + c, (uint32_t)sizeof(QChunk), full + s);
#pragma unroll
        for (int g = 0; g < RPL; ++g)
            bulk_g2s(&stage[s].ref[g][0][0][0], rsrc + g * group_stride + (size_t)c * (W4 * NPLANE * GROUP), STAGE_REF_BYTES, full + s);
    };
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES && s < p.n_chunks; ++s) issue(s, s);
    }

    // carry-save accumulators of the match words (weights 1 and 2) and packed counters: acc = M4 | V << 16 where M4 counts
    // the weight-4 carries (already multiplied by 4); accI (--acgt): mismatches on segregating sites
    uint32_t ones[RPL][QPW], twos[RPL][QPW], acc[RPL][QPW], accI[MODE == 1 ? RPL : 1][MODE == 1 ? QPW : 1] = {};
    int32_t tot[LONG ? 3 : 1][LONG ? RPL : 1][LONG ? QPW : 1] = {};
#pragma unroll
    for (int g = 0; g < RPL; ++g)
#pragma unroll
        for (int j = 0; j < QPW; ++j) {
            ones[g][j] = 0; twos[g][j] = 0; acc[g][j] = 0;
            if constexpr (MODE == 1) accI[g][j] = 0;
            if constexpr (LONG) { tot[0][g][j] = 0; tot[1][g][j] = 0; tot[2][g][j] = 0; }
        }
    auto cntM4 = [&](int g, int j) { return (int32_t)(acc[g][j] & 0xffffu) + (LONG ? tot[0][LONG ? g : 0][LONG ? j : 0] : 0); };
    auto cntV = [&](int g, int j) { return (int32_t)(acc[g][j] >> 16) + (LONG ? tot[LONG ? 1 : 0][LONG ? g : 0][LONG ? j : 0] : 0); };
    auto cntI = [&](int g, int j) { return (int32_t)(accI[MODE == 1 ? g : 0][MODE == 1 ? j : 0]) + (LONG ? tot[LONG ? 2 : 0][LONG ? g : 0][LONG ? j : 0] : 0); };

    const int q0 = warp * QPW;
    bool warp_dead = false;
    uint32_t executed = 0;
    int c = 0;
    for (; c < p.n_chunks; ++c) {
        const int s = c % STAGES;
        mbar_wait(full + s, (uint32_t)((c / STAGES) & 1));
        const Stage& st = stage[s];
        if (!warp_dead) {
            ++executed;
#pragma unroll
            for (int u = 0; u < W4; ++u) {
                uint4 rb[RPL], r0[RPL], r1[RPL];
#pragma unroll
                for (int g = 0; g < RPL; ++g) { rb[g] = st.ref[g][u][0][lane]; r0[g] = st.ref[g][u][1][lane]; r1[g] = st.ref[g][u][2][lane]; }
#pragma unroll
                for (int j = 0; j < QPW; ++j) {
                    const uint4 qb = st.q.pl[u][q0 + j][0], qx = st.q.pl[u][q0 + j][1], qy = st.q.pl[u][q0 + j][2];
                    uint4 qi = make_uint4(0u, 0u, 0u, 0u);
                    if constexpr (MODE == 1) qi = st.q.pl[u][q0 + j][3];
#pragma unroll
                    for (int g = 0; g < RPL; ++g) {
                        const uint32_t V0 = rb[g].x & qb.x, V1 = rb[g].y & qb.y, V2 = rb[g].z & qb.z, V3 = rb[g].w & qb.w;
                        const uint32_t M0 = keep_eq(keep_eq(V0, r0[g].x, qx.x), r1[g].x, qy.x);
                        const uint32_t M1 = keep_eq(keep_eq(V1, r0[g].y, qx.y), r1[g].y, qy.y);
                        const uint32_t M2 = keep_eq(keep_eq(V2, r0[g].z, qx.z), r1[g].z, qy.z);
                        const uint32_t M3 = keep_eq(keep_eq(V3, r0[g].w, qx.w), r1[g].w, qy.w);
                        uint32_t o = ones[g][j], ta, tb, f, t;
                        csa(o, ta, o, M0, M1);
                        csa(o, tb, o, M2, M3);
                        csa(t, f, twos[g][j], ta, tb);
                        ones[g][j] = o; twos[g][j] = t;
                        uint32_t a = acc[g][j];
                        a = imad(__popc(f), k4, a);
                        a = imad(__popc(V0), k64k, a);
                        a = imad(__popc(V1), k64k, a);
                        a = imad(__popc(V2), k64k, a);
                        a = imad(__popc(V3), k64k, a);
                        acc[g][j] = a;
                        if constexpr (MODE == 1) {
                            // mismatches on the segregating (idx) sites: valid, not equal, inside the mask
                            const uint32_t I0 = lop3<0x20>(V0, M0, qi.x), I1 = lop3<0x20>(V1, M1, qi.y);
                            const uint32_t I2 = lop3<0x20>(V2, M2, qi.z), I3 = lop3<0x20>(V3, M3, qi.w);
                            accI[MODE == 1 ? g : 0][MODE == 1 ? j : 0] += __popc(I0) + __popc(I1) + __popc(I2) + __popc(I3);
                        }
                    }
                }
            }
            if constexpr (LONG) if ((c + 1) % FLUSH_CHUNKS == 0) {
#pragma unroll
                for (int g = 0; g < RPL; ++g)
#pragma unroll
                    for (int j = 0; j < QPW; ++j) {
                        tot[0][LONG ? g : 0][LONG ? j : 0] += acc[g][j] & 0xffffu; tot[LONG ? 1 : 0][LONG ? g : 0][LONG ? j : 0] += acc[g][j] >> 16;
                        acc[g][j] = 0;
                        if constexpr (MODE == 1) { tot[LONG ? 2 : 0][LONG ? g : 0][LONG ? j : 0] += accI[MODE == 1 ? g : 0][MODE == 1 ? j : 0]; accI[MODE == 1 ? g : 0][MODE == 1 ? j : 0] = 0; }
                    }
            }
            if (p.early_abort && (c % ABORT_EVERY) == ABORT_EVERY - 1) {
                // exact: mismatches only grow along the genome (ambiguity sites add to valid, never to acgt), and the ACGT
                // count can grow by at most what is left of the query.  96 = the most the carry-save registers can hold.
                bool dead = true;
#pragma unroll
                for (int g = 0; g < RPL; ++g) {
                    if (!in_range[g]) continue;
#pragma unroll
                    for (int j = 0; j < QPW; ++j) {
                        if (qt * QTILE + q0 + j >= p.n_slots) continue;          // padding slot
                        const Threshold t = s_thr[q0 + j];
                        const int32_t m_ub = cntM4(g, j) + 96;
                        if (!t.full || (cntV(g, j) - m_ub < t.dmax && m_ub + (int32_t)st.q.rem[q0 + j] >= t.root_s0)) dead = false;
                    }
                }
                warp_dead = __all_sync(0xffffffffu, dead);
            }
        }
        const int all_dead = p.early_abort ? __syncthreads_and(warp_dead ? 1 : 0) : (__syncthreads(), 0);
        if (all_dead) break;
        if (threadIdx.x == 0 && c + STAGES < p.n_chunks) issue(c + STAGES, s);
    }
    if (lane == 0 && p.work) atomicAdd(p.work, (unsigned long long)executed * (QPW * RPL * W4));
    if (c < p.n_chunks) {
        // whole block aborted: drain the bulk copies still in flight before the shared memory is released
        for (int cc = c + 1; cc < p.n_chunks && cc < c + STAGES; ++cc) mbar_wait(full + cc % STAGES, (uint32_t)((cc / STAGES) & 1));
    }

    // ---- epilogue: exact contribution of the ambiguity codes, then the four counts of every pair go to the dense score
    // tile of the round on the rank that owns the query (coalesced 512 B per warp and query).  The replay kernel
    // assembles the 6-tuple (src/nearest.c:466-470, 499-501) for the pairs it admits.  Aborted pairs get a sentinel whose
    // mismatch count can never pass the admission test.
#pragma unroll
    for (int j = 0; j < QPW; ++j) {
        const int slot = qt * QTILE + q0 + j;
        if (slot >= p.n_slots) continue;
        const bool is_cons = slot == p.n_slots - 1;
#pragma unroll
        for (int g = 0; g < RPL; ++g) {
            if (!in_range[g]) continue;
            uint4 v;
            if (warp_dead) {
                v = (MODE == 0) ? make_uint4(0u, 0u, 0u, 0x7fffffffu) : make_uint4(0u, 0x7fffffffu, 0u, 0u);
            } else {
                const int32_t M = cntM4(g, j) + __popc(ones[g][j]) + 2 * __popc(twos[g][j]);
                const int32_t V = cntV(g, j);
                if (MODE == 0) v = pair_epilogue_default(p, p.qamb_off, p.qamb_ent, slot, group0 + g, lane, local[g], M, V);
                else v = make_uint4((uint32_t)M, (uint32_t)V, (uint32_t)cntI(g, j), 0u);
            }
            pair_store(p, owner, row0 + q0 + j, is_cons, local[g], v);
        }
    }
}

cudaError_t paircount_setup()
{
    cudaError_t e;
#define SETUP_(K) e = cudaFuncSetAttribute(K, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PAIR_SMEM); if (e != cudaSuccess) return e;
    SETUP_((k_paircount<0, false>)) SETUP_((k_paircount<1, false>)) SETUP_((k_paircount<0, true>)) SETUP_((k_paircount<1, true>))
#undef SETUP_
    return cudaSuccess;
}

cudaError_t launch_paircount(const PairParams& p, int acgt_mode, cudaStream_t st)
{
    if (p.local_end <= p.local_begin) return cudaSuccess;
    const int64_t per = GROUP * RPL;
    const int64_t g0 = p.local_begin / per, g1 = (p.local_end + per - 1) / per;
    const int64_t blocks = (g1 - g0) * p.n_qtiles;
    if (blocks > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
    const bool lng = (int64_t)p.Wp * 32 > 65535;
    const uint32_t k4 = 4u, k64k = 65536u;       // passed at run time so that the compiler keeps IMAD (no shift/add on the ALU pipe)
    if (acgt_mode) {
        if (lng) k_paircount<1, true><<<(unsigned)blocks, PAIR_THREADS, PAIR_SMEM, st>>>(p, k4, k64k);
        else k_paircount<1, false><<<(unsigned)blocks, PAIR_THREADS, PAIR_SMEM, st>>>(p, k4, k64k);
    } else {
        if (lng) k_paircount<0, true><<<(unsigned)blocks, PAIR_THREADS, PAIR_SMEM, st>>>(p, k4, k64k);
        else k_paircount<0, false><<<(unsigned)blocks, PAIR_THREADS, PAIR_SMEM, st>>>(p, k4, k64k);
    }
    return cudaGetLastError();
}

}  // namespace uvg
