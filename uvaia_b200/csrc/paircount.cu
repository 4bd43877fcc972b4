// K2: pair-count kernels for sm_100a.  Replace the reference's scoring calls
//   biomcmc_pairwise_score_matches_truncated_idx   biomcmc alignment.c:855-884   (default mode, x3 per pair)
//   quick_pairwise_score_acgt_and_valid            src/fastaseq.c:585-596        (--acgt mode)
//   queue_distance_to_consensus                    src/nearest.c:428-433         (k_consensus)
//   the 6-tuple assembly of queue_update_min_heaps src/nearest.c:466-470, 499-501
// by closed-form bit-plane arithmetic: per 32-site column word and (query, reference) pair
//   P = (ra&qa)|(rc&qc)|(rg&qg)|(rt&qt)      sites whose IUPAC sets intersect     -> partial
//   V = rv & qv                               both valid                           -> valid
//   Y = P & ~(r_onehot & q_onehot)            intersecting but not a plain base    -> acgt = P - Y   (rare words only)
//   E = q_ambiguous & (r == q)                identical ambiguity codes            -> text = acgt + E (rare words only)
// Lane = 2 references (one of each of the block's two groups; 32 references interleaved per column word -> a group's
// chunk is one contiguous 4 KB run), warp = 8 queries whose planes are broadcast from shared memory, block = 64 queries
// x 64 references.  Both operands are staged by TMA bulk copies (cp.async.bulk + mbarrier) through a 4-deep ring, so
// the inner loop issues only LDS, LOP3 (ALU pipe), POPC (XU pipe) and IMAD (FMA pipe).
#include "common.cuh"
#include "kernels.h"

namespace uvg {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ uint32_t plane_any(const uint4& r) { return r.x | r.y | r.z | r.w; }
__device__ __forceinline__ uint32_t plane_onehot(const uint4& r)
{
    uint32_t multi = (r.x & r.y) | (r.z & r.w) | ((r.x | r.y) & (r.z | r.w));
    return plane_any(r) & ~multi;
}
__device__ __forceinline__ uint32_t plane_overlap(const uint4& r, const uint4& q)
{
    return (r.x & q.x) | (r.y & q.y) | (r.z & q.z) | (r.w & q.w);
}
__device__ __forceinline__ uint32_t plane_differ(const uint4& r, const uint4& q)
{
    return (r.x ^ q.x) | (r.y ^ q.y) | (r.z ^ q.z) | (r.w ^ q.w);
}

// One pipeline stage: a query chunk (64 queries x WC words) and the same WC words of the block's 64 references.
struct __align__(16) Stage {
    QChunk q;
    uint4  ref[RPL][WC][GROUP];
};
constexpr uint32_t STAGE_REF_BYTES = WC * GROUP * sizeof(uint4);
constexpr size_t PAIR_SMEM = STAGES * sizeof(Stage) + 64 /*barriers*/ + QTILE * sizeof(Threshold);
constexpr int FLUSH_CHUNKS = 2040 / WC;        // packed 16-bit counters are flushed before they can overflow

// a*b+c on the FMA pipe (IMAD): keeps the accumulations off the ALU pipe that issues the LOP3s
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// LONG = alignments of 65,536 columns or more (counters may exceed 16 bits): packed counters are flushed into
// 32-bit totals every FLUSH_CHUNKS chunks.
template <int MODE, bool LONG>
__global__ void __launch_bounds__(PAIR_THREADS, 2) k_paircount(const PairParams p, const uint32_t k1, const uint32_t k64k)
{
    extern __shared__ __align__(128) uint8_t smem[];
    Stage* stage = reinterpret_cast<Stage*>(smem);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * sizeof(Stage));
    Threshold* s_thr = reinterpret_cast<Threshold*>(smem + STAGES * sizeof(Stage) + 64);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int qt = blockIdx.x % p.n_qtiles;
    const int64_t group0 = (p.local_begin / (GROUP * RPL) + blockIdx.x / p.n_qtiles) * RPL;
    const QChunk* qsrc = p.qdata + (size_t)qt * p.n_chunks;
    const uint4* rsrc = p.db + (size_t)group0 * p.Wp * GROUP;
    const size_t group_stride = (size_t)p.Wp * GROUP;                // uint4 elements between consecutive groups
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
    {
        const uint4* tsrc = reinterpret_cast<const uint4*>(p.thr + (size_t)qt * QTILE);
        uint4* tdst = reinterpret_cast<uint4*>(s_thr);
        for (int i = threadIdx.x; i < (int)(QTILE * sizeof(Threshold) / 16); i += blockDim.x) tdst[i] = tsrc[i];
    }
    __syncthreads();
    auto issue = [&](int c, int s) {
        mbar_expect_tx(full + s, (uint32_t)(sizeof(QChunk) + RPL * STAGE_REF_BYTES));
        bulk_g2s(&stage[s].q, qsrc + c, (uint32_t)sizeof(QChunk), full + s);
#pragma unroll
        for (int g = 0; g < RPL; ++g)
            bulk_g2s(&stage[s].ref[g][0][0], rsrc + g * group_stride + (size_t)c * WC * GROUP, STAGE_REF_BYTES, full + s);
    };
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES && s < p.n_chunks; ++s) issue(s, s);
    }

    uint32_t accPV[RPL][QPW], accYE[RPL][QPW];                 // packed: P | V << 16  and  Y | E << 16
    int32_t tot[LONG ? 4 : 1][RPL][QPW];
#pragma unroll
    for (int g = 0; g < RPL; ++g)
#pragma unroll
        for (int j = 0; j < QPW; ++j) {
            accPV[g][j] = 0; accYE[g][j] = 0;
#pragma unroll
            for (int k = 0; k < (LONG ? 4 : 1); ++k) tot[k][g][j] = 0;
        }
    auto flush = [&]() {
        if (LONG) {
#pragma unroll
            for (int g = 0; g < RPL; ++g)
#pragma unroll
                for (int j = 0; j < QPW; ++j) {
                    tot[0][g][j] += accPV[g][j] & 0xffffu; tot[LONG ? 1 : 0][g][j] += accPV[g][j] >> 16;
                    tot[LONG ? 2 : 0][g][j] += accYE[g][j] & 0xffffu; tot[LONG ? 3 : 0][g][j] += accYE[g][j] >> 16;
                    accPV[g][j] = 0; accYE[g][j] = 0;
                }
        }
    };
    auto cntP = [&](int g, int j) { return (int32_t)(accPV[g][j] & 0xffffu) + (LONG ? tot[0][g][j] : 0); };
    auto cntV = [&](int g, int j) { return (int32_t)(accPV[g][j] >> 16) + (LONG ? tot[LONG ? 1 : 0][g][j] : 0); };
    auto cntY = [&](int g, int j) { return (int32_t)(accYE[g][j] & 0xffffu) + (LONG ? tot[LONG ? 2 : 0][g][j] : 0); };
    auto cntE = [&](int g, int j) { return (int32_t)(accYE[g][j] >> 16) + (LONG ? tot[LONG ? 3 : 0][g][j] : 0); };

    const uint32_t* ambp = p.refamb + (size_t)group0 * p.ambw;
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
            uint32_t ambbits = 0;
            if (MODE == 0) {
#pragma unroll
                for (int g = 0; g < RPL; ++g) ambbits |= (__ldg(ambp + g * p.ambw + ((c * WC) >> 5)) >> ((c * WC) & 31)) & ((1u << WC) - 1u);
            }
#pragma unroll
            for (int wl = 0; wl < WC; ++wl) {
                uint4 r[RPL];
                uint32_t rv[RPL];
#pragma unroll
                for (int g = 0; g < RPL; ++g) { r[g] = st.ref[g][wl][lane]; rv[g] = plane_any(r[g]); }
                const bool slow = (MODE == 0) && ((((st.q.slow[wl] >> warp) | (ambbits >> wl)) & 1u) != 0u);
                if (!slow) {
#pragma unroll
                    for (int j = 0; j < QPW; ++j) {
                        const uint4 q = st.q.acgt[wl][q0 + j];
                        const uint2 vo = st.q.vo[wl][q0 + j];
#pragma unroll
                        for (int g = 0; g < RPL; ++g) {
                            const uint32_t P = plane_overlap(r[g], q);
                            const uint32_t V = rv[g] & vo.x;
                            accPV[g][j] = imad(__popc(V), k64k, imad(__popc(P), k1, accPV[g][j]));
                            if (MODE == 1) accYE[g][j] = imad(__popc(V & ~P & vo.y), k1, accYE[g][j]);   // mismatches on idx sites
                        }
                    }
                } else {
#pragma unroll
                    for (int g = 0; g < RPL; ++g) {
                        const uint32_t rone = plane_onehot(r[g]);
#pragma unroll
                        for (int j = 0; j < QPW; ++j) {
                            const uint4 q = st.q.acgt[wl][q0 + j];
                            const uint2 vo = st.q.vo[wl][q0 + j];
                            const uint32_t P = plane_overlap(r[g], q);
                            const uint32_t V = rv[g] & vo.x;
                            accPV[g][j] += __popc(P) + (__popc(V) << 16);
                            accYE[g][j] += __popc(P & ~(rone & vo.y)) + (__popc(vo.x & ~vo.y & ~plane_differ(r[g], q)) << 16);
                        }
                    }
                }
            }
            if (LONG && (c + 1) % FLUSH_CHUNKS == 0) flush();
            if (p.early_abort) {
                // exact: a pair that can no longer satisfy the admission staircase is never a survivor.  Mismatches so
                // far only grow along the genome and the best reachable ACGT count is what is left of the query.
                bool dead = true;
#pragma unroll
                for (int g = 0; g < RPL; ++g) {
                    if (!in_range[g]) continue;
#pragma unroll
                    for (int j = 0; j < QPW; ++j) {
                        const Threshold& t = s_thr[q0 + j];
                        if (!t.full) { dead = false; continue; }
                        const int32_t dsofar = (MODE == 0) ? (cntV(g, j) - cntP(g, j) + cntY(g, j)) : (cntV(g, j) - cntP(g, j));
                        if (dsofar < t.dmax && cntP(g, j) + (int32_t)st.q.rem[q0 + j] >= t.bar[dsofar >> (t.full - 1)]) dead = false;
                    }
                }
                warp_dead = __all_sync(0xffffffffu, dead);
            }
        }
        const int all_dead = __syncthreads_and(warp_dead ? 1 : 0);
        if (all_dead) break;
        if (threadIdx.x == 0 && c + STAGES < p.n_chunks) issue(c + STAGES, s);
    }
    if (lane == 0 && p.work) atomicAdd(p.work, (unsigned long long)executed);
    if (c < p.n_chunks) {
        // whole block aborted: drain the bulk copies still in flight before the shared memory is released
        for (int cc = c + 1; cc < p.n_chunks && cc < c + STAGES; ++cc) mbar_wait(full + cc % STAGES, (uint32_t)((cc / STAGES) & 1));
    }

    // ---- epilogue: the four counts of every pair go to the dense score tile of the round (coalesced 512 B per warp
    // and query); the replay kernel assembles the 6-tuple (src/nearest.c:466-470, 499-501) for the pairs it admits.
    // Aborted pairs get a sentinel whose mismatch count can never pass the admission test.
#pragma unroll
    for (int g = 0; g < RPL; ++g) {
        if (!in_range[g]) continue;
        uint4* out = p.scores + p.score_col0 + (local[g] - p.local_begin);
#pragma unroll
        for (int j = 0; j < QPW; ++j) {
            const int q = qt * QTILE + q0 + j;
            if (q >= p.n_query) continue;
            uint4 v;
            if (warp_dead) {
                v = (MODE == 0) ? make_uint4(0u, 0u, 0u, 0x7fffffffu) : make_uint4(0u, 0x7fffffffu, 0u, 0u);
            } else if (MODE == 0) {
                const int32_t A = cntP(g, j) - cntY(g, j);
                v = make_uint4((uint32_t)A, (uint32_t)(A + cntE(g, j)), (uint32_t)cntP(g, j), (uint32_t)cntV(g, j));
            } else {
                v = make_uint4((uint32_t)cntP(g, j), (uint32_t)cntV(g, j), (uint32_t)cntY(g, j), 0u);
            }
            out[(size_t)q * p.score_stride] = v;
        }
    }
}

// ---- consensus stage: one warp per group -------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) k_consensus(const uint4* __restrict__ db, const uint4* __restrict__ cons_acgt,
                                                   const uint2* __restrict__ cons_vo, int Wp, int64_t local_begin, int64_t local_end,
                                                   int32_t* __restrict__ c0, int32_t* __restrict__ ci)
{
    const int lane = threadIdx.x & 31;
    const int64_t wid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t group = local_begin / GROUP + wid;
    const int64_t local = group * GROUP + lane;
    if (group * GROUP >= local_end) return;
    const uint4* rp = db + (size_t)group * Wp * GROUP + lane;
    int32_t m = 0, v = 0;
#pragma unroll 4
    for (int w = 0; w < Wp; ++w) {
        const uint4 r = __ldg(rp + (size_t)w * GROUP);
        const uint4 q = __ldg(cons_acgt + w);
        const uint2 vo = __ldg(cons_vo + w);
        const uint32_t P = plane_overlap(r, q);
        v += __popc(plane_any(r) & vo.x);
        if (MODE == 0) m += __popc(P & plane_onehot(r) & vo.y);       // cons == r and r is a plain base
        else m += __popc(P);
    }
    if (local >= local_begin && local < local_end) { c0[local] = m; ci[local] = v - m; }
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
    const uint32_t k1 = 1u, k64k = 65536u;
    if (acgt_mode) {
        if (lng) k_paircount<1, true><<<(unsigned)blocks, PAIR_THREADS, PAIR_SMEM, st>>>(p, k1, k64k);
        else k_paircount<1, false><<<(unsigned)blocks, PAIR_THREADS, PAIR_SMEM, st>>>(p, k1, k64k);
    } else {
        if (lng) k_paircount<0, true><<<(unsigned)blocks, PAIR_THREADS, PAIR_SMEM, st>>>(p, k1, k64k);
        else k_paircount<0, false><<<(unsigned)blocks, PAIR_THREADS, PAIR_SMEM, st>>>(p, k1, k64k);
    }
    return cudaGetLastError();
}

cudaError_t launch_consensus(const uint4* db, const uint4* cons_acgt, const uint2* cons_vo, int Wp, int acgt_mode,
                             int64_t local_begin, int64_t local_end, int32_t* c0, int32_t* ci, cudaStream_t st)
{
    if (local_end <= local_begin) return cudaSuccess;
    const int64_t g0 = local_begin / GROUP, g1 = (local_end + GROUP - 1) / GROUP;
    const int64_t blocks = (g1 - g0 + 7) / 8;
    if (acgt_mode) k_consensus<1><<<(unsigned)blocks, 256, 0, st>>>(db, cons_acgt, cons_vo, Wp, local_begin, local_end, c0, ci);
    else k_consensus<0><<<(unsigned)blocks, 256, 0, st>>>(db, cons_acgt, cons_vo, Wp, local_begin, local_end, c0, ci);
    return cudaGetLastError();
}

}  // namespace uvg
