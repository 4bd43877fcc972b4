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
// Lane = reference (32 references of a group interleaved per column word -> one coalesced 512 B request per warp and
// word), warp = 8 queries whose planes are broadcast from shared memory, block = 64 queries x 32 references.
// Query chunks are staged by TMA bulk copies (cp.async.bulk + mbarrier) through a 3-deep ring.
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

// strictly better than the threshold item in the order of compare_q_item_score (src/min_heap.c:41-47)
__device__ __forceinline__ bool lex_greater(const int32_t (&s)[6], const Threshold& t)
{
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        if (s[i] != t.s[i]) return s[i] > t.s[i];
    }
    return false;
}

constexpr size_t PAIR_SMEM = STAGES * sizeof(QChunk) + 64 /*barriers*/ + QTILE * sizeof(Threshold);

template <int MODE>
__global__ void __launch_bounds__(PAIR_THREADS, 2) k_paircount(const PairParams p)
{
    extern __shared__ __align__(128) uint8_t smem[];
    QChunk* stage = reinterpret_cast<QChunk*>(smem);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * sizeof(QChunk));
    Threshold* s_thr = reinterpret_cast<Threshold*>(smem + STAGES * sizeof(QChunk) + 64);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int qt = blockIdx.x % p.n_qtiles;
    const int64_t group = p.local_begin / GROUP + blockIdx.x / p.n_qtiles;
    const int64_t local = group * GROUP + lane;
    const bool in_range = local >= p.local_begin && local < p.local_end;
    const QChunk* qsrc = p.qdata + (size_t)qt * p.n_chunks;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(full + s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (threadIdx.x < QTILE) s_thr[threadIdx.x] = p.thr[qt * QTILE + threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES && s < p.n_chunks; ++s) {
            mbar_expect_tx(full + s, (uint32_t)sizeof(QChunk));
            bulk_g2s(stage + s, qsrc + s, (uint32_t)sizeof(QChunk), full + s);
        }
    }

    int32_t accP[QPW], accV[QPW], accY[QPW], accE[QPW];
#pragma unroll
    for (int j = 0; j < QPW; ++j) { accP[j] = 0; accV[j] = 0; accY[j] = 0; accE[j] = 0; }

    const uint4* rp = p.db + (size_t)group * p.Wp * GROUP + lane;
    const uint32_t* ambp = p.refamb + (size_t)group * p.ambw;
    const int q0 = warp * QPW;
    bool warp_dead = false;
    int c = 0;
    for (; c < p.n_chunks; ++c) {
        const int s = c % STAGES;
        mbar_wait(full + s, (uint32_t)((c / STAGES) & 1));
        const QChunk& ch = stage[s];
        if (!warp_dead) {
            const uint32_t ambbits = (__ldg(ambp + ((c * WC) >> 5)) >> ((c * WC) & 31)) & 0xffffu;
            const uint4* rw = rp + (size_t)(c * WC) * GROUP;
#pragma unroll 4
            for (int wl = 0; wl < WC; ++wl) {
                const uint4 r = __ldg(rw + (size_t)wl * GROUP);
                const uint32_t rv = plane_any(r);
                const bool slow = (MODE == 0) && ((((ch.slow[wl] >> warp) | (ambbits >> wl)) & 1u) != 0u);
                if (!slow) {
#pragma unroll
                    for (int j = 0; j < QPW; ++j) {
                        const uint4 q = ch.acgt[wl][q0 + j];
                        const uint2 vo = ch.vo[wl][q0 + j];
                        const uint32_t P = plane_overlap(r, q);
                        const uint32_t V = rv & vo.x;
                        accP[j] += __popc(P);
                        accV[j] += __popc(V);
                        if (MODE == 1) accY[j] += __popc(V & ~P & vo.y);     // mismatches on the idx sites
                    }
                } else {
                    const uint32_t rone = plane_onehot(r);
#pragma unroll
                    for (int j = 0; j < QPW; ++j) {
                        const uint4 q = ch.acgt[wl][q0 + j];
                        const uint2 vo = ch.vo[wl][q0 + j];
                        const uint32_t P = plane_overlap(r, q);
                        const uint32_t V = rv & vo.x;
                        accP[j] += __popc(P);
                        accV[j] += __popc(V);
                        accY[j] += __popc(P & ~(rone & vo.y));
                        accE[j] += __popc(vo.x & ~vo.y & ~plane_differ(r, q));
                    }
                }
            }
            if (p.early_abort) {
                // exact: a pair whose ACGT matches cannot reach the query's k-th best first key is never a survivor
                bool dead = true;
                if (in_range) {
#pragma unroll
                    for (int j = 0; j < QPW; ++j) {
                        const Threshold& t = s_thr[q0 + j];
                        // mismatches so far (monotone along the genome) and the best ACGT count still reachable
                        const int32_t dsofar = (MODE == 0) ? (accV[j] - accP[j] + accY[j]) : (accV[j] - accP[j]);
                        if (!t.full || (dsofar < t.dmax && accP[j] + (int32_t)ch.rem[q0 + j] >= t.s[0])) dead = false;
                    }
                }
                warp_dead = __all_sync(0xffffffffu, dead);
            }
        }
        const int all_dead = __syncthreads_and(warp_dead ? 1 : 0);
        if (all_dead) break;
        if (threadIdx.x == 0 && c + STAGES < p.n_chunks) {
            mbar_expect_tx(full + s, (uint32_t)sizeof(QChunk));
            bulk_g2s(stage + s, qsrc + c + STAGES, (uint32_t)sizeof(QChunk), full + s);
        }
    }
    if (c < p.n_chunks) {
        // whole block aborted: drain the bulk copies still in flight before the shared memory is released
        for (int cc = c + 1; cc < p.n_chunks && cc < c + STAGES; ++cc) mbar_wait(full + cc % STAGES, (uint32_t)((cc / STAGES) & 1));
        return;
    }
    // ---- epilogue: 6-tuple, threshold test, warp-aggregated survivor append ----
    const int32_t non_n = in_range ? p.non_n[local] : 0;
    const int32_t c0 = in_range ? p.c0[local] : 0;
    const int32_t ci = in_range ? p.ci[local] : 0;
    const int64_t n_launch = p.local_end - p.local_begin;
#pragma unroll
    for (int j = 0; j < QPW; ++j) {
        const int q = qt * QTILE + q0 + j;
        int32_t sc[6];
        if (MODE == 0) {
            const int32_t A = accP[j] - accY[j];
            sc[0] = A; sc[1] = A + accE[j]; sc[2] = accP[j]; sc[3] = accV[j]; sc[4] = A - c0; sc[5] = non_n;
        } else {
            const int32_t M = accV[j] - accP[j];
            sc[0] = accP[j]; sc[1] = accV[j]; sc[2] = accP[j] - c0; sc[3] = non_n; sc[4] = M - accY[j]; sc[5] = accY[j];
        }
        bool pass = false;
        if (in_range && q < p.n_query && !warp_dead) {
            const Threshold& t = s_thr[q0 + j];
            const int32_t D = (MODE == 0) ? (sc[3] - sc[0]) : (sc[1] - sc[0]);
            pass = !t.full || (D < t.dmax && lex_greater(sc, t));
            if (p.dbg_counts) {
                int32_t* o = p.dbg_counts + ((size_t)q * n_launch + (local - p.local_begin)) * 4;
                if (MODE == 0) { o[0] = sc[0]; o[1] = sc[1]; o[2] = sc[2]; o[3] = sc[3]; }
                else { o[0] = accP[j]; o[1] = accV[j]; o[2] = accY[j]; o[3] = 0; }
            }
        }
        const uint32_t m = __ballot_sync(0xffffffffu, pass);
        if (m) {
            unsigned long long base = 0;
            if (lane == 0) {
                base = atomicAdd(p.surv_count, (unsigned long long)__popc(m));
                atomicAdd(p.qcount + q, __popc(m));
            }
            base = __shfl_sync(0xffffffffu, base, 0);
            if (pass) {
                const unsigned long long slot = base + __popc(m & ((1u << lane) - 1u));
                if (slot < (unsigned long long)p.surv_cap) {
                    Survivor sv;
                    sv.q = q; sv.r = (int32_t)(local + p.global_minus_local);
#pragma unroll
                    for (int k = 0; k < 6; ++k) sv.s[k] = sc[k];
                    p.surv[slot] = sv;
                    p.surv_ci[slot] = ci;
                }
            }
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
    cudaError_t e = cudaFuncSetAttribute(k_paircount<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PAIR_SMEM);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(k_paircount<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PAIR_SMEM);
}

cudaError_t launch_paircount(const PairParams& p, int acgt_mode, cudaStream_t st)
{
    if (p.local_end <= p.local_begin) return cudaSuccess;
    const int64_t g0 = p.local_begin / GROUP, g1 = (p.local_end + GROUP - 1) / GROUP;
    const int64_t blocks = (g1 - g0) * p.n_qtiles;
    if (blocks > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
    if (acgt_mode) k_paircount<1><<<(unsigned)blocks, PAIR_THREADS, PAIR_SMEM, st>>>(p);
    else k_paircount<0><<<(unsigned)blocks, PAIR_THREADS, PAIR_SMEM, st>>>(p);
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
