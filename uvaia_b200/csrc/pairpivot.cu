// K2p: pair-count kernel in PIVOT form for sm_100a (default scoring mode; the general kernel is paircount.cu).
// Same contract as k_paircount<0,false> -- the four counts of biomcmc_pairwise_score_matches_truncated_idx
// (biomcmc alignment.c:855-884) of every (query slot, reference) pair of a round, untruncated, stored into the dense
// score tile of the query's owner -- with a cheaper inner loop.
//
// All queries of one search are close to a pivot sequence P (per site the plain base most queries hold; uvg_set_queries).
// For a site where the query holds P's base, "query == reference" is "reference == P", which does not depend on the query:
//     mc  (reference plane, derived once per query set by k_derive_pivot)  = r.base & (r == P)
//     qeq (query plane)                                                    = q.base & (q == P)
//     M' = mc & qeq        V = r.base & q.base                             (2 LOP3 per pair-word instead of 3)
// Query sites with a plain base != P (the query's own SNPs, a few dozen per genome) are outside M'; they are "fix"
// entries (slot, word, base, bit mask) that travel with the query chunk: the loop ANDs the mask with the one-hot plane
// word of that base of each staged reference (one conflict-free LDS.32) and, on the rare lanes that hit, adds the
// popcount to the match counter.  Entries that overflow a chunk's fix block and all ambiguity codes go to the epilogue
// (pair_epilogue.cuh).
// Both popcount streams go through one 3:2 compressor per two words (ones register + one carry word of weight 2):
//     4 LOP3 (ALU) + 1 POPC (XU) + 1 IMAD (FMA) per pair and 32-site word
// which balances the ALU and XU pipes at the measured 4:1 LOP3:POPC rate of sm_100 (profiles/int_pipes_b200_r01.json);
// the general kernel needs 4.5 LOP3 + 1.25 POPC.
// Shape: block = 64 query slots x 64 references, 16 warps x 4 slots, lane = 2 references, 64 registers, 2 blocks per
// SM (32 warps per SM).  Staging: one TMA bulk copy per operand into a 6-deep ring with full/empty mbarriers (no
// __syncthreads in the loop: warps drift up to the ring depth apart); one elected thread refills a stage two chunks
// after it was consumed.
#include "common.cuh"
#include "kernels.h"
#include "pair_epilogue.cuh"
#include "sm100_prims.cuh"

namespace uvg {

struct __align__(16) RefUnit2 {
    uint4    base[GROUP];
    uint4    mc[GROUP];
    uint32_t onehot[4][4][GROUP];                       // [word][base A,C,G,T][lane]
};
static_assert(sizeof(RefUnit2) == UNIT2_BYTES, "derived database record");
struct __align__(16) Stage2 {
    QChunk2  q;
    RefUnit2 ref[RPL][W4];
};
constexpr uint32_t STAGE2_REF_BYTES = W4 * UNIT2_BYTES;
constexpr size_t PIVOT_SMEM = STAGES2 * sizeof(Stage2) + 2 * STAGES2 * sizeof(uint64_t);
constexpr int REFILL_LAG = 2;
constexpr int QW2 = PIVOT_QW;                           // query slots per warp (register tile QW2 x RPL pairs per lane)
constexpr int PIVOT_THREADS = PIVOT_WARPS * 32;

__global__ void __launch_bounds__(PIVOT_THREADS, 1024 / PIVOT_THREADS) k_pairpivot(const PairParams p, const uint32_t k2, const uint32_t k128k)
{
    extern __shared__ __align__(128) uint8_t smem[];
    Stage2* stage = reinterpret_cast<Stage2*>(smem);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES2 * sizeof(Stage2));
    uint64_t* empty = full + STAGES2;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int qt = blockIdx.x % p.n_qtiles;
    const int64_t group0 = (p.local_begin / (GROUP * RPL) + blockIdx.x / p.n_qtiles) * RPL;
    const QChunk2* qsrc = p.qdata2 + (size_t)qt * p.n_chunks;
    const size_t group_stride = (size_t)(p.Wp / 4) * UNIT2_BYTES;            // bytes between consecutive groups
    const uint8_t* rsrc = p.db2 + (size_t)group0 * group_stride;
    const int owner = qt % p.world, row0 = (qt / p.world) * QTILE;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES2; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, PIVOT_THREADS / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int c, int s) {
        mbar_expect_tx(full + s, (uint32_t)(sizeof(QChunk2) + RPL * STAGE2_REF_BYTES));
        bulk_g2s(&stage[s].q, qsrc + c, (uint32_t)sizeof(QChunk2), full + s);
#pragma unroll
        for (int g = 0; g < RPL; ++g)
            bulk_g2s(&stage[s].ref[g][0], rsrc + g * group_stride + (size_t)c * STAGE2_REF_BYTES, STAGE2_REF_BYTES, full + s);
    };
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES2 && s < p.n_chunks; ++s) issue(s, s);
    }

    // per pair: carry-save "ones" registers of the valid and the match stream and one packed counter
    // acc = (#weight-2 carries of M) * 2 | (#weight-2 carries of V) * 2 << 16
    uint32_t oV[RPL][QW2], oM[RPL][QW2], acc[RPL][QW2];
#pragma unroll
    for (int g = 0; g < RPL; ++g)
#pragma unroll
        for (int j = 0; j < QW2; ++j) { oV[g][j] = 0; oM[g][j] = 0; acc[g][j] = 0; }

    const int q0 = warp * QW2;
    for (int c = 0; c < p.n_chunks; ++c) {
        const int s = c % STAGES2;
        mbar_wait(full + s, (uint32_t)((c / STAGES2) & 1));
        const Stage2& st = stage[s];
#pragma unroll
        for (int u = 0; u < W4; ++u) {
            uint4 rb[RPL], rm[RPL];
#pragma unroll
            for (int g = 0; g < RPL; ++g) { rb[g] = st.ref[g][u].base[lane]; rm[g] = st.ref[g][u].mc[lane]; }
#pragma unroll
            for (int j = 0; j < QW2; ++j) {
                const uint4 qb = st.q.pl[u][q0 + j][0], qe = st.q.pl[u][q0 + j][1];
#pragma unroll
                for (int g = 0; g < RPL; ++g) {
                    const uint32_t V0 = rb[g].x & qb.x, V1 = rb[g].y & qb.y, V2 = rb[g].z & qb.z, V3 = rb[g].w & qb.w;
                    const uint32_t M0 = rm[g].x & qe.x, M1 = rm[g].y & qe.y, M2 = rm[g].z & qe.z, M3 = rm[g].w & qe.w;
                    uint32_t ov = oV[g][j], om = oM[g][j], cv0, cv1, cm0, cm1;
                    csa(ov, cv0, ov, V0, V1);
                    csa(ov, cv1, ov, V2, V3);
                    csa(om, cm0, om, M0, M1);
                    csa(om, cm1, om, M2, M3);
                    oV[g][j] = ov; oM[g][j] = om;
                    uint32_t a = acc[g][j];
                    a = imad(__popc(cv0), k128k, a);
                    a = imad(__popc(cm0), k2, a);
                    a = imad(__popc(cv1), k128k, a);
                    a = imad(__popc(cm1), k2, a);
                    acc[g][j] = a;
                }
            }
        }
        // fix entries of this warp's slots in this chunk (usually 0..3): query base != pivot base at the masked bits
        {
            const uint32_t fb = st.q.fix_begin[warp], fe = st.q.fix_begin[warp + 1];
            const uint8_t* r0 = reinterpret_cast<const uint8_t*>(&st.ref[0][0]) + lane * 4;
            for (uint32_t e = fb; e < fe; ++e) {
                const uint2 ent = st.q.fix[e];
                const uint32_t off = ent.y & 0xffffu, jj = ent.y >> 16;
#pragma unroll
                for (int g = 0; g < RPL; ++g) {
                    const uint32_t hit = *reinterpret_cast<const uint32_t*>(r0 + g * (W4 * UNIT2_BYTES) + off) & ent.x;
                    if (hit) {
                        const uint32_t n = __popc(hit);
#pragma unroll
                        for (int j = 0; j < QW2; ++j) if (jj == (uint32_t)j) acc[g][j] += n;
                    }
                }
            }
        }
        // release the stage; the elected thread refills the stage consumed REFILL_LAG chunks ago (everybody is done with it
        // by now, or will be within a few hundred cycles)
        __syncwarp();
        if (lane == 0) {
            mbar_arrive(empty + s);
            if (warp == 0) {
                const int cr = c - REFILL_LAG, cn = cr + STAGES2;
                if (cr >= 0 && cn < p.n_chunks) {
                    const int sr = cr % STAGES2;
                    mbar_wait(empty + sr, (uint32_t)((cr / STAGES2) & 1));
                    issue(cn, sr);
                }
            }
        }
        __syncwarp();
    }
    if (lane == 0 && p.work) atomicAdd(p.work, (unsigned long long)p.n_chunks * (QW2 * RPL));

    // ---- epilogue: plane counts -> exact counts -> score tile of the owner ----
    bool in_range[RPL];
    int64_t local[RPL];
#pragma unroll
    for (int g = 0; g < RPL; ++g) {
        local[g] = (group0 + g) * GROUP + lane;
        in_range[g] = local[g] >= p.local_begin && local[g] < p.local_end;
    }
#pragma unroll
    for (int j = 0; j < QW2; ++j) {
        const int slot = qt * QTILE + q0 + j;
        if (slot >= p.n_slots) continue;
        const bool is_cons = slot == p.n_slots - 1;
#pragma unroll
        for (int g = 0; g < RPL; ++g) {
            if (!in_range[g]) continue;
            const int32_t M = (int32_t)(acc[g][j] & 0xffffu) + __popc(oM[g][j]);
            const int32_t V = (int32_t)(acc[g][j] >> 16) + __popc(oV[g][j]);
            const uint4 v = pair_epilogue_default(p, p.qfix_off, p.qfix_ent, slot, group0 + g, lane, local[g], M, V);
            pair_store(p, owner, row0 + q0 + j, is_cons, local[g], v);
        }
    }
}

// Derived database of the pivot form for the references [begin, end): per (group, unit) a RefUnit2 record with
//   base (copy), mc = base & ~(b0 ^ P.b0) & ~(b1 ^ P.b1), and the four one-hot planes of every word.
// One thread per (group, unit, lane): 3 coalesced 16-byte loads, 2 16-byte + 16 coalesced 4-byte stores.
// pivot[unit] = (b0, b1) planes of the pivot sequence, 4 words each.
__global__ void __launch_bounds__(256) k_derive_pivot(const uint4* __restrict__ db, uint8_t* __restrict__ db2, const uint4* __restrict__ pivot,
                                                      int64_t begin, int64_t end, int64_t group_first, int units, int Wp)
{
    const int lane = threadIdx.x & 31;
    const int64_t t = (int64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);        // (group, unit) index of this warp
    const int64_t group = group_first + t / units;
    const int unit = (int)(t % units);
    const int64_t ordinal = group * GROUP + lane;
    if (ordinal < begin || ordinal >= end) return;
    const uint4 rb = db[db_unit_index(group, unit, 0, lane, Wp)];
    const uint4 r0 = db[db_unit_index(group, unit, 1, lane, Wp)];
    const uint4 r1 = db[db_unit_index(group, unit, 2, lane, Wp)];
    const uint4 p0 = __ldg(pivot + 2 * unit), p1 = __ldg(pivot + 2 * unit + 1);
    uint4 mc;
    mc.x = keep_eq(keep_eq(rb.x, r0.x, p0.x), r1.x, p1.x);
    mc.y = keep_eq(keep_eq(rb.y, r0.y, p0.y), r1.y, p1.y);
    mc.z = keep_eq(keep_eq(rb.z, r0.z, p0.z), r1.z, p1.z);
    mc.w = keep_eq(keep_eq(rb.w, r0.w, p0.w), r1.w, p1.w);
    RefUnit2* rec = reinterpret_cast<RefUnit2*>(db2 + db2_unit_bytes(group, unit, Wp));
    rec->base[lane] = rb;
    rec->mc[lane] = mc;
    const uint32_t b[4] = {rb.x, rb.y, rb.z, rb.w}, x[4] = {r0.x, r0.y, r0.z, r0.w}, y[4] = {r1.x, r1.y, r1.z, r1.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        rec->onehot[k][0][lane] = b[k] & ~x[k] & ~y[k];      // A
        rec->onehot[k][1][lane] = b[k] & x[k] & ~y[k];       // C
        rec->onehot[k][2][lane] = b[k] & ~x[k] & y[k];       // G
        rec->onehot[k][3][lane] = b[k] & x[k] & y[k];        // T
    }
}

cudaError_t pairpivot_setup()
{
    return cudaFuncSetAttribute(k_pairpivot, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PIVOT_SMEM);
}

cudaError_t launch_pairpivot(const PairParams& p, cudaStream_t st)
{
    if (p.local_end <= p.local_begin) return cudaSuccess;
    const int64_t per = GROUP * RPL;
    const int64_t g0 = p.local_begin / per, g1 = (p.local_end + per - 1) / per;
    const int64_t blocks = (g1 - g0) * p.n_qtiles;
    if (blocks > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
    // multipliers passed at run time so that the compiler keeps IMAD on the FMA pipe (no shift/add on the ALU pipe);
    // a carry word has weight 2: matches count 2 per set bit, valid sites 2 << 16
    k_pairpivot<<<(unsigned)blocks, PIVOT_THREADS, PIVOT_SMEM, st>>>(p, 2u, 2u << 16);
    return cudaGetLastError();
}

cudaError_t launch_derive_pivot(const uint4* db, uint8_t* db2, const uint4* pivot, int64_t begin, int64_t end, int Wp, cudaStream_t st)
{
    if (end <= begin) return cudaSuccess;
    const int64_t g0 = begin / GROUP, g1 = (end + GROUP - 1) / GROUP;
    const int units = Wp / 4;
    const int64_t warps = (g1 - g0) * units;
    const int64_t blocks = (warps + 7) / 8;
    if (blocks > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
    k_derive_pivot<<<(unsigned)blocks, 256, 0, st>>>(db, db2, pivot, begin, end, g0, units, Wp);
    return cudaGetLastError();
}

}  // namespace uvg
