// K2p: pair-count kernel in PIVOT form for sm_100a (default scoring mode; the general kernel is paircount.cu).
// Same contract as k_paircount<0,false> -- the four counts of biomcmc_pairwise_score_matches_truncated_idx
// (biomcmc alignment.c:855-884) of every (query slot, reference) pair of a round, untruncated, stored into the dense
// score tile of the query's owner -- with a cheaper inner loop.
//
// All queries of one search are close to a pivot sequence P (per site the plain base most queries hold; uvg_set_queries).
// For a site where the query holds P's base, "query == reference" is "reference == P", which does not depend on the query:
//     mc (reference plane, derived once per query set by k_derive_pivot) = r.base & (r == P)
// The query planes are stored INVERTED: nq = no plain base here (or trimmed away / padding), ne = not P's base (nq is a
// subset of ne), so that
//     valid  = popc(r.base) - sum popc(r.base & nq)        match' = popc(r.mc) - sum popc(r.mc & ne)
// with per-reference totals (k_derive_pivot) as first terms and sums that only get contributions from 128-site units in
// which the query has an N, a gap, an ambiguity code or a SNP: a (slot, unit) pair whose `ne` words are all zero is
// skipped (QChunk2::active, warp-uniform branch).  Per executed pair-word: 2 LOP3 for the two ANDs, one 3:2 compressor
// per two words and stream (ones register + one carry word of weight 2), IMAD adds the popcounts:
//     4 LOP3 (ALU) + 1 POPC (XU) + 1 IMAD (FMA) per pair and 32-site word
// which balances the ALU and XU pipes at the measured 4:1 LOP3:POPC rate of sm_100 (profiles/int_pipes_b200_r01.json);
// the general kernel needs 4.5 LOP3 + 1.25 POPC and cannot skip.
// Query sites with a plain base != P (the query's own SNPs, a few dozen per genome) are outside match'; they are "fix"
// entries (slot, site, base) that travel with the query chunk and are resolved against column masks of the derived
// database (per site and base one word: bit l = reference l of the group holds that base): a warp fetches the masks of
// all its entries of a chunk with one load (lane = entry x group) while it counts the chunk, almost always finds them
// all zero, and otherwise bumps a per-pair counter in shared memory.  Entries that overflow a chunk's fix block and all
// ambiguity codes go to the epilogue (pair_epilogue.cuh).
// Shape: block = 64 query slots x 64 references, 16 warps x 4 slots, lane = 2 references, 64 registers, 2 blocks per
// SM (32 warps per SM).  Staging: TMA bulk copies (query chunk + the base/mc planes of the two reference groups) into an
// 8-deep ring with full/empty mbarriers (no __syncthreads in the loop: warps drift up to the ring depth apart); one
// elected thread refills a stage two chunks after it was consumed.
#include "common.cuh"
#include "kernels.h"
#include "pair_epilogue.cuh"
#include "sm100_prims.cuh"

namespace uvg {

// staged part of a chunk record: the base and mc planes of the block's two groups (the column masks stay in global memory /
// L2: only fix entries read them)
struct __align__(16) RefPlanes2 {
    uint4 base[GROUP];
    uint4 mc[GROUP];
};
static_assert(sizeof(RefPlanes2) == UNIT2_PLANES, "planes of one (group, unit) of the derived database");
struct __align__(16) Stage2 {
    QChunk2    q;
    RefPlanes2 ref[RPL][W4];
    uint32_t   cnt[RPL][W4][GROUP];      // unit totals of every reference: popc(mc) | popc(base) << 16
};
static_assert(sizeof(RefPlanes2) * RPL * W4 == CHUNK2_PLANES && sizeof(Stage2) == sizeof(QChunk2) + CHUNK2_STAGED, "staged prefix of a chunk record");
constexpr size_t PIVOT_FIXHIT_BYTES = QTILE * GROUP * RPL * sizeof(uint16_t);       // per (slot, reference of the block): matches found by fix entries
constexpr size_t PIVOT_SMEM = STAGES2 * sizeof(Stage2) + PIVOT_FIXHIT_BYTES + 2 * STAGES2 * sizeof(uint64_t);
#ifndef PIVOT_ROTATE
#define PIVOT_ROTATE 0          // 1: the warps take turns refilling the ring (chunk c: warp c % 16) -- measured 1.6x SLOWER; 0: warp 0 refills
#endif
#ifndef PIVOT_COPIES
#define PIVOT_COPIES 1          // bulk copies per stage for the reference side: 1 (4.5 KB) or 4 (1 KB per (group, unit) + the counts)
#endif
#ifndef PIVOT_LAG
#define PIVOT_LAG 2
#endif
constexpr int REFILL_LAG = PIVOT_LAG;

constexpr int QW2 = PIVOT_QW;                           // query slots per warp (register tile QW2 x RPL pairs per lane)
constexpr int PIVOT_THREADS = PIVOT_WARPS * 32;

__global__ void __launch_bounds__(PIVOT_THREADS, 2) k_pairpivot(const PairParams p, const uint32_t k2, const uint32_t k128k)
{
    extern __shared__ __align__(128) uint8_t smem[];
    Stage2* stage = reinterpret_cast<Stage2*>(smem);
    uint16_t* fixhit = reinterpret_cast<uint16_t*>(smem + STAGES2 * sizeof(Stage2));   // [slot][RPL * GROUP]: written and read by one thread each
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES2 * sizeof(Stage2) + PIVOT_FIXHIT_BYTES);
    uint64_t* empty = full + STAGES2;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int qt = blockIdx.x % p.n_qtiles;
    const int64_t group0 = (p.local_begin / (GROUP * RPL) + blockIdx.x / p.n_qtiles) * RPL;
    const QChunk2* qsrc = p.qdata2 + (size_t)qt * p.n_chunks;
    const uint8_t* rsrc = p.db2 + (size_t)(group0 / RPL) * (size_t)p.n_chunks * CHUNK2_BYTES;      // chunk records of this block's pair of groups
    const int owner = qt % p.world, row0 = (qt / p.world) * QTILE;
    // shared-memory addresses as plain integers: nothing in the loop goes through a generic -> shared conversion
    uint32_t stage_a = smem_u32(stage);
    asm volatile("mov.u32 %0, %0;" : "+r"(stage_a));                         // opaque: kept in a register instead of being re-derived (S2R + LEA) per use
    const uint32_t full_a = stage_a + (uint32_t)(STAGES2 * sizeof(Stage2) + PIVOT_FIXHIT_BYTES), empty_a = full_a + STAGES2 * 8u;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES2; ++s) { mbar_init(full + s, 1); mbar_init(empty + s, PIVOT_THREADS / 32); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < (int)(PIVOT_FIXHIT_BYTES / sizeof(uint4)); i += PIVOT_THREADS) reinterpret_cast<uint4*>(fixhit)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    // one stage = the query chunk + the planes of the pair's chunk record: two TMA bulk copies
    auto issue = [&](int c, int s) {
        const uint32_t bar = full_a + (uint32_t)s * 8u, dst = stage_a + (uint32_t)s * (uint32_t)sizeof(Stage2);
        mbar_expect_tx_a(bar, (uint32_t)(sizeof(QChunk2) + CHUNK2_STAGED));
        bulk_g2s_a(dst, qsrc + c, (uint32_t)sizeof(QChunk2), bar);
        const uint8_t* src = rsrc + (size_t)c * CHUNK2_BYTES;
#if PIVOT_COPIES == 1
        bulk_g2s_a(dst + (uint32_t)sizeof(QChunk2), src, (uint32_t)CHUNK2_STAGED, bar);
#else
#pragma unroll
        for (int k = 0; k < RPL * W4; ++k)
            bulk_g2s_a(dst + (uint32_t)sizeof(QChunk2) + (uint32_t)k * UNIT2_PLANES, src + (size_t)k * UNIT2_PLANES, (uint32_t)UNIT2_PLANES, bar);
        bulk_g2s_a(dst + (uint32_t)sizeof(QChunk2) + CHUNK2_PLANES, src + CHUNK2_PLANES, (uint32_t)CHUNK2_COUNTS, bar);
#endif
    };
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES2 && s < p.n_chunks; ++s) issue(s, s);
    }

    // per pair: carry-save "ones" registers of the two COMPLEMENT streams and one packed counter.  The query planes are
    // stored inverted (nq = no plain base here, ne = not the pivot's base; nq is a subset of ne):
    //     valid   = popc(r.base) - popc(r.base & nq)        match' = popc(r.mc) - popc(r.mc & ne)
    // The first terms are per-reference totals (k_derive_pivot); the second terms are zero wherever the query is complete
    // and equal to the pivot, so a (slot, unit) whose `ne` words are all zero is SKIPPED (uniform branch: the planes
    // are warp-wide broadcasts).  acc = (#weight-2 carries of r.mc & ne) * 2 | (#weight-2 carries of r.base & nq) * 2 << 16
    uint32_t oV[RPL][QW2], oM[RPL][QW2], acc[RPL][QW2];
#pragma unroll
    for (int g = 0; g < RPL; ++g)
#pragma unroll
        for (int j = 0; j < QW2; ++j) { oV[g][j] = 0; oM[g][j] = 0; acc[g][j] = 0; }

    // the slots of the tile this warp counts (uvg_set_queries balances the warps of a tile by work)
    // kept as byte offsets of the slot's planes inside QChunk2::pl[u], two per register (the loop is short of registers)
    static_assert(QW2 == 4 && sizeof(uint4) * 2 * QTILE <= 65536, "slot offsets are packed as 4 x 16 bits");
    uint32_t so[2];
    {
        const uint8_t* ws = p.wslot + ((size_t)qt * PIVOT_WARPS + warp) * QW2;
        so[0] = (uint32_t)ws[0] * 32u | ((uint32_t)ws[1] * 32u) << 16;
        so[1] = (uint32_t)ws[2] * 32u | ((uint32_t)ws[3] * 32u) << 16;
        asm volatile("mov.u32 %0, %0;" : "+r"(so[0]));                       // opaque: unpacked where used, not hoisted into four spilled registers
        asm volatile("mov.u32 %0, %0;" : "+r"(so[1]));
    }
    auto slot_off = [&](int j) -> uint32_t { return (j & 1) ? so[j >> 1] >> 16 : so[j >> 1] & 0xffffu; };
    auto slot_of = [&](int j) -> int { return (int)(slot_off(j) >> 5); };
    // resolve fetched column masks: a set bit = that reference holds the query's (non-pivot) base at the site, a match the
    // planes did not see.  Almost always no lane holds a non-zero mask.
    auto resolve = [&](uint32_t ent_l, uint32_t col_l) {
        uint32_t any = __ballot_sync(0xffffffffu, col_l != 0u);
        while (any) {
            const int src = __ffs(any) - 1;
            any &= any - 1u;
            const uint32_t col = __shfl_sync(0xffffffffu, col_l, src), ent = __shfl_sync(0xffffffffu, ent_l, src);
            fixhit[(ent >> 16) * (RPL * GROUP) + (src % RPL) * GROUP + lane] += (uint16_t)((col >> lane) & 1u);   // branch-free; this thread owns the counter
        }
    };
#if !PIVOT_SKIP2
    uint32_t executed = 0;                                                   // (slot, unit) iterations that were not skipped
#endif
    for (int c = 0; c < p.n_chunks; ++c) {
        const int s = c % STAGES2;
        mbar_wait_a(full_a + (uint32_t)s * 8u, (uint32_t)((c / STAGES2) & 1));
        const Stage2& st = stage[s];
        // fix entries of this warp's slots in this chunk (usually 0..3): lane l fetches the column mask of entry l / RPL for
        // reference group l % RPL (one load for the whole warp, consumed after the counting work of the chunk)
        const uint32_t fb = st.q.fix_begin[warp], fe = st.q.fix_begin[warp + 1];
        const uint8_t* gl0 = rsrc + (size_t)c * CHUNK2_BYTES;                  // fix entries: byte offsets inside this chunk record
        uint32_t my_ent = 0, my_col = 0;
        if (fb + lane / RPL < fe) {
            my_ent = st.q.fix[fb + lane / RPL];
            my_col = __ldg(reinterpret_cast<const uint32_t*>(gl0 + (lane % RPL) * (W4 * UNIT2_MASKS) + (my_ent & 0xffffu)));
        }
#if PIVOT_SKIP2
        // what this warp has to do in the chunk: per unit u and slot j either nothing (the query is complete and equal to
        // the pivot over the unit), or one add from the counts table (the query holds no plain base in the unit: all-N), or
        // up to two halves of 64 sites counted from the planes
        const uint32_t hmask = st.q.half[warp], nmask = st.q.fulln[warp];
#pragma unroll
        for (int u = 0; u < W4; ++u) {
            const uint32_t nu = (nmask >> (u * 4)) & 15u;
            if (nu) {
                uint32_t ct[RPL];
#pragma unroll
                for (int g = 0; g < RPL; ++g) ct[g] = st.cnt[g][u][lane];
#pragma unroll
                for (int j = 0; j < QW2; ++j)
                    if ((nu >> j) & 1u) {
#pragma unroll
                        for (int g = 0; g < RPL; ++g) acc[g][j] += ct[g];
                    }
            }
            const uint32_t hu = (hmask >> (u * 8)) & 255u;
            if (hu == 0u) continue;
            uint4 rb[RPL], rm[RPL];
#pragma unroll
            for (int g = 0; g < RPL; ++g) { rb[g] = st.ref[g][u].base[lane]; rm[g] = st.ref[g][u].mc[lane]; }
#pragma unroll
            for (int j = 0; j < QW2; ++j) {
                const uint32_t hj = (hu >> (j * 2)) & 3u;
                if (hj == 0u) continue;
                const uint2* qp = reinterpret_cast<const uint2*>(reinterpret_cast<const uint8_t*>(&st.q.pl[u][0][0]) + slot_off(j));
                if (hj & 1u) {
                    const uint2 qb = qp[0], qe = qp[2];                      // words 0, 1 of the two complement planes
#pragma unroll
                    for (int g = 0; g < RPL; ++g) {
                        uint32_t ov = oV[g][j], om = oM[g][j], cv, cm;
                        csa(ov, cv, ov, rb[g].x & qb.x, rb[g].y & qb.y);
                        csa(om, cm, om, rm[g].x & qe.x, rm[g].y & qe.y);
                        oV[g][j] = ov; oM[g][j] = om;
                        acc[g][j] = imad(__popc(cm), k2, imad(__popc(cv), k128k, acc[g][j]));
                    }
                }
                if (hj & 2u) {
                    const uint2 qb = qp[1], qe = qp[3];                      // words 2, 3
#pragma unroll
                    for (int g = 0; g < RPL; ++g) {
                        uint32_t ov = oV[g][j], om = oM[g][j], cv, cm;
                        csa(ov, cv, ov, rb[g].z & qb.x, rb[g].w & qb.y);
                        csa(om, cm, om, rm[g].z & qe.x, rm[g].w & qe.y);
                        oV[g][j] = ov; oM[g][j] = om;
                        acc[g][j] = imad(__popc(cm), k2, imad(__popc(cv), k128k, acc[g][j]));
                    }
                }
            }
        }
#else
#pragma unroll
        for (int u = 0; u < W4; ++u) {
            // which of this warp's slots have work in this unit (the others are complete and equal to the pivot over these
            // 128 sites); nothing to do for any of them: not even the reference words are read
            const uint32_t act = st.q.active[u][warp];
            if (act == 0u) continue;
            uint4 rb[RPL], rm[RPL];
#pragma unroll
            for (int g = 0; g < RPL; ++g) { rb[g] = st.ref[g][u].base[lane]; rm[g] = st.ref[g][u].mc[lane]; }
#pragma unroll
            for (int j = 0; j < QW2; ++j) {
                if (!((act >> j) & 1u)) continue;
                const uint4* qp = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(&st.q.pl[u][0][0]) + slot_off(j));
                const uint4 qb = qp[0], qe = qp[1];
                ++executed;
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
#endif
        if (fe > fb) {
            resolve(my_ent, my_col);
            for (uint32_t e0 = fb + 32 / RPL; e0 < fe; e0 += 32 / RPL) {    // more than 16 entries (rare)
                my_col = 0;
                if (e0 + lane / RPL < fe) {
                    my_ent = st.q.fix[e0 + lane / RPL];
                    my_col = __ldg(reinterpret_cast<const uint32_t*>(gl0 + (lane % RPL) * (W4 * UNIT2_MASKS) + (my_ent & 0xffffu)));
                }
                resolve(my_ent, my_col);
            }
        }
        // release the stage.  Refill: the stage consumed REFILL_LAG chunks ago gets the chunk STAGES2 further on.  The warps take
        // turns (chunk c: warp c % 16), so that no warp carries the producer's instructions on top of its counting work -- with
        // one fixed producer warp that warp sets the pace of the block and the other fifteen wait for it at the head of the ring.
        __syncwarp();
        if (lane == 0) {
            mbar_arrive_a(empty_a + (uint32_t)s * 8u);
#if PIVOT_ROTATE
            if (warp == (c % PIVOT_WARPS)) {
#else
            if (warp == 0) {
#endif
                const int cr = c - REFILL_LAG, cn = cr + STAGES2;
                if (cr >= 0 && cn < p.n_chunks) {
                    const int sr = cr % STAGES2;
                    mbar_wait_a(empty_a + (uint32_t)sr * 8u, (uint32_t)((cr / STAGES2) & 1));
                    issue(cn, sr);
                }
            }
        }
        __syncwarp();
    }
    // p.work counts (pair, 4-word unit) iterations per lane.  The half-unit variant does not count in the kernel (no register
    // to spare): the host knows the masks and adds halves x references per launch (engine.cu, run_round)
#if !PIVOT_SKIP2
    if (lane == 0 && p.work) atomicAdd(p.work, (unsigned long long)executed * RPL);
#endif

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
        const int sjj = slot_of(j);
        const int slot = qt * QTILE + sjj;
        if (slot >= p.n_slots) continue;
        const bool is_cons = slot == p.n_slots - 1;
#pragma unroll
        for (int g = 0; g < RPL; ++g) {
            if (!in_range[g]) continue;
            const int32_t M = p.mc_cnt[local[g]] - (int32_t)((acc[g][j] & 0xffffu) + __popc(oM[g][j])) + (int32_t)fixhit[sjj * (RPL * GROUP) + g * GROUP + lane];
            const int32_t V = p.rb_cnt[local[g]] - (int32_t)((acc[g][j] >> 16) + __popc(oV[g][j]));
            const uint4 v = pair_epilogue_default(p, p.qfix_off, p.qfix_ent, slot, group0 + g, lane, local[g], M, V);
            pair_store(p, owner, row0 + sjj, is_cons, local[g], v);
        }
    }
}

// Derived database of the pivot form for the references [begin, end): per (group, unit), inside the chunk record of the
// group's pair (common.cuh): base (copy), mc = base & ~(b0 ^ P.b0) & ~(b1 ^ P.b1), and the column masks of every word.
// One thread per (group, unit, lane): 3 coalesced 16-byte loads, 2 16-byte + 16 coalesced 4-byte stores.
// pivot[unit] = (b0, b1) planes of the pivot sequence, 4 words each.
__global__ void __launch_bounds__(256) k_derive_pivot(const uint4* __restrict__ db, uint8_t* __restrict__ db2, const uint4* __restrict__ pivot,
                                                      int32_t* __restrict__ rb_cnt, int32_t* __restrict__ mc_cnt,
                                                      int64_t begin, int64_t end, int64_t group_first, int units, int Wp)
{
    const int lane = threadIdx.x & 31;
    const int64_t t = (int64_t)blockIdx.x * (blockDim.x / 32) + (threadIdx.x >> 5);        // (group, unit) index of this warp
    const int64_t group = group_first + t / units;
    const int unit = (int)(t % units);
    const int64_t ordinal = group * GROUP + lane;
    if (group * GROUP >= end) return;                                                       // whole warp
    const bool resident = ordinal < end;                     // lanes below `begin` were derived before: they only vote
    const bool fresh = resident && ordinal >= begin;
    uint4 rb = make_uint4(0u, 0u, 0u, 0u), r0 = rb, r1 = rb;
    if (resident) {
        rb = db[db_unit_index(group, unit, 0, lane, Wp)];
        r0 = db[db_unit_index(group, unit, 1, lane, Wp)];
        r1 = db[db_unit_index(group, unit, 2, lane, Wp)];
    }
    const uint4 p0 = __ldg(pivot + 2 * unit), p1 = __ldg(pivot + 2 * unit + 1);
    uint4 mc;
    mc.x = keep_eq(keep_eq(rb.x, r0.x, p0.x), r1.x, p1.x);
    mc.y = keep_eq(keep_eq(rb.y, r0.y, p0.y), r1.y, p1.y);
    mc.z = keep_eq(keep_eq(rb.z, r0.z, p0.z), r1.z, p1.z);
    mc.w = keep_eq(keep_eq(rb.w, r0.w, p0.w), r1.w, p1.w);
    RefPlanes2* rec = reinterpret_cast<RefPlanes2*>(db2 + db2_planes_bytes(group, unit, Wp));
    uint32_t* colmask = reinterpret_cast<uint32_t*>(db2 + db2_masks_bytes(group, unit, Wp));     // [word][base A,C,G,T][site in word]
    if (fresh) {
        rec->base[lane] = rb;
        rec->mc[lane] = mc;
        reinterpret_cast<uint32_t*>(db2 + db2_counts_bytes(group, unit, Wp))[lane] =
            (uint32_t)(__popc(mc.x) + __popc(mc.y) + __popc(mc.z) + __popc(mc.w)) | (uint32_t)(__popc(rb.x) + __popc(rb.y) + __popc(rb.z) + __popc(rb.w)) << 16;
        atomicAdd(rb_cnt + ordinal, __popc(rb.x) + __popc(rb.y) + __popc(rb.z) + __popc(rb.w));
        atomicAdd(mc_cnt + ordinal, __popc(mc.x) + __popc(mc.y) + __popc(mc.z) + __popc(mc.w));
    }
    // column masks: 32 x 32 bit transposes of the one-hot planes (ballots); lane s keeps and stores the mask of site s.
    // The whole word is rewritten: the bits of the references derived earlier do not change.
    const uint32_t b[4] = {rb.x, rb.y, rb.z, rb.w}, x[4] = {r0.x, r0.y, r0.z, r0.w}, y[4] = {r1.x, r1.y, r1.z, r1.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const uint32_t oh[4] = {b[k] & ~x[k] & ~y[k], b[k] & x[k] & ~y[k], b[k] & ~x[k] & y[k], b[k] & x[k] & y[k]};      // A, C, G, T
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            uint32_t mine = 0;
#pragma unroll
            for (int sbit = 0; sbit < 32; ++sbit) {
                const uint32_t m = __ballot_sync(0xffffffffu, (oh[a] >> sbit) & 1u);
                if (lane == sbit) mine = m;
            }
            colmask[(k * 4 + a) * 32 + lane] = mine;
        }
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

cudaError_t launch_derive_pivot(const uint4* db, uint8_t* db2, const uint4* pivot, int32_t* rb_cnt, int32_t* mc_cnt, int64_t begin, int64_t end,
                                int Wp, cudaStream_t st)
{
    if (end <= begin) return cudaSuccess;
    const int64_t g0 = begin / GROUP, g1 = (end + GROUP - 1) / GROUP;
    const int units = Wp / 4;
    const int64_t warps = (g1 - g0) * units;
    const int64_t blocks = (warps + 7) / 8;
    if (blocks > 0x7fffffffLL) return cudaErrorInvalidConfiguration;
    k_derive_pivot<<<(unsigned)blocks, 256, 0, st>>>(db, db2, pivot, rb_cnt, mc_cnt, begin, end, g0, units, Wp);
    return cudaGetLastError();
}

}  // namespace uvg
