// K2p: pair-count kernel in PIVOT form for sm_100a (default scoring mode; the general kernel is paircount.cu).
// Same contract as k_paircount<0,false> -- the four counts of biomcmc_pairwise_score_matches_truncated_idx
// (biomcmc alignment.c:855-884) of every (query slot, reference) pair of a round, untruncated, stored into the dense
// score tile of the query's owner -- with a much cheaper inner loop.
//
// All queries of one search are close to a pivot sequence P (per site the plain base most queries hold; uvg_set_queries).
// For a site where the query holds P's base, "query == reference" is "reference == P", which does not depend on the query:
//     mc (reference plane, derived once per query set by k_derive_pivot) = r.base & (r == P)
// The query plane is used INVERTED: nq = no plain base here (or trimmed away / padding), so that
//     valid  = popc(r.base) - sum popc(r.base & nq)        match' = popc(r.mc) - sum popc(r.mc & nq)
// with per-reference totals (k_derive_pivot) as first terms and sums that only get contributions from sites where the query
// has an N, a gap or an ambiguity code.  match' treats EVERY plain base of the query as the pivot's; a base that differs (a
// SNP, a few dozen per genome) is settled by its fix entry after the loop.  Per (slot, 128-site unit) a warp does one of
// three things (QWarpHead):
//   * nothing: the query holds a plain base everywhere in the unit (~75 % of the units of a SARS-CoV-2 query set);
//   * one add from the counts table of the derived database: the query holds no plain base in the unit (inside an N run,
//     trimmed away), so both sums are the reference's own unit totals (~12 %);
//   * count up to two halves of 64 sites from the planes (~14 % of the units): per pair and half 4 AND + 4 POPC + 4 IMAD
//         = 2 AND (ALU) + 2 POPC (XU) + 2 IMAD (FMA) per counted pair and 32-site word.
//     (The first half of round 2 used a carry-save step, 4 LOP3 + 1 POPC + 1 IMAD, balanced for a dense loop; this loop is
//     mostly the walk, which keeps the ALU pipe busy, and without the "ones" registers a lane holds four references.)
// After the main loop every warp resolves the fix entries of its four slots against the column masks of the derived database
// (per site and base one word: bit l = reference l of the group holds that base): minus the references that hold the pivot's
// base, plus those that hold the query's; loads for 32 entries at a time, all independent, then a 32 x 32 bit transpose across
// the warp.  The query's ambiguity codes go the same way; the references' codes are looked up per reference in the epilogue.
// Shape: block = 60 query slots ("pivot tile") x 128 references, 16 warps: warp 0 is the PRODUCER (it keeps the staging ring
// full and counts nothing), warps 1..15 count 4 slots each, lane = PRPL = 4 references (what a warp does per chunk besides
// counting -- wait, head, walk, release -- is paid once per 128 references); 64 registers, 2 blocks per SM.  Staging: a ring
// of STAGES2 stages, each filled by two TMA bulk copies -- the (pivot tile, chunk) record of the query stream (variable
// length: 128-byte head + 8 bytes per counted half) and the planes + counts of the block's chunk record (9 KB) -- with
// full/empty mbarriers (no __syncthreads in the loop: warps drift up to the ring depth apart).  7 stages = 80 KB per block: the
// shared memory of the two blocks of an SM is carved out of the L1, and a deeper ring (205 KB per SM) measured 15 % slower.
// Why a warp that only produces: a bulk copy costs its issuing warp ~200 cycles; a warp that issued the copies on top of an
// equal share of the counting set the pace of the block, and the other fifteen waited for it at the head of the ring (r02
// profile: 20 % of all warp time in the full-barrier wait; rotating the duty among the warps was 1.6x slower still).
#include "common.cuh"
#include "kernels.h"
#include "pair_epilogue.cuh"
#include "sm100_prims.cuh"

namespace uvg {

// staged part of a chunk record of the derived database
struct __align__(16) RefPlanes2 {           // half-major: a warp's 8-byte loads of one half are consecutive (no bank conflicts)
    uint2 base[2][GROUP];                   // [half: words 0,1 | words 2,3][reference of the group]
    uint2 mc[2][GROUP];
};
static_assert(sizeof(RefPlanes2) == UNIT2_PLANES, "planes of one (group, unit) of the derived database");
struct __align__(16) Stage2 {
    uint8_t    q[QREC_MAX];              // QWarpHead[16] | items
    RefPlanes2 ref[PRPL][W4];
    uint32_t   cnt[PRPL][W4][GROUP];     // unit totals of every reference: popc(mc) | popc(base) << 16
};
static_assert(sizeof(RefPlanes2) * PRPL * W4 == CHUNK2_PLANES && sizeof(Stage2) == QREC_MAX + CHUNK2_STAGED, "ring stage");
constexpr int PIVOT_MAX_CHUNKS = 256;                                  // < 65,536 columns
constexpr size_t PIVOT_SMEM = STAGES2 * sizeof(Stage2) + 2 * STAGES2 * sizeof(uint64_t) + (PIVOT_MAX_CHUNKS + 4) * sizeof(uint32_t);
#ifndef PIVOT_UNROLL_UNITS
#define PIVOT_UNROLL_UNITS 0    // 1: the loop over the two units of a chunk is unrolled (twice the loop's instruction footprint: measured slower)
#endif
#ifndef PIVOT_PRODUCER_SLEEP
#define PIVOT_PRODUCER_SLEEP 0   // ns the producer thread sleeps between two tests of a stage that is still in use
#endif
#ifndef PIVOT_UNIFORM_HEAD
#define PIVOT_UNIFORM_HEAD 1
#endif
#ifndef PIVOT_CONSUMER_SLEEP
#define PIVOT_CONSUMER_SLEEP 0  // ns a counting warp sleeps between two tests of a stage that has not arrived yet (0 = try_wait)
#endif
#ifndef PIVOT_WAIT_HINT
#define PIVOT_WAIT_HINT 0       // ns a counting warp may sleep inside the try_wait for its stage (0 = the default time limit)
#endif
#ifndef PIVOT_GRANULE
#define PIVOT_GRANULE 0         // bytes per bulk copy of the reference side of a stage (0 = one copy of 4.5 KB; smaller copies measured slower)
#endif
constexpr int QW2 = PIVOT_QW;                           // query slots per warp (register tile QW2 x PRPL pairs per lane)
constexpr int PIVOT_THREADS = PIVOT_WARPS * 32;

__global__ void __launch_bounds__(PIVOT_THREADS, 2) k_pairpivot(const PairParams p, const uint32_t k2, const uint32_t k128k)
{
    extern __shared__ __align__(128) uint8_t smem[];
    Stage2* stage = reinterpret_cast<Stage2*>(smem);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pt = blockIdx.x % p.n_ptiles;                                  // pivot tile: slots [pt * 60, pt * 60 + 60)
    const int64_t group0 = (p.local_begin / (GROUP * PRPL) + blockIdx.x / p.n_ptiles) * PRPL;
    const uint32_t* qoff = p.qoff + (size_t)pt * (p.n_chunks + 1);
    const uint8_t* rsrc = p.db2 + (size_t)(group0 / PRPL) * (size_t)p.n_chunks * CHUNK2_BYTES;      // chunk records of this block's pair of groups
    // shared-memory addresses as plain integers: nothing in the loop goes through a generic -> shared conversion
    uint32_t stage_a = smem_u32(stage);
    asm volatile("mov.u32 %0, %0;" : "+r"(stage_a));                         // opaque: kept in a register instead of being re-derived (S2R + LEA) per use
    const uint32_t full_a = stage_a + (uint32_t)(STAGES2 * sizeof(Stage2)), empty_a = full_a + STAGES2 * 8u;
    // offsets of the tile's records in shared memory: the producer thread must not wait for a global load per chunk
    uint32_t* s_qoff = reinterpret_cast<uint32_t*>(smem + STAGES2 * sizeof(Stage2) + 2 * STAGES2 * sizeof(uint64_t));
    for (int i = threadIdx.x; i <= p.n_chunks; i += PIVOT_THREADS) s_qoff[i] = __ldg(qoff + i);

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES2; ++s) { mbar_init_a(full_a + s * 8u, 1); mbar_init_a(empty_a + s * 8u, PIVOT_WARPS - 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    // one stage = the query record of the chunk (exactly its bytes) + planes and counts of the pair's chunk record.  (Two
    // chunks per stage and barrier round trip were measured 4 % slower: the stages get coarser, the warps drift less freely.)
    auto issue_at = [&](int c, int s, uint32_t o0, uint32_t o1) {
        const uint32_t bar = full_a + (uint32_t)s * 8u, dst = stage_a + (uint32_t)s * (uint32_t)sizeof(Stage2);
        mbar_expect_tx_a(bar, (o1 - o0) + (uint32_t)CHUNK2_STAGED);
        bulk_g2s_a(dst, p.qstream + o0, o1 - o0, bar);
#if PIVOT_GRANULE == 0
        bulk_g2s_a(dst + (uint32_t)QREC_MAX, rsrc + (size_t)c * CHUNK2_BYTES, (uint32_t)CHUNK2_STAGED, bar);
#else
        static_assert(CHUNK2_STAGED % PIVOT_GRANULE == 0 && PIVOT_GRANULE % 16 == 0, "granule");
#pragma unroll
        for (int k = 0; k < CHUNK2_STAGED / PIVOT_GRANULE; ++k)
            bulk_g2s_a(dst + (uint32_t)QREC_MAX + k * PIVOT_GRANULE, rsrc + (size_t)c * CHUNK2_BYTES + k * PIVOT_GRANULE, (uint32_t)PIVOT_GRANULE, bar);
#endif
    };
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES2 && s < p.n_chunks; ++s) issue_at(s, s, s_qoff[s], s_qoff[s + 1]);
    }

    // per pair one packed counter: what has to be SUBTRACTED from the reference's totals,
    //     acc = sum popc(r.mc & ne) | sum popc(r.base & nq) << 16           (+ the unit totals of all-N units)
    // Every counted word gets its own POPC (per pair and half 4 AND + 4 POPC + 4 IMAD: ALU 4, XU 4, FMA 4).  The carry-save
    // step of the first half of the round (8 LOP3 + 2 POPC + 2 IMAD) suited a dense loop; this loop is mostly the walk, which
    // keeps the ALU pipe busy while the XU pipe idles -- and without the "ones" registers a lane holds PRPL = 4 references.
    uint32_t acc[PRPL][QW2];
#pragma unroll
    for (int g = 0; g < PRPL; ++g)
#pragma unroll
        for (int j = 0; j < QW2; ++j) acc[g][j] = 0;

    const uint32_t k1 = k2 >> 1, k64k = k128k >> 1;                          // run-time multipliers 1 and 1 << 16: IMAD stays on the FMA pipe
    if (warp == 0) {
        // ---- producer warp: keeps the ring full.  Chunk c goes into stage c % STAGES2 as soon as the fifteen counting warps
        // have released what the stage held before; it has no slots of its own, so nobody ever waits for its counting work ----
        if (lane == 0) {
            int s = 0; uint32_t ph = 0;                                      // stage / phase of the chunk being issued
            for (int c = STAGES2; c < p.n_chunks; ++c) {
#if PIVOT_PRODUCER_SLEEP
                while (!mbar_test_a(empty_a + (uint32_t)s * 8u, ph)) __nanosleep(PIVOT_PRODUCER_SLEEP);
#else
                mbar_wait_a(empty_a + (uint32_t)s * 8u, ph);                 // use c / STAGES2 - 1 of the stage is over
#endif
                issue_at(c, s, s_qoff[c], s_qoff[c + 1]);
                if (++s == STAGES2) { s = 0; ph ^= 1u; }
            }
        }
        return;
    }
    int s = 0;                                                               // stage of chunk c = c % STAGES2, ph = (c / STAGES2) & 1
    uint32_t ph = 0;
    for (int c = 0; c < p.n_chunks; ++c) {
#if PIVOT_CONSUMER_SLEEP
        while (!mbar_test_a(full_a + (uint32_t)s * 8u, ph)) __nanosleep(PIVOT_CONSUMER_SLEEP);
#elif PIVOT_WAIT_HINT
        mbar_wait_hint_a(full_a + (uint32_t)s * 8u, ph, PIVOT_WAIT_HINT);
#else
        mbar_wait_a(full_a + (uint32_t)s * 8u, ph);
#endif
        const Stage2& st = stage[s];
        // what this warp has to do in the chunk, and where its items start
        uint2 hd = reinterpret_cast<const uint2*>(st.q)[warp];
#if PIVOT_UNIFORM_HEAD
        // the head is the same for all lanes; a broadcast from lane 0 lets the compiler KNOW it, so that the branches on its bits
        // are uniform branches without convergence barriers (BSSY / BSYNC were ~10 % of the instructions of the loop)
        hd.x = __shfl_sync(0xffffffffu, hd.x, 0); hd.y = __shfl_sync(0xffffffffu, hd.y, 0);
#endif
        const uint32_t hmask = hd.x & 0xffffu, nmask = (hd.x >> 16) & 0xffu;
        const uint2* item = reinterpret_cast<const uint2*>(st.q + QHEAD_BYTES) + (hd.y & 0xffffu);
#if PIVOT_UNROLL_UNITS
#pragma unroll
#else
#pragma unroll 1                                                             // one copy of the unit body: half the loop's instruction footprint
#endif
        for (int u = 0; u < W4; ++u) {
            const uint32_t nu = (nmask >> (u * 4)) & 15u;
            if (nu) {                                                        // all-N units: the reference's own unit totals
                uint32_t ct[PRPL];
#pragma unroll
                for (int g = 0; g < PRPL; ++g) ct[g] = st.cnt[g][u][lane];
#pragma unroll
                for (int j = 0; j < QW2; ++j)
                    if ((nu >> j) & 1u) {
#pragma unroll
                        for (int g = 0; g < PRPL; ++g) acc[g][j] += ct[g];
                    }
            }
            const uint32_t hu = (hmask >> (u * 8)) & 255u;
            if (hu == 0u) continue;
            // the loop over the slots is unrolled (a slot's counters are registers), the loops over units and halves are not: with
            // four references per lane the fully unrolled walk was 16.6 KB of SASS and half of all warp samples waited for
            // instructions (profiles/ncu_pairpivot_r02av_*); a unit or half only moves the shared-memory address
#pragma unroll
            for (int j = 0; j < QW2; ++j) {
                const uint32_t hj = (hu >> (j * 2)) & 3u;
                if (hj == 0u) continue;
#pragma unroll 1
                for (int h = 0; h < 2; ++h) {
                    if (!((hj >> h) & 1u)) continue;
                    const uint2 q = *item++;                                 // nq words 2h, 2h+1
#pragma unroll
                    for (int g = 0; g < PRPL; ++g) {
                        // words 2h, 2h+1 of the reference's two planes: 8 bytes each, transient registers
                        const uint2 rb = st.ref[g][u].base[h][lane];
                        const uint2 rm = st.ref[g][u].mc[h][lane];
                        acc[g][j] = imad(__popc(rm.x & q.x), k1, imad(__popc(rm.y & q.y), k1,
                                    imad(__popc(rb.x & q.x), k64k, imad(__popc(rb.y & q.y), k64k, acc[g][j]))));
                    }
                }
            }
        }
        // release the stage to the producer warp
        __syncwarp();
        if (lane == 0) mbar_arrive_a(empty_a + (uint32_t)s * 8u);
        if (++s == STAGES2) { s = 0; ph ^= 1u; }
    }

    // ---- sparse corrections from the column masks of the derived database (per site and base one word: bit l = reference l of
    // the group holds that base there), per slot of the warp:
    //   fix entries    -- plain bases that differ from the pivot: a reference that holds the same base is one more match;
    //   ambiguity codes -- the site is outside the query planes: a reference with a plain base there makes the pair valid, and
    //                      partial when the code contains that base (alignment.c:871-877; never acgt / text: the codes differ).
    // Lane = entry for the loads (32 entries x 1..4 independent 4-byte loads in flight, L2 hits or the 8 KB mask part of the
    // chunk records), then the 32 x 32 bit matrix (entry x reference) is transposed across the warp (5 butterfly steps), so
    // that every lane holds the 32 entries' bits of ITS reference and one POPC counts them: ~2 instructions per entry and
    // group instead of a serial bit test per entry (r01: ~18 per entry in the epilogue, 12x the main loop on config 4). ----
    // slots of the pivot tile this warp counted (-1: none), broadcast from lane 0 like everything below that is the same for
    // all lanes (list bounds, owners): the compiler then keeps it in uniform registers and the loops need no convergence barriers
    int slot_of[QW2];
    {
        static_assert(QW2 == 4, "four slot bytes in one word");
        const uint32_t ws4 = __shfl_sync(0xffffffffu, __ldg(reinterpret_cast<const uint32_t*>(p.wslot) + (size_t)pt * PIVOT_WARPS + warp), 0);
#pragma unroll
        for (int j = 0; j < QW2; ++j) {
            const uint32_t b = (ws4 >> (j * 8)) & 0xffu;
            const int slot = pt * PIVOT_SLOTS + (int)b;
            slot_of[j] = (b == 0xffu || slot >= p.n_slots) ? -1 : slot;
        }
    }
    // matches by fix entries, partial by ambiguity codes; the valid sites the codes add are folded into the counter at once
    // (it holds what is SUBTRACTED from the reference's total, modulo 2^16: alignments of the pivot form have < 65,536 columns).
    // Corrections and epilogue run in passes of TG = 2 reference groups, so that the per-pair correction registers stay few.
    constexpr int TG = 2;
    static_assert(PRPL % TG == 0, "passes of the tail");
    auto transpose32 = [&](uint32_t x) -> uint32_t {
#pragma unroll
        for (int j = 16; j >= 1; j >>= 1) {
            const uint32_t m = j == 16 ? 0x0000ffffu : j == 8 ? 0x00ff00ffu : j == 4 ? 0x0f0f0f0fu : j == 2 ? 0x33333333u : 0x55555555u;
            const uint32_t y = __shfl_xor_sync(0xffffffffu, x, j);
            x = (lane & j) ? ((x & ~m) | ((y >> j) & m)) : ((x & m) | ((y << j) & ~m));
        }
        return x;
    };
    auto mask_words = [&](uint32_t site) -> const uint8_t* {                     // column masks [base A,C,G,T][site in word] of the site's word, first group
        const uint32_t w = site >> 5;
        return rsrc + (size_t)(w / WC) * CHUNK2_BYTES + CHUNK2_STAGED + ((w % WC) / 4) * UNIT2_MASKS + (((w & 3u) * 4u) * 32u + (site & 31u)) * 4u;
    };
#pragma unroll
    for (int gp = 0; gp < PRPL; gp += TG) {
        uint32_t fh[TG][QW2], fp[TG][QW2];
#pragma unroll
        for (int j = 0; j < QW2; ++j) {
#pragma unroll
            for (int g = 0; g < TG; ++g) { fh[g][j] = 0; fp[g][j] = 0; }
            const int slot = slot_of[j];
            if (slot < 0) continue;
            uint32_t e1 = __shfl_sync(0xffffffffu, __ldg(p.snp_off + slot + 1), 0);
            for (uint32_t e0 = __shfl_sync(0xffffffffu, __ldg(p.snp_off + slot), 0); e0 < e1; e0 += 32) {
                const bool have = e0 + lane < e1;
                const uint32_t ent = have ? __ldg(p.snp_ent + e0 + lane) : 0u;
                const uint8_t* mrec = rsrc + (size_t)(ent >> 16) * CHUNK2_BYTES + CHUNK2_STAGED + ((ent >> 9) & 1u) * UNIT2_MASKS + gp * (W4 * UNIT2_MASKS);
                const uint32_t oq = (ent & 0x1ffu) * 4u, op = ((ent & 0x19fu) | ((ent >> 5) & 0x60u)) * 4u;       // same word and site, the query's / the pivot's base
                uint32_t mq[TG], mp[TG];
#pragma unroll
                for (int g = 0; g < TG; ++g) {
                    mq[g] = have ? __ldg(reinterpret_cast<const uint32_t*>(mrec + g * (W4 * UNIT2_MASKS) + oq)) : 0u;
                    mp[g] = have ? __ldg(reinterpret_cast<const uint32_t*>(mrec + g * (W4 * UNIT2_MASKS) + op)) : 0u;
                }
#pragma unroll
                for (int g = 0; g < TG; ++g) {
                    // the planes counted the site as "query holds the pivot's base": take back the references that hold it, add those
                    // that hold the query's base
                    if (__ballot_sync(0xffffffffu, mq[g] != 0u)) fh[g][j] += __popc(transpose32(mq[g]));
                    fh[g][j] -= __popc(transpose32(mp[g]));
                }
            }
            e1 = __shfl_sync(0xffffffffu, __ldg(p.qamb_off + slot + 1), 0);
            for (uint32_t e0 = __shfl_sync(0xffffffffu, __ldg(p.qamb_off + slot), 0); e0 < e1; e0 += 32) {
                const bool have = e0 + lane < e1;
                const uint32_t ent = have ? __ldg(p.qamb_ent + e0 + lane) : 0u, code = ent >> 28;
                const uint8_t* mrec = mask_words(ent & AMB_SITE_MASK) + gp * (W4 * UNIT2_MASKS);
#pragma unroll
                for (int g = 0; g < TG; ++g) {
                    uint32_t vm = 0, pm = 0;
                    if (have) {
#pragma unroll
                        for (int a = 0; a < 4; ++a) {
                            const uint32_t m = __ldg(reinterpret_cast<const uint32_t*>(mrec + g * (W4 * UNIT2_MASKS) + a * 128));
                            vm |= m;
                            pm |= (code >> a) & 1u ? m : 0u;
                        }
                    }
                    acc[gp + g][j] -= (uint32_t)__popc(transpose32(vm)) << 16;
                    fp[g][j] += __popc(transpose32(pm));
                }
            }
        }

        // ---- epilogue: plane counts -> exact counts -> score tile of the owner (the four counts of alignment.c:866-879).  The
        // sites where the REFERENCE holds an ambiguity code are outside its planes: its (short) list is walked once per reference
        // and every entry is looked up in the 4-bit codes of the warp's four slots (pair_epilogue.cuh does the same per pair). ----
#pragma unroll
        for (int g = 0; g < TG; ++g) {
            const int64_t local = (group0 + gp + g) * GROUP + lane;
            if (local < p.local_begin || local >= p.local_end) continue;
            int32_t dv[QW2], dp[QW2], dt[QW2];
#pragma unroll
            for (int j = 0; j < QW2; ++j) { dv[j] = 0; dp[j] = 0; dt[j] = 0; }
            const uint32_t re1 = p.ramb_off[local + 1];
            for (uint32_t e = p.ramb_off[local]; e < re1; ++e) {
                const uint32_t ent = __ldg(p.ramb_ent + e), site = ent & AMB_SITE_MASK, rc = ent >> 28;
#pragma unroll
                for (int j = 0; j < QW2; ++j) {
                    if (slot_of[j] < 0) continue;
                    const uint32_t qc = (__ldg(p.qnib + (size_t)slot_of[j] * p.qnib_stride + (site >> 3)) >> ((site & 7u) * 4u)) & 15u;
                    dv[j] += qc != 0u; dp[j] += (qc & rc) != 0u; dt[j] += qc == rc;
                }
            }
            const int32_t mc_total = p.mc_cnt[local], rb_total = p.rb_cnt[local];
#pragma unroll
            for (int j = 0; j < QW2; ++j) {
                const int slot = slot_of[j];
                if (slot < 0) continue;
                const int32_t M = mc_total - (int32_t)(acc[gp + g][j] & 0xffffu) + (int32_t)fh[g][j];
                const int32_t V = (int32_t)(((uint32_t)rb_total - (acc[gp + g][j] >> 16)) & 0xffffu);
                // acgt, text, partial (+ the query's ambiguity codes), valid
                const uint4 v = make_uint4((uint32_t)M, (uint32_t)(M + dt[j]), (uint32_t)(M + dp[j]) + fp[g][j], (uint32_t)(V + dv[j]));
                const int qt = slot / QTILE;                                 // uniform: one division per slot and reference group
                pair_store(p, qt % p.world, (qt / p.world) * QTILE + slot % QTILE, slot == p.n_slots - 1, local, v);
            }
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
        rec->base[0][lane] = make_uint2(rb.x, rb.y); rec->base[1][lane] = make_uint2(rb.z, rb.w);
        rec->mc[0][lane] = make_uint2(mc.x, mc.y); rec->mc[1][lane] = make_uint2(mc.z, mc.w);
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
    const int64_t per = GROUP * PRPL;
    const int64_t g0 = p.local_begin / per, g1 = (p.local_end + per - 1) / per;
    const int64_t blocks = (g1 - g0) * p.n_ptiles;
    if (blocks > 0x7fffffffLL || p.n_chunks > PIVOT_MAX_CHUNKS) return cudaErrorInvalidConfiguration;
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
