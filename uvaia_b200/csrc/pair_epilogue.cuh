// Epilogue shared by the two pair-count kernels (paircount.cu: general 3-plane kernel; pairpivot.cu: pivot kernel).
// Input: the plane counts of one (query slot, reference) pair -- M = sites where both hold the same plain base (as far as
// the planes of the calling kernel know), V = sites where both hold a plain base.  The epilogue adds the exact contribution
// of the sites that are outside the planes and stores the four counts of the default mode
//     acgt, text, partial, valid                      (biomcmc alignment.c:866-879)
// into the dense score tile of the rank that owns the query (peer-mapped memory over NVLink when sharded).
//   (b) sites where the REFERENCE holds an ambiguity code (sparse list of the reference): look the query's 4-bit code up;
//   (c) "fix" entries of the QUERY (sparse list of the slot): a site and the query's 4-bit set there.
//       - an ambiguity code (2..3 bases): the site is outside the query planes; if the reference holds a plain base it
//         is valid, and partial when the sets overlap (alignment.c:871-877; never acgt/text: the codes differ);
//       - pivot kernel only: a plain base that DIFFERS from the pivot sequence.  The site is inside the `valid` plane but
//         outside the match plane (the match plane only knows "equals the pivot"), so a reference holding the same base
//         adds one to acgt, text and partial.
#pragma once
#include "common.cuh"

namespace uvg {

__device__ __forceinline__ uint4 pair_epilogue_default(const PairParams& p, const uint32_t* __restrict__ fix_off, const uint32_t* __restrict__ fix_ent,
                                                       int slot, int64_t group, int lane, int64_t local, int32_t M, int32_t V,
                                                       const bool query_list = true)
{
    const uint32_t* dbw = reinterpret_cast<const uint32_t*>(p.db);
    int32_t dm = 0, dv = 0, dp = 0, dt = 0;
    const uint32_t re1 = p.ramb_off[local + 1];
    for (uint32_t e = p.ramb_off[local]; e < re1; ++e) {
        const uint32_t ent = __ldg(p.ramb_ent + e), site = ent & AMB_SITE_MASK, rc = ent >> 28;
        const uint32_t qc = (__ldg(p.qnib + (size_t)slot * p.qnib_stride + (site >> 3)) >> ((site & 7u) * 4u)) & 15u;
        dv += qc != 0u; dp += (qc & rc) != 0u; dt += qc == rc;
    }
    const uint32_t qe0 = query_list ? fix_off[slot] : 0u, qe1 = query_list ? fix_off[slot + 1] : 0u;   // the pivot kernel resolves (c) itself
    const size_t lane_base = db_unit_index(group, 0, 0, lane, p.Wp) * 4;
#pragma unroll 4
    for (uint32_t e = qe0; e < qe1; ++e) {
        const uint32_t ent = __ldg(fix_ent + e), site = ent & AMB_SITE_MASK, qc = ent >> 28;
        const uint32_t w = site >> 5, b = site & 31u;
        const size_t base = lane_base + (size_t)(w >> 2) * (NPLANE * GROUP * 4) + (w & 3u);
        const uint32_t rb = (__ldg(dbw + base) >> b) & 1u;
        const uint32_t x = (__ldg(dbw + base + GROUP * 4) >> b) & 1u, y = (__ldg(dbw + base + 2 * GROUP * 4) >> b) & 1u;
        const uint32_t hit = rb & (qc >> (x + 2u * y));
        const bool ambiguous = (qc & (qc - 1u)) != 0u;              // uniform over the warp
        if (ambiguous) { dv += rb; dp += hit; }
        else dm += hit;
    }
    M += dm;
    return make_uint4((uint32_t)M, (uint32_t)(M + dt), (uint32_t)(M + dp), (uint32_t)(V + dv));
}

// store the counts of one pair: query rows go to the owner's tile, the consensus row and the per-reference record to every rank
__device__ __forceinline__ void pair_store(const PairParams& p, int owner, int row, bool is_cons, int64_t local, uint4 v)
{
    const int64_t col = p.col0 + (local - p.local_begin);
    if (!is_cons) {
        p.out[owner][(size_t)row * p.row_stride + col] = v;
    } else {
        const uint4 meta = make_uint4((uint32_t)p.non_n[local], 0u, 0u, 0u);
        for (int r = 0; r < p.world; ++r) {
            p.out[r][(size_t)p.rows_cap * p.row_stride + col] = v;
            p.out[r][(size_t)(p.rows_cap + 1) * p.row_stride + col] = meta;
        }
    }
}

}  // namespace uvg
