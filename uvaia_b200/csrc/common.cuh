// Shared layout definitions of the B200 n-best search engine (kernels + host).  See DESIGN.md sections 3-5.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace uvg {

// ---- packed database layout -------------------------------------------------------------------------------
// A site is a 4-bit IUPAC set (A=1,C=2,G=4,T=8; 0 = invalid: N,O,X,?,- and unknown bytes; biomcmc alignment.c:906-931
// with the skip rule of :870).  The database keeps, per reference and 32-site column word, THREE bit-planes
//     plane 0  "base"  bit i set <=> site 32*w+i is a plain base (one-hot code A, C, G or T)
//     plane 1  "b0"    bit i set <=> the plain base is C or T
//     plane 2  "b1"    bit i set <=> the plain base is G or T
// so two plain bases are equal iff both b0 and b1 agree (3 LOP3 per pair and word give the valid AND the match mask).
// Sites holding an ambiguity code (M R W S Y K V H D B: 2-3 bases) are rare (~1.7 per SARS-CoV-2 genome); they are
// kept out of the planes (treated as invalid there) and listed per reference as (site, code) entries; the pair-count
// kernel adds their exact contribution in its epilogue (paircount.cu).
// Four consecutive column words of one plane form one uint4 "unit"; 32 consecutive references form a GROUP; the
// database is
//     uint4 db[group][Wp/4][NPLANE][32]      (lane = reference within group)
// so that (a) a lane reads 128 sites of one plane of its reference with one conflict-free 128-bit shared-memory load
// and (b) one chunk (WC = 8 column words) of a group is a single contiguous 3 KB run for a TMA bulk copy.
constexpr int GROUP = 32;
constexpr int WC = 8;             // column words per staged chunk
constexpr int W4 = WC / 4;        // uint4 units per chunk
constexpr int NPLANE = 3;
constexpr int QTILE = 64;         // query slots per thread block (8 warps x QPW)
constexpr int QPW = 8;            // queries per warp (register tile)
constexpr int PAIR_THREADS = 256;
constexpr int STAGES = 6;
constexpr int RPL = 2;            // references per lane (a block scores 2 groups = 64 references)
constexpr int MAX_RANKS = 8;      // GPUs of one box
constexpr uint32_t AMB_SITE_MASK = 0x0fffffffu;       // ambiguity entry = site | code << 28

// One staged chunk of a query tile: WC column words of 64 query slots.  Contiguous in global memory so that one
// cp.async.bulk (TMA 1-D bulk copy) moves it into shared memory.
struct __align__(16) QChunk {
    uint4    pl[W4][QTILE][4];    // [unit][slot][plane]: 0 base (plain base, inside the trimmed range), 1 b0, 2 b1,
                                  //                      3 --acgt mode: segregating-site (idx) mask
    uint32_t rem[QTILE];          // plain bases of the query in the chunks AFTER this one (upper bound of matches to come)
};
static_assert(sizeof(QChunk) % 16 == 0, "bulk copy size must be a multiple of 16");

// ---- pivot form (pairpivot.cu) ----------------------------------------------------------------------------------
// All queries of a search are close to one "pivot" sequence (per site: the plain base most queries hold).  With
//     mc  = sites where the REFERENCE holds the pivot base         (derived database plane, one pass per query set)
//     qeq = sites where the QUERY holds the pivot base
// the match word of a pair is mc & qeq and the valid word r.base & q.base: two ANDs instead of three LOP3 with five
// inputs.  The kernel works on the COMPLEMENT nq = no plain base here (every plain base counts as the pivot's; one that
// differs is settled by its fix entry): valid = popc(r.base) - sum popc(r.base & nq), match' = popc(r.mc) - sum popc(r.mc &
// nq), with per-reference totals as first terms and sums that only get contributions where the query has an N, a gap or a
// code; every counted word gets its own POPC (2 AND + 2 POPC + 2 IMAD per counted pair-word).
// Derived database: per (PRPL = 4 groups = the 128 references of a thread block, chunk of WC column words) one contiguous
// 25 KB record
//     planes   [group of the block][unit][ uint2 base[half][32] | uint2 mc[half][32] ]       8 KB  \ staged by TMA bulk copies
//     counts   [group of the block][unit][32 references] uint32: popc(mc unit) | popc(base unit) << 16   1 KB /
//     colmasks [group of the block][unit][4 words][4 bases A,C,G,T][32 sites] uint32        16 KB: read by fix entries only
// (planes half-major: words 0,1 of the 32 references, then words 2,3 -- a warp's 8-byte loads of one half are consecutive).
// The counts serve the units in which a query holds no plain base at all (inside an N run, trimmed away): there
// popc(r.base & nq) and popc(r.mc & nq) are the reference's own unit totals -- one add instead of 8 AND + 8 POPC.
#ifndef PIVOT_STAGES
#define PIVOT_STAGES 7
#endif
constexpr int STAGES2 = PIVOT_STAGES;
constexpr int UNIT2_PLANES = 2 * GROUP * 16;                         // base + mc of one (group, unit): 1 KB
constexpr int UNIT2_MASKS = 4 * 4 * GROUP * 4;                       // column masks of one (group, unit): 2 KB
#ifndef PIVOT_RPL
#define PIVOT_RPL 4
#endif
constexpr int PRPL = PIVOT_RPL;                                      // reference groups per thread block of the pivot form = references per lane:
                                                                     // what a warp does per chunk besides counting (wait, head, walk, release) is
                                                                     // paid once per PRPL x 32 references
constexpr int CHUNK2_PLANES = PRPL * W4 * UNIT2_PLANES;              // 8 KB
constexpr int CHUNK2_COUNTS = PRPL * W4 * GROUP * 4;                 // 1 KB
constexpr int CHUNK2_STAGED = CHUNK2_PLANES + CHUNK2_COUNTS;         // what a ring stage holds of a chunk record
constexpr int CHUNK2_BYTES = CHUNK2_STAGED + PRPL * W4 * UNIT2_MASKS; // 25 KB per (PRPL groups, chunk)
constexpr int PIVOT_QW = 4;                                          // query slots per counting warp of k_pairpivot
constexpr int PIVOT_WARPS = 16;                                      // warp 0 only fills the staging ring, warps 1..15 count
constexpr int PIVOT_SLOTS = (PIVOT_WARPS - 1) * PIVOT_QW;            // 60 query slots per thread block ("pivot tile"): consecutive slots,
                                                                     // independent of the 64-slot tiles that own the score rows
// Query side of the pivot form: per (pivot tile of 60 slots, chunk) one variable-length record of a byte stream (PairParams::qstream,
// offsets in PairParams::qoff), staged by one TMA bulk copy of exactly its size:
//     QWarpHead head[16 warps]                          what each warp has to do in the chunk
//     uint2     item[n]                                 per half (64 sites) that is counted from the planes: words 2h, 2h+1 of the
//                                                       COMPLEMENT plane of the slot -- nq: no plain base here, or trimmed /
//                                                       padding.  (A plain base that differs from the pivot's is settled by its
//                                                       fix entry, so "not the pivot's base" is the same plane.)  The items of a
//                                                       warp are contiguous, in the order the kernel walks them (unit, slot, half);
//                                                       a record is padded to a multiple of 16 bytes (bulk copy)
// Only ~16 % of the (slot, half) pairs of a SARS-CoV-2 query set need an item: the record averages ~0.8 KB where the dense
// planes of the tile took 4 KB of every ring stage.
struct __align__(8) QWarpHead {
    uint16_t half;        // bit u * 8 + j * 2 + h: words 2h, 2h+1 of unit u of the warp's slot j are counted from the planes (an item)
    uint8_t  fulln;       // bit u * 4 + j: slot j holds no plain base in unit u (inside an N run, trimmed): the counts table answers
    uint8_t  n_items;
    uint16_t item0;       // index of the warp's first item in the record
    uint16_t pad;
};
constexpr int QHEAD_BYTES = PIVOT_WARPS * (int)sizeof(QWarpHead);    // 128
constexpr int QITEMS_MAX = QTILE * W4 * 2;                           // every half of every slot: 256
constexpr int QITEM_BYTES = 8;                                       // words 2h, 2h+1 of the slot's complement plane
constexpr int QREC_MAX = QHEAD_BYTES + QITEMS_MAX * QITEM_BYTES;     // 2176: what a ring stage reserves
// A query site holding a plain base that differs from the pivot is outside the planes' match count; a reference holding
// the same base there is one more match.  These "fix" entries (~the SNPs of the query, a few dozen per genome) are resolved
// after the main loop against the column masks of the derived database: entry = chunk << 16 | unit << 9 | word << 7 |
// base (A,C,G,T) << 5 | site in word; per slot a range of PairParams::snp_ent.
static_assert(W4 == 2 && PIVOT_WARPS == 16 && sizeof(QWarpHead) == 8 && PIVOT_SLOTS <= QTILE, "QWarpHead bit layout");
__host__ __device__ inline size_t db2_chunk_bytes(int64_t group, int unit, int Wp)        // record of the chunk that holds (group, unit)
{
    return ((size_t)(group / PRPL) * (size_t)(Wp / WC) + (size_t)(unit / W4)) * CHUNK2_BYTES;
}
__host__ __device__ inline size_t db2_planes_bytes(int64_t group, int unit, int Wp)       // base plane of (group, unit); mc follows at + GROUP * 16
{
    return db2_chunk_bytes(group, unit, Wp) + (size_t)((group % PRPL) * W4 + unit % W4) * UNIT2_PLANES;
}
__host__ __device__ inline size_t db2_masks_bytes(int64_t group, int unit, int Wp)        // column masks of (group, unit)
{
    return db2_chunk_bytes(group, unit, Wp) + CHUNK2_STAGED + (size_t)((group % PRPL) * W4 + unit % W4) * UNIT2_MASKS;
}
__host__ __device__ inline size_t db2_counts_bytes(int64_t group, int unit, int Wp)       // unit totals of the 32 references of (group, unit)
{
    return db2_chunk_bytes(group, unit, Wp) + CHUNK2_PLANES + (size_t)((group % PRPL) * W4 + unit % W4) * GROUP * 4;
}

// Heap slot (32 B): the reference's q_item (src/min_heap.h:14-18) with name -> ordinal.
struct __align__(16) HeapItem {
    int32_t s[6];
    int32_t r;
    int32_t pad;
};

// Per-query bounds for the warp-granular early abort of the pair-count kernel, written by the replay kernel once the
// heap is full.  A pair with D = valid - acgt mismatches is admitted only if D < max_incompatible = D(root) + 1
// (src/nearest.c:488-496, 506-508) where the root is a kept item; every kept item was itself admitted under that
// rule, so by induction D(c) < dmax := 1 + max D over the items kept NOW holds for every pair admitted from now on.
// And an admitted pair is strictly better than the root (min_heap.c:97), whose first key never decreases, so its ACGT
// count is >= root_s0.  Both bounds only tighten with time: a stale snapshot is always safe.
struct __align__(16) Threshold {
    int32_t full;                 // 0: heap not full (nothing can be discarded)
    int32_t dmax;
    int32_t root_s0;
    int32_t pad;
};

struct PairParams {
    const uint4*     db;          // whole local database
    const QChunk*    qdata;       // [n_qtiles][n_chunks]
    const Threshold* thr;         // indexed like the score rows: owner * rows_cap + row
    const uint32_t*  qnib;        // 4-bit codes of every query slot (trimmed), qnib_stride words per slot, 8 sites per word
    const uint32_t*  qamb_off;    // [n_slots + 1] ambiguity entries of the query slots
    const uint32_t*  qamb_ent;
    // pivot form (null when the general kernel runs)
    const uint8_t*   db2;         // derived database: 3 KB records (base, mc, one-hot planes)
    const int32_t*   rb_cnt;      // per local reference: popcount of its base plane (all words)
    const int32_t*   mc_cnt;      //                      popcount of its mc plane
    const uint8_t*   qstream;     // records of every (query tile, chunk), see QWarpHead
    const uint32_t*  qoff;        // [n_ptiles][n_chunks + 1] byte offsets of the records (multiples of 16)
    const uint8_t*   wslot;       // [n_ptiles][PIVOT_WARPS][PIVOT_QW]: which slots of the pivot tile each warp counts (balanced by work; 0xff: none)
    const uint32_t*  snp_off;     // [n_slots + 1] fix entries of every slot: plain bases that differ from the pivot
    const uint32_t*  snp_ent;
    const uint32_t*  ramb_off;    // [local refs + 1] ambiguity entries of the references
    const uint32_t*  ramb_ent;
    const int32_t*   non_n;       // per local reference
    int32_t qnib_stride;
    int32_t Wp, n_chunks, n_qtiles;
    int32_t n_ptiles;             // pivot tiles: ceil(n_slots / PIVOT_SLOTS)
    int32_t n_slots;              // real queries + 1: the last slot is the consensus pseudo-query (nearest.c:428-433)
    int64_t local_begin, local_end;   // local reference range of this launch
    // dense score tile(s): slot q of query tile qt goes to rank (qt % world), row (qt / world) * QTILE + q % QTILE; the
    // consensus slot goes to row rows_cap of EVERY rank and the per-reference meta record (non_n) to row rows_cap + 1.
    uint4*   out[MAX_RANKS];      // tile base on every rank (peer-mapped pointers when world > 1)
    int64_t  row_stride;          // uint4 elements between rows
    int64_t  col0;                // column of local_begin in the destination tiles
    int32_t  world, rows_cap;
    int32_t  early_abort;
    unsigned long long* work;     // executed (pair, 4-word unit) iterations per lane: x 4 words x 32 lanes = pair-words (early abort / skipped units: < dense)
};

__host__ __device__ inline size_t db_unit_index(int64_t group, int unit, int plane, int lane, int Wp)
{
    return (((size_t)group * (size_t)(Wp / 4) + (size_t)unit) * NPLANE + (size_t)plane) * GROUP + (size_t)lane;
}

}  // namespace uvg
