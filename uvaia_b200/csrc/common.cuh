// Shared layout definitions of the B200 n-best search engine (kernels + host).  See DESIGN.md sections 3-5.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace uvg {

// ---- packed database layout -------------------------------------------------------------------------------
// A site is a 4-bit IUPAC set (A=1,C=2,G=4,T=8; 0 = invalid: N,O,X,?,- and unknown bytes; biomcmc alignment.c:906-931
// with the skip rule of :870).  32 consecutive sites of ONE reference form a "column word" of four bit-planes
// (uint4: x=A, y=C, z=G, w=T; bit i = site 32*w+i).  32 consecutive references form a GROUP; the database is
//     uint4 db[group][Wp][32]      (lane = reference within group)
// so a warp reading column word w of its group issues one fully coalesced 512-byte request of 128-bit loads.
constexpr int GROUP = 32;
constexpr int WC = 8;             // column words per staged chunk
constexpr int QTILE = 64;         // queries per thread block (8 warps x QPW)
constexpr int QPW = 8;            // queries per warp (register tile)
constexpr int PAIR_THREADS = 256;
constexpr int STAGES = 4;
constexpr int RPL = 2;            // references per lane (a block scores 2 groups = 64 references)
constexpr int NBAR = 16;          // steps of the admission staircase kept per query

// One staged chunk of a query tile: WC column words of 64 queries.  Contiguous in global memory so that one
// cp.async.bulk (TMA 1-D bulk copy) moves it into shared memory.
struct __align__(16) QChunk {
    uint4    acgt[WC][QTILE];     // query bit-planes (zero outside the trimmed range / beyond nchar)
    uint2    vo[WC][QTILE];       // x = valid plane (a|c|g|t); y = default mode: one-hot plane, --acgt mode: idx site mask
    uint32_t slow[WC];            // bit k: warp k has an ambiguous (multi-base) query site in this word
    uint32_t rem[QTILE];          // max ACGT matches query can still collect in the chunks AFTER this one
};
static_assert(sizeof(QChunk) % 16 == 0, "bulk copy size must be a multiple of 16");

// Heap slot (32 B): the reference's q_item (src/min_heap.h:14-18) with name -> ordinal.
struct __align__(16) HeapItem {
    int32_t s[6];
    int32_t r;
    int32_t pad;
};

// Per-query threshold snapshot read by the pair-count kernel (written by the replay kernel once the heap is full).
// A pair c with D(c) = valid - acgt mismatches can be admitted at some later time only if the heap root r' at that time
// has D(r') >= D(c) (src/nearest.c:488-496: D < max_incompatible = D(root)+1) and c is strictly better than r'
// (min_heap.c:97).  Every future root is either an item kept NOW or was itself admitted under the same rule, so by
// induction:  c is ever admissible  =>  some item h kept now has D(h) >= D(c) and h < c.   Hence with
//   bar[b] = min ACGT count (first key) over kept items with D >= b << sh      (non-decreasing in b; bar[0] = root)
//   dmax   = 1 + max D over kept items
// a pair with D >= dmax, or with ACGT count < bar[D >> sh], or not strictly better than the root, is never admitted.
struct __align__(16) Threshold {
    int32_t s[6];                 // heap root (worst kept item)
    int32_t full;                 // 0: heap not full (every pair survives); else 1 + sh
    int32_t dmax;
    int32_t bar[NBAR];
};
static_assert(sizeof(Threshold) == 96, "Threshold layout");

struct PairParams {
    const uint4*    db;           // whole local database
    const uint32_t* refamb;       // [group][ambw] bit w%32 of word w/32: some reference of the group is ambiguous in column word w
    const QChunk*   qdata;        // [n_qtiles][n_chunks]
    const Threshold* thr;         // [n_qtiles*QTILE]
    int32_t Wp, n_chunks, ambw, n_qtiles, n_query;
    int64_t local_begin, local_end;   // local reference range of this launch
    int64_t global_minus_local;       // global ordinal = local index + this
    uint4*   scores;              // dense output: scores[q * score_stride + score_col0 + (local - local_begin)]
    int64_t  score_stride, score_col0;   //   default mode (acgt, text, partial, valid); --acgt mode (match, valid, mism_idx, 0)
    int32_t early_abort;
    unsigned long long* work;     // executed (warp, chunk) iterations (early abort makes this < the dense count)
};

__host__ __device__ inline int64_t db_word_index(int64_t group, int w, int lane, int Wp)
{
    return (group * Wp + w) * GROUP + lane;
}

}  // namespace uvg
