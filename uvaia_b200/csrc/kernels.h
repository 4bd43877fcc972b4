// Host-callable launchers of the CUDA kernels (one translation unit each: encode.cu, paircount.cu, replay.cu).
#pragma once
#include "common.cuh"

namespace uvg {

// ---- encode.cu (K1 + synthetic rows) ----
void build_luts(uint8_t lut[3][256]);
cudaError_t upload_luts();
cudaError_t launch_count_non_n(const uint8_t* d_text, int64_t n, int nchar, int min_non_n, int32_t* d_non_n, uint8_t* d_acc, cudaStream_t st);
cudaError_t launch_scan_accept(const uint8_t* d_acc, int64_t n, int64_t dst_base, int64_t* d_dst, int64_t* d_nacc, cudaStream_t st);
cudaError_t launch_encode(const uint8_t* d_text, int64_t n, int nchar, int W, int Wp, int ambw, int acgt_mode, const int64_t* d_dst,
                          const int32_t* d_non_n_src, uint4* db, uint32_t* refamb, int32_t* non_n_db, cudaStream_t st);
cudaError_t launch_unpack_ref(const uint4* db, int64_t local, int nchar, int Wp, uint8_t* d_out, cudaStream_t st);
cudaError_t launch_synth_rows(const uint8_t* d_clades, int n_clades, int nchar, uint64_t seed, int64_t first, int64_t n,
                              double n_frac_mean, double ambiguous_frac, uint8_t* d_out, cudaStream_t st);

// ---- paircount.cu (K2) ----
// consensus stage: per local reference c0 (matches) and ci (mismatches) against the consensus pseudo-query
cudaError_t launch_consensus(const uint4* db, const uint4* cons_acgt, const uint2* cons_vo, int Wp, int acgt_mode,
                             int64_t local_begin, int64_t local_end, int32_t* c0, int32_t* ci, cudaStream_t st);
cudaError_t launch_paircount(const PairParams& p, int acgt_mode, cudaStream_t st);
cudaError_t paircount_setup();       // opt in to the dynamic shared memory size

// ---- replay.cu (K3) ----
constexpr int MAX_SEG = 16;
struct ReplayParams {
    const uint4* scores;          // dense counts of the round: row = query (q_first + block), column = position in the round
    int64_t row_stride;           // uint4 elements between rows
    // a row is nseg segments of seg_stride positions (one per source rank); segment g holds seg_cnt[g] references whose
    // global ordinals start at seg_global[g].  Single GPU: one segment.
    int32_t nseg, seg_stride;
    int32_t seg_cnt[MAX_SEG];
    int64_t seg_global[MAX_SEG];
    const int32_t* non_n; const int32_t* c0; const int32_t* ci;      // per position of the round
    HeapItem* heaps; int32_t* heap_n; int32_t* max_inc;               // [n_query_total][K+1], [..], [..]
    int32_t* batch_id; int32_t* max_inc_batch;                        // H1 bookkeeping per query
    Threshold* thr;                                                   // [n_qtiles*QTILE]
    uint8_t* ever_best;                                               // per GLOBAL ordinal
    unsigned long long* counters;                                     // [0]=admitted, [1]=H1 flags
    int32_t q_first, n_query;                                         // queries replayed by this launch (rows)
    int32_t K, acgt_mode, pool;
};
cudaError_t launch_replay(const ReplayParams& p, cudaStream_t st);
cudaError_t replay_setup(int K);

}  // namespace uvg
