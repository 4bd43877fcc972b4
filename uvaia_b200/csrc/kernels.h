// Host-callable launchers of the CUDA kernels (one translation unit each: encode.cu, paircount.cu, replay.cu).
#pragma once
#include "common.cuh"

namespace uvg {

// ---- encode.cu (K1 + synthetic rows) ----
void build_luts(uint8_t lut[3][256]);
cudaError_t upload_luts();
cudaError_t launch_count_non_n(const uint8_t* d_text, int64_t n, int nchar, int min_non_n, int acgt_mode, int32_t* d_non_n, uint8_t* d_acc,
                               int32_t* d_namb, cudaStream_t st);
// d_totals[0] = accepted rows, d_totals[1] = their ambiguity entries; fills ramb_off[dst_base .. dst_base + accepted]
cudaError_t launch_scan_accept(const uint8_t* d_acc, const int32_t* d_namb, int64_t n, int64_t dst_base, uint32_t amb_base, int64_t* d_dst,
                               uint32_t* ramb_off, int64_t* d_totals, cudaStream_t st);
cudaError_t launch_encode(const uint8_t* d_text, int64_t n, int nchar, int Wp, int acgt_mode, const int64_t* d_dst,
                          const int32_t* d_non_n_src, uint4* db, const uint32_t* ramb_off, uint32_t* ramb_ent, int32_t* non_n_db, cudaStream_t st);
cudaError_t launch_unpack_ref(const uint4* db, const uint32_t* ramb_off, const uint32_t* ramb_ent, int64_t local, int nchar, int Wp,
                              uint8_t* d_out, cudaStream_t st);
cudaError_t launch_synth_rows(const uint8_t* d_clades, int n_clades, int nchar, uint64_t seed, int64_t first, int64_t n,
                              double n_frac_mean, double ambiguous_frac, uint8_t* d_out, cudaStream_t st);

// ---- paircount.cu (K2) ----
cudaError_t launch_paircount(const PairParams& p, int acgt_mode, cudaStream_t st);
cudaError_t paircount_setup();       // opt in to the dynamic shared memory size

// ---- pairpivot.cu (K2 in pivot form: default mode, < 65,536 columns, no early abort) ----
cudaError_t launch_pairpivot(const PairParams& p, cudaStream_t st);
cudaError_t pairpivot_setup();
// derived planes (base, mc) of the resident references [begin, end) from the packed database and the pivot planes
// rb_cnt / mc_cnt (zeroed by the caller for [begin, end)): popcount of the base plane / of the mc plane per reference
cudaError_t launch_derive_pivot(const uint4* db, uint8_t* db2, const uint4* pivot, int32_t* rb_cnt, int32_t* mc_cnt, int64_t begin, int64_t end,
                                int Wp, cudaStream_t st);

// ---- replay.cu (K3) ----
struct ReplayParams {
    const uint4* tile;            // this rank's score tile of the round: row = owned query row, column = position in the round
    int64_t row_stride;           // uint4 elements between rows
    // a row is nseg segments of seg_stride positions (one per source rank, in rank = file order); segment g holds
    // seg_cnt[g] references whose global ordinals start at seg_global[g].  Single GPU: one segment.
    int32_t nseg, seg_stride;
    int32_t seg_cnt[MAX_RANKS];
    int64_t seg_global[MAX_RANKS];
    int32_t rows_cap;             // row rows_cap = consensus counts, row rows_cap + 1 = per-reference record (non_n)
    HeapItem* heaps; int32_t* heap_n; int32_t* max_inc;               // per owned row: [rows_cap][K+1], [rows_cap], [rows_cap]
    int32_t* batch_id; int32_t* max_inc_batch;                        // H1 bookkeeping per owned row: first ordinal of the reference's batch of the
                                                                      //     row's last candidate (-1: none), max_incompatible at the start of that batch
    Threshold* thr;                                                   // [rows_cap] thresholds of the owned rows (output)
    uint8_t* ever_best;                                               // per GLOBAL ordinal
    unsigned long long* counters;                                     // [0]=admitted, [1]=H1 flags
    int32_t n_query;                                                  // real queries (all ranks)
    int32_t world, rank;                                              // row -> query: ((row / QTILE) * world + rank) * QTILE + row % QTILE
    int32_t K, acgt_mode, pool;
    // the reference's batches that touch this launch: sorted first ordinals, bounds[0] <= first ordinal of the launch, the last
    // entry >= its end (--pool accepted references per batch, a batch also ends with its file: nearest.c:245-286)
    const int64_t* bounds; int32_t n_bounds;
    // speculative pass of a literal round (see engine.cu run_round): admissions are marked in ever_round[ordinal - ord_round0]
    // instead of ever_best; literal launches return at once when *spec_flags == 0
    uint8_t* ever_round; int64_t ord_round0;
    const unsigned long long* spec_flags;
    // the launch replays the global ordinals [ord_begin, ord_end) of the round (literal mode: one launch per batch piece)
    int64_t ord_begin, ord_end;
    // ---- literal emulation of the reference's batch-wide consensus limit (hazard H1; single GPU) ----
    int32_t literal;              // 0: pool-independent closed form + H1 flags; 1: the reference's result for this --pool
    int32_t batch_start;          // the launch starts a batch: the limit is the maximum the previous launch accumulated
    const int32_t* b_prev;        // max over the queries of max_incompatible after the previous launch
    const int32_t* b_carry_in;    // limit of the batch in progress (written by the previous launch)
    int32_t* b_carry_out;         // ... for the next launch
    int32_t* b_acc;               // this launch's accumulator (atomicMax of every query's final max_incompatible)
    int32_t* b_zero;              // the accumulator of the NEXT launch (zeroed here)
    // packed references of every segment's rank (peer-mapped when sharded), for the prefix recount of truncated pairs:
    // the reference at position i of segment g is row seg_local0 + i of that rank's database
    const uint4* seg_db[MAX_RANKS]; const uint32_t* seg_ramb_off[MAX_RANKS]; const uint32_t* seg_ramb_ent[MAX_RANKS];
    int64_t seg_local0;
    const QChunk* qcons; int32_t cons_qi; const uint32_t* cons_nib;
    const uint32_t* camb; int32_t n_camb;
    int32_t Wp;
};
cudaError_t launch_replay(const ReplayParams& p, cudaStream_t st);
// speculation of a literal round: commit the speculative pass if it raised no flag (spec[1] == 0) ...
cudaError_t launch_spec_commit(const unsigned long long* spec, const uint8_t* ever_round, int64_t n, int64_t ord_round0, uint8_t* ever_best,
                               unsigned long long* counters, cudaStream_t st);
// ... else put the state of the round's start back and prepare the batch-wide limits of the first literal launch
struct SpecRestore {
    const unsigned long long* spec;
    const uint4* snap_own; uint4* own; int64_t own_vec;               // heaps | heap_n | max_inc block, in 16-byte units
    const int32_t* snap_batch_id; int32_t* batch_id; const int32_t* snap_mib; int32_t* mib; int32_t rows;
    const int32_t* snap_max_inc;                                      // inside snap_own
    int32_t n_query;                                                  // rows of real queries (single GPU: row = query)
    int32_t open_batch;                                               // first ordinal of the batch in progress at the round's start (-1: a batch starts there)
    int32_t* b_prev; int32_t* b_carry; int32_t* b_acc;                // what the first literal launch reads / accumulates into
};
cudaError_t launch_spec_restore(const SpecRestore& r, cudaStream_t st);
// final tables of `rows` heaps (slot swap + stable descending sort), as uvg_item rows [rows][K]
cudaError_t launch_finalise(const HeapItem* heaps, const int32_t* heap_n, int K, int rows, void* out, cudaStream_t st);
cudaError_t replay_setup(int K);

}  // namespace uvg
