/* uvaia_b200.h -- C ABI of the B200-native n-best database search (drop-in for the scoring/heap seam of the
 * reference program `uvaia`, src/nearest.c).  Plain C: pointers and sizes only, no torch / C++ types.
 *
 * The reference has no plugin or FFI interface: `uvaia` is a monolith whose main() calls
 *     queue_distance_to_consensus (query_t, queue_t, int ir)            src/nearest.c:428-433
 *     queue_update_min_heaps     (query_t, int iq, queue_t, int ir)    src/nearest.c:442-510
 *     heap_insert (heap_t, q_item) / heap_finalise_heap_qsort          src/min_heap.c:93-117, 149-158
 *     quick_count_sequence_non_N (char*, size_t)                       src/fastaseq.c:642-648
 *     create_query_indices (query_t)                                   src/fastaseq.c:732-777
 * inside two OpenMP loops (src/nearest.c:293-299).  The entry points below are what a C host bound to this
 * library calls INSTEAD of those functions; INTEGRATION.md shows the binding in nearest.c.
 *
 * Conventions: every function returns 0 on success, a negative UVG_E* code otherwise; uvg_last_error() gives the
 * message (the reference's biomcmc_error() prints and exits, lowlevel.c:39-52 -- the host decides here).
 * One host thread drives one context; a context owns one CUDA device and its streams.
 * There is NO CPU fallback: without a usable CUDA device uvg_create() fails with UVG_ENODEV.
 */
#ifndef UVAIA_B200_H
#define UVAIA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UVG_ABI_VERSION 3

enum { UVG_OK = 0, UVG_EINVAL = -1, UVG_ENODEV = -2, UVG_ENOMEM = -3, UVG_ECUDA = -4, UVG_ESTATE = -5, UVG_ENCCL = -6 };

enum { UVG_PAIR_AUTO = 0, UVG_PAIR_GENERAL = 1, UVG_PAIR_PIVOT = 2 };

typedef struct uvg_ctx uvg_ctx;

/* One kept neighbour = the reference's q_item (src/min_heap.h:14-18) with the strdup'ed name replaced by the
 * ordinal of the reference sequence among the ACCEPTED references, in file order (the host keeps the names). */
typedef struct uvg_item {
    int32_t score[6];   /* default: acgt, text, partial, valid, acgt_unique, valid_ref_sites  (nearest.c:499-501)
                           --acgt : match, valid, match_unique, valid_ref_sites, dist_consensus, dist_unique (:466-470) */
    int64_t ref;        /* global ordinal of the accepted reference */
} uvg_item;

typedef struct uvg_params {
    int32_t abi_version;    /* UVG_ABI_VERSION */
    int32_t nchar;          /* alignment columns (query->aln->nchar) */
    int32_t trim;           /* --trim, clamped by the library like fastaseq.c:708-712 */
    int32_t nbest;          /* --nbest; heap capacity is max(2, nbest) (min_heap.c:58) */
    int32_t acgt_mode;      /* --acgt */
    int32_t pool;           /* --pool: accepted references per batch of the reference program (nearest.c:81,250).  The reference's
                               consensus stage is truncated at a batch-wide limit (nearest.c:290-291,432), so its table can depend
                               on this value (hazard H1); the search reproduces the table of THIS pool size */
    int32_t device;         /* CUDA ordinal */
    int32_t first_round;    /* references in the first scoring round per rank (0 = default 1024) */
    int32_t max_round;      /* largest round per rank (0 = default 65536).  Single GPU: also capped so that a score tile stays
                               <= 3 GiB.  Sharded: never capped -- (first_round, max_round) is the row distribution of
                               uvg_shard_of; uvg_set_queries fails with UVG_ENOMEM if the tiles of a query set do not fit */
    int32_t early_abort;    /* 1 = warp-granular abort of pairs that cannot beat the k-th best (exact, perf only) */
    int32_t rank, world;    /* row sharding: this context holds tile `rank` of every round (world<=1: single GPU) */
    int32_t pair_kernel;    /* UVG_PAIR_AUTO (0): pivot form of the pair-count kernel when the mode and the query set allow it
                               (default mode, < 65,536 columns, no early abort, queries close to one pivot sequence), else the
                               general 3-plane kernel; UVG_PAIR_GENERAL / UVG_PAIR_PIVOT force one (PIVOT only skips the
                               closeness test).  Results are identical either way. */
    int32_t closed_form;    /* 0 (default): literal emulation of the reference's batches -- the table is the reference's for this
                               --pool and these file boundaries (uvg_mark_file_end); costs one replay launch per batch.
                               1: pool-independent closed form (= the reference at --pool 1) in one replay launch per round;
                               uvg_stats.n_h1_flags counts the pairs the reference might have scored differently (0 => the
                               table is the reference's for every --pool).  Sharded contexts (world > 1) emulate the batches
                               literally for query sets of one tile (at most 63 queries: rank 0 owns them all, truncated
                               references are read from the rank that holds them over peer memory) and run this form for
                               larger query sets. */
    int64_t db_capacity;    /* references to reserve HBM for (0 = grow on demand) */
    int32_t nominal_tail;   /* sharded: 1 = the last round keeps its nominal size (the first ranks take what is left), for hosts that
                               deal the rows out while they stream the files and do not know the total: uvg_shard_of with
                               n_total = -1 is their distribution.  0 = the rounds taper off towards the end of the database and share it evenly (needs the total up front) */
    int32_t reserved2;      /* must be 0 */
} uvg_params;

typedef struct uvg_query_info {
    int32_t n_query, nchar, trim;           /* trim after clamping */
    int32_t n_idx, n_idx_m, n_idx_c;        /* sizes of the three site sets (stderr line fastaseq.c:774) */
} uvg_query_info;

typedef struct uvg_stats {
    int64_t n_refs;             /* references scored so far */
    int64_t n_rounds;
    int64_t n_admitted;         /* successful heap insertions */
    int64_t n_h1_flags;         /* closed form: pairs the reference MAY have scored differently for this --pool (hazard H1);
                                   literal emulation: pairs that reached the heap with truncated consensus counts */
    int64_t pair_kernel_launches, other_kernel_launches;
    double  pair_kernel_ms;     /* CUDA-event time of the pair-count launches of the last uvg_search */
    double  replay_ms;          /* ... of bucket+replay */
    double  search_ms;          /* ... whole uvg_search */
    double  encode_ms;          /* ... of the encoder launches of the last uvg_push_refs* */
    int64_t warp_chunks_executed; /* (warp, chunk) iterations the pair-count kernel ran in the last uvg_search: x 8 words x 16 pairs x 32 lanes
                                     = executed pair-words (early abort makes this smaller than queries x refs x words) */
    int64_t max_mismatch;       /* "Highest number of ACGT mismatches" (nearest.c:324): max over queries of max_incompatible-1 */
    int64_t pivot_fix_entries;  /* pivot form of the pair-count kernel in use: fix entries over all query slots; -1: general kernel */
    double  set_queries_host_ms, set_queries_device_ms;   /* last uvg_set_queries: host preparation / allocations + uploads */
} uvg_stats;

/* ---- life cycle ---- */
int  uvg_create(const uvg_params* p, uvg_ctx** out);
void uvg_destroy(uvg_ctx* ctx);
const char* uvg_last_error(const uvg_ctx* ctx);        /* ctx may be NULL: error of the last failed uvg_create */
int  uvg_abi_version(void);

/* ---- queries: replaces create_query_indices + the per-query strings read by queue_update_min_heaps ----
 * text = n_query rows of nchar bytes, ALREADY filtered (utils.c:10-48) and ordered (fastaseq.c:779-795) by the
 * host; case-insensitive.  Builds consensus / idx / idx_m / idx_c, the trimmed bit-planes of every query and the
 * consensus pseudo-query on the device, and empties all heaps (max_incompatible = nchar, nearest.c:375,387). */
int uvg_set_queries(uvg_ctx* ctx, const uint8_t* text, int32_t n_query, uvg_query_info* info);
/* -k / --keep_resolved: replaces exclude_redundant_query_sequences (fastaseq.c:797-841; called at nearest.c:219-224 and
 * ball.c:190).  text = the filtered and ordered query rows as for uvg_set_queries; keep[i] = 1 if query i stays, *n_kept
 * (may be NULL) = how many.  keep_more_resolved: of two equivalent queries drop the one with fewer usable sites (what
 * `uvaia -k` asks for); 0 drops the more resolved one (uvaiaball's default).  The caller compacts the rows and calls
 * uvg_set_queries with the survivors (the reference rebuilds its indices the same way, nearest.c:222).  nchar, trim and
 * acgt_mode as in uvg_params (trim is clamped the same way).  Runs on host threads and needs no context: the walk over
 * the pairs is sequential in the reference and short (first differing site ends a pair). */
int uvg_prune_queries(const uint8_t* text, int32_t n_query, int32_t nchar, int32_t trim, int32_t acgt_mode,
                      int32_t keep_more_resolved, uint8_t* keep, int32_t* n_kept);

/* ---- references: replaces quick_count_sequence_non_N + the ref filter (nearest.c:263-270) + the char2bit lookups.
 * text = n rows of nchar bytes (host memory; pinned memory makes the copy asynchronous).  Each row gets
 * non_n_out[i] = #bytes not in "NnXx-?Oo." over the whole row and accepted_out[i] = (non_n >= min_non_n);
 * accepted rows are appended, in order, to the device-resident packed database (4-bit IUPAC sets as three
 * bit-planes + a sparse list of ambiguity codes, 32 references interleaved per 16-byte unit: DESIGN.md 3).  Either output pointer may be NULL.
 * *n_accepted (may be NULL) receives the number appended. */
int uvg_push_refs(uvg_ctx* ctx, const uint8_t* text, int64_t n, int32_t min_non_n,
                  int32_t* non_n_out, uint8_t* accepted_out, int64_t* n_accepted);
/* Same, with text already in device memory (bench "inputs resident in HBM"; outputs stay host pointers). */
int uvg_push_refs_device(uvg_ctx* ctx, const uint8_t* d_text, int64_t n, int32_t min_non_n,
                         int32_t* non_n_out, uint8_t* accepted_out, int64_t* n_accepted);
/* Forget the packed references scored so far but keep heaps and ordinals (streaming hosts: push, search, drop). */
int uvg_drop_refs(uvg_ctx* ctx);
/* Forget every reference and restart ordinals at 0 (heaps are emptied too). */
int uvg_clear_refs(uvg_ctx* ctx);
int64_t uvg_db_count(const uvg_ctx* ctx);              /* accepted references resident on this rank */
/* The references accepted so far end a -r file: the reference's batch ends with its file and the next file starts a new
 * batch (nearest.c:245-286).  n_accepted_so_far = accepted references of all files up to here, or -1 = what this context
 * has been given (single GPU).  Call it after the last uvg_push_refs of each file, before the next push. */
int uvg_mark_file_end(uvg_ctx* ctx, int64_t n_accepted_so_far);

/* ---- packed reference cache (no counterpart in the reference, which re-reads the FASTA text on every run) ----
 * The packed references leave and enter the device as they are: planes of whole groups of 32 references, non_n and the
 * ambiguity lists, behind a 64-byte header that names nchar and the mode they were encoded for.  A host keeps the blob
 * (a file, pinned memory) and the names; the next search skips reading, parsing and encoding the text and moves 0.375
 * bytes per site to the device instead of 1.  `first` must be a multiple of 32 and [first, first+n) resident;
 * uvg_import_packed appends like uvg_push_refs (accepted rows only: the -A filter was applied when the blob was made) and
 * needs the resident count to be a multiple of 32.  Pinned blobs (uvg_host_alloc_pinned) copy at full PCIe speed. */
int uvg_packed_bytes(uvg_ctx* ctx, int64_t first, int64_t n, size_t* bytes);
int uvg_export_packed(uvg_ctx* ctx, int64_t first, int64_t n, void* blob, size_t capacity);
int uvg_import_packed(uvg_ctx* ctx, const void* blob, size_t bytes, int64_t* n_refs);

/* ---- search: replaces the two OpenMP loops nearest.c:293-299 (consensus stage, per-query heap updates) ----
 * Scores every resident reference not scored yet against every query, in file order, in rounds.  Bit-exact with
 * the reference's sequential admission (order-dependent max_incompatible rule and first-come tie-break). */
int uvg_search(uvg_ctx* ctx);
/* Pipelined hosts: uvg_search_async enqueues the rounds for the references pushed so far and returns at once, so that
 * the next uvg_push_refs (host->device copy + encoder, on their own streams) runs under them; uvg_wait blocks until
 * everything enqueued has finished (uvg_search = uvg_search_async + uvg_wait).  Results and statistics wait implicitly. */
int uvg_search_async(uvg_ctx* ctx);
int uvg_wait(uvg_ctx* ctx);
/* Empty all heaps and mark every resident reference as not scored (repeat a search over the same database). */
int uvg_reset_search(uvg_ctx* ctx);

/* ---- results: replaces heap_finalise_heap_qsort + the reads in save_distance_table (nearest.c:513-547) ----
 * out must hold max(2,nbest) items; rows come in the reference's output order (heap array after the slot0<->slot n
 * swap, then qsort with compare_q_item: glibc's qsort is a stable merge sort, so rows that compare equal keep their heap
 * order -- heaps of up to 2,048 slots are rank-sorted on the device at the end of the search, larger ones by libc qsort on
 * the host); *n = rows (hazard H2: never the NULL-name rows). */
int uvg_get_heap(uvg_ctx* ctx, int32_t iq, uvg_item* out, int32_t* n);
/* Every query's table in one call (one device->host copy of the sorted tables): out holds n_query x max(2,nbest)
 * items, the rows of query iq start at out + iq * max(2,nbest); n[iq] = rows of query iq. */
int uvg_get_all_heaps(uvg_ctx* ctx, uvg_item* out, int32_t* n);
/* flags[i] = 1 if accepted reference ordinal first+i ever entered any heap (is_best, nearest.c:303-314).  Sharded contexts:
 * COLLECTIVE -- the first call after a search combines the flags of all ranks (NCCL all-reduce), so every rank calls it. */
int uvg_get_ever_best(uvg_ctx* ctx, int64_t first, int64_t n, uint8_t* flags);
int uvg_get_stats(uvg_ctx* ctx, uvg_stats* out);

/* ---- row sharding over several GPUs (one context per GPU / process) ----
 * Round j covers global ordinals [S_j, S_j + world*T_j) and rank g holds [S_j + g*T_j, S_j + (g+1)*T_j), with
 * T_1 = first_round, T_{j+1} = min(4*T_j, max_round); once every rank has less than two nominal rounds left, a round takes
 * half of what is left per rank (rounded up to 64, not below 4,096; the rest when that is at most 4,096), so that the tail
 * is shared evenly and the last replay, which no pair counting hides, is short.  uvg_shard_of maps a global ordinal to
 * (rank, local index); n_total = accepted references in all (the value given to uvg_set_global_count). */
int uvg_shard_of(int64_t ordinal, int64_t n_total, int32_t world, int32_t first_round, int32_t max_round, int32_t* rank, int64_t* local);
int uvg_nccl_unique_id(void* out128);                          /* 128 bytes, made by rank 0, sent to all by the host */
int uvg_nccl_init(uvg_ctx* ctx, const void* id128);            /* collective over all `world` contexts */
/* In sharded mode every rank declares how many accepted references exist in total before uvg_search. */
int uvg_set_global_count(uvg_ctx* ctx, int64_t n_total);
/* Pipelined sharded hosts: every rank pushes its rows round by round and calls uvg_search_async after each round; only
 * rounds that end at or below `n_resident` (accepted references, counted over all ranks in file order) are searched by
 * that call -- the next round's rows may still be crossing PCIe on some rank.  -1 = no limit (the default).  Every rank
 * passes the same values in the same order (the rounds are collective). */
int uvg_set_search_limit(uvg_ctx* ctx, int64_t n_resident);

/* ---- radius search: the reference program `uvaiaball` (src/ball.c) ----
 * Which reference sequences are within `dist` differences of SOME query: replaces seq_ball_against_query_structure
 * (src/fastaseq.c:660-696) inside the OpenMP loop ball.c:240-244 and the reference filter ball.c:201,216-222.  The ball
 * distance compares upper-cased BYTES over the site sets of create_query_indices (two usable bytes that differ count one;
 * --acgt: only A, C, G, T are usable), so this path works on the sequence text, not on the packed planes, and has its own
 * small context.  Host flow (ball.c:172-194): filter the queries (utils.c:10-48), order them (fastaseq.c:779-795), ALWAYS
 * prune them with uvg_prune_queries (keep_more_resolved = the -k flag), then uvg_ball_set_queries. */
typedef struct uvg_ball uvg_ball;
typedef struct uvg_ball_params {
    int32_t abi_version;    /* UVG_ABI_VERSION */
    int32_t nchar;
    int32_t trim;           /* --trim, clamped like fastaseq.c:708-712 */
    int32_t dist;           /* -d / --distance (the program's default is 5); clamped to (nchar - 2 trim) / 10, fastaseq.c:713-715 */
    int32_t acgt_mode;      /* -x / --acgt */
    int32_t device;
} uvg_ball_params;
int  uvg_ball_create(const uvg_ball_params* p, uvg_ball** out);
void uvg_ball_destroy(uvg_ball* b);
const char* uvg_ball_last_error(const uvg_ball* b);   /* b may be NULL: error of the last failed uvg_ball_create */
int32_t uvg_ball_radius(const uvg_ball* b);           /* dist after clamping */
/* text = n_query rows of nchar bytes, filtered, ordered and pruned by the host; builds consensus / idx / idx_m / idx_c */
int uvg_ball_set_queries(uvg_ball* b, const uint8_t* text, int32_t n_query, uvg_query_info* info);
/* text = n reference rows of nchar bytes (host memory, any case).  keep[i] = 0: too ambiguous (valid_ref_sites < min_non_n;
 * uvaiaball uses min_non_n = (int)(nchar * A), ball.c:201), 1: outside the ball, 2: within `dist` of some query (the rows
 * uvaiaball writes to <prefix>.aln.xz, in file order).  non_n_out (may be NULL) = valid_ref_sites of every row. */
int uvg_ball_push_refs(uvg_ball* b, const uint8_t* text, int64_t n, int32_t min_non_n, uint8_t* keep, int32_t* non_n_out);
/* The same classification for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO counterpart in the reference.  d_text: n rows of nchar
 * bytes in DEVICE memory; d_keep [n] and d_non_n [n] (may be NULL) are DEVICE buffers of the caller and receive what
 * uvg_ball_push_refs returns in keep / non_n_out.  The call allocates nothing (no cudaMalloc / cudaFree): the scratch between its
 * two kernels is a file-scope device array of 8 bytes x UVG_BALL_NOALLOC_MAX_ROWS = 512 KiB that every GPU which loads the library
 * holds until the process exits.  n > UVG_BALL_NOALLOC_MAX_ROWS is rejected (UVG_EINVAL, message in uvg_ball_last_error): split the
 * rows over several calls.  All calls on a device share the array: calls that overlap on the GPU (contexts of the same device used
 * from several threads) race; a call returns when its kernels have finished.  uvg_ball_set_queries (which does allocate) comes first. */
#define UVG_BALL_NOALLOC_MAX_ROWS 65536
int uvg_ball_classify_device_noalloc(uvg_ball* b, const uint8_t* d_text, int64_t n, int32_t min_non_n, uint8_t* d_keep, int32_t* d_non_n);

/* ---- bench / test utilities (not part of the reference seam) ---- */
/* Device generator of SARS-CoV-2-shaped alignment rows (same model as uvaia_b200/synth.py; counter-based RNG keyed
 * by (seed, ordinal)).  clades = n_clades rows of nchar bytes (host).  Writes n rows for ordinals first.. into d_text. */
int uvg_synth_rows_device(uvg_ctx* ctx, const uint8_t* clades, int32_t n_clades, uint64_t seed, int64_t first,
                          int64_t n, double n_frac_mean, double ambiguous_frac, uint8_t* d_text);
void* uvg_device_alloc(uvg_ctx* ctx, size_t bytes);
void  uvg_device_free(uvg_ctx* ctx, void* p);
int   uvg_memcpy_d2h(uvg_ctx* ctx, void* dst, const void* d_src, size_t bytes);
int   uvg_memcpy_h2d(uvg_ctx* ctx, void* d_dst, const void* src, size_t bytes);
void* uvg_host_alloc_pinned(size_t bytes);
void  uvg_host_free_pinned(void* p);
/* Raw per-pair counts of queries x resident references [first, first+n): out[(iq*n + i)*4 + k], k = the four
 * counters of the mode (default: acgt,text,partial,valid; --acgt: match,valid,mism_idx,0).  Unit-test hook. */
int uvg_debug_pair_counts(uvg_ctx* ctx, int64_t first, int64_t n, int32_t* out);
/* Same for the query rows [q_first, q_first+q_count) only: out[((iq-q_first)*n + i)*4 + k]. */
int uvg_debug_pair_counts_rows(uvg_ctx* ctx, int32_t q_first, int32_t q_count, int64_t first, int64_t n, int32_t* out);
/* Per-reference values of resident references [first, first+n): valid_ref_sites (non_n), consensus-stage ACGT matches
 * c0 and mismatches ci (queue_distance_to_consensus, nearest.c:428-433, untruncated).  Pointers may be NULL. */
int uvg_debug_ref_meta(uvg_ctx* ctx, int64_t first, int64_t n, int32_t* non_n, int32_t* c0, int32_t* ci);
/* Unpack resident reference `local` back to its 4-bit codes (nchar bytes, values 0..15).  Unit-test hook. */
int uvg_debug_unpack_ref(uvg_ctx* ctx, int64_t local, uint8_t* codes_out);

#ifdef __cplusplus
}
#endif
#endif /* UVAIA_B200_H */
