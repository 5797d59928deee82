/* kvmini_b200.h - C ABI of libkvmini_b200.so (sm_100a).
 *
 * Drop-in boundary for the run-analysis hot path of Siryoos/kserve-vllm-mini.  The
 * reference has no FFI of its own: its boundary is the
 * CLI + three importable function sets, all Python.  The Python host in
 * kserve-vllm-mini_b200/ keeps those names and calls the entry points below through
 * ctypes; each entry point cites the reference code whose arithmetic it replaces.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer owned by the caller (torch tensors are only
 *    buffers) unless the parameter name ends in _host;
 *  - `stream` is a cudaStream_t passed as void*; calls are stream-ordered and do not
 *    synchronise unless stated;
 *  - no entry point allocates: scratch comes from the caller (`workspace`, sized by
 *    the matching *_workspace_bytes) - workspaces need 256-byte alignment;
 *  - return value: KVM_OK or a negative KVM_ERR_*; kvm_last_error() gives the text
 *    (thread-local);
 *  - fp64 columns use NaN for "missing"; int64 columns are plain counts / codes.
 */
#ifndef KVMINI_B200_H
#define KVMINI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KVM_ABI_VERSION 1

#define KVM_OK 0
#define KVM_ERR_INVALID (-1)   /* bad argument */
#define KVM_ERR_CUDA (-2)      /* CUDA runtime error; text in kvm_last_error() */
#define KVM_ERR_WORKSPACE (-3) /* workspace_bytes too small */

const char* kvm_last_error(void);
int kvm_abi_version(void);
/* Number of SMs of the current device (grids are sized as multiples of it). */
int kvm_sm_count(void);

/* ------------------------------------------------------------------ request log */

/* Columnar request log.  Replaces the list-of-dicts of analyze.py:34-56. */
typedef struct kvm_request_log {
  const double* start_ms;
  const double* latency_ms;
  const double* ttfb_ms;            /* NaN = missing */
  const double* tllt_ms;            /* NaN = missing */
  const int64_t* status;            /* success == 200 (analyze.py:127 `str(int(status))=="200"`) */
  const int64_t* completion_tokens; /* 0 where the reference's `or 0` applies */
  const int64_t* total_tokens;
  int64_t n;
} kvm_request_log;

/* Row flags written by kvm_log_scan (one byte per row). */
#define KVM_F_SUCCESS 0x01 /* status == 200 */
#define KVM_F_COLD 0x02    /* analyze.py:402-419 */
#define KVM_F_TOKENS 0x04  /* completion_tokens > 0 */
#define KVM_F_GEN 0x08     /* row contributes to generation / per-token sets (analyze.py:143) */

/* The seven disjoint "base sets" of (value, row-predicate) the analysis ranks.
 * Every percentile multiset of analyze.py is a union of these:
 *   latency all = LAT_COLD|LAT_WARM, ttfb all = TTFB_COLD|TTFB_WARM, ... */
enum {
  KVM_SET_LAT_COLD = 0, /* latency_ms | success & cold & !nan   (analyze.py:437) */
  KVM_SET_LAT_WARM = 1,
  KVM_SET_TTFB_COLD = 2, /* ttfb_ms | success & cold & !nan     (analyze.py:438-442) */
  KVM_SET_TTFB_WARM = 3,
  KVM_SET_TLLT = 4,   /* tllt_ms | success & !nan               (analyze.py:131,150) */
  KVM_SET_GEN = 5,    /* tllt-ttfb | success & both !nan & tokens>0 (analyze.py:143-145) */
  KVM_SET_PERTOK = 6, /* (tllt-ttfb)/tokens, same rows            (analyze.py:146) */
  KVM_LOG_NSETS = 7
};
/* value columns of the request log (digit parameters are per column) */
enum { KVM_COL_LAT = 0, KVM_COL_TTFB = 1, KVM_COL_TLLT = 2, KVM_COL_GEN = 3, KVM_COL_PERTOK = 4,
       KVM_LOG_NCOLS = 5 };

#define KVM_MAX_SETS 8
#define KVM_MAX_COLS 8

/* Counts / ranges of the base sets of one select problem. Keys are the order-preserving
 * uint64 image of the fp64 value (sign-flip encoding). */
typedef struct kvm_set_info {
  int64_t count[KVM_MAX_SETS];
  uint64_t kmin[KVM_MAX_COLS]; /* per value column, over all of its sets */
  uint64_t kmax[KVM_MAX_COLS];
  int32_t set_col[KVM_MAX_SETS]; /* column of each base set */
  int32_t nsets;
  int32_t ncols;
} kvm_set_info;

/* Result of the fused first pass over the log.
 * Replaces analyze.py:183-189 (window), :402-419 (cold/warm), :496-514 (counts, token
 * sums), cost_estimator.py:117-124,293-308 and the NaN filters of :68,:89. */
typedef struct kvm_log_stats {
  int64_t n_total;
  int64_t n_success;
  int64_t n_cold;         /* rows classified cold (any status) */
  int64_t n_cold_success; /* cold & success */
  int64_t sum_completion_tokens; /* over success rows */
  int64_t sum_total_tokens;      /* over success rows */
  int64_t sum_total_tokens_cold; /* over success & cold rows (warm = total - cold) */
  int64_t n_gen_rows;            /* rows passing analyze.py:143 (incl. NaN differences) */
  double min_start_ms;           /* over ALL rows (analyze.py:185) */
  double max_end_ms;             /* max(start_ms+latency_ms) over ALL rows (analyze.py:187) */
  double ttfb_sum_hi;            /* double-double sum of ttfb over success & !nan (analyze.py:581) */
  double ttfb_sum_lo;
  double lat_sum_hi; /* double-double sum of latency over success rows (cost_calculator.py:48) */
  double lat_sum_lo;
  kvm_set_info sets; /* the seven base sets / five columns above */
} kvm_log_stats;

size_t kvm_log_scan_workspace_bytes(void);

/* K1+K7.  One pass over all seven columns.  `cold_times` = sorted cold-start times in
 * seconds (may be NULL when n_cold_times == 0); a row is cold iff some c has
 * c <= start_ms/1000 <= c + cold_window_s.  When `cold_mask` is non-NULL it is a per-row
 * 0/1 byte array that REPLACES that classification (cost_estimator.py:108-111 reads
 * `is_cold_start` from requests_classified.csv).  Writes flags[n] (may be NULL) and *stats. */
int kvm_log_scan(const kvm_request_log* log_host, const double* cold_times, int32_t n_cold_times,
                 double cold_window_s, const uint8_t* cold_mask, uint8_t* flags, kvm_log_stats* stats,
                 void* workspace, size_t workspace_bytes, void* stream);

/* Multi-GPU (log sharded by time range): shard statistics are combined
 * on the device with two all-reduces issued by the host between these calls -
 * sums[kvm_log_stats_pack_sums()] with SUM, mins[kvm_log_stats_pack_mins()] with MIN (maxima
 * travel complemented).  The double-double sums are NOT combined here: they are gathered and
 * added in rank order by the host (deterministic).  No reference counterpart (the reference
 * is single-process); replaces analyze.py:183-189,496-514 across shards. */
int32_t kvm_log_stats_pack_sums(void);
int32_t kvm_log_stats_pack_mins(void);
int kvm_log_stats_pack(const kvm_log_stats* stats, int64_t* sums, int64_t* mins, void* stream);
int kvm_log_stats_unpack(kvm_log_stats* stats, const int64_t* sums, const int64_t* mins, void* stream);
/* One-collective form: pack a shard's statistics (sums, mins, the two double-double sums) into
 * kvm_log_stats_pack_words() int64 words, all-gather them, and combine the `world` records
 * locally - integer fields exactly, the double-double sums in rank order. */
int32_t kvm_log_stats_pack_words(void);
int kvm_log_stats_pack_words_into(const kvm_log_stats* stats, int64_t* packed, void* stream);
int kvm_log_stats_combine(kvm_log_stats* stats, const int64_t* gathered, int32_t world, void* stream);

/* ------------------------------------------------------------------ exact selection */

/* How a query turns (n, p) into a 0-based rank. */
enum {
  KVM_RANK_LERP_LO = 0,     /* floor(clamp(p,0,1)*(n-1))   analyze.py:75-76 */
  KVM_RANK_LERP_HI = 1,     /* ceil (clamp(p,0,1)*(n-1))   analyze.py:77   */
  KVM_RANK_ABSOLUTE = 2,    /* rank given                                   */
  KVM_RANK_UPPER_MEDIAN = 3,/* n/2  `sorted(v)[len(v)//2]`  collector.py:333 */
  KVM_RANK_FLOOR_PN = 4,    /* int(p*n)      runners/backends/tgi/invoke.sh:208-210 */
  KVM_RANK_FLOOR_PNM1 = 5   /* int(p*(n-1))  scripts/fairness_dual_tenant.py:59-65 */
};

#define KVM_MAX_QUERIES 48

typedef struct kvm_query {
  uint32_t set_mask; /* union of base sets (bit b = set b); all of one column */
  int32_t rank_kind;
  double p;
  int64_t rank; /* KVM_RANK_ABSOLUTE only */
} kvm_query;

#define KVM_Q_OK 0
#define KVM_Q_EMPTY 1        /* multiset empty: value = NaN */
#define KVM_Q_OUT_OF_RANGE 2 /* rank >= n (reference would raise IndexError) */
#define KVM_Q_FAILED 3       /* the select as a whole failed (status word != KVM_OK): value = NaN */

typedef struct kvm_query_result {
  double value; /* the order statistic */
  int64_t n;    /* size of the multiset (NaNs excluded) */
  int64_t rank; /* resolved 0-based rank */
  int32_t status;
  int32_t pad;
} kvm_query_result;

/* Value histograms produced as a by-product of a select problem (K5 folded into K2): up to
 * KVM_MAX_HISTS histograms of num_bins (<= KVM_HIST_MAX_BINS) uniform bins each, over one value
 * column and a union of its base sets; bin index = min(nb-1, int((v-min)/w)), w=(max-min)/nb,
 * exactly as analyze.py:104-111 (min/max = the column's range in kvm_set_info). */
#define KVM_MAX_HISTS 4
#define KVM_HIST_MAX_BINS 64
typedef struct kvm_hist_spec {
  int32_t nhist;
  int32_t num_bins;
  int32_t col[KVM_MAX_HISTS];
  uint32_t set_mask[KVM_MAX_HISTS];
} kvm_hist_spec;

/* Optional by-products of kvm_log_select (all device pointers; zero / NULL = not wanted). */
typedef struct kvm_log_extras {
  int32_t num_bins;        /* > 0: the four histograms of analyze.py:149-152 (ttfb, tllt, generation, per-token) */
  int32_t pad;
  int64_t* hist_counts;    /* [4][num_bins], in a buffer of KVM_MAX_HISTS*KVM_HIST_MAX_BINS int64 */
  double series_window_ms; /* > 0: extension A12 window series, binned in the histogram pass */
  int64_t series_max_bins;
  int64_t* series;         /* [3][series_max_bins]: requests, successes, completion tokens */
  int64_t* series_dropped; /* [1]: rows whose window index >= series_max_bins */
  uint64_t* digit_records; /* [n] scratch, optional: the histogram pass stores each row's five
                              fine-bin indices (12 bits each) so the compact pass streams 9 B/row
                              instead of re-reading and re-keying four columns */
} kvm_log_extras;

/* Workspace for a select problem whose candidate arena holds `arena_capacity` keys.
 * kvm_select_default_arena(n) is a size that fits unless the data is pathologically
 * concentrated; on overflow the select returns KVM_ERR_ARENA in the status word and
 * the caller retries with capacity = total number of elements. */
size_t kvm_select_workspace_bytes(int64_t arena_capacity);
int64_t kvm_select_default_arena(int64_t n);
#define KVM_ERR_ARENA (-4) /* candidate arena too small: retry with a larger one */

/* Fused multi-rank exact select over the request log (K2+K6, with K5 and K8 riding on the
 * same two passes when `extras` asks for them).  Replaces the 25 `percentile()` calls of one
 * analyze.py run (analyze.py:59-81, call sites :159-176, :447-451, :574-583) and its four
 * compute_histograms calls (:149-152).  `stats` is the DEVICE struct kvm_log_scan wrote,
 * `flags` its flag bytes; `extras` is a host struct (may be NULL).  Stream-ordered;
 * results[nq] land in device memory.  After the stream is synchronised,
 * kvm_select_status(workspace) tells whether the arena overflowed. */
int kvm_log_select(const kvm_request_log* log_host, const uint8_t* flags,
                   const kvm_log_stats* stats, const kvm_query* queries_host, int32_t nq,
                   kvm_query_result* results, const kvm_log_extras* extras_host, void* workspace,
                   size_t workspace_bytes, int64_t arena_capacity, void* stream);

/* ---- one-pass form (csrc/kvm_fused.cu).  kvm_log_scan and the select's histogram pass in ONE
 * kernel: 56 B/row read once, flags + digit records written, statistics, the seven digit
 * histograms and the window series produced together.  The digit parameters come from a 512-row
 * sample drawn identically by every CTA; the two end bins catch what lies outside the sampled
 * range and are resolved exactly by the select (plan_preset marks them, resolve spans their
 * buckets with the true range), so the results are exact for any input.  Needs every column
 * 16-byte aligned (128-bit loads), flags 2-byte and extras->digit_records (required) 16-byte
 * aligned, n >= 1: otherwise KVM_ERR_UNSUPPORTED and the caller uses kvm_log_scan + kvm_log_select.
 * The window series is binned relative to the minimum start_ms of the sample + first + last row;
 * that reference is stored (as double bits) in extras->series_dropped[2] (the buffer needs three
 * int64 here): when it differs from stats->min_start_ms (a log that is not time-ordered), the
 * series must be recomputed with kvm_log_series.
 * Stage order: fused_pass -> plan_preset -> resolve -> {refine_hist -> refine}* -> compact -> finish;
 * kvm_log_analyze_fused chains them.  Replaces analyze.py:59-189,402-514 like scan + select. */
size_t kvm_log_analyze_fused_scan_workspace_bytes(void);
int kvm_log_fused_pass(const kvm_request_log* log_host, const double* cold_times, int32_t n_cold_times,
                       double cold_window_s, uint8_t* flags, kvm_log_stats* stats,
                       const kvm_log_extras* extras_host, void* scan_workspace, size_t scan_workspace_bytes,
                       void* select_workspace, size_t select_workspace_bytes, int64_t arena_capacity,
                       void* stream);
/* Sharded logs (one rank per GPU): the digit histograms of the ranks are summed, so every rank must
 * bin with the SAME parameters.  kvm_log_fused_sample writes a shard's sample record
 * (kvm_log_fused_preset_words() u64 words), the ranks exchange the records,
 * kvm_log_fused_preset_combine folds them (records `stride_words` apart) into the agreed hull, and
 * kvm_log_fused_pass_preset runs the pass with it (any alignment, empty shards included; the series
 * reference is the minimum over all ranks' samples).  No reference counterpart (one process). */
int32_t kvm_log_fused_preset_words(void);
int kvm_log_fused_sample(const kvm_request_log* log_host, uint64_t* record, void* stream);
int kvm_log_fused_preset_combine(const uint64_t* records, int32_t world, int64_t stride_words, uint64_t* preset,
                                 void* stream);
int kvm_log_fused_pass_preset(const kvm_request_log* log_host, const double* cold_times, int32_t n_cold_times,
                              double cold_window_s, uint8_t* flags, kvm_log_stats* stats,
                              const kvm_log_extras* extras_host, void* scan_workspace, size_t scan_workspace_bytes,
                              void* select_workspace, size_t select_workspace_bytes, int64_t arena_capacity,
                              const uint64_t* preset, void* stream);
int kvm_select_plan_preset(const kvm_set_info* info, const kvm_query* queries_host, int32_t nq,
                           const kvm_hist_spec* hists_host, void* workspace, size_t workspace_bytes,
                           int64_t arena_capacity, int32_t refine_rounds, int64_t refine_min_keys,
                           void* stream);
int kvm_log_analyze_fused(const kvm_request_log* log_host, const double* cold_times, int32_t n_cold_times,
                          double cold_window_s, uint8_t* flags, kvm_log_stats* stats,
                          const kvm_query* queries_host, int32_t nq, kvm_query_result* results,
                          const kvm_log_extras* extras_host, void* scan_workspace, size_t scan_workspace_bytes,
                          void* select_workspace, size_t select_workspace_bytes, int64_t arena_capacity,
                          int32_t refine_rounds, int64_t refine_min_keys, void* stream);

/* Generic single-column variant.  Set membership of element i:
 *   mode 0: every element is in set 0
 *   mode 1: set_ids[i] (uint8; 255 = excluded)
 *   mode 2: status[i] == 200 -> set 0, else excluded   (the 16 B/record percentile path)
 * NaN values are always excluded.  Replaces analyze.py:59-81 for one list. */
enum { KVM_MEMBER_ALL = 0, KVM_MEMBER_SET_IDS = 1, KVM_MEMBER_STATUS = 2 };

int kvm_column_select(const double* values, int64_t n, int32_t member_mode, const void* member,
                      int32_t nsets, const kvm_query* queries_host, int32_t nq,
                      kvm_query_result* results, void* workspace, size_t workspace_bytes,
                      int64_t arena_capacity, void* stream);

/* Fast form of kvm_column_select for a few quantiles of ONE multiset (member mode ALL or STATUS,
 * up to 8 queries of rank kinds that carry a p, 2^18 <= n < 2^31): the 16 B/record percentile
 * path of SURVEY.md section 8d in one pass and an eighth.  A 4096-element stratified sample
 * estimates the key range; ONE pass over the data histograms 13-bit digits relative to it and
 * writes them as 2-byte records; a pass over those records fetches the ~0.1 % of values whose
 * digit holds a requested rank; one CTA per query ranks them.  Exact whenever it answers;
 * otherwise (a rank falls outside the sampled range, heavy duplicates, too few members) the
 * status word - the first int32 of the workspace, stream-ordered - is KVM_ERR_UNSUPPORTED and
 * the caller runs kvm_column_select.  Arguments outside the envelope are refused on the host
 * with KVM_ERR_UNSUPPORTED.  Replaces analyze.py:59-81. */
int kvm_column_select_sampled(const double* values, int64_t n, int32_t member_mode, const void* member,
                              const kvm_query* queries_host, int32_t nq, kvm_query_result* results,
                              void* workspace, size_t workspace_bytes, void* stream);
size_t kvm_column_select_sampled_workspace_bytes(int64_t n);
int64_t kvm_column_select_sampled_min_n(void);

/* No counterpart in the reference.  kvm_column_select for callers that cannot give the
 * library a workspace because they must not (or cannot) allocate device memory - e.g. a
 * latency monitor sharing the GPU with an inference server that has claimed the whole HBM
 * pool, where a cudaMalloc would fail or eat into the server's KV cache.  The scratch is a
 * fixed array inside the library image (about 9.5 MB per GPU that loads the library, held
 * until exit), which bounds the input: n <= kvm_column_select_static_max_n() (1,048,576
 * values), member modes KVM_MEMBER_ALL / KVM_MEMBER_STATUS only; larger or multi-set inputs
 * are rejected with KVM_ERR_INVALID and a message.  Inputs and `results[nq]` are still the
 * caller's device buffers.  Calls cudaMalloc / cudaFree never.
 *  - The first call on each device looks up the scratch's device address (that device must
 *    be current); make that first call before starting any stream capture.  The lookup is
 *    guarded by a host mutex, so two threads making a device's first call together is safe;
 *    with lazy module loading the array itself materialises on a device at first use there.
 *  - All calls on a device share the one scratch: calls whose GPU work can overlap (two
 *    streams, two threads) race.  Serialise them or use kvm_column_select.
 *  - The scratch keeps the previous call's contents; the call clears what it needs.
 *  - There is no workspace, hence no status word to read: only bit 0 of a query's set_mask
 *    is looked at (one set), which bounds the candidates of all queries together by n - the
 *    size of the scratch arena - so the arena cannot overflow; should a select fail all the
 *    same, every result carries status KVM_Q_FAILED and value NaN. */
int kvm_column_select_static(const double* values, int64_t n, int32_t member_mode, const void* member,
                             const kvm_query* queries_host, int32_t nq, kvm_query_result* results,
                             void* stream);
int64_t kvm_column_select_static_max_n(void);

/* Reads back the status word of the last select run in `workspace` (synchronises the
 * stream): KVM_OK or KVM_ERR_ARENA. */
int kvm_select_status(const void* workspace, void* stream);

/* ---- staged form of the same select, for multi-GPU runs where the host all-reduces
 * between the stages.  Stage order:
 *   column_stats -> [allreduce set_info: counts sum, kmin min, kmax max] -> plan ->
 *   hist -> [allreduce hist: sum] -> resolve -> {refine_hist -> [allreduce hist2: sum] ->
 *   refine}* -> compact -> [allgather arena+counts, allreduce edge counters: sum] -> finish
 */
int kvm_column_stats(const double* values, int64_t n, int32_t member_mode, const void* member,
                     int32_t nsets, kvm_set_info* info, void* workspace, size_t workspace_bytes,
                     void* stream);
size_t kvm_column_stats_workspace_bytes(void);
/* device addresses inside a select workspace, so the host can hand them to NCCL */
uint64_t* kvm_select_hist_ptr(void* workspace);   /* [KVM_MAX_SETS][kvm_select_bins()] u64 */
uint64_t* kvm_select_edge_ptr(void* workspace);   /* [KVM_MAX_HISTS][KVM_HIST_MAX_BINS] u64 edge counters */
int32_t kvm_select_bins(void);
uint64_t* kvm_select_arena_ptr(void* workspace);  /* candidate keys */
int64_t* kvm_select_bucket_counts_ptr(void* workspace); /* [KVM_MAX_QUERIES] local counts */
int64_t* kvm_select_bucket_offsets_ptr(void* workspace); /* [KVM_MAX_QUERIES+1] arena offsets */
kvm_set_info* kvm_select_info_ptr(void* workspace);      /* set info block inside the workspace */

/* one small record per rank after the compact pass: [bucket cursors | edge counters | dropped];
 * pack -> all-gather -> unpack (cursors per rank into counts_all[world][KVM_MAX_QUERIES], the
 * rest summed over ranks back into the workspace / *dropped) */
int32_t kvm_select_post_words(void);
int kvm_select_post_pack(const void* workspace, const int64_t* dropped, int64_t* record, void* stream);
int kvm_select_post_unpack(void* workspace, const int64_t* gathered, int32_t world, int64_t* counts_all,
                           int64_t* dropped, void* stream);

/* sets the status word to KVM_ERR_ARENA when the buckets hold more than `capacity_keys` keys in
 * total (the size of the slots the candidates were exchanged through), so that the finish does
 * not read past them; the caller retries with larger slots */
int kvm_select_require_capacity(void* workspace, int64_t capacity_keys, void* stream);

int kvm_select_plan(const kvm_set_info* info, const kvm_query* queries_host, int32_t nq,
                    const kvm_hist_spec* hists_host, void* workspace, size_t workspace_bytes,
                    int64_t arena_capacity, void* stream);
int kvm_column_select_hist(const double* values, int64_t n, int32_t member_mode, const void* member,
                           void* workspace, void* stream);
int kvm_log_select_hist(const kvm_request_log* log_host, const uint8_t* flags,
                        const kvm_log_stats* stats, const kvm_log_extras* extras_host,
                        void* workspace, void* stream);
int kvm_select_resolve(void* workspace, void* stream);

/* Refine rounds (large logs).  A level-1 bucket holds ~n/2048 keys of its column; at 1e8..1e9
 * rows that is too many to exchange between ranks and to rank inside one CTA.  A refine round
 * narrows every bucket to the 1/2048 sub-range that holds the requested rank:
 *   refine_hist  pass over the digit records: per bucket, histogram of the next 11 key bits
 *                (values are fetched only for the ~3% of rows that hit a bucket)
 *   [allreduce kvm_select_hist2_ptr: sum, kvm_select_hist2_words() u64]
 *   refine       walk each query's bucket histogram, replace the buckets by the sub-ranges
 * kvm_select_plan_refine tells the device how many rounds the host will launch
 * (`refine_rounds`, 0..4) and the total bucket population (`refine_min_keys`) above which a
 * round does work; below it both kernels return at once, so the host may launch them
 * unconditionally (all ranks take the same decision from the all-reduced histograms).
 * The compact pass after a refine round must be the record-driven one (digit_records != NULL).
 * kvm_log_select runs one round by itself when n >= KVM_REFINE_MIN_ROWS and digit records
 * are given.  No counterpart in the reference (it sorts each list). */
#define KVM_REFINE_MIN_ROWS (1ll << 25)
#define KVM_REFINE_MIN_KEYS (1ll << 19)
int kvm_select_plan_refine(const kvm_set_info* info, const kvm_query* queries_host, int32_t nq,
                           const kvm_hist_spec* hists_host, void* workspace, size_t workspace_bytes,
                           int64_t arena_capacity, int32_t refine_rounds, int64_t refine_min_keys,
                           void* stream);
int kvm_log_select_refine_hist(const kvm_request_log* log_host, const uint8_t* flags,
                               const uint64_t* digit_records, void* workspace, void* stream);
int kvm_select_refine(void* workspace, void* stream);
uint64_t* kvm_select_hist2_ptr(void* workspace); /* [KVM_MAX_QUERIES][kvm_select_bins()] u64 */
int64_t kvm_select_hist2_words(void);

int kvm_column_select_compact(const double* values, int64_t n, int32_t member_mode,
                              const void* member, void* workspace, void* stream);
int kvm_log_select_compact(const kvm_request_log* log_host, const uint8_t* flags,
                           const uint64_t* digit_records /* from the hist pass, or NULL */,
                           void* workspace, void* stream);
/* Finish from `nsrc` arenas (own + gathered peers), each laid out with the same bucket
 * offsets; src_counts[s*KVM_MAX_QUERIES + b] = keys source s holds for bucket b. */
int kvm_select_finish(void* workspace, const uint64_t* const* arenas_host, int32_t nsrc,
                      const int64_t* src_counts, kvm_query_result* results,
                      int64_t* hist_counts /* [KVM_MAX_HISTS*KVM_HIST_MAX_BINS] capacity, histogram h at
                                              h*num_bins; zeroed here; NULL = no histograms */,
                      void* stream);

/* ---- peer-memory exchange (multi-GPU, one NVSwitch node).  No counterpart in the reference.
 * Every rank owns an "exchange buffer" of the same layout that the host has mapped into all
 * ranks (`peers[r]` = device address of rank r's buffer in THIS process; peers[rank] is the
 * own one).  The sharded analysis replaces its collectives by these stream-ordered kernels:
 *   put / put_candidates  stores into every rank's buffer over NVLink
 *   add_u64               system-scope 64-bit atomic adds of the non-zero words (histograms)
 *   barrier               per-pair arrival counters in the flag block at `flags_offset`
 *                         (kvm_peer_flag_bytes() bytes, zeroed once by the host before first
 *                         use): releases everything this rank stored before, then waits for
 *                         the same from every rank.  The epoch is kept on the device, so the
 *                         call can be captured in a CUDA graph and replayed.  A peer that does
 *                         not arrive within the time-out (60 s; env KVM_PEER_TIMEOUT_S or
 *                         kvm_peer_set_timeout_s; the value is read when the call is ENQUEUED, so a
 *                         captured graph keeps its own) makes the waiting rank write 1 + that
 *                         peer's rank into word [KVM_PEER_MAX_WORLD+1] of EVERY rank's flag block
 *                         instead of hanging the GPU.  The word is sticky: later barriers on the
 *                         buffer return at once, and every rank - the late one too - must treat the
 *                         step's results as void (the host checks the word before it reports).
 * Buffers that peers accumulate into must be zeroed by their owner BEFORE the owner's previous
 * barrier in stream order (the peers' adds only start after they have passed that barrier). */
#define KVM_PEER_MAX_WORLD 16
int64_t kvm_peer_flag_bytes(void);
int kvm_peer_put(const void* src, int64_t bytes, int64_t dst_offset, void* const* peers_host,
                 int32_t world, int32_t rank, void* stream);
int kvm_peer_add_u64(const uint64_t* src, int64_t words, int64_t dst_offset, void* const* peers_host,
                     int32_t world, int32_t rank, void* stream);
int kvm_peer_barrier(int64_t flags_offset, void* const* peers_host, int32_t world, int32_t rank,
                     void* stream);
/* sets the barrier time-out for later calls (seconds > 0; <= 0 only queries); returns the old one */
double kvm_peer_set_timeout_s(double seconds);
/* after the compact pass: the filled part of every candidate bucket of `workspace` goes to
 * offset slot_offset (+ the bucket's arena offset) of every rank's buffer; more than
 * slot_capacity candidates (global bucket sizes, so all ranks agree) -> status KVM_ERR_ARENA */
int kvm_peer_put_candidates(void* workspace, int64_t slot_offset, int64_t slot_capacity,
                            void* const* peers_host, int32_t world, int32_t rank, void* stream);

/* ------------------------------------------------------------------ nearest element (F2)
 * Index of the member element whose value is closest to `target`; ties -> lowest index; NaN
 * values never win; index -1 / distance NaN when no member qualifies.  Replaces the exemplar
 * search of report_generator.py:437-454 (`sorted((abs(latency - p95), trace_id))[0]` over the
 * status-200 rows): one 16 B/row pass instead of a CSV re-parse and a sort of tuples. */
typedef struct {
  int64_t index;
  double distance; /* |values[index] - target| */
} kvm_nearest_result;
size_t kvm_column_nearest_workspace_bytes(void);
int kvm_column_nearest(const double* values, int64_t n, int32_t member_mode, const void* member,
                       double target, kvm_nearest_result* result /* device */, void* workspace,
                       size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------ CSV ingest (host only)
 * SURVEY.md section 8f F1.  Columnar fast path for `read_requests_csv` (analyze.py:34-56) and the
 * requests_classified.csv writer (analyze.py:597-604); no device work.  A cell that is a plain
 * decimal literal is parsed here (std::from_chars: correctly rounded, same double as float());
 * "" / "NaN" are reported as such; every other spelling is reported IRREGULAR and left to the
 * caller's float().  Files this parser does not model carry flags and the caller uses the
 * csv module instead. */
typedef struct kvm_csv kvm_csv;
enum { KVM_CSV_CELL_NUMBER = 0, KVM_CSV_CELL_EMPTY = 1, KVM_CSV_CELL_NAN_TEXT = 2, KVM_CSV_CELL_IRREGULAR = 3 };
#define KVM_CSV_RAGGED 1      /* a row has more or fewer fields than the header */
#define KVM_CSV_LONE_CR 2     /* '\r' used as a row terminator / inside an unquoted field */
#define KVM_CSV_BARE_QUOTE 4  /* a '"' that csv.writer would not have produced */
#define KVM_CSV_NEEDS_CSV_MODULE (KVM_CSV_RAGGED | KVM_CSV_LONE_CR | KVM_CSV_BARE_QUOTE)
#define KVM_ERR_UNSUPPORTED (-5) /* input outside what the fast path models: use the fallback the text names */
int kvm_csv_open(const char* path, kvm_csv** out);   /* mmap + header + quote-aware row index */
int kvm_csv_close(kvm_csv* csv);
int64_t kvm_csv_rows(const kvm_csv* csv);            /* data rows (blank lines skipped like csv.reader) */
int32_t kvm_csv_ncols(const kvm_csv* csv);
int32_t kvm_csv_flags(const kvm_csv* csv);
const char* kvm_csv_col_name(const kvm_csv* csv, int32_t i);
int32_t kvm_csv_find_col(const kvm_csv* csv, const char* name);  /* last column of that name, -1 if none */
/* column `col` of every row: value[i] (NaN unless kind[i]==NUMBER) and kind[i] (KVM_CSV_CELL_*) */
int kvm_csv_parse_f64(const kvm_csv* csv, int32_t col, double* value, uint8_t* kind, int32_t threads);
/* several columns in one pass over the file (ncols <= 16) */
int kvm_csv_parse_columns(const kvm_csv* csv, const int32_t* cols, int32_t ncols, double* const* values,
                          uint8_t* const* kinds, int32_t threads);
/* parsed text of one cell -> dst (no terminator); returns its length, -1 if out of range */
int64_t kvm_csv_cell(const kvm_csv* csv, int64_t row, int32_t col, char* dst, int64_t cap);
/* out[i] = 1 iff the cell equals one of the options (cost_estimator.py:108-111 is_cold_start) */
int kvm_csv_match_any(const kvm_csv* csv, int32_t col, const char* const* options, int32_t noptions,
                      uint8_t* out, int32_t threads);
/* requests_classified.csv: all columns (+ is_cold_start), coerced columns re-serialised as
 * repr(float); KVM_ERR_UNSUPPORTED when a coerced cell is IRREGULAR (the file is then incomplete) */
int kvm_csv_write_classified(const kvm_csv* csv, const char* out_path, const uint8_t* is_cold,
                             const int32_t* coerced_cols, int32_t ncoerced, int32_t threads);

/* ------------------------------------------------------------------ power-sample ingest (host only)
 * SURVEY.md section 8f F3.  Columnar fast path for `load_power_samples` (energy/collector.py:152-177:
 * power.json = {"samples": [{"ts" | "ts_s": t, "watts": w | null}, ...]} or the bare list) and for
 * Prometheus query / query_range responses (analyze.py:219-247, energy/collector.py:203-251:
 * "values": [[t, "v"], ...] per series): the text is scanned once, in parallel, into two fp64
 * columns per series (timestamp, value; NaN = null) - no object per sample.  Numbers are parsed
 * with std::from_chars (the same double as Python's float()).  Text the scanner does not model
 * returns KVM_ERR_UNSUPPORTED and the caller falls back to the json module. */
typedef struct kvm_samples kvm_samples;
int kvm_power_json_parse(const char* path, int32_t threads, kvm_samples** out);
int kvm_prom_result_parse(const char* text, int64_t len, int32_t threads, kvm_samples** out);
int32_t kvm_samples_series(const kvm_samples* s);
int64_t kvm_samples_count(const kvm_samples* s, int32_t series);
int kvm_samples_copy(const kvm_samples* s, int32_t series, double* ts_host, double* values_host);
void kvm_samples_free(kvm_samples* s);

/* ------------------------------------------------------------------ histograms / series */

/* K5: the four 20-bin value histograms of analyze.py:84-122,149-152 in one pass.
 * counts[4][num_bins] int64 in the order ttfb, tllt, generation, per-token; bin index is
 * min(num_bins-1, int((v-min)/w)) with w=(max-min)/num_bins, bit-exact.  min/max come
 * from `stats`.  Degenerate cases (empty, min==max) are resolved by the host from stats. */
int kvm_log_histograms(const kvm_request_log* log_host, const uint8_t* flags,
                       const kvm_log_stats* stats, int32_t num_bins, int64_t* counts,
                       void* workspace, size_t workspace_bytes, void* stream);
size_t kvm_log_histograms_workspace_bytes(int32_t num_bins);

/* Generic K5 for one column (used by `compute_histograms`, analyze.py:84-122):
 * NaNs skipped; lo/hi given by the caller. */
int kvm_column_histogram(const double* values, int64_t n, double lo, double hi, int32_t num_bins,
                         int64_t* counts, void* workspace, size_t workspace_bytes, void* stream);

/* K8 (extension A12; no reference function): per window of `window_ms` starting at
 * stats->min_start_ms: requests, successes, completion tokens of successes.
 * series[3][max_bins] int64; rows whose bin >= max_bins are counted in *dropped. */
int kvm_log_series(const kvm_request_log* log_host, const uint8_t* flags,
                   const kvm_log_stats* stats, double window_ms, int64_t max_bins, int64_t* series,
                   int64_t* dropped, void* stream);

/* ------------------------------------------------------------------ segmented (sweep) analysis */

/* K4: the numeric core of analyze.py for `runs` independent logs stored back to back
 * (CSR: run r = rows [run_offsets[r], run_offsets[r+1])).  A persistent CTA takes one run at a
 * time and reads its columns ONCE (sampled digit range, shared-memory histograms, one 8-byte
 * digit record per row into `scratch`), then finds the candidates from the records; all other
 * intermediate state stays in shared memory.  Replaces the per-run loop of
 * grid-sweep.sh:103-174 (one analyze.py process per run).  Per run it writes the same
 * kvm_log_stats as kvm_log_scan, the order statistics of the nq queries (results[run*nq + q])
 * and the four num_bins-bin histograms (hist_counts[(run*4 + h)*num_bins + b]; h = ttfb, tllt,
 * generation, per-token).
 * cold_offsets/cold_times: CSR of sorted cold-start times per run (cold_offsets NULL = none).
 * run_status[run] = 1 when the run's candidates did not fit on chip: analyse that run with
 * the flat entry points instead (stats[run] is still valid).  flags[n] is scratch + output;
 * scratch: kvm_seg_analyze_scratch_bytes(n) bytes of device memory (contents undefined after).
 * One launch; stream-ordered; allocates nothing. */
size_t kvm_seg_analyze_scratch_bytes(int64_t n);
int kvm_seg_analyze(const kvm_request_log* log_host, const int64_t* run_offsets, int64_t runs,
                    const int64_t* cold_offsets, const double* cold_times, double cold_window_s,
                    const kvm_query* queries_host, int32_t nq, int32_t num_bins, uint8_t* flags,
                    void* scratch, size_t scratch_bytes, kvm_log_stats* stats,
                    kvm_query_result* results, int64_t* hist_counts, int32_t* run_status, void* stream);

/* The machine-filling streaming kernels (one-pass scan, compact) leave `slots` CTAs out of their
 * 2-per-SM grid, so that many SMs keep half their registers and shared memory free for the few-CTA
 * kernels (plan, resolve, finish) of another analysis in flight on another stream.  Default 16
 * (environment KVM_TAIL_SLOTS); 0 = full grids.  Process-wide; call before capturing graphs. */
int kvm_set_tail_slots(int32_t slots);

/* Debug aid: microseconds since the first CTA entered the last one-pass kernel that used this scan
 * workspace - [0] last CTA set up (sample + parameters), [1] last CTA done streaming, [2] last CTA
 * flushed its histograms, [3] statistics written.  Synchronises (one small copy). */
int kvm_log_fused_phase_us(const void* scan_workspace, double* out4);

/* Debug aid: nanoseconds the segmented kernel spent per phase, summed over all runs and CTAs since
 * the last reset - [0] sample + set-up, [1] stream, [2] plan / resolve, [3] records pass,
 * [4] finish + emit.  Synchronises the device. */
int kvm_seg_phase_times(uint64_t* out5, int32_t reset);

/* ------------------------------------------------------------------ sort */

/* K3: stable ascending sort of (key, payload) fp64 pairs by key (LSD radix, 8-bit digits).
 * Replaces `ss.sort(key=lambda s: s.ts)` of energy/collector.py:141 for traces that arrive
 * out of order.  Synchronises the stream once (constant digits are skipped).  Out-of-place. */
size_t kvm_sort_pairs_workspace_bytes(int64_t m);
int kvm_sort_pairs_f64(const double* keys_in, const double* vals_in, double* keys_out,
                       double* vals_out, int64_t m, void* workspace, size_t workspace_bytes,
                       void* stream);

/* ------------------------------------------------------------------ energy */

/* Whole-trace facts integrate_energy needs (energy/collector.py:296-347). */
typedef struct kvm_power_stats {
  int64_t m;
  int64_t n_valid;         /* samples with non-null watts */
  int64_t n_outside;       /* ts < t0 or ts > t1, nulls included   (collector.py:345) */
  int64_t n_outside_valid; /* of those, non-null                   (collector.py:327-331) */
  int64_t n_unsorted;      /* #i with ts[i+1] < ts[i]; 0 = already ascending */
  double ts_min, ts_max;   /* over all samples                     (collector.py:299-300,313-314) */
  double step_sum_hi;      /* double-double sum of ts[i+1]-ts[i] in load order (collector.py:338-341) */
  double step_sum_lo;
} kvm_power_stats;

size_t kvm_energy_workspace_bytes(int64_t m);

/* One pass over (ts, watts).  outside_ids (may be NULL) receives 0 for non-null samples
 * outside [t0,t1] and 255 otherwise - the set-id array kvm_column_select ranks for the
 * upper-median idle power (collector.py:333). */
int kvm_power_scan(const double* ts, const double* watts, int64_t m, double t0, double t1,
                   uint8_t* outside_ids, kvm_power_stats* stats, void* workspace,
                   size_t workspace_bytes, void* stream);

/* K9: trapezoidal Wh over up to 8 windows in one pass; ts ascending.  windows_host = k pairs
 * (t0,t1) on the host; out_dd[k][2] = double-double (hi,lo) on the device.
 * Replaces trapezoidal_wh, energy/collector.py:133-149. */
int kvm_trapezoid_wh(const double* ts, const double* watts, int64_t m, const double* windows_host,
                     int32_t k, double* out_dd, void* workspace, size_t workspace_bytes, void* stream);

/* K10 (extension A16): inclusive double-double prefix of the segment energies of a trace,
 * then per request interval [start_ms, start_ms+latency_ms]/1000 the trapezoidal energy of
 * the samples inside it (membership clip, like trapezoidal_wh).  accumulate != 0 adds to
 * energy_wh[] (several GPU traces are summed in call order - PromQL `sum(...)`,
 * collector.py:45). */
int kvm_energy_prefix(const double* ts, const double* watts, int64_t m, double* p_hi, double* p_lo,
                      void* workspace, size_t workspace_bytes, void* stream);
int kvm_interval_energy(const double* ts, const double* watts, const double* p_hi,
                        const double* p_lo, int64_t m, const double* start_ms,
                        const double* latency_ms, int64_t n, double* energy_wh, int32_t accumulate,
                        void* stream);
/* All traces (one per GPU of the served model; PromQL `sum(...)`, collector.py:45) in one pass over
 * the intervals: per-trace energies are added in trace order in a register, the interval columns
 * are read once and energy_wh[] written once.  The two boundary searches per interval and trace
 * start from an interpolated guess (traces are sampled regularly) and gallop: time-ordered
 * intervals stream through ts / prefix instead of bisecting 27 levels each. */
/* Multi-GPU energy (one trace per rank; no reference counterpart - the reference is one process
 * and lets PromQL add the GPUs, collector.py:45): out[i] = ((s_0[i] + s_1[i]) + s_2[i]) + ... with
 * s_r = slices + r*stride - the per-interval energies of the ranks added in RANK ORDER, never by a
 * floating-point all-reduce (kserve-vllm-mini_b200/dist_energy.py). */
int kvm_ordered_sum_f64(const double* slices, int32_t world, int64_t n, int64_t stride, double* out, void* stream);

/* Diagnostic: out[i] = the library's division-free evaluation of x[i] / 3600.0 (the trapezoid and
 * prefix kernels use it for collector.py:146); must equal the IEEE quotient bit for bit. */
int kvm_selfcheck_div3600(const double* x, double* out, int64_t n, void* stream);

#define KVM_MAX_TRACES 8
typedef struct kvm_power_trace {
  const double* ts;
  const double* watts;
  const double* p_hi; /* kvm_energy_prefix */
  const double* p_lo;
  int64_t m;
} kvm_power_trace;
int kvm_interval_energy_multi(const kvm_power_trace* traces_host, int32_t ntraces, const double* start_ms,
                              const double* latency_ms, int64_t n, double* energy_wh, int32_t accumulate,
                              void* stream);

#ifdef __cplusplus
}
#endif
#endif /* KVMINI_B200_H */
