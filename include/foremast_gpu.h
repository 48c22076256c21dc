/*
 * foremast_gpu.h -- C ABI of libforemast_gpu, the B200 (sm_100a) anomaly-scoring library.
 *
 * The library replaces ONE path of Foremast: the foremast-brain scoring pass that sits between
 * the Elasticsearch job document and Prometheus.  The reference has no FFI for
 * that path -- the seam is the ES document written by foremast-service
 * (foremast-service/pkg/search/elasticsearchstore.go:58-87, "Status: initial") and read back by
 * SearchByID (foremast-service/cmd/manager/main.go:227-255).  Every entry point below therefore
 * cites the reference artefact whose job it takes over.  All paths are relative to the
 * reference checkout.
 *
 * Rules of the ABI (cgo-safe):
 *   - plain C, POD structs, fixed-width integers, no callbacks, no C++ exceptions across the edge;
 *   - every function returns int: 0 = FM_OK, negative = FM_E_* (mirrors the (value, code, reason)
 *     triples of foremast-service/pkg/search/elasticsearchstore.go:94-98, 0 = ok);
 *   - input pointers are borrowed for the duration of the call only;
 *   - results are library-owned opaque handles, copied out by accessor calls, freed by *_free;
 *   - every entry point sets the CUDA device of its context itself (goroutines hop OS threads);
 *   - there is NO CPU fallback: without a usable sm_100 device fm_init fails with FM_E_NO_DEVICE.
 */
#ifndef FOREMAST_GPU_H
#define FOREMAST_GPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FM_ABI_VERSION 1

/* ---------------------------------------------------------------- error codes */
enum {
  FM_OK = 0,
  FM_E_INVALID = -1,     /* bad argument (NULL, zero size, inconsistent CSR)            */
  FM_E_NOMEM = -2,       /* host or device allocation failed                            */
  FM_E_CUDA = -3,        /* CUDA runtime error; text in fm_last_error                   */
  FM_E_NO_DEVICE = -4,   /* no sm_100 GPU visible -- the library has no CPU path        */
  FM_E_ALIGN = -5,       /* device CSR offsets must be multiples of 4 elements (16 B)   */
  FM_E_STATE = -6,       /* call order violated (e.g. add after seal)                   */
  FM_E_RANGE = -7,       /* index / size outside what the call supports                 */
  FM_E_UNSUPPORTED = -8, /* valid request the build cannot serve                        */
  FM_E_COMM = -9,        /* multi-GPU exchange failed                                   */
  FM_E_TOO_LARGE = -10   /* more items than a fixed-capacity entry point can take       */
};

/* ---------------------------------------------------------------- vocabulary */

/* Historical-model algorithms.  ML_ALGORITHM (deploy/foremast/3_brain/foremast-brain.yaml:24-25,
 * default moving_average_all) and the model list of docs/guides/design.md:68-71.  A bit mask so
 * that "all scorers enabled" (BASELINE.json north_star) is one pass. */
#define FM_ALG_MA_ALL     (1u << 0) /* "Moving Average", global mean/std of the model window */
#define FM_ALG_MA_ROLLING (1u << 1) /* rolling-W mean/std band from fp64 prefix sums          */
#define FM_ALG_EWMA       (1u << 2) /* "Exponential Smoothing"                                */
#define FM_ALG_DES        (1u << 3) /* "Double Exponential Smoothing" (Holt linear)           */
#define FM_ALG_HW         (1u << 4) /* "Holt-Winters", additive, period fm_params.period      */
#define FM_N_ALG 5
#define FM_ALG_ALL 0x1Fu

/* ML_PAIRWISE_ALGORITHM, foremast-brain/README.md:34 */
enum { FM_PW_NONE = 0, FM_PW_ALL = 1, FM_PW_ANY = 2, FM_PW_MANN_WHITE = 3, FM_PW_WILCOXON = 4, FM_PW_KRUSKAL = 5 };

/* ML_BOUND, foremast-brain/README.md:24 ("upper bound, lower bound or upper and lower bound");
 * values 1 and 3 appear in foremast-brain.yaml:30-31,54-55.  Bit mask (oracle/DECISIONS.md D1). */
enum { FM_BOUND_UPPER = 1, FM_BOUND_LOWER = 2, FM_BOUND_BOTH = 3 };

/* metric_type rows of foremast-brain.yaml:34-73 (+ two extra slots used by the 8-metric configs). */
enum {
  FM_MT_OTHER = 0, FM_MT_ERROR5XX = 1, FM_MT_ERROR4XX = 2, FM_MT_LATENCY = 3, FM_MT_CPU = 4,
  FM_MT_MEMORY = 5, FM_MT_TPS = 6, FM_MT_GC = 7, FM_MT_THREADS = 8
};
#define FM_MAX_METRIC_TYPES 16

/* Verdicts = the three terminal statuses the brain writes into the ES document,
 * foremast-service/pkg/converter/converter.go:16-21.  Ordered so that "worse" is larger. */
enum { FM_V_HEALTH = 0 /* completed_health */, FM_V_UNKNOWN = 1 /* completed_unknown */, FM_V_UNHEALTH = 2 /* completed_unhealth */ };

/* per-unit status bits (fm_series_score.status) */
#define FM_S_UNKNOWN      (1u << 0) /* no current data, or no model window long enough (design.md:25)   */
#define FM_S_ANOMALY      (1u << 1) /* at least min_anomaly_points current points outside a band        */
#define FM_S_DIFFERS      (1u << 2) /* pairwise tests say baseline and current differ (design.md:35)    */
#define FM_S_MW_RUN       (1u << 3) /* Mann-Whitney had enough points and was evaluated                 */
#define FM_S_WILCOXON_RUN (1u << 4)
#define FM_S_KRUSKAL_RUN  (1u << 5)
#define FM_S_MODEL_BASE   (1u << 6) /* model window was the baseline segment (no usable history)        */
#define FM_S_ALG_SHIFT    8         /* bit (8+a) set: algorithm a was evaluated for this unit           */

/* ---------------------------------------------------------------- parameters */

/* One row of the threshold table, foremast-brain.yaml:26-73 (threshold / bound / min_lower_bound). */
typedef struct fm_metric_rule {
  double threshold;        /* k: band half-width in standard deviations */
  double min_lower_bound;  /* lower band is clamped to at least this    */
  uint32_t bound;          /* FM_BOUND_*                                */
  uint32_t reserved;
} fm_metric_rule;

/* Mirrors the brain's environment block, foremast-brain.yaml:21-81 and foremast-brain/README.md:20-38.
 * fm_params_default() fills the documented defaults; oracle/DECISIONS.md lists the undocumented ones. */
typedef struct fm_params {
  uint32_t algorithms;          /* FM_ALG_* mask; ML_ALGORITHM                                         */
  uint32_t pairwise;            /* FM_PW_*;  ML_PAIRWISE_ALGORITHM                                     */
  double pairwise_threshold;    /* ML_PAIRWISE_THRESHOLD: reject "same distribution" when p < this     */
  double differ_scale;          /* threshold multiplier when distributions differ (design.md:35)       */
  uint32_t min_mann_white;      /* MIN_MANN_WHITE_DATA_POINTS  = 20, foremast-brain.yaml:74-75         */
  uint32_t min_wilcoxon;        /* MIN_WILCOXON_DATA_POINTS    = 20, :76-77                            */
  uint32_t min_kruskal;         /* MIN_KRUSKAL_DATA_POINTS     = 5,  :78-79                            */
  uint32_t min_historical;      /* MIN_HISTORICAL_DATA_POINT_TO_MEASURE, README.md:23                  */
  uint32_t min_anomaly_points;  /* fail-fast: 1 (design.md:43)                                         */
  uint32_t ma_window;           /* W of FM_ALG_MA_ROLLING                                              */
  uint32_t period;              /* Holt-Winters season length in samples (1440 = one day at 60 s)      */
  uint32_t band_algorithm;      /* ONE FM_ALG_* bit whose per-point band is materialised, or 0         */
  double alpha, beta, gamma;    /* smoothing constants (level, trend, season)                          */
  fm_metric_rule rules[FM_MAX_METRIC_TYPES]; /* indexed by metric_type; row 0 = the un-typed default   */
} fm_params;

/* ---------------------------------------------------------------- data */

/* One scoring unit = one (app, metric, pod) series-window: up to three segments stored back to
 * back in `values` -- historical | baseline | current -- exactly the three range queries of one
 * job alias (historicalConfig / baselineConfig / currentConfig, pkg/models/models.go:108-110).
 * Samples are fp32; timestamps are implicit (start + i*step, step = 60 s,
 * foremast-barrelman/pkg/client/metrics/metricsquery.go:63).
 *
 * Ragged CSR: unit u owns values[offsets[u] .. offsets[u]+n_hist+n_base+n_cur).  offsets has
 * n_units+1 entries, is non-decreasing, and gaps (padding) are allowed.  Device-resident CSR must
 * have every offset a multiple of 4 elements (16 B) and 32 readable elements after the last unit. */
typedef struct fm_csr {
  uint64_t n_units;
  const float *values;
  const int64_t *offsets;      /* [n_units+1], in elements                    */
  const uint32_t *seg_lens;    /* [3*n_units]: n_hist, n_base, n_cur per unit */
  const uint32_t *service_id;  /* [n_units], dense ids < n_services           */
  const uint8_t *metric_type;  /* [n_units], FM_MT_*                          */
} fm_csr;

typedef struct fm_series_meta {
  uint32_t service_id;
  uint8_t metric_type;
  uint8_t reserved[3];
} fm_series_meta;

/* Everything the scorers produce for one unit.  Integer statistics are exact (twice the rank
 * sums so that half-integer average ranks stay integral); doubles are fp64 end to end. */
typedef struct fm_series_score {
  uint32_t status;                 /* FM_S_* */
  uint32_t n_anom[FM_N_ALG];       /* anomalous current points per algorithm               */
  uint32_t first_idx[FM_N_ALG];    /* index in the current segment of the first one, or ~0 */
  uint32_t w_n;                    /* Wilcoxon: number of non-zero differences             */
  int64_t mw_2u1;                  /* 2*U of the baseline sample                           */
  int64_t mw_2r1;                  /* 2*(rank sum of the baseline sample), pooled average ranks */
  int64_t mw_ties;                 /* sum(t^3 - t) over pooled tie groups                  */
  int64_t w_2wplus, w_2wminus;     /* 2*W+, 2*W- (current minus baseline)                  */
  int64_t w_ties;                  /* sum(t^3 - t) over tie groups of |d|                  */
  double p_mw, p_w, p_k;           /* two-sided normal-approximation p-values (1 if not run) */
  double kw_h;                     /* Kruskal-Wallis H, tie-corrected                      */
  double k_eff;                    /* threshold actually applied                           */
  double ma_mean, ma_std;          /* FM_ALG_MA_ALL model                                  */
  double sigma[FM_N_ALG];          /* band sigma per algorithm (residual sigma for smoothers) */
} fm_series_score;

/* Per-service verdict = what the worker writes to the ES document's `status`
 * (converter.go:16-21) plus the counts used to build `reason`. */
typedef struct fm_verdict {
  uint32_t service_id;
  uint32_t verdict;            /* FM_V_* */
  uint32_t n_units;
  uint32_t n_unknown;
  uint32_t n_anom_units;
  uint32_t n_anom_points;
  uint32_t worst_unit;         /* global unit index with most anomalous points (lowest index on ties), ~0 if none */
  uint32_t worst_unit_points;
} fm_verdict;

/* Synthetic workload generator; bit-identical on device and in oracle/. */
typedef struct fm_synth_config {
  uint64_t n_units;
  uint32_t n_hist, n_base, n_cur;  /* nominal segment lengths                                   */
  uint32_t units_per_service;      /* service_id = unit / units_per_service                     */
  uint32_t n_metric_types;         /* metric_type = 1 + (unit % n_metric_types), <= 8           */
  uint32_t seed;
  uint32_t flags;                  /* FM_SYNTH_*                                                */
  uint64_t first_unit;             /* global index of unit 0 (sharding across GPUs)             */
} fm_synth_config;
#define FM_SYNTH_RAGGED    (1u << 0) /* ~2 % of units lose samples, ~0.1 % have no current window */
#define FM_SYNTH_ANOMALIES (1u << 1) /* ~1 % spike, ~0.5 % level shift in the current segment     */

typedef struct fm_config {
  int32_t device;           /* CUDA ordinal, -1 = current device */
  uint32_t flags;
} fm_config;

typedef struct fm_ctx fm_ctx;
typedef struct fm_batch fm_batch;
typedef struct fm_result fm_result;
typedef struct fm_model fm_model;

/* ---------------------------------------------------------------- life cycle */

/* Takes the place of the brain process start-up (deploy/foremast/3_brain/foremast-brain.yaml:16-20). */
int fm_init(const fm_config *cfg, fm_ctx **out);
void fm_shutdown(fm_ctx *ctx);
int fm_version(void);
const char *fm_strerror(int code);
/* Text of the last failure on this context ("reason" of elasticsearchstore.go:94-98).  Safe while other
 * threads use the context: fm_last_error returns a copy that belongs to the calling thread (valid until
 * that thread calls it again); fm_last_error_copy writes into the caller's buffer (what a cgo shim uses). */
const char *fm_last_error(const fm_ctx *ctx);
int fm_last_error_copy(const fm_ctx *ctx, char *buf, uint32_t cap);
/* Documented defaults of foremast-brain.yaml:24-79 + oracle/DECISIONS.md. */
void fm_params_default(fm_params *p);
int fm_device_info(const fm_ctx *ctx, int32_t *sm_count, uint64_t *hbm_bytes, char *name, uint32_t name_cap);

/* CUDA-event stopwatch on the context's stream (all library work is ordered on it). */
int fm_timer_start(fm_ctx *ctx);
int fm_timer_stop(fm_ctx *ctx, float *ms);

/* Pinned host memory for callers that want zero-copy staging of large CSR batches. */
int fm_host_alloc(fm_ctx *ctx, uint64_t bytes, void **out);
void fm_host_free(fm_ctx *ctx, void *p);

/* ---------------------------------------------------------------- ingest */

/* Fast path from a Prometheus query_range response body (matrix) to fp32 series: the step the brain
 * performs after fetching the URLs built by pkg/prometheus/prometheushelper.go:13-43; response shape
 * data.result[].values[[ts,"v"],...] as read by foremast-dashboard/src/services/ApiService.js:36-39.
 * One pass, no allocation: series[i] describes result i (byte span of its "metric" label object in
 * `json`, first timestamp, step, and the slice values[first_value .. first_value+n_values)).
 * With too small buffers it returns FM_E_RANGE and the required counts in n_series / n_values
 * (call once with caps 0 to size).  Host code only, and reentrant (no shared state, no context):
 * a worker parses one response per thread -- a response is one query of one job, so a scoring
 * cycle has thousands of them. */
typedef struct fm_prom_series {
  uint64_t metric_off, metric_len;
  double t0, step;             /* first timestamp; smallest gap between neighbouring samples (or the hint) */
  uint64_t first_value;
  uint32_t n_values;
  uint32_t reserved;           /* grid slots t0 + k*step up to the last sample that have NO sample (missing
                                  scrapes; 0 = the series is regular), 0xFFFFFFFF if timestamps do not ascend */
} fm_prom_series;
int fm_parse_prometheus_matrix(const char *json, uint64_t len, fm_prom_series *series, uint32_t series_cap,
                               uint32_t *n_series, float *values, uint64_t values_cap, uint64_t *n_values);
/* query_range omits a missing scrape instead of sending NaN, while the scorers are index-based (Wilcoxon
 * pairs sample i with sample i, the Holt-Winters phase is index mod period, anomaly times are t0 + i*step,
 * foremast-barrelman/pkg/client/metrics/metricsquery.go:63).  With FM_PROM_GRID every sample is placed at
 * round((ts - t0) / step) -- step_hint (the `step` of the query URL, prometheushelper.go:13-43) or, when
 * <= 0, the smallest gap of the series -- and the holes are NaN (dropped later = the old behaviour, but now
 * visible in `reserved`), or the previous sample with FM_PROM_FILL, which keeps index and time aligned. */
#define FM_PROM_GRID (1u << 0)
#define FM_PROM_FILL (1u << 1)
int fm_parse_prometheus_matrix_grid(const char *json, uint64_t len, uint32_t flags, double step_hint, fm_prom_series *series,
                                    uint32_t series_cap, uint32_t *n_series, float *values, uint64_t values_cap,
                                    uint64_t *n_values);
/* Response bodies straight into a batch, no host language in the loop: the three query_range answers of ONE
 * job alias (historicalConfig / baselineConfig / currentConfig entry, main.go:50-81; any may be NULL) become
 * one unit per current series -- baseline series of the same rank, first historical series -- exactly the
 * pairing of the worker (foremast_b200/brain.py::build_units).  _many parses n_jobs such triples on a thread
 * pool (n_threads = 0: all cores; the scanner is reentrant) and appends them in job order; units_per_job[i]
 * = units added for job i, or 0xFFFFFFFF if its responses do not parse (the others are still added). */
typedef struct fm_prom_job {
  fm_series_meta meta;
  uint32_t reserved;
  double step;                 /* query step in seconds, <= 0: take it from the data */
  const char *hist_json, *base_json, *cur_json;
  uint64_t hist_len, base_len, cur_len;
} fm_prom_job;
int fm_batch_add_prometheus(fm_batch *b, const fm_prom_job *job, uint32_t flags, uint32_t *n_units_added);
int fm_batch_add_prometheus_many(fm_batch *b, const fm_prom_job *jobs, uint32_t n_jobs, uint32_t flags, uint32_t n_threads,
                                 uint32_t *units_per_job);

/* Incremental ingest of one job's series: what the brain does after fetching the three
 * query_range matrices named by currentConfig / baselineConfig / historicalConfig
 * (foremast-service/cmd/manager/main.go:50-81).  Non-finite samples are dropped (gaps),
 * -0.0 is stored as +0.0 (DECISIONS D11). */
int fm_batch_begin(fm_ctx *ctx, uint64_t n_units_hint, uint64_t n_points_hint, fm_batch **out);
int fm_batch_add_series(fm_batch *b, const fm_series_meta *meta,
                        const float *hist, uint32_t n_hist,
                        const float *base, uint32_t n_base,
                        const float *cur, uint32_t n_cur);
/* Pads, uploads (H2D) and freezes the batch. */
int fm_batch_seal(fm_batch *b);
/* Bulk ingest of a host CSR; copied (and re-aligned if needed) during the call.  The bulk paths
 * (this one, fm_batch_from_device, fm_score_host) take `values` as they are: the packer passes
 * finite samples only, i.e. it has already dropped the gaps that fm_batch_add_series drops. */
int fm_batch_from_csr(fm_ctx *ctx, const fm_csr *host_csr, fm_batch **out);
/* Wrap a CSR that already lives in this device's HBM (borrowed until fm_batch_free). */
int fm_batch_from_device(fm_ctx *ctx, const fm_csr *device_csr, fm_batch **out);
/* Generate the synthetic workload of SURVEY 8(d) directly in HBM. */
int fm_batch_synth(fm_ctx *ctx, const fm_synth_config *cfg, fm_batch **out);
/* Multi-GPU shards: global index of this batch's unit 0 and the global service count, so that
 * every shard fills the same dense per-service vector (the reference scales the same way, with
 * shared-nothing brain pods, docs/guides/design.md:37-41). */
int fm_batch_set_shard(fm_batch *b, uint64_t first_unit, uint32_t n_services);
int fm_batch_info(const fm_batch *b, uint64_t *n_units, uint64_t *n_values, uint32_t *n_services);
/* Copy the packed CSR back to host buffers (tests, CPU baseline); any pointer may be NULL. */
int fm_batch_download(const fm_batch *b, float *values, int64_t *offsets, uint32_t *seg_lens,
                      uint32_t *service_id, uint8_t *metric_type);
void fm_batch_free(fm_batch *b);

/* ---------------------------------------------------------------- scoring (a2-a11) */

/* The scoring pass itself: pairwise tests -> threshold adjust -> band models -> anomaly flags ->
 * per-service verdicts (docs/guides/design.md:35,43).  Asynchronous on the context's stream;
 * accessors synchronise.  Rank tests work on segments of up to FM_RANK_MAX_SEGMENT samples (the
 * reference's baseline / current windows are 30-1440 one-minute points): a unit whose baseline or
 * current segment is longer has Mann-Whitney and Kruskal-Wallis not run, a unit with more than that many
 * PAIRS has Wilcoxon not run (status bits FM_S_*_RUN stay clear, p = 1, exactly like a unit with too
 * few points, oracle/DECISIONS.md D17) and everything else of the unit and of the batch is scored as
 * usual.  History length is only bounded by memory. */
#define FM_RANK_MAX_SEGMENT 2048
int fm_score(fm_ctx *ctx, const fm_batch *b, const fm_params *p, fm_result **out);
/* One-shot, host in / host out: chunked H2D overlapped with scoring, verdicts copied back.
 * `verdicts` has room for n_services entries. */
int fm_score_host(fm_ctx *ctx, const fm_csr *host_csr, const fm_params *p,
                  uint32_t n_services, fm_verdict *verdicts);
int fm_result_sync(fm_result *r);
/* status `completed_health|completed_unhealth|completed_unknown` per service (converter.go:16-21). */
int fm_result_verdicts(fm_result *r, fm_verdict *out, uint64_t cap, uint64_t *n);
/* The verdicts of services [first_service, first_service + count) only: after fm_result_merge every rank
 * holds all counters, and a rank of a sharded sweep usually writes back only the documents of its own shard. */
int fm_result_verdicts_range(fm_result *r, uint64_t first_service, uint64_t count, fm_verdict *out);
int fm_result_series(fm_result *r, uint64_t first_unit, uint64_t count, fm_series_score *out);
/* upper/lower band of the current segment of one unit, for the foremastbrain:*_upper/_lower
 * gauges (foremast-dashboard/src/config/metrics.js:20-31).  Needs params.band_algorithm. */
int fm_result_bands(fm_result *r, uint64_t unit, float *upper, float *lower, uint32_t cap, uint32_t *n);
/* Continuous-mode model cache: the brain keeps fitted models between cycles
 * (MAX_CACHE_SIZE, foremast-brain/README.md:30) while continuous / hpa jobs are re-armed with fresh
 * START_TIME / END_TIME every cycle (foremast-service/cmd/manager/main.go:59-63;
 * foremast-barrelman/pkg/controller/Barrelman.go:552-565).  fm_model_fit runs Holt-Winters over the
 * model window of every unit once and keeps (level, trend, residual sigma, phase, seasonals) in
 * HBM -- (period + 4) doubles per unit.  fm_model_score then scores a batch that carries only the
 * NEW current samples of the same units, in the same order, from that state, and advances the
 * state, so each cycle costs the new window, not the 7-day history.  sigma stays the one of the
 * fit.  Units whose model window was shorter than two periods are `unknown`. */
int fm_model_fit(fm_ctx *ctx, const fm_batch *history, const fm_params *p, fm_model **out);
int fm_model_score(fm_ctx *ctx, fm_model *m, const fm_batch *current, fm_result **out);
void fm_model_free(fm_model *m);

/* Compact list of the anomalous current points of ONE algorithm, the device-side form of the
 * anomaly pairs barrelman decodes as [t0, v0, t1, v1, ...]
 * (foremast-barrelman/pkg/controller/DeploymentController.go:443-452; t = current start + idx*step).
 * build: per-unit counts -> exclusive scan -> one emit pass over the units that have anomalies. */
int fm_result_anomalies_build(fm_result *r, uint32_t algorithm, uint64_t *n_points);
int fm_result_anomalies_unit(fm_result *r, uint64_t unit, uint32_t *cur_idx, float *values, uint32_t cap, uint32_t *n);
/* offsets has n_units+1 entries; cur_idx / values have room for cap >= n_points entries. */
int fm_result_anomalies_all(fm_result *r, uint64_t *offsets, uint32_t *cur_idx, float *values, uint64_t cap);
/* fm_scan_counts_static -- NO COUNTERPART IN THE REFERENCE (the brain is a Python process that
 * allocates freely); it exists for hosts that must not allocate GPU memory while they score.
 *
 * Who that is: a worker that sized its device arena once at start-up and runs under a fixed memory
 * budget next to other tenants of the GPU; a scoring loop with a latency bound, where cudaMalloc /
 * cudaFree may stall on the driver's allocator lock or synchronise the whole device; and code that
 * is not allowed to fail for lack of memory half-way through a sweep.  Every other entry point of
 * this library allocates result buffers; this one allocates nothing -- no cudaMalloc, cudaFree or
 * stream-ordered pool call -- it clears a fixed scratch array that is part of the library image and
 * launches one kernel.
 *
 * What it computes: the exclusive prefix sum that fm_result_anomalies_build uses to lay out the
 * compact anomaly list.  d_counts[n] (device, uint32: anomalous points per unit) ->
 * d_offsets[n+1] (device, uint64; d_offsets[n] = total).  Both buffers are the caller's.
 *
 * Limits and rules:
 *  - n <= FM_SCAN_STATIC_MAX_COUNTS (16,777,216); more returns FM_E_TOO_LARGE and does nothing.
 *    The scratch is 64 KiB of device memory per GPU that has loaded the library, held until exit.
 *  - Runs on the CURRENT device, on `cuda_stream` (a cudaStream_t, NULL = default stream),
 *    asynchronously: synchronise the stream before reading d_offsets.
 *  - The first call on each device looks up the scratch array's device address
 *    (cudaGetSymbolAddress); make that first call before starting any stream capture.
 *  - All calls on a device share the one scratch array: two calls whose GPU work can overlap
 *    (different streams, or different threads) race and produce garbage.  Serialise them. */
#define FM_SCAN_STATIC_MAX_COUNTS 16777216ull
int fm_scan_counts_static(const uint32_t *d_counts, uint64_t n, uint64_t *d_offsets, void *cuda_stream);

/* ---------------------------------------------------------------- multi-GPU merge
 *
 * The reference scales with shared-nothing brain pods that only meet in Elasticsearch
 * (docs/guides/design.md:37-41).  Here every GPU scores its shard of the series (one context per GPU, in
 * one process or many) and the dense per-service counters are merged by one NCCL all-reduce pair
 * (u32[4S] summed, u64[S] maxed, one launch) that is ordered on the context's own stream:
 *
 *     rank 0: fm_comm_unique_id(id)  -> hand the 128 bytes to every rank (the host's own channel)
 *     all   : fm_comm_init(ctx, id, rank, world)            -- collective, once per context
 *     sweep : fm_batch_set_shard(b, first_unit, n_services_total); fm_score(ctx, b, p, &r);
 *             fm_result_merge(r);                            -- collective, asynchronous, no host sync
 *             fm_result_verdicts(r, ...)                     -- every rank gets the merged verdicts
 *
 * NCCL is loaded at run time (libnccl.so.2); a host that never calls fm_comm_* does not need it.
 * Failures return FM_E_COMM with the NCCL text in fm_last_error. */
#define FM_COMM_ID_BYTES 128
int fm_comm_unique_id(void *id_out, uint32_t cap);
int fm_comm_init(fm_ctx *ctx, const void *unique_id, int32_t rank, int32_t world);
int fm_comm_info(const fm_ctx *ctx, int32_t *rank, int32_t *world);
void fm_comm_destroy(fm_ctx *ctx); /* also done by fm_shutdown */
/* In-place merge of this result's per-service counters across the ranks of the context's communicator
 * (a no-op without one).  Call it on every rank, in the same order of results. */
int fm_result_merge(fm_result *r);

/* Device pointers of the dense per-service counters, for hosts that merge them by other means.  They are
 * written by fm_score on the context's PRIVATE non-blocking stream and read again, on that same stream, by
 * fm_result_verdicts: a caller that touches them on another stream must call fm_result_sync first and must
 * finish (synchronise) its own work on them before calling fm_result_verdicts.
 * counts = uint32[4*n_services] (n_units, n_unknown, n_anom_units, n_anom_points), summed;
 * worst  = uint64[n_services]   ((points << 32) | ~unit), maxed. */
int fm_result_service_buffers(fm_result *r, uint32_t **counts, uint64_t **worst, uint32_t *n_services);
/* Device time of this fm_score in milliseconds per stage: {set-up (allocation, clears), Wilcoxon
 * rank kernel, Mann-Whitney/Kruskal rank kernel + p-values, band kernels, verdict reduce}; CUDA
 * events on the context's stream; cap >= 5. */
int fm_result_timings(fm_result *r, float *ms, uint32_t cap);
/* Free results before the batch they were scored from (accessors read the batch's offsets). */
void fm_result_free(fm_result *r);

#ifdef __cplusplus
}
#endif
#endif /* FOREMAST_GPU_H */
