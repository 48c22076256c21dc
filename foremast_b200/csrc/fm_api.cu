// fm_api.cu -- the C ABI of libforemast_gpu (include/foremast_gpu.h): contexts, ingest into the
// ragged CSR layout in HBM, the scoring pass, result accessors.  No torch types, no exceptions
// across the edge, no CPU scoring path: every score comes from the kernels in this directory.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "fm_common.cuh"
#include "fm_internal.h"

// ------------------------------------------------------------------ objects

struct fm_ctx {
  int device = 0;
  int sm_count = 0;
  uint64_t hbm_bytes = 0;
  char name[256] = {0};
  cudaStream_t stream = nullptr;       // compute
  cudaStream_t copy_stream = nullptr;  // H2D staging of fm_score_host
  cudaEvent_t timer[2] = {nullptr, nullptr};
  std::mutex mu;         // GPU submission
  std::mutex err_mu;     // last_error only: set_err may run with or without `mu` held
  std::string last_error;
  fm_comm_state comm;    // multi-GPU merge (fm_comm.cu)
  // small-job arena of fm_score_host: one device slab + one pinned slab, made on the first small call, so that a
  // canary job (a handful of series) costs one H2D, one clear, the launches and one D2H -- no allocation, no events
  unsigned char *arena_dev = nullptr, *arena_host = nullptr;
  DevParams arena_params;       // what the slab's DevParams currently holds
  bool arena_params_valid = false;
  bool arena_clean = false;     // the result part of the slab is zero (cleared behind the previous call's download)
  cudaEvent_t arena_done = nullptr;  // the previous call's verdicts have reached the host slab
};

namespace {
constexpr uint64_t kArenaUnits = 1024, kArenaValues = 256 * 1024, kArenaServices = 1024;
constexpr uint64_t a16(uint64_t x) { return (x + 255) & ~255ull; }
// device slab layout (byte offsets); the host slab mirrors [params .. values] for the single upload, then the verdicts
constexpr uint64_t kOffParams = 0;
constexpr uint64_t kOffOffsets = a16(sizeof(DevParams));   // params | per-unit arrays | values: ONE upload, which ends
constexpr uint64_t kOffSeg = kOffOffsets + a16((kArenaUnits + 1) * 8);   // where the job's samples end
constexpr uint64_t kOffSid = kOffSeg + a16(3 * kArenaUnits * 4);
constexpr uint64_t kOffMt = kOffSid + a16(kArenaUnits * 4);
constexpr uint64_t kOffValues = kOffMt + a16(kArenaUnits);
constexpr uint64_t kOffScores = kOffValues + a16((kArenaValues + 32) * 4);  // scores | counts | worst are cleared together
constexpr uint64_t kOffCounts = kOffScores + a16((kArenaUnits + 1) * sizeof(fm_series_score));
constexpr uint64_t kOffWorst = kOffCounts + a16((4 * kArenaServices + 4) * 4);
constexpr uint64_t kOffUstat = kOffWorst + a16((kArenaServices + 1) * 8);
constexpr uint64_t kOffUpts = kOffUstat + a16((kArenaUnits + 1) * 4);
constexpr uint64_t kOffVerdicts = kOffUpts + a16((kArenaUnits + 1) * 4);
constexpr uint64_t kArenaBytes = kOffVerdicts + a16((kArenaServices + 1) * sizeof(fm_verdict));
}  // namespace

// std::allocator value-initialises on resize(); the bulk ingest path reserves hundreds of MB that its threads then
// fill completely, so zeroing them first (one thread, fresh pages) was a third of its time
template <class T>
struct DefaultInitAlloc : std::allocator<T> {
  template <class U> struct rebind { using other = DefaultInitAlloc<U>; };
  template <class U> void construct(U *p) noexcept { ::new (static_cast<void *>(p)) U; }
  template <class U, class... A> void construct(U *p, A &&...a) { ::new (static_cast<void *>(p)) U(std::forward<A>(a)...); }
};

struct fm_batch {
  fm_ctx *ctx = nullptr;
  bool sealed = false, owns = true;
  uint64_t n_units = 0, n_values = 0;  // n_values: allocated elements incl. 32 pad
  uint32_t n_services = 0, max_hist = 0, max_base = 0, max_cur = 0;
  uint32_t rank_mask_w = 0, rank_mask_m = 0;  // sort-size classes present (fm_rank_class); 0 = not known (one launch for all)
  uint64_t first_unit = 0;
  float *d_values = nullptr;
  int64_t *d_offsets = nullptr;
  uint32_t *d_seg = nullptr, *d_sid = nullptr;
  uint8_t *d_mt = nullptr;
  // host staging of the incremental path
  std::vector<float, DefaultInitAlloc<float>> h_values;
  std::vector<int64_t> h_offsets;
  std::vector<uint32_t> h_seg, h_sid;
  std::vector<uint8_t> h_mt;
};

struct fm_result {
  fm_ctx *ctx = nullptr;
  const fm_batch *batch = nullptr;
  uint64_t n_units = 0;
  uint32_t n_services = 0;
  uint32_t band_algorithm = 0;
  DevParams *d_params = nullptr;
  fm_series_score *d_scores = nullptr;
  uint32_t *d_unit_status = nullptr, *d_unit_points = nullptr, *d_counts = nullptr;
  uint32_t *d_order = nullptr;  // k_band's unit order (fm_launch_band_order); not in the small-job arena
  uint32_t *d_rank_classes = nullptr;  // class sizes + unit lists of a batch with several rank sort sizes (fm_launch_rank)
  unsigned long long *d_worst = nullptr;
  fm_verdict *d_verdicts = nullptr;
  float *d_band_up = nullptr, *d_band_lo = nullptr;
  // compact anomaly list (fm_result_anomalies_build)
  fm_params params;
  unsigned long long *d_anom_off = nullptr;
  uint32_t *d_anom_idx = nullptr;
  float *d_anom_val = nullptr;
  uint64_t anom_total = 0;
  uint32_t anom_alg = 0;
  bool from_model = false;  // scored by fm_model_score: the batch holds no model window, so no emit pass
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // start, pre-rank, post-Wilcoxon, post-rank, post-band, post-verdict
};

struct fm_model {
  fm_ctx *ctx = nullptr;
  uint64_t n_units = 0;
  fm_params params;
  DevParams *d_params = nullptr;
  double *d_state = nullptr;  // [n_units][period + 4]
};

namespace {

int set_err(fm_ctx *ctx, int code, const std::string &msg) {
  if (ctx) {
    std::lock_guard<std::mutex> g(ctx->err_mu);
    ctx->last_error = msg;
  }
  return code;
}

// a NULL / malformed argument: the stale text of an earlier failure must not be what fm_last_error shows next
int bad_arg(const fm_ctx *ctx, const char *fn) { return set_err(const_cast<fm_ctx *>(ctx), FM_E_INVALID, std::string(fn) + ": NULL or inconsistent argument"); }

#define FM_CUDA(ctx, expr)                                                                      \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      cudaGetLastError();                                                                       \
      return set_err((ctx), _e == cudaErrorMemoryAllocation ? FM_E_NOMEM : FM_E_CUDA,           \
                     std::string(#expr) + ": " + cudaGetErrorString(_e));                       \
    }                                                                                           \
  } while (0)

DevBatch dev_view(const fm_batch *b) {
  DevBatch d;
  d.n_units = b->n_units;
  d.values = b->d_values;
  d.offsets = b->d_offsets;
  d.seg_lens = b->d_seg;
  d.service_id = b->d_sid;
  d.metric_type = b->d_mt;
  d.first_unit = b->first_unit;
  return d;
}

void resolve_params(const fm_params *p, DevParams *d) {
  memset(d, 0, sizeof(*d));
  d->algorithms = p->algorithms;
  d->pairwise = p->pairwise;
  d->pairwise_threshold = p->pairwise_threshold;
  d->differ_scale = p->differ_scale;
  d->min_mann_white = p->min_mann_white;
  d->min_wilcoxon = p->min_wilcoxon;
  d->min_kruskal = p->min_kruskal;
  d->min_historical = p->min_historical;
  d->min_anomaly_points = p->min_anomaly_points;
  d->ma_window = p->ma_window ? p->ma_window : 1;
  d->period = p->period;
  d->band_algorithm = p->band_algorithm;
  d->alpha = p->alpha;
  d->beta = p->beta;
  d->gamma = p->gamma;
  {  // powers of the Holt-Winters level/trend transition matrix: with k samples per lane,
     // lane[l] = A^(k*l) (identity for lane 0) and pow[i] = A^(k*2^i) -- once for k_hw (k = period/32), once for
     // k_hw_tm (k = period / fm_hw_tm_lanes(period)), which also takes the chunk weights A^(k-1-i) g
    const double a = p->alpha, ab = p->alpha * p->beta;
    const double A[4] = {1.0 - a, 1.0 - a, -ab, 1.0 - ab};
    auto mul = [](const double *x, const double *y, double *o) {
      const double r[4] = {x[0] * y[0] + x[1] * y[2], x[0] * y[1] + x[1] * y[3], x[2] * y[0] + x[3] * y[2], x[2] * y[1] + x[3] * y[3]};
      for (int i = 0; i < 4; ++i) o[i] = r[i];
    };
    auto powers = [&](uint32_t k, double *lane, double *pw) {
      double Ak[4] = {1.0, 0.0, 0.0, 1.0};
      for (uint32_t i = 0; i < k; ++i) mul(Ak, A, Ak);
      double P[4] = {1.0, 0.0, 0.0, 1.0};
      for (int l = 0; l < 32; ++l) {  // A^(k*l)
        for (int i = 0; i < 4; ++i) lane[4 * l + i] = P[i];
        mul(P, Ak, P);
      }
      double Q[4] = {Ak[0], Ak[1], Ak[2], Ak[3]};
      for (int i = 0; i < 5; ++i) {   // A^(k*2^i)
        for (int j = 0; j < 4; ++j) pw[4 * i + j] = Q[j];
        mul(Q, Q, Q);
      }
    };
    powers(p->period >= 32 ? p->period / 32 : 1, d->hw_lane, d->hw_pow);
    d->hw_tl = fm_hw_tm_lanes(p->period);
    d->hw_tk = d->hw_tl ? p->period / d->hw_tl : 1;
    powers(d->hw_tk, d->hw_tlane, d->hw_tpow);
    double v[2] = {a, ab};  // A^(k-1-i) g for i = k-1 .. 0
    for (int i = 0; i < 64; ++i) d->hw_w[2 * i] = d->hw_w[2 * i + 1] = 0.0;
    for (int i = (int)std::min<uint32_t>(d->hw_tk, 64u) - 1; i >= 0; --i) {
      d->hw_w[2 * i] = v[0];
      d->hw_w[2 * i + 1] = v[1];
      const double nl = A[0] * v[0] + A[1] * v[1], nb = A[2] * v[0] + A[3] * v[1];
      v[0] = nl;
      v[1] = nb;
    }
  }
  for (int i = 0; i < FM_MAX_METRIC_TYPES; ++i) {
    const fm_metric_rule &r = p->rules[i].bound ? p->rules[i] : p->rules[0];  // D16
    d->rules[i].threshold = r.threshold;
    d->rules[i].min_lower_bound = r.min_lower_bound;
    d->rules[i].bound = r.bound;
  }
}

int alloc_device_csr(fm_batch *b, uint64_t n_units, uint64_t n_values) {
  fm_ctx *ctx = b->ctx;
  FM_CUDA(ctx, cudaMalloc(&b->d_values, n_values * sizeof(float)));
  FM_CUDA(ctx, cudaMalloc(&b->d_offsets, (n_units + 1) * sizeof(int64_t)));
  FM_CUDA(ctx, cudaMalloc(&b->d_seg, (3 * n_units + 1) * sizeof(uint32_t)));
  FM_CUDA(ctx, cudaMalloc(&b->d_sid, (n_units + 1) * sizeof(uint32_t)));
  FM_CUDA(ctx, cudaMalloc(&b->d_mt, n_units + 1));
  b->n_units = n_units;
  b->n_values = n_values;
  return FM_OK;
}

// upload a host CSR whose offsets are already multiples of 4
int upload_aligned(fm_batch *b, const fm_csr *h, uint64_t used) {
  fm_ctx *ctx = b->ctx;
  const uint64_t n = h->n_units;
  int rc = alloc_device_csr(b, n, used + 32);
  if (rc) return rc;
  cudaStream_t st = ctx->stream;
  if (used) FM_CUDA(ctx, cudaMemcpyAsync(b->d_values, h->values, used * sizeof(float), cudaMemcpyHostToDevice, st));
  FM_CUDA(ctx, cudaMemsetAsync(b->d_values + used, 0, 32 * sizeof(float), st));
  if (n && h->offsets)
    FM_CUDA(ctx, cudaMemcpyAsync(b->d_offsets, h->offsets, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, st));
  else  // an empty batch may come without arrays: its one offset is 0
    FM_CUDA(ctx, cudaMemsetAsync(b->d_offsets, 0, sizeof(int64_t), st));
  if (n) {
    FM_CUDA(ctx, cudaMemcpyAsync(b->d_seg, h->seg_lens, 3 * n * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    if (h->service_id)
      FM_CUDA(ctx, cudaMemcpyAsync(b->d_sid, h->service_id, n * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    else
      FM_CUDA(ctx, cudaMemsetAsync(b->d_sid, 0, n * sizeof(uint32_t), st));
    if (h->metric_type)
      FM_CUDA(ctx, cudaMemcpyAsync(b->d_mt, h->metric_type, n, cudaMemcpyHostToDevice, st));
    else
      FM_CUDA(ctx, cudaMemsetAsync(b->d_mt, 0, n, st));
  }
  FM_CUDA(ctx, cudaStreamSynchronize(st));  // inputs are borrowed for the call only
  return FM_OK;
}

// host-side scan of a CSR: validity, alignment, maxima
int scan_host_csr(fm_ctx *ctx, const fm_csr *h, bool *aligned, uint64_t *used, fm_batch *b) {
  *aligned = true;
  uint32_t mh = 0, mb = 0, mc = 0, ms = 0, kw = 0, km = 0;
  for (uint64_t u = 0; u < h->n_units; ++u) {
    const int64_t o = h->offsets[u], o1 = h->offsets[u + 1];
    const uint64_t len = (uint64_t)h->seg_lens[3 * u] + h->seg_lens[3 * u + 1] + h->seg_lens[3 * u + 2];
    if (o < 0 || o1 < o || (uint64_t)(o1 - o) < len) return set_err(ctx, FM_E_INVALID, "CSR offsets inconsistent with seg_lens");
    if (o & 3) *aligned = false;
    mh = std::max(mh, h->seg_lens[3 * u]);
    mb = std::max(mb, h->seg_lens[3 * u + 1]);
    mc = std::max(mc, h->seg_lens[3 * u + 2]);
    kw |= 1u << fm_rank_class(std::min(h->seg_lens[3 * u + 1], h->seg_lens[3 * u + 2]));
    km |= 1u << fm_rank_class(std::max(h->seg_lens[3 * u + 1], h->seg_lens[3 * u + 2]));
    if (h->service_id) ms = std::max(ms, h->service_id[u]);
  }
  b->rank_mask_w = kw; b->rank_mask_m = km;
  *used = h->n_units ? (uint64_t)h->offsets[h->n_units] : 0;
  if (h->n_units && (h->offsets[h->n_units] & 3)) *used = (*used + 3) & ~3ull;
  b->max_hist = mh; b->max_base = mb; b->max_cur = mc;
  b->n_services = h->n_units ? ms + 1 : 0;
  return FM_OK;
}

void free_batch_device(fm_batch *b) {
  if (b->owns) {
    cudaFree(b->d_values); cudaFree(b->d_offsets); cudaFree(b->d_seg); cudaFree(b->d_sid); cudaFree(b->d_mt);
  }
  b->d_values = nullptr; b->d_offsets = nullptr; b->d_seg = nullptr; b->d_sid = nullptr; b->d_mt = nullptr;
}

}  // namespace

std::mutex &fm_ctx_mutex(fm_ctx *ctx) { return ctx->mu; }
fm_comm_state *fm_ctx_comm(fm_ctx *ctx) { return &ctx->comm; }
int fm_ctx_device(const fm_ctx *ctx) { return ctx->device; }
cudaStream_t fm_ctx_stream(const fm_ctx *ctx) { return ctx->stream; }
int fm_set_error(fm_ctx *ctx, int code, const std::string &msg) { return set_err(ctx, code, msg); }

// ------------------------------------------------------------------ life cycle

extern "C" {

int fm_version(void) { return FM_ABI_VERSION; }

const char *fm_strerror(int code) {
  switch (code) {
    case FM_OK: return "ok";
    case FM_E_INVALID: return "invalid argument";
    case FM_E_NOMEM: return "out of memory";
    case FM_E_CUDA: return "CUDA error";
    case FM_E_NO_DEVICE: return "no sm_100 GPU available (libforemast_gpu has no CPU path)";
    case FM_E_ALIGN: return "device CSR offsets must be multiples of 4 elements and values 16-byte aligned";
    case FM_E_STATE: return "call order violated";
    case FM_E_RANGE: return "index or size out of range";
    case FM_E_UNSUPPORTED: return "unsupported request";
    case FM_E_COMM: return "multi-GPU exchange failed";
    case FM_E_TOO_LARGE: return "more items than this fixed-capacity entry point can take (see its limit in foremast_gpu.h)";
    default: return "unknown error";
  }
}

// The text is copied under the error lock into a buffer of the CALLING thread, so the pointer stays valid
// (until this thread asks again) even while other threads fail on the same context.
const char *fm_last_error(const fm_ctx *ctx) {
  if (!ctx) return "no context";
  static thread_local std::string copy;
  fm_ctx *c = const_cast<fm_ctx *>(ctx);
  std::lock_guard<std::mutex> g(c->err_mu);
  copy = c->last_error;
  return copy.c_str();
}

int fm_last_error_copy(const fm_ctx *ctx, char *buf, uint32_t cap) {
  if (!ctx || !buf || cap == 0) return bad_arg(ctx, __func__);
  fm_ctx *c = const_cast<fm_ctx *>(ctx);
  std::lock_guard<std::mutex> g(c->err_mu);
  snprintf(buf, cap, "%s", c->last_error.c_str());
  return FM_OK;
}

// Documented defaults: deploy/foremast/3_brain/foremast-brain.yaml:24-79; the rest oracle/DECISIONS.md.
void fm_params_default(fm_params *p) {
  if (!p) return;
  memset(p, 0, sizeof(*p));
  p->algorithms = FM_ALG_MA_ALL;
  p->pairwise = FM_PW_ALL;
  p->pairwise_threshold = 0.05;
  p->differ_scale = 0.5;
  p->min_mann_white = 20;
  p->min_wilcoxon = 20;
  p->min_kruskal = 5;
  p->min_historical = 10;
  p->min_anomaly_points = 1;
  p->ma_window = 30;
  p->period = 1440;
  p->alpha = 0.3;
  p->beta = 0.1;
  p->gamma = 0.1;
  p->rules[FM_MT_OTHER] = fm_metric_rule{2.0, 0.0, FM_BOUND_UPPER, 0};
  p->rules[FM_MT_ERROR5XX] = fm_metric_rule{2.0, 0.0, FM_BOUND_UPPER, 0};
  p->rules[FM_MT_ERROR4XX] = fm_metric_rule{3.0, 0.0, FM_BOUND_UPPER, 0};
  p->rules[FM_MT_LATENCY] = fm_metric_rule{10.0, 0.0, FM_BOUND_BOTH, 0};
  p->rules[FM_MT_CPU] = fm_metric_rule{5.0, 0.0, FM_BOUND_UPPER, 0};
  p->rules[FM_MT_MEMORY] = fm_metric_rule{5.0, 0.0, FM_BOUND_UPPER, 0};
}

int fm_init(const fm_config *cfg, fm_ctx **out) {
  if (!out) return FM_E_INVALID;
  *out = nullptr;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
    cudaGetLastError();
    return FM_E_NO_DEVICE;
  }
  int dev = cfg ? cfg->device : -1;
  if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) return FM_E_NO_DEVICE;
  if (dev >= count) return FM_E_NO_DEVICE;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return FM_E_NO_DEVICE;
  if (prop.major != 10) return FM_E_NO_DEVICE;  // kernels are built for sm_100a only
  fm_ctx *ctx = new (std::nothrow) fm_ctx;
  if (!ctx) return FM_E_NOMEM;
  ctx->device = dev;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->hbm_bytes = prop.totalGlobalMem;
  snprintf(ctx->name, sizeof(ctx->name), "%s", prop.name);
  if (cudaSetDevice(dev) != cudaSuccess || cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) {
    delete ctx;
    return FM_E_CUDA;
  }
  // keep freed result buffers in the stream-ordered pool: steady-state fm_score allocates nothing new
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
    uint64_t thresh = UINT64_MAX;
    cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh);
  }
  *out = ctx;
  return FM_OK;
}

void fm_shutdown(fm_ctx *ctx) {
  if (!ctx) return;
  fm_comm_destroy(ctx);
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->arena_dev) cudaFree(ctx->arena_dev);
  if (ctx->arena_host) cudaFreeHost(ctx->arena_host);
  if (ctx->arena_done) cudaEventDestroy(ctx->arena_done);
  for (auto &e : ctx->timer)
    if (e) cudaEventDestroy(e);
  cudaStreamDestroy(ctx->stream);
  cudaStreamDestroy(ctx->copy_stream);
  delete ctx;
}

int fm_device_info(const fm_ctx *ctx, int32_t *sm_count, uint64_t *hbm_bytes, char *name, uint32_t name_cap) {
  if (!ctx) return bad_arg(ctx, __func__);
  if (sm_count) *sm_count = ctx->sm_count;
  if (hbm_bytes) *hbm_bytes = ctx->hbm_bytes;
  if (name && name_cap) snprintf(name, name_cap, "%s", ctx->name);
  return FM_OK;
}

int fm_timer_start(fm_ctx *ctx) {
  if (!ctx) return bad_arg(ctx, __func__);
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  for (auto &e : ctx->timer)
    if (!e) FM_CUDA(ctx, cudaEventCreate(&e));
  FM_CUDA(ctx, cudaEventRecord(ctx->timer[0], ctx->stream));
  return FM_OK;
}

int fm_timer_stop(fm_ctx *ctx, float *ms) {
  if (!ctx || !ms || !ctx->timer[0]) return bad_arg(ctx, __func__);
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  FM_CUDA(ctx, cudaEventRecord(ctx->timer[1], ctx->stream));
  FM_CUDA(ctx, cudaEventSynchronize(ctx->timer[1]));
  FM_CUDA(ctx, cudaEventElapsedTime(ms, ctx->timer[0], ctx->timer[1]));
  return FM_OK;
}

int fm_host_alloc(fm_ctx *ctx, uint64_t bytes, void **out) {
  if (!ctx || !out) return bad_arg(ctx, __func__);
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  FM_CUDA(ctx, cudaMallocHost(out, bytes ? bytes : 1));
  return FM_OK;
}

void fm_host_free(fm_ctx *ctx, void *p) {
  if (!ctx || !p) return;
  std::lock_guard<std::mutex> g(ctx->mu);
  cudaSetDevice(ctx->device);
  cudaFreeHost(p);
}

// ------------------------------------------------------------------ ingest

int fm_batch_begin(fm_ctx *ctx, uint64_t n_units_hint, uint64_t n_points_hint, fm_batch **out) {
  if (!ctx || !out) return bad_arg(ctx, __func__);
  fm_batch *b = new (std::nothrow) fm_batch;
  if (!b) return FM_E_NOMEM;
  b->ctx = ctx;
  try {
    b->h_values.reserve(n_points_hint + 4 * n_units_hint + 32);
    b->h_offsets.reserve(n_units_hint + 1);
    b->h_seg.reserve(3 * n_units_hint);
    b->h_sid.reserve(n_units_hint);
    b->h_mt.reserve(n_units_hint);
  } catch (...) {
    delete b;
    return FM_E_NOMEM;
  }
  *out = b;
  return FM_OK;
}

int fm_batch_add_series(fm_batch *b, const fm_series_meta *meta, const float *hist, uint32_t n_hist,
                        const float *base, uint32_t n_base, const float *cur, uint32_t n_cur) {
  if (!b || !meta || (n_hist && !hist) || (n_base && !base) || (n_cur && !cur)) return bad_arg(b ? b->ctx : nullptr, __func__);
  if (b->sealed) return set_err(b->ctx, FM_E_STATE, "batch already sealed");
  try {
    b->h_offsets.push_back((int64_t)b->h_values.size());
    const float *seg[3] = {hist, base, cur};
    const uint32_t len[3] = {n_hist, n_base, n_cur};
    for (int s = 0; s < 3; ++s) {
      uint32_t kept = 0;
      for (uint32_t i = 0; i < len[s]; ++i) {
        const float x = seg[s][i];
        if (!std::isfinite(x)) continue;       // a gap (D11)
        b->h_values.push_back(x + 0.0f);       // -0.0 -> +0.0
        ++kept;
      }
      b->h_seg.push_back(kept);
    }
    while (b->h_values.size() & 3) b->h_values.push_back(0.0f);
    b->h_sid.push_back(meta->service_id);
    b->h_mt.push_back(meta->metric_type);
  } catch (...) {
    return set_err(b->ctx, FM_E_NOMEM, "host staging allocation failed");
  }
  return FM_OK;
}

}  // extern "C"

int fm_batch_reserve_units(fm_batch *b, uint64_t n, const uint32_t *seg_lens, const uint32_t *service_id,
                           const uint8_t *metric_type, float **values_out, uint64_t *rel_off) {
  if (!b || (n && (!seg_lens || !values_out || !rel_off))) return bad_arg(b ? b->ctx : nullptr, __func__);
  if (b->sealed) return set_err(b->ctx, FM_E_STATE, "batch already sealed");
  try {
    const uint64_t start = b->h_values.size();
    uint64_t pos = start;
    for (uint64_t k = 0; k < n; ++k) {
      rel_off[k] = pos - start;
      b->h_offsets.push_back((int64_t)pos);
      for (int s = 0; s < 3; ++s) b->h_seg.push_back(seg_lens[3 * k + s]);
      b->h_sid.push_back(service_id ? service_id[k] : 0u);
      b->h_mt.push_back(metric_type ? metric_type[k] : (uint8_t)0);
      pos += ((uint64_t)seg_lens[3 * k] + seg_lens[3 * k + 1] + seg_lens[3 * k + 2] + 3) & ~3ull;
    }
    b->h_values.resize(pos);  // NOT zero-filled (DefaultInitAlloc): the caller writes every sample and the padding
    *values_out = b->h_values.data() + start;
  } catch (...) {
    return set_err(b->ctx, FM_E_NOMEM, "host staging allocation failed");
  }
  return FM_OK;
}

extern "C" {

int fm_batch_seal(fm_batch *b) {
  if (!b) return bad_arg(b ? b->ctx : nullptr, __func__);
  if (b->sealed) return set_err(b->ctx, FM_E_STATE, "batch already sealed");
  fm_ctx *ctx = b->ctx;
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  b->h_offsets.push_back((int64_t)b->h_values.size());  // the end offset; taken back if the upload fails
  fm_csr h;
  h.n_units = b->h_sid.size();
  h.values = b->h_values.data();
  h.offsets = b->h_offsets.data();
  h.seg_lens = b->h_seg.data();
  h.service_id = b->h_sid.data();
  h.metric_type = b->h_mt.data();
  bool aligned;
  uint64_t used;
  int rc = scan_host_csr(ctx, &h, &aligned, &used, b);
  if (rc == FM_OK) rc = upload_aligned(b, &h, used);
  if (rc) {  // the batch stays open and consistent: a retry, or more add_series calls, start from a clean CSR
    free_batch_device(b);
    b->h_offsets.pop_back();
    return rc;
  }
  b->sealed = true;
  std::vector<float, DefaultInitAlloc<float>>().swap(b->h_values);
  return FM_OK;
}

int fm_batch_from_csr(fm_ctx *ctx, const fm_csr *h, fm_batch **out) {
  if (!ctx || !h || !out || (h->n_units && (!h->values || !h->offsets || !h->seg_lens))) return bad_arg(ctx, __func__);
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  fm_batch *b = new (std::nothrow) fm_batch;
  if (!b) return FM_E_NOMEM;
  b->ctx = ctx;
  bool aligned;
  uint64_t used;
  int rc = scan_host_csr(ctx, h, &aligned, &used, b);
  if (rc == FM_OK) {
    if (aligned) {
      rc = upload_aligned(b, h, h->n_units ? (uint64_t)h->offsets[h->n_units] : 0);
    } else {  // re-pack so that every unit starts on a 16-byte boundary
      try {
        std::vector<float> v;
        std::vector<int64_t> o(h->n_units + 1);
        v.reserve(used + 4 * h->n_units);
        for (uint64_t u = 0; u < h->n_units; ++u) {
          o[u] = (int64_t)v.size();
          const uint64_t len = (uint64_t)h->seg_lens[3 * u] + h->seg_lens[3 * u + 1] + h->seg_lens[3 * u + 2];
          v.insert(v.end(), h->values + h->offsets[u], h->values + h->offsets[u] + len);
          while (v.size() & 3) v.push_back(0.0f);
        }
        o[h->n_units] = (int64_t)v.size();
        fm_csr h2 = *h;
        h2.values = v.data();
        h2.offsets = o.data();
        rc = upload_aligned(b, &h2, v.size());
      } catch (...) {
        rc = set_err(ctx, FM_E_NOMEM, "host re-pack allocation failed");
      }
    }
  }
  if (rc) { free_batch_device(b); delete b; return rc; }
  b->sealed = true;
  *out = b;
  return FM_OK;
}

int fm_batch_from_device(fm_ctx *ctx, const fm_csr *d, fm_batch **out) {
  if (!ctx || !d || !out || !d->values || !d->offsets || !d->seg_lens || !d->service_id || !d->metric_type) return bad_arg(ctx, __func__);
  if (reinterpret_cast<uintptr_t>(d->values) & 15) return set_err(ctx, FM_E_ALIGN, "values pointer is not 16-byte aligned");
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  fm_batch *b = new (std::nothrow) fm_batch;
  if (!b) return FM_E_NOMEM;
  b->ctx = ctx;
  b->owns = false;
  b->n_units = d->n_units;
  b->d_values = const_cast<float *>(d->values);
  b->d_offsets = const_cast<int64_t *>(d->offsets);
  b->d_seg = const_cast<uint32_t *>(d->seg_lens);
  b->d_sid = const_cast<uint32_t *>(d->service_id);
  b->d_mt = const_cast<uint8_t *>(d->metric_type);
  uint32_t *d_stats = nullptr, h_stats[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  cudaError_t e = cudaMallocAsync(&d_stats, sizeof(h_stats), ctx->stream);
  if (e == cudaSuccess) e = cudaMemsetAsync(d_stats, 0, sizeof(h_stats), ctx->stream);
  if (e == cudaSuccess) {
    fm_launch_batch_stats(dev_view(b), d_stats, ctx->stream);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(h_stats, d_stats, sizeof(h_stats), cudaMemcpyDeviceToHost, ctx->stream);
  int64_t end_off = 0;
  if (e == cudaSuccess) e = cudaMemcpyAsync(&end_off, d->offsets + d->n_units, sizeof(end_off), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  b->n_values = (uint64_t)(end_off > 0 ? end_off : 0) + 32;  // the ABI requires 32 readable elements of slack
  if (d_stats) cudaFreeAsync(d_stats, ctx->stream);
  if (e != cudaSuccess) {
    delete b;
    cudaGetLastError();
    return set_err(ctx, FM_E_CUDA, std::string("batch statistics: ") + cudaGetErrorString(e));
  }
  if (h_stats[4]) { delete b; return set_err(ctx, FM_E_ALIGN, "device CSR offsets must be multiples of 4 elements"); }
  if (h_stats[5]) { delete b; return set_err(ctx, FM_E_INVALID, "device CSR units overlap"); }
  b->max_hist = h_stats[0]; b->max_base = h_stats[1]; b->max_cur = h_stats[2];
  b->rank_mask_w = h_stats[6]; b->rank_mask_m = h_stats[7];
  b->n_services = d->n_units ? h_stats[3] + 1 : 0;
  b->sealed = true;
  *out = b;
  return FM_OK;
}

int fm_batch_synth(fm_ctx *ctx, const fm_synth_config *cfg, fm_batch **out) {
  if (!ctx || !cfg || !out || cfg->n_metric_types == 0 || cfg->n_metric_types > 8 || cfg->units_per_service == 0) return bad_arg(ctx, __func__);
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  fm_batch *b = new (std::nothrow) fm_batch;
  if (!b) return FM_E_NOMEM;
  b->ctx = ctx;
  const uint64_t stride = ((uint64_t)cfg->n_hist + cfg->n_base + cfg->n_cur + 3) & ~3ull;
  int rc = alloc_device_csr(b, cfg->n_units, cfg->n_units * stride + 32);
  if (rc) { free_batch_device(b); delete b; return rc; }
  fm_launch_synth(*cfg, stride, b->d_values, b->d_offsets, b->d_seg, b->d_sid, b->d_mt, ctx->stream);
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
    free_batch_device(b); delete b;
    cudaGetLastError();
    return set_err(ctx, FM_E_CUDA, std::string("synthetic generator: ") + cudaGetErrorString(e));
  }
  b->max_hist = cfg->n_hist; b->max_base = cfg->n_base; b->max_cur = cfg->n_cur;
  b->first_unit = cfg->first_unit;
  b->n_services = cfg->n_units ? (uint32_t)((cfg->first_unit + cfg->n_units - 1) / cfg->units_per_service) + 1 : 0;
  b->sealed = true;
  *out = b;
  return FM_OK;
}

int fm_batch_set_shard(fm_batch *b, uint64_t first_unit, uint32_t n_services) {
  if (!b) return bad_arg(b ? b->ctx : nullptr, __func__);
  if (n_services < b->n_services) return set_err(b->ctx, FM_E_RANGE, "n_services smaller than the largest service id + 1");
  b->first_unit = first_unit;
  b->n_services = n_services;
  return FM_OK;
}

int fm_batch_info(const fm_batch *b, uint64_t *n_units, uint64_t *n_values, uint32_t *n_services) {
  if (!b) return bad_arg(b ? b->ctx : nullptr, __func__);
  if (n_units) *n_units = b->n_units;
  if (n_values) *n_values = b->n_values;
  if (n_services) *n_services = b->n_services;
  return FM_OK;
}

int fm_batch_download(const fm_batch *b, float *values, int64_t *offsets, uint32_t *seg_lens, uint32_t *service_id, uint8_t *metric_type) {
  if (!b || !b->sealed) return bad_arg(b ? b->ctx : nullptr, __func__);
  fm_ctx *ctx = b->ctx;
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  const uint64_t n = b->n_units;
  if (values && b->n_values) FM_CUDA(ctx, cudaMemcpy(values, b->d_values, b->n_values * sizeof(float), cudaMemcpyDeviceToHost));
  if (offsets) FM_CUDA(ctx, cudaMemcpy(offsets, b->d_offsets, (n + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost));
  if (seg_lens && n) FM_CUDA(ctx, cudaMemcpy(seg_lens, b->d_seg, 3 * n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  if (service_id && n) FM_CUDA(ctx, cudaMemcpy(service_id, b->d_sid, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  if (metric_type && n) FM_CUDA(ctx, cudaMemcpy(metric_type, b->d_mt, n, cudaMemcpyDeviceToHost));
  return FM_OK;
}

void fm_batch_free(fm_batch *b) {
  if (!b) return;
  if (b->ctx) {
    std::lock_guard<std::mutex> g(b->ctx->mu);
    cudaSetDevice(b->ctx->device);
    cudaStreamSynchronize(b->ctx->stream);
    free_batch_device(b);
  }
  delete b;
}

// ------------------------------------------------------------------ scoring

constexpr uint64_t kMinOrderedUnits = 131072;  // the order costs ~10 us per launch (its kernel + one more dependent launch): not worth it on small batches

static void result_release(fm_result *r) {
  if (!r->ctx) { delete r; return; }
  cudaStream_t st = r->ctx->stream;
  void *ptrs[] = {r->d_params, r->d_scores, r->d_unit_status, r->d_unit_points, r->d_counts, r->d_order, r->d_rank_classes,
                  r->d_worst, r->d_verdicts, r->d_band_up, r->d_band_lo, r->d_anom_off, r->d_anom_idx, r->d_anom_val};
  for (void *p : ptrs)
    if (p) cudaFreeAsync(p, st);
  for (auto &e : r->ev)
    if (e) cudaEventDestroy(e);
  delete r;
}

static int validate_params(fm_ctx *ctx, const fm_params *p) {
  if (p->algorithms & ~FM_ALG_ALL) return set_err(ctx, FM_E_INVALID, "unknown algorithm bits");
  if ((p->algorithms & FM_ALG_HW) && (p->period < 32 || p->period > 2048))
    return set_err(ctx, FM_E_UNSUPPORTED, "Holt-Winters period must be between 32 and 2048 samples");
  if (p->pairwise > FM_PW_KRUSKAL) return set_err(ctx, FM_E_INVALID, "unknown pairwise algorithm");
  if (p->band_algorithm && (!(p->band_algorithm & p->algorithms) || (p->band_algorithm & (p->band_algorithm - 1))))
    return set_err(ctx, FM_E_INVALID, "band_algorithm must be one enabled FM_ALG_* bit");
  return FM_OK;
}

#define FM_R(expr)                                                                              \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      cudaGetLastError();                                                                       \
      return set_err(ctx, _e == cudaErrorMemoryAllocation ? FM_E_NOMEM : FM_E_CUDA,             \
                     std::string(#expr) + ": " + cudaGetErrorString(_e));                       \
    }                                                                                           \
  } while (0)

// result buffers for a batch of n units / ns services, cleared, parameters resolved and uploaded
static int result_alloc(fm_ctx *ctx, const fm_batch *b, const fm_params *p, fm_result *r) {
  r->ctx = ctx;
  r->batch = b;
  r->n_units = b->n_units;
  r->n_services = b->n_services;
  r->band_algorithm = p->band_algorithm;
  r->params = *p;
  cudaStream_t st = ctx->stream;
  const uint64_t n = b->n_units, ns = b->n_services;
  for (auto &e : r->ev) FM_R(cudaEventCreate(&e));
  FM_R(cudaMallocAsync(&r->d_params, sizeof(DevParams), st));
  FM_R(cudaMallocAsync(&r->d_scores, (n + 1) * sizeof(fm_series_score), st));
  FM_R(cudaMallocAsync(&r->d_unit_status, (n + 1) * sizeof(uint32_t), st));
  FM_R(cudaMallocAsync(&r->d_unit_points, (n + 1) * sizeof(uint32_t), st));
  static const bool no_order = getenv("FOREMAST_GPU_NO_ORDER") != nullptr;  // A/B switch for scripts/ragged_time.py
  if (n >= kMinOrderedUnits && !no_order) FM_R(cudaMallocAsync(&r->d_order, (n + 1) * sizeof(uint32_t), st));
  const bool mixed = ((b->rank_mask_w & (b->rank_mask_w - 1u)) | (b->rank_mask_m & (b->rank_mask_m - 1u))) != 0;
  if (mixed && p->pairwise != FM_PW_NONE) FM_R(cudaMallocAsync(&r->d_rank_classes, (2 * n + 64) * sizeof(uint32_t), st));
  FM_R(cudaMallocAsync(&r->d_counts, (4 * ns + 4) * sizeof(uint32_t), st));
  FM_R(cudaMallocAsync(&r->d_worst, (ns + 1) * sizeof(unsigned long long), st));
  FM_R(cudaMallocAsync(&r->d_verdicts, (ns + 1) * sizeof(fm_verdict), st));
  if (p->band_algorithm) {
    FM_R(cudaMallocAsync(&r->d_band_up, b->n_values * sizeof(float), st));
    FM_R(cudaMallocAsync(&r->d_band_lo, b->n_values * sizeof(float), st));
  }
  FM_R(cudaEventRecord(r->ev[0], st));
  DevParams hp;
  resolve_params(p, &hp);
  FM_R(cudaMemcpyAsync(r->d_params, &hp, sizeof(hp), cudaMemcpyHostToDevice, st));  // pageable: staged before return
  FM_R(cudaMemsetAsync(r->d_scores, 0, (n + 1) * sizeof(fm_series_score), st));
  FM_R(cudaMemsetAsync(r->d_counts, 0, (4 * ns + 4) * sizeof(uint32_t), st));
  FM_R(cudaMemsetAsync(r->d_worst, 0, (ns + 1) * sizeof(unsigned long long), st));
  return FM_OK;
}

// launch the scoring kernels over units [i0, i1) of the batch (results land at the same indices)
static int launch_range(fm_ctx *ctx, fm_result *r, const fm_batch *b, const fm_params *p, uint64_t i0, uint64_t i1,
                        bool timed, bool small = false) {
  cudaStream_t st = ctx->stream;
  const bool do_rank = p->pairwise != FM_PW_NONE && b->max_base > 0 && b->max_cur > 0;
  DevBatch db = dev_view(b);
  db.n_units = i1 - i0;
  db.offsets += i0;
  db.seg_lens += 3 * i0;
  db.service_id += i0;
  db.metric_type += i0;
  db.first_unit += i0;
  fm_series_score *scores = r->d_scores + i0;
  uint32_t *ustat = r->d_unit_status + i0, *upts = r->d_unit_points + i0;
  if (timed) FM_R(cudaEventRecord(r->ev[1], st));
  if (do_rank)
    // (class scratch: 64 words of class sizes and cursors, then 2 n unit indices; a range uses the 2 (i1 - i0) at 2 i0)
    FM_R(fm_launch_rank(db, r->d_params, p->pairwise, b->max_base, b->max_cur, scores, ctx->sm_count, st,
                        timed ? r->ev[2] : nullptr, b->rank_mask_w, b->rank_mask_m, r->d_rank_classes,
                        r->d_rank_classes ? r->d_rank_classes + 64 + 2 * i0 : nullptr));
  else if (timed)
    FM_R(cudaEventRecord(r->ev[2], st));
  if (timed) FM_R(cudaEventRecord(r->ev[3], st));
  const bool band_in_kband = (p->band_algorithm & (FM_ALG_MA_ALL | FM_ALG_MA_ROLLING | FM_ALG_EWMA | FM_ALG_DES)) != 0;
  if (r->d_order && i1 - i0 >= kMinOrderedUnits && (p->algorithms & (FM_ALG_MA_ALL | FM_ALG_MA_ROLLING | FM_ALG_EWMA | FM_ALG_DES))) {
    FM_R(fm_launch_band_order(db, r->d_order + i0, st));  // ragged batches: about equally long units per warp
    db.order = r->d_order + i0;
  }
  FM_R(fm_launch_band(db, r->d_params, p->algorithms, do_rank, scores, ustat, upts,
                      band_in_kband ? r->d_band_up : nullptr, band_in_kband ? r->d_band_lo : nullptr, st));
  if (p->algorithms & FM_ALG_HW)
    FM_R(fm_launch_hw(db, r->d_params, p->period, scores, ustat, upts,
                      p->band_algorithm == FM_ALG_HW ? r->d_band_up : nullptr,
                      p->band_algorithm == FM_ALG_HW ? r->d_band_lo : nullptr, ctx->sm_count, st));
  if (timed) FM_R(cudaEventRecord(r->ev[4], st));
  if (small) fm_launch_verdict_small(db, ustat, upts, r->n_services, r->d_counts, r->d_worst, r->d_verdicts, st);
  else fm_launch_verdict_reduce(db, ustat, upts, r->n_services, r->d_counts, r->d_worst, st);
  FM_R(cudaGetLastError());
  if (timed) FM_R(cudaEventRecord(r->ev[5], st));
  return FM_OK;
}
#undef FM_R

static int score_locked(fm_ctx *ctx, const fm_batch *b, const fm_params *p, fm_result **out) {
  int rc = validate_params(ctx, p);
  if (rc) return rc;
  fm_result *r = new (std::nothrow) fm_result;
  if (!r) return FM_E_NOMEM;
  rc = result_alloc(ctx, b, p, r);
  if (rc == FM_OK) rc = launch_range(ctx, r, b, p, 0, b->n_units, true);
  if (rc) { result_release(r); return rc; }
  *out = r;
  return FM_OK;
}

int fm_score(fm_ctx *ctx, const fm_batch *b, const fm_params *p, fm_result **out) {
  if (!ctx || !b || !p || !out) return bad_arg(ctx, __func__);
  if (!b->sealed) return set_err(ctx, FM_E_STATE, "batch is not sealed");
  if (b->ctx != ctx) return set_err(ctx, FM_E_INVALID, "batch belongs to another context");
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  return score_locked(ctx, b, p, out);
}

// ---- continuous-mode model cache (Holt-Winters state kept in HBM between cycles)

int fm_model_fit(fm_ctx *ctx, const fm_batch *b, const fm_params *p, fm_model **out) {
  if (!ctx || !b || !p || !out) return bad_arg(ctx, __func__);
  if (!b->sealed || b->ctx != ctx) return set_err(ctx, FM_E_STATE, "batch is not a sealed batch of this context");
  if (p->period < 32 || p->period > 2048) return set_err(ctx, FM_E_UNSUPPORTED, "Holt-Winters period must be between 32 and 2048 samples");
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  fm_model *m = new (std::nothrow) fm_model;
  if (!m) return FM_E_NOMEM;
  m->ctx = ctx;
  m->n_units = b->n_units;
  m->params = *p;
  m->params.algorithms = FM_ALG_HW;
  m->params.pairwise = FM_PW_NONE;
  m->params.band_algorithm = 0;
  cudaError_t e = cudaMalloc(&m->d_params, sizeof(DevParams));
  if (e == cudaSuccess) e = cudaMalloc(&m->d_state, (b->n_units * (uint64_t)(p->period + 4) + 1) * sizeof(double));
  DevParams hp;
  resolve_params(&m->params, &hp);
  if (e == cudaSuccess) e = cudaMemcpyAsync(m->d_params, &hp, sizeof(hp), cudaMemcpyHostToDevice, ctx->stream);
  if (e == cudaSuccess) e = fm_launch_hw_fit(dev_view(b), m->d_params, p->period, m->d_state, ctx->sm_count, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  if (e != cudaSuccess) {
    cudaGetLastError();
    cudaFree(m->d_params); cudaFree(m->d_state);
    delete m;
    return set_err(ctx, e == cudaErrorMemoryAllocation ? FM_E_NOMEM : FM_E_CUDA, std::string("model fit: ") + cudaGetErrorString(e));
  }
  *out = m;
  return FM_OK;
}

int fm_model_score(fm_ctx *ctx, fm_model *m, const fm_batch *b, fm_result **out) {
  if (!ctx || !m || !b || !out) return bad_arg(ctx, __func__);
  if (!b->sealed || b->ctx != ctx || m->ctx != ctx) return set_err(ctx, FM_E_STATE, "batch / model do not belong to this context");
  if (b->n_units != m->n_units) return set_err(ctx, FM_E_INVALID, "the current batch must hold the same units, in the same order, as the fitted one");
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  fm_result *r = new (std::nothrow) fm_result;
  if (!r) return FM_E_NOMEM;
  int rc = result_alloc(ctx, b, &m->params, r);
  r->from_model = true;
  cudaStream_t st = ctx->stream;
  cudaError_t e = cudaSuccess;
  if (rc == FM_OK) {
    DevBatch db = dev_view(b);
    for (int i = 1; i <= 3 && e == cudaSuccess; ++i) e = cudaEventRecord(r->ev[i], st);
    if (e == cudaSuccess) e = fm_launch_hw_resume(db, r->d_params, m->params.period, m->d_state, r->d_scores, r->d_unit_status,
                                                  r->d_unit_points, ctx->sm_count, st);
    if (e == cudaSuccess) e = cudaEventRecord(r->ev[4], st);
    if (e == cudaSuccess) {
      fm_launch_verdict_reduce(db, r->d_unit_status, r->d_unit_points, r->n_services, r->d_counts, r->d_worst, st);
      e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaEventRecord(r->ev[5], st);
    if (e != cudaSuccess) {
      cudaGetLastError();
      rc = set_err(ctx, FM_E_CUDA, std::string("model score: ") + cudaGetErrorString(e));
    }
  }
  if (rc) { result_release(r); return rc; }
  *out = r;
  return FM_OK;
}

void fm_model_free(fm_model *m) {
  if (!m) return;
  {
    std::lock_guard<std::mutex> g(m->ctx->mu);
    cudaSetDevice(m->ctx->device);
    cudaStreamSynchronize(m->ctx->stream);
    cudaFree(m->d_params);
    cudaFree(m->d_state);
  }
  delete m;
}

int fm_result_sync(fm_result *r) {
  if (!r) return bad_arg(r ? r->ctx : nullptr, __func__);
  fm_ctx *ctx = r->ctx;
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  FM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FM_OK;
}

int fm_result_verdicts(fm_result *r, fm_verdict *out, uint64_t cap, uint64_t *n) {
  if (!r || !out) return bad_arg(r ? r->ctx : nullptr, __func__);
  fm_ctx *ctx = r->ctx;
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (n) *n = r->n_services;
  if (cap < r->n_services) return set_err(ctx, FM_E_RANGE, "verdict buffer too small");
  if (r->n_services == 0) return FM_OK;
  // finalised at read time so that a cross-GPU merge of the counters can happen in between
  fm_launch_verdict_finalize(r->n_services, r->d_counts, r->d_worst, r->d_verdicts, 0u, ctx->stream);
  FM_CUDA(ctx, cudaGetLastError());
  FM_CUDA(ctx, cudaMemcpyAsync(out, r->d_verdicts, r->n_services * sizeof(fm_verdict), cudaMemcpyDeviceToHost, ctx->stream));
  FM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FM_OK;
}

// A slice of the verdict vector: after a cross-GPU merge every rank holds all counters, but a rank usually
// reports only the services of its own shard.
int fm_result_verdicts_range(fm_result *r, uint64_t first_service, uint64_t count, fm_verdict *out) {
  if (!r || (!out && count)) return bad_arg(r ? r->ctx : nullptr, __func__);
  fm_ctx *ctx = r->ctx;
  if (first_service > r->n_services || count > r->n_services - first_service)
    return set_err(ctx, FM_E_RANGE, "service range outside the result");
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (count == 0) return FM_OK;
  fm_launch_verdict_finalize((uint32_t)count, r->d_counts + 4 * first_service, r->d_worst + first_service,
                             r->d_verdicts + first_service, (uint32_t)first_service, ctx->stream);
  FM_CUDA(ctx, cudaGetLastError());
  FM_CUDA(ctx, cudaMemcpyAsync(out, r->d_verdicts + first_service, count * sizeof(fm_verdict), cudaMemcpyDeviceToHost, ctx->stream));
  FM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FM_OK;
}

int fm_result_series(fm_result *r, uint64_t first_unit, uint64_t count, fm_series_score *out) {
  if (!r || (!out && count)) return bad_arg(r ? r->ctx : nullptr, __func__);
  fm_ctx *ctx = r->ctx;
  if (first_unit > r->n_units || count > r->n_units - first_unit) return set_err(ctx, FM_E_RANGE, "unit range outside the batch");
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  if (count)
    FM_CUDA(ctx, cudaMemcpyAsync(out, r->d_scores + first_unit, count * sizeof(fm_series_score), cudaMemcpyDeviceToHost, ctx->stream));
  FM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FM_OK;
}

int fm_result_bands(fm_result *r, uint64_t unit, float *upper, float *lower, uint32_t cap, uint32_t *n) {
  if (!r || !upper || !lower) return bad_arg(r ? r->ctx : nullptr, __func__);
  fm_ctx *ctx = r->ctx;
  if (!r->d_band_up) return set_err(ctx, FM_E_STATE, "bands were not requested (params.band_algorithm == 0)");
  if (unit >= r->n_units) return set_err(ctx, FM_E_RANGE, "unit outside the batch");
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  int64_t off;
  uint32_t seg[3];
  FM_CUDA(ctx, cudaMemcpyAsync(&off, r->batch->d_offsets + unit, sizeof(off), cudaMemcpyDeviceToHost, ctx->stream));
  FM_CUDA(ctx, cudaMemcpyAsync(seg, r->batch->d_seg + 3 * unit, sizeof(seg), cudaMemcpyDeviceToHost, ctx->stream));
  FM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (n) *n = seg[2];
  if (cap < seg[2]) return set_err(ctx, FM_E_RANGE, "band buffer too small");
  const int64_t c0 = off + seg[0] + seg[1];
  if (seg[2]) {
    FM_CUDA(ctx, cudaMemcpyAsync(upper, r->d_band_up + c0, seg[2] * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    FM_CUDA(ctx, cudaMemcpyAsync(lower, r->d_band_lo + c0, seg[2] * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    FM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return FM_OK;
}

// ---- compact anomaly list (a10)

int fm_result_anomalies_build(fm_result *r, uint32_t algorithm, uint64_t *n_points) {
  if (!r) return bad_arg(r ? r->ctx : nullptr, __func__);
  fm_ctx *ctx = r->ctx;
  int a = -1;
  for (int i = 0; i < FM_N_ALG; ++i)
    if (algorithm == (1u << i)) a = i;
  if (a < 0 || !(r->params.algorithms & algorithm))
    return set_err(ctx, FM_E_INVALID, "algorithm must be one FM_ALG_* bit that was enabled in fm_score");
  if (r->from_model) return set_err(ctx, FM_E_UNSUPPORTED, "anomaly lists are not available for results of fm_model_score");
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  for (void *p : {(void *)r->d_anom_off, (void *)r->d_anom_idx, (void *)r->d_anom_val})
    if (p) cudaFreeAsync(p, st);
  r->d_anom_off = nullptr; r->d_anom_idx = nullptr; r->d_anom_val = nullptr;
  r->anom_total = 0; r->anom_alg = 0;
  const uint64_t n = r->n_units;
  unsigned long long *scratch = nullptr;
  FM_CUDA(ctx, cudaMallocAsync(&r->d_anom_off, (n + 1) * sizeof(unsigned long long), st));
  FM_CUDA(ctx, cudaMallocAsync(&scratch, fm_scan_scratch_bytes(n), st));
  cudaError_t e = fm_launch_scan_counts(&r->d_scores->n_anom[a], sizeof(fm_series_score), n, r->d_anom_off, scratch, st);
  unsigned long long total = 0;
  if (e == cudaSuccess) e = cudaMemcpyAsync(&total, r->d_anom_off + n, sizeof(total), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFreeAsync(scratch, st);
  if (e != cudaSuccess) { cudaGetLastError(); return set_err(ctx, FM_E_CUDA, std::string("anomaly offsets: ") + cudaGetErrorString(e)); }
  FM_CUDA(ctx, cudaMallocAsync(&r->d_anom_idx, (total + 1) * sizeof(uint32_t), st));
  FM_CUDA(ctx, cudaMallocAsync(&r->d_anom_val, (total + 1) * sizeof(float), st));
  if (total) {
    const DevBatch db = dev_view(r->batch);
    if (algorithm == FM_ALG_HW)
      FM_CUDA(ctx, fm_launch_hw_emit(db, r->d_params, r->params.period, r->d_scores, r->d_anom_off, r->d_anom_idx, r->d_anom_val, ctx->sm_count, st));
    else
      FM_CUDA(ctx, fm_launch_band_emit(db, r->d_params, algorithm, r->d_scores, r->d_anom_off, r->d_anom_idx, r->d_anom_val, st));
    FM_CUDA(ctx, cudaStreamSynchronize(st));
  }
  r->anom_total = total;
  r->anom_alg = algorithm;
  if (n_points) *n_points = total;
  return FM_OK;
}

int fm_result_anomalies_unit(fm_result *r, uint64_t unit, uint32_t *cur_idx, float *values, uint32_t cap, uint32_t *n) {
  if (!r || !n) return bad_arg(r ? r->ctx : nullptr, __func__);
  fm_ctx *ctx = r->ctx;
  if (!r->d_anom_off) return set_err(ctx, FM_E_STATE, "call fm_result_anomalies_build first");
  if (unit >= r->n_units) return set_err(ctx, FM_E_RANGE, "unit outside the batch");
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  unsigned long long lim[2];
  FM_CUDA(ctx, cudaMemcpyAsync(lim, r->d_anom_off + unit, sizeof(lim), cudaMemcpyDeviceToHost, ctx->stream));
  FM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const uint64_t cnt = lim[1] - lim[0];
  *n = (uint32_t)cnt;
  if (cnt > cap) return set_err(ctx, FM_E_RANGE, "anomaly buffer too small");
  if (cnt && cur_idx) FM_CUDA(ctx, cudaMemcpyAsync(cur_idx, r->d_anom_idx + lim[0], cnt * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  if (cnt && values) FM_CUDA(ctx, cudaMemcpyAsync(values, r->d_anom_val + lim[0], cnt * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  FM_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return FM_OK;
}

int fm_result_anomalies_all(fm_result *r, uint64_t *offsets, uint32_t *cur_idx, float *values, uint64_t cap) {
  if (!r) return bad_arg(r ? r->ctx : nullptr, __func__);
  fm_ctx *ctx = r->ctx;
  if (!r->d_anom_off) return set_err(ctx, FM_E_STATE, "call fm_result_anomalies_build first");
  if (cap < r->anom_total) return set_err(ctx, FM_E_RANGE, "anomaly buffer too small");
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  cudaStream_t st = ctx->stream;
  if (offsets) FM_CUDA(ctx, cudaMemcpyAsync(offsets, r->d_anom_off, (r->n_units + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
  if (r->anom_total && cur_idx) FM_CUDA(ctx, cudaMemcpyAsync(cur_idx, r->d_anom_idx, r->anom_total * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
  if (r->anom_total && values) FM_CUDA(ctx, cudaMemcpyAsync(values, r->d_anom_val, r->anom_total * sizeof(float), cudaMemcpyDeviceToHost, st));
  FM_CUDA(ctx, cudaStreamSynchronize(st));
  return FM_OK;
}

int fm_result_service_buffers(fm_result *r, uint32_t **counts, uint64_t **worst, uint32_t *n_services) {
  if (!r) return bad_arg(r ? r->ctx : nullptr, __func__);
  if (counts) *counts = r->d_counts;
  if (worst) *worst = reinterpret_cast<uint64_t *>(r->d_worst);
  if (n_services) *n_services = r->n_services;
  return FM_OK;
}

// Cross-GPU merge of the per-service counters, in place, ordered on the context's stream between the
// scoring kernels and the finalize pass of fm_result_verdicts: no host synchronisation in between.
int fm_result_merge(fm_result *r) {
  if (!r) return bad_arg(r ? r->ctx : nullptr, __func__);
  fm_ctx *ctx = r->ctx;
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  return fm_comm_merge_counters(ctx, r->d_counts, r->d_worst, r->n_services);
}

int fm_result_timings(fm_result *r, float *ms, uint32_t cap) {
  if (!r || !ms || cap < 5) return bad_arg(r ? r->ctx : nullptr, __func__);
  fm_ctx *ctx = r->ctx;
  std::lock_guard<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  FM_CUDA(ctx, cudaEventSynchronize(r->ev[5]));
  for (int i = 0; i < 5; ++i) FM_CUDA(ctx, cudaEventElapsedTime(&ms[i], r->ev[i], r->ev[i + 1]));
  return FM_OK;
}

void fm_result_free(fm_result *r) {
  if (!r) return;
  std::lock_guard<std::mutex> g(r->ctx->mu);
  cudaSetDevice(r->ctx->device);
  result_release(r);
}

// One-shot host-in / host-out scoring (the call a Go worker makes per sweep).  The upload is
// chunked on a second stream and each chunk is scored as soon as it has landed, so PCIe and the
// kernels overlap; with pinned input (fm_host_alloc) the call runs at PCIe speed.
int fm_score_host(fm_ctx *ctx, const fm_csr *h, const fm_params *p, uint32_t n_services, fm_verdict *verdicts) {
  if (!ctx || !h || !p || !verdicts || (h->n_units && (!h->values || !h->offsets || !h->seg_lens))) return bad_arg(ctx, __func__);
  std::unique_lock<std::mutex> g(ctx->mu);
  FM_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = validate_params(ctx, p);
  if (rc) return rc;
  fm_batch tb;  // transient, device arrays from the stream-ordered pool
  tb.ctx = ctx;
  tb.owns = false;
  bool aligned;
  uint64_t used;
  rc = scan_host_csr(ctx, h, &aligned, &used, &tb);
  if (rc) return rc;
  if (n_services < tb.n_services) return set_err(ctx, FM_E_RANGE, "n_services smaller than the largest service id + 1");
  tb.n_services = n_services;
  if (!aligned || h->n_units == 0) {
    // unaligned input has to be re-packed first: plain path
    g.unlock();
    fm_batch *b = nullptr;
    rc = fm_batch_from_csr(ctx, h, &b);
    if (rc) return rc;
    b->n_services = n_services;
    fm_result *r = nullptr;
    rc = fm_score(ctx, b, p, &r);
    if (rc == FM_OK) {
      uint64_t n = 0;
      rc = fm_result_verdicts(r, verdicts, n_services, &n);
      fm_result_free(r);
    }
    fm_batch_free(b);
    return rc;
  }
  const uint64_t n = h->n_units, n_vals = (uint64_t)h->offsets[n];
  cudaStream_t st = ctx->stream, cs = ctx->copy_stream;
  if (n <= kArenaUnits && n_vals <= kArenaValues && n_services <= kArenaServices && !p->band_algorithm) {
    // ---- small job (the reference's own shape: one deployment, a handful of series, 10-30 minute windows,
    //      foremast-barrelman/pkg/controller/DeploymentController.go:48): latency, not throughput
    if (!ctx->arena_dev) {
      FM_CUDA(ctx, cudaMalloc(&ctx->arena_dev, kArenaBytes));
      FM_CUDA(ctx, cudaMallocHost(&ctx->arena_host, kArenaBytes));
      FM_CUDA(ctx, cudaEventCreateWithFlags(&ctx->arena_done, cudaEventDisableTiming));
    }
    unsigned char *dv = ctx->arena_dev, *hs = ctx->arena_host;
    DevParams hp;
    resolve_params(p, &hp);
    uint64_t up0 = kOffOffsets;
    if (!ctx->arena_params_valid || memcmp(&hp, &ctx->arena_params, sizeof(hp)) != 0) {
      memcpy(hs + kOffParams, &hp, sizeof(hp));  // new parameters ride along with the samples
      ctx->arena_params = hp;
      ctx->arena_params_valid = true;
      up0 = kOffParams;
    }
    memcpy(hs + kOffValues, h->values, n_vals * 4);
    memset(hs + kOffValues + n_vals * 4, 0, 32 * 4);
    memcpy(hs + kOffOffsets, h->offsets, (n + 1) * 8);
    memcpy(hs + kOffSeg, h->seg_lens, 3 * n * 4);
    if (h->service_id) memcpy(hs + kOffSid, h->service_id, n * 4); else memset(hs + kOffSid, 0, n * 4);
    if (h->metric_type) memcpy(hs + kOffMt, h->metric_type, n); else memset(hs + kOffMt, 0, n);
    // one upload: from the first byte that changed (the parameters, or the per-unit arrays behind them) to the end
    // of the job's samples -- the per-unit arrays travel whole (25 KB), which is cheaper than a second copy call
    FM_CUDA(ctx, cudaMemcpyAsync(dv + up0, hs + up0, kOffValues + (n_vals + 32) * 4 - up0, cudaMemcpyHostToDevice, st));
    // the result part of the slab is cleared BEHIND a call's download, off the next call's critical path
    if (!ctx->arena_clean) FM_CUDA(ctx, cudaMemsetAsync(dv + kOffScores, 0, kOffUstat - kOffScores, st));
    ctx->arena_clean = false;
    tb.d_values = reinterpret_cast<float *>(dv + kOffValues);
    tb.d_offsets = reinterpret_cast<int64_t *>(dv + kOffOffsets);
    tb.d_seg = reinterpret_cast<uint32_t *>(dv + kOffSeg);
    tb.d_sid = reinterpret_cast<uint32_t *>(dv + kOffSid);
    tb.d_mt = dv + kOffMt;
    tb.n_units = n;
    tb.n_values = n_vals + 32;
    tb.sealed = true;
    fm_result sr;  // lives on this frame: nothing in it is owned
    sr.ctx = ctx;
    sr.batch = &tb;
    sr.n_units = n;
    sr.n_services = n_services;
    sr.params = *p;
    sr.d_params = reinterpret_cast<DevParams *>(dv + kOffParams);
    sr.d_scores = reinterpret_cast<fm_series_score *>(dv + kOffScores);
    sr.d_counts = reinterpret_cast<uint32_t *>(dv + kOffCounts);
    sr.d_worst = reinterpret_cast<unsigned long long *>(dv + kOffWorst);
    sr.d_unit_status = reinterpret_cast<uint32_t *>(dv + kOffUstat);
    sr.d_unit_points = reinterpret_cast<uint32_t *>(dv + kOffUpts);
    sr.d_verdicts = reinterpret_cast<fm_verdict *>(dv + kOffVerdicts);
    rc = launch_range(ctx, &sr, &tb, p, 0, n, false, /*small=*/true);  // its verdict kernel folds and finalises in one launch
    if (rc) { cudaStreamSynchronize(st); return rc; }
    if (n_services)
      FM_CUDA(ctx, cudaMemcpyAsync(hs + kOffVerdicts, sr.d_verdicts, n_services * sizeof(fm_verdict), cudaMemcpyDeviceToHost, st));
    FM_CUDA(ctx, cudaEventRecord(ctx->arena_done, st));
    const cudaError_t ce = cudaMemsetAsync(dv + kOffScores, 0, kOffUstat - kOffScores, st);  // for the next call
    FM_CUDA(ctx, cudaEventSynchronize(ctx->arena_done));
    ctx->arena_clean = ce == cudaSuccess;
    if (n_services) memcpy(verdicts, hs + kOffVerdicts, n_services * sizeof(fm_verdict));
    return FM_OK;
  }
  fm_result *r = new (std::nothrow) fm_result;
  if (!r) return FM_E_NOMEM;
  std::vector<cudaEvent_t> evs;
  cudaEvent_t ev_alloc = nullptr;
  auto cleanup = [&]() {
    cudaStreamSynchronize(cs);
    cudaStreamSynchronize(st);
    for (void *q : {(void *)tb.d_values, (void *)tb.d_offsets, (void *)tb.d_seg, (void *)tb.d_sid, (void *)tb.d_mt})
      if (q) cudaFreeAsync(q, st);
    for (auto e : evs) cudaEventDestroy(e);
    if (ev_alloc) cudaEventDestroy(ev_alloc);
    result_release(r);
  };
#define FM_H(expr)                                                                              \
  do {                                                                                          \
    cudaError_t _e = (expr);                                                                    \
    if (_e != cudaSuccess) {                                                                    \
      cudaGetLastError();                                                                       \
      cleanup();                                                                                \
      return set_err(ctx, _e == cudaErrorMemoryAllocation ? FM_E_NOMEM : FM_E_CUDA,             \
                     std::string(#expr) + ": " + cudaGetErrorString(_e));                       \
    }                                                                                           \
  } while (0)
  FM_H(cudaMallocAsync(&tb.d_values, (n_vals + 32) * sizeof(float), st));
  FM_H(cudaMallocAsync(&tb.d_offsets, (n + 1) * sizeof(int64_t), st));
  FM_H(cudaMallocAsync(&tb.d_seg, (3 * n + 1) * sizeof(uint32_t), st));
  FM_H(cudaMallocAsync(&tb.d_sid, (n + 1) * sizeof(uint32_t), st));
  FM_H(cudaMallocAsync(&tb.d_mt, n + 1, st));
  tb.n_units = n;
  tb.n_values = n_vals + 32;
  tb.sealed = true;
  FM_H(cudaMemsetAsync(tb.d_values + n_vals, 0, 32 * sizeof(float), st));
  if (!h->service_id) FM_H(cudaMemsetAsync(tb.d_sid, 0, n * sizeof(uint32_t), st));
  if (!h->metric_type) FM_H(cudaMemsetAsync(tb.d_mt, 0, n, st));
  rc = result_alloc(ctx, &tb, p, r);
  if (rc) { cleanup(); return rc; }
  // the copy stream may only touch the new buffers once the pool allocations are ordered before it
  FM_H(cudaEventCreateWithFlags(&ev_alloc, cudaEventDisableTiming));
  FM_H(cudaEventRecord(ev_alloc, st));
  FM_H(cudaStreamWaitEvent(cs, ev_alloc, 0));
  uint64_t target_bytes = 256ull << 20;  // ~256 MiB of samples per chunk
  if (const char *env = getenv("FOREMAST_GPU_CHUNK_BYTES")) target_bytes = std::max<uint64_t>(4096, strtoull(env, nullptr, 10));
  // the per-unit arrays (29 B per unit) go up once, ahead of the samples: four copies instead of
  // four per chunk, and the sample stream then runs back to back
  FM_H(cudaMemcpyAsync(tb.d_offsets, h->offsets, (n + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, cs));
  FM_H(cudaMemcpyAsync(tb.d_seg, h->seg_lens, 3 * n * sizeof(uint32_t), cudaMemcpyHostToDevice, cs));
  if (h->service_id) FM_H(cudaMemcpyAsync(tb.d_sid, h->service_id, n * sizeof(uint32_t), cudaMemcpyHostToDevice, cs));
  if (h->metric_type) FM_H(cudaMemcpyAsync(tb.d_mt, h->metric_type, n, cudaMemcpyHostToDevice, cs));
  uint64_t i0 = 0;
  while (i0 < n) {
    uint64_t i1 = i0 + 1;
    while (i1 < n && (uint64_t)(h->offsets[i1] - h->offsets[i0]) * sizeof(float) < target_bytes) ++i1;
    if (n - i1 < 64) i1 = n;
    const uint64_t v0 = (uint64_t)h->offsets[i0], v1 = (uint64_t)h->offsets[i1];
    if (v1 > v0) FM_H(cudaMemcpyAsync(tb.d_values + v0, h->values + v0, (v1 - v0) * sizeof(float), cudaMemcpyHostToDevice, cs));
    cudaEvent_t ev;
    FM_H(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    evs.push_back(ev);
    FM_H(cudaEventRecord(ev, cs));
    FM_H(cudaStreamWaitEvent(st, ev, 0));
    rc = launch_range(ctx, r, &tb, p, i0, i1, false);
    if (rc) { cleanup(); return rc; }
    i0 = i1;
  }
  fm_launch_verdict_finalize(n_services, r->d_counts, r->d_worst, r->d_verdicts, 0u, st);
  FM_H(cudaGetLastError());
  if (n_services)
    FM_H(cudaMemcpyAsync(verdicts, r->d_verdicts, n_services * sizeof(fm_verdict), cudaMemcpyDeviceToHost, st));
  FM_H(cudaStreamSynchronize(st));
#undef FM_H
  cleanup();
  return FM_OK;
}

}  // extern "C"
