/*
 * fm_synth.c -- CPU twin of the device synthetic-workload generator (SURVEY.md section 8(d),
 * oracle/DECISIONS.md "Synthetic workload").  TEST INFRASTRUCTURE ONLY.  Every operation is a
 * single IEEE fp32 / integer step in a fixed order so that the CUDA generator (compiled with
 * --fmad=false) produces the same bits; series names follow the recording rules of
 * deploy/foremast/2_barrelman/metrics-rules-default.yaml:62-151 only in spirit (5xx, 4xx,
 * latency, cpu, memory, tps, gc, threads).
 */
#include "fm_oracle.h"

#include <math.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { uint32_t v[4]; } u32x4;

static inline u32x4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  u32x4 o = {{c0, c1, c2, c3}};
  return o;
}

static inline float u01(uint32_t w) { return (float)(w >> 8) * 5.9604644775390625e-8f; }

static const float kBase[9] = {1.0f, 0.2f, 1.0f, 180.0f, 0.45f, 6.0e8f, 50.0f, 0.02f, 40.0f};
static const float kSigma[9] = {0.1f, 0.15f, 0.5f, 40.0f, 0.08f, 3.0e7f, 12.0f, 0.01f, 6.0f};

uint64_t fmo_synth_stride(const fm_synth_config *cfg) {
  uint64_t n = (uint64_t)cfg->n_hist + cfg->n_base + cfg->n_cur;
  return (n + 3) & ~(uint64_t)3;
}

static void unit_lens(const fm_synth_config *cfg, uint64_t gu, uint32_t lens[3]) {
  lens[0] = cfg->n_hist; lens[1] = cfg->n_base; lens[2] = cfg->n_cur;
  if (cfg->flags & FM_SYNTH_RAGGED) {
    u32x4 s = philox4x32_10((uint32_t)gu, (uint32_t)(gu >> 32), 1u, 0xAu, cfg->seed, 0x466F7265u);
    if (s.v[0] % 50u == 0u)
      for (int k = 0; k < 3; ++k)
        if (lens[k] > 0) {
          uint32_t lose = (s.v[1] >> (8 * k)) % 16u;
          lens[k] = lens[k] > lose ? lens[k] - lose : 1u;
        }
    if (s.v[0] % 1000u == 7u) lens[2] = 0;
    /* a whole service whose new pods have not reported yet: no current window in any of its series */
    uint64_t svc = gu / cfg->units_per_service;
    u32x4 w = philox4x32_10((uint32_t)svc, (uint32_t)(svc >> 32), 2u, 0xAu, cfg->seed, 0x466F7265u);
    if (w.v[0] % 500u == 3u) lens[2] = 0;
  }
}

static inline float irwin_hall(u32x4 q) {
  return (((u01(q.v[0]) + u01(q.v[1])) + (u01(q.v[2]) + u01(q.v[3]))) - 2.0f) * 1.7320508f;
}

/* t = unit time of the sample; baseline and current windows are CONCURRENT (old pods and new pods of one
 * deployment over the same minutes), so a current sample shares t -- season, drift and the load-driven
 * noise term -- with the baseline sample beside it and adds a smaller pod-private term. */
static float sample(const fm_synth_config *cfg, uint64_t gu, uint32_t mt, u32x4 r, uint32_t t, int in_cur, uint32_t cur_idx) {
  float base_m = kBase[mt], sigma_m = kSigma[mt];
  float phi = u01(r.v[0]);
  float drift = (u01(r.v[1]) - 0.5f) * base_m * 1.0e-5f;
  u32x4 q = philox4x32_10((uint32_t)gu, (uint32_t)(gu >> 32), t, 0xBu, cfg->seed, 0x466F7265u);
  float noise = irwin_hall(q);
  if (in_cur) {
    u32x4 w = philox4x32_10((uint32_t)gu, (uint32_t)(gu >> 32), t, 0xCu, cfg->seed, 0x466F7265u);
    noise = noise + 0.25f * irwin_hall(w);
  }
  noise = fminf(1.25f, fmaxf(-1.25f, noise));  /* bounded noise: k >= 3 bands stay quiet without an injected anomaly */
  float ph = (float)(t % 1440u) * (1.0f / 1440.0f) + phi;
  ph = ph - floorf(ph);
  float tri = 1.0f - 4.0f * fabsf(ph - 0.5f);
  float season = base_m * (1.0f + 0.02f * tri);
  float x = season + drift * (float)t;
  x = x + sigma_m * noise;
  if (mt == FM_MT_ERROR5XX) x = 0.0f;  /* a healthy service answers no 5xx at all */
  if (in_cur && (cfg->flags & FM_SYNTH_ANOMALIES)) {
    if (r.v[2] % 400u == 0u && cfg->n_cur > 0 && cur_idx == r.v[3] % cfg->n_cur) x = x * 20.0f + 5.0f * base_m;
    if (r.v[2] % 800u == 1u) x = x * 1.5f + sigma_m;
    if (mt == FM_MT_LATENCY && r.v[2] % 8u == 5u) x = x * 1.04f;  /* benign shift: differs, still inside the k = 10 band */
  }
  if (!(x > 0.0f)) x = 0.0f;
  if (mt == FM_MT_ERROR5XX || mt == FM_MT_ERROR4XX) x = rintf(x * 1000.0f) * 0.001f;
  return x;
}

int fmo_synth(const fm_synth_config *cfg, float *values, int64_t *offsets, uint32_t *seg_lens,
              uint32_t *service_id, uint8_t *metric_type, int threads) {
  if (!cfg || !values || !offsets || !seg_lens || cfg->n_metric_types == 0 || cfg->n_metric_types > 8 ||
      cfg->units_per_service == 0)
    return FM_E_INVALID;
  uint64_t stride = fmo_synth_stride(cfg);
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
#else
  threads = 1;
#endif
  memset(values + cfg->n_units * stride, 0, 32 * sizeof(float));
#pragma omp parallel for schedule(static) num_threads(threads)
  for (int64_t u = 0; u < (int64_t)cfg->n_units; ++u) {
    uint64_t gu = cfg->first_unit + (uint64_t)u;
    uint32_t mt = 1u + (uint32_t)(gu % cfg->n_metric_types);
    uint32_t lens[3];
    unit_lens(cfg, gu, lens);
    offsets[u] = (int64_t)((uint64_t)u * stride);
    seg_lens[3 * u] = lens[0]; seg_lens[3 * u + 1] = lens[1]; seg_lens[3 * u + 2] = lens[2];
    if (service_id) service_id[u] = (uint32_t)(gu / cfg->units_per_service);
    if (metric_type) metric_type[u] = (uint8_t)mt;
    u32x4 r = philox4x32_10((uint32_t)gu, (uint32_t)(gu >> 32), 0u, 0xAu, cfg->seed, 0x466F7265u);
    float *v = values + (uint64_t)u * stride;
    uint64_t k = 0;
    for (uint32_t i = 0; i < lens[0]; ++i) v[k++] = sample(cfg, gu, mt, r, i, 0, 0);
    for (uint32_t i = 0; i < lens[1]; ++i) v[k++] = sample(cfg, gu, mt, r, cfg->n_hist + i, 0, 0);
    for (uint32_t i = 0; i < lens[2]; ++i) v[k++] = sample(cfg, gu, mt, r, cfg->n_hist + i, 1, i);
    for (; k < stride; ++k) v[k] = 0.0f;
  }
  offsets[cfg->n_units] = (int64_t)(cfg->n_units * stride);
  return 0;
}
