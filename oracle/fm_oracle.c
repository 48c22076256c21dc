/*
 * fm_oracle.c -- scalar CPU restatement of the foremast-brain scoring pass (see fm_oracle.h:
 * TEST INFRASTRUCTURE ONLY, PARITY UNPINNED).  Spec: oracle/DECISIONS.md.
 *
 * Build: gcc -O3 -march=native -ffp-contract=off -fopenmp (oracle/Makefile).  -ffp-contract=off
 * matters: every fused multiply-add is an explicit fma() so that the CUDA path, compiled with
 * --fmad=false, reproduces the same roundings.
 */
#include "fm_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define FMO_NONE 0xFFFFFFFFu

int fmo_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

/* Defaults: deploy/foremast/3_brain/foremast-brain.yaml:24-79 (documented) and DECISIONS D2,D5,D6,D10. */
void fmo_params_default(fm_params *p) {
  memset(p, 0, sizeof(*p));
  p->algorithms = FM_ALG_MA_ALL;        /* ML_ALGORITHM=moving_average_all, brain.yaml:24-25 */
  p->pairwise = FM_PW_ALL;
  p->pairwise_threshold = 0.05;
  p->differ_scale = 0.5;
  p->min_mann_white = 20;               /* brain.yaml:74-75 */
  p->min_wilcoxon = 20;                 /* brain.yaml:76-77 */
  p->min_kruskal = 5;                   /* brain.yaml:78-79 */
  p->min_historical = 10;
  p->min_anomaly_points = 1;
  p->ma_window = 30;
  p->period = 1440;
  p->band_algorithm = 0;
  p->alpha = 0.3;
  p->beta = 0.1;
  p->gamma = 0.1;
  /* brain.yaml:26-73: threshold / bound / min_lower_bound rows */
  p->rules[FM_MT_OTHER] = (fm_metric_rule){2.0, 0.0, FM_BOUND_UPPER, 0};
  p->rules[FM_MT_ERROR5XX] = (fm_metric_rule){2.0, 0.0, FM_BOUND_UPPER, 0};
  p->rules[FM_MT_ERROR4XX] = (fm_metric_rule){3.0, 0.0, FM_BOUND_UPPER, 0};
  p->rules[FM_MT_LATENCY] = (fm_metric_rule){10.0, 0.0, FM_BOUND_BOTH, 0};
  p->rules[FM_MT_CPU] = (fm_metric_rule){5.0, 0.0, FM_BOUND_UPPER, 0};
  p->rules[FM_MT_MEMORY] = (fm_metric_rule){5.0, 0.0, FM_BOUND_UPPER, 0};
}

static const fm_metric_rule *rule_for(const fm_params *p, uint32_t mt) {
  uint32_t i = mt < FM_MAX_METRIC_TYPES ? mt : 0;
  if (p->rules[i].bound == 0) i = 0;
  return &p->rules[i];
}

/* ------------------------------------------------------------------ rank statistics
 * Mann-Whitney / Wilcoxon / Kruskal-Wallis, foremast-brain/README.md:34-38; design.md:91. */

/* order-preserving map of a finite float onto uint32 */
static inline uint32_t fkey(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

static void sort_u64(uint64_t *a, uint64_t *tmp, uint32_t n) {
  /* bottom-up merge sort, insertion sort on runs of 16 */
  for (uint32_t lo = 0; lo < n; lo += 16) {
    uint32_t hi = lo + 16 < n ? lo + 16 : n;
    for (uint32_t i = lo + 1; i < hi; ++i) {
      uint64_t v = a[i];
      uint32_t j = i;
      while (j > lo && a[j - 1] > v) { a[j] = a[j - 1]; --j; }
      a[j] = v;
    }
  }
  uint64_t *src = a, *dst = tmp;
  for (uint32_t w = 16; w < n; w *= 2) {
    for (uint32_t lo = 0; lo < n; lo += 2 * w) {
      uint32_t mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
      uint32_t i = lo, j = mid, k = lo;
      while (i < mid && j < hi) dst[k++] = (src[j] < src[i]) ? src[j++] : src[i++];
      while (i < mid) dst[k++] = src[i++];
      while (j < hi) dst[k++] = src[j++];
    }
    uint64_t *t = src; src = dst; dst = t;
  }
  if (src != a) memcpy(a, src, (size_t)n * sizeof(uint64_t));
}

/* keys are (value_key << 1 | tag); groups are runs of equal value_key.
 * returns sum over groups of cnt_tag0 * (2s + t + 1), and *ties = sum(t^3 - t). */
static int64_t rank_groups(const uint64_t *k, uint32_t n, int64_t *ties) {
  int64_t r2 = 0, tt = 0;
  uint32_t s = 0;
  while (s < n) {
    uint64_t v = k[s] >> 1;
    uint32_t e = s, c0 = 0;
    while (e < n && (k[e] >> 1) == v) { c0 += (uint32_t)(~k[e] & 1u); ++e; }
    int64_t t = (int64_t)(e - s);
    r2 += (int64_t)c0 * (2 * (int64_t)s + t + 1);
    tt += t * t * t - t;
    s = e;
  }
  *ties = tt;
  return r2;
}

static double p_mann_whitney(int64_t u1x2, uint32_t n1, uint32_t n2, int64_t ties) {
  double nn = (double)n1 * (double)n2, N = (double)n1 + (double)n2;
  double var = nn / 12.0 * ((N + 1.0) - (double)ties / (N * (N - 1.0)));
  if (!(var > 0.0)) return 1.0;
  int64_t u2x2 = 2 * (int64_t)n1 * (int64_t)n2 - u1x2;
  int64_t umax = u1x2 > u2x2 ? u1x2 : u2x2;
  double z = ((double)(umax - (int64_t)n1 * (int64_t)n2) * 0.5 - 0.5) / sqrt(var);
  double p = erfc(z * 0.70710678118654752440);
  return p > 1.0 ? 1.0 : p;
}

static double p_kruskal(int64_t r1x2, uint32_t n1, uint32_t n2, int64_t ties, double *h_out) {
  double N = (double)n1 + (double)n2;
  int64_t Ni = (int64_t)n1 + (int64_t)n2;
  int64_t r2x2 = Ni * (Ni + 1) - r1x2;
  double r1 = 0.5 * (double)r1x2, r2 = 0.5 * (double)r2x2;
  double c = 1.0 - (double)ties / (N * N * N - N);
  if (!(c > 0.0)) { *h_out = 0.0; return 1.0; }
  double h = 12.0 / (N * (N + 1.0)) * (r1 * r1 / (double)n1 + r2 * r2 / (double)n2) - 3.0 * (N + 1.0);
  h = h / c;
  if (!(h > 0.0)) h = 0.0;
  *h_out = h;
  return erfc(sqrt(h * 0.5));
}

static double p_wilcoxon(int64_t wp2, int64_t wm2, uint32_t n, int64_t ties) {
  if (n == 0) return 1.0;
  double nd = (double)n;
  double se2 = nd * (nd + 1.0) * (2.0 * nd + 1.0) / 24.0 - (double)ties / 48.0;
  if (!(se2 > 0.0)) return 1.0;
  int64_t tmin = wp2 < wm2 ? wp2 : wm2;
  double z = ((double)tmin * 0.5 - nd * (nd + 1.0) / 4.0) / sqrt(se2);
  double p = erfc(fabs(z) * 0.70710678118654752440);
  return p > 1.0 ? 1.0 : p;
}

/* ------------------------------------------------------------------ band models
 * design.md:68-71 (Moving Average, Exponential / Double Exponential Smoothing, Holt-Winters);
 * threshold / bound / min_lower_bound semantics README.md:24-25, brain.yaml:26-33. */

typedef struct {
  double k, mlb;
  uint32_t bound;
} band_rule;

static inline int outside(double x, double up, double lo, uint32_t bound) {
  return ((bound & 1u) && x > up) || ((bound & 2u) && x < lo);
}

static inline void band(double centre, double sigma, const band_rule *r, double *up, double *lo) {
  *up = fma(r->k, sigma, centre);
  double l = fma(-r->k, sigma, centre);
  *lo = l < r->mlb ? r->mlb : l;
}

typedef struct {
  double *S;      /* Holt-Winters seasonals, period entries */
  double *P1, *P2; /* rolling prefix sums, ns+nc+1 entries  */
  uint64_t *keys, *tmp;
  size_t cap_S, cap_P1, cap_P2, cap_K, cap_T;
} scratch_t;

static int grow(void **p, size_t *cap, size_t need, size_t elt) {
  if (need <= *cap) return 0;
  void *q = realloc(*p, need * elt);
  if (!q) return -1;
  *p = q;
  *cap = need;
  return 0;
}

static inline void note(fm_series_score *o, int a, uint32_t j) {
  if (o->n_anom[a]++ == 0) o->first_idx[a] = j;
}

static int score_unit(const fm_params *p, const float *hist, uint32_t nh, const float *base,
                      uint32_t nb, const float *cur, uint32_t nc, uint32_t mt, fm_series_score *o,
                      float *bu, float *bl, scratch_t *sc) {
  memset(o, 0, sizeof(*o));
  for (int a = 0; a < FM_N_ALG; ++a) o->first_idx[a] = FMO_NONE;
  o->p_mw = o->p_w = o->p_k = 1.0;

  /* ---- 1. pairwise (README.md:32-38; gated by MIN_*_DATA_POINTS brain.yaml:74-79) */
  int want_mw = p->pairwise == FM_PW_ALL || p->pairwise == FM_PW_ANY || p->pairwise == FM_PW_MANN_WHITE;
  int want_w = p->pairwise == FM_PW_ALL || p->pairwise == FM_PW_ANY || p->pairwise == FM_PW_WILCOXON;
  int want_k = p->pairwise == FM_PW_ALL || p->pairwise == FM_PW_ANY || p->pairwise == FM_PW_KRUSKAL;
  uint32_t m = nb < nc ? nb : nc;
  /* D17: rank tests look at segments of at most FM_RANK_MAX_SEGMENT samples (Wilcoxon: pairs) */
  int short_enough = nb <= FM_RANK_MAX_SEGMENT && nc <= FM_RANK_MAX_SEGMENT;
  int run_mw = want_mw && nb >= p->min_mann_white && nc >= p->min_mann_white && m > 0 && short_enough;
  int run_w = want_w && m >= p->min_wilcoxon && m > 0 && m <= FM_RANK_MAX_SEGMENT;
  int run_k = want_k && nb >= p->min_kruskal && nc >= p->min_kruskal && m > 0 && short_enough;
  int n_run = 0, n_rej = 0;
  if (run_mw || run_k) {
    uint32_t N = nb + nc;
    if (grow((void **)&sc->keys, &sc->cap_K, N, 8)) return FM_E_NOMEM;
    if (grow((void **)&sc->tmp, &sc->cap_T, N, 8)) return FM_E_NOMEM;
    for (uint32_t i = 0; i < nb; ++i) sc->keys[i] = ((uint64_t)fkey(base[i]) << 1);
    for (uint32_t i = 0; i < nc; ++i) sc->keys[nb + i] = ((uint64_t)fkey(cur[i]) << 1) | 1u;
    sort_u64(sc->keys, sc->tmp, N);
    int64_t ties;
    int64_t r1x2 = rank_groups(sc->keys, N, &ties);
    o->mw_2r1 = r1x2;
    o->mw_2u1 = r1x2 - (int64_t)nb * ((int64_t)nb + 1);
    o->mw_ties = ties;
    if (run_mw) {
      o->status |= FM_S_MW_RUN;
      o->p_mw = p_mann_whitney(o->mw_2u1, nb, nc, ties);
      ++n_run;
      n_rej += o->p_mw < p->pairwise_threshold;
    }
    if (run_k) {
      o->status |= FM_S_KRUSKAL_RUN;
      o->p_k = p_kruskal(r1x2, nb, nc, ties, &o->kw_h);
      ++n_run;
      n_rej += o->p_k < p->pairwise_threshold;
    }
  }
  if (run_w) {
    if (grow((void **)&sc->keys, &sc->cap_K, m, 8)) return FM_E_NOMEM;
    if (grow((void **)&sc->tmp, &sc->cap_T, m, 8)) return FM_E_NOMEM;
    uint32_t n = 0;
    for (uint32_t i = 0; i < m; ++i) {
      float d = cur[i] - base[i];
      if (d == 0.0f) continue;
      uint32_t mag;
      float ad = fabsf(d);
      memcpy(&mag, &ad, 4);
      sc->keys[n++] = ((uint64_t)mag << 1) | (d < 0.0f ? 1u : 0u);
    }
    sort_u64(sc->keys, sc->tmp, n);
    int64_t ties;
    int64_t wp2 = rank_groups(sc->keys, n, &ties);
    o->w_n = n;
    o->w_2wplus = wp2;
    o->w_2wminus = (int64_t)n * ((int64_t)n + 1) - wp2;
    o->w_ties = ties;
    o->status |= FM_S_WILCOXON_RUN;
    o->p_w = p_wilcoxon(wp2, o->w_2wminus, n, ties);
    ++n_run;
    n_rej += o->p_w < p->pairwise_threshold;
  }
  int differs = 0;
  if (p->pairwise == FM_PW_ALL) differs = n_run > 0 && n_rej == n_run;
  else if (p->pairwise != FM_PW_NONE) differs = n_rej > 0;
  if (differs) o->status |= FM_S_DIFFERS;

  /* ---- 2. threshold (design.md:35 "it will be lower threshold") */
  const fm_metric_rule *mr = rule_for(p, mt);
  band_rule r = {differs ? mr->threshold * p->differ_scale : mr->threshold, mr->min_lower_bound, mr->bound};
  o->k_eff = r.k;

  /* ---- 3. model window (D15) */
  const float *src = NULL;
  uint32_t ns = 0;
  if (nh >= p->min_historical && nh > 0) { src = hist; ns = nh; }
  else if (nb >= p->min_historical && nb > 0) { src = base; ns = nb; o->status |= FM_S_MODEL_BASE; }
  if (!src || nc == 0) { o->status |= FM_S_UNKNOWN; return 0; }
  int want_band_alg = (bu && bl) ? (int)p->band_algorithm : 0;

  /* ---- MA_ALL and MA_ROLLING share the shifted sums */
  double x0 = (double)src[0];
  if (p->algorithms & FM_ALG_MA_ALL) {
    double s1 = 0.0, s2 = 0.0;
    for (uint32_t i = 0; i < ns; ++i) {
      double d = (double)src[i] - x0;
      s1 += d;
      s2 = fma(d, d, s2);
    }
    double m1 = s1 / (double)ns, mean = x0 + m1;
    double var = fma(-m1, m1, s2 / (double)ns);
    if (!(var > 0.0)) var = 0.0;
    double sd = sqrt(var), up, lo;
    band(mean, sd, &r, &up, &lo);
    o->ma_mean = mean;
    o->ma_std = sd;
    o->sigma[0] = sd;
    o->status |= 1u << (FM_S_ALG_SHIFT + 0);
    for (uint32_t j = 0; j < nc; ++j) {
      if (outside((double)cur[j], up, lo, r.bound)) note(o, 0, j);
      if (want_band_alg == FM_ALG_MA_ALL) { bu[j] = (float)up; bl[j] = (float)lo; }
    }
  }
  if (p->algorithms & FM_ALG_MA_ROLLING) {
    uint32_t W = p->ma_window ? p->ma_window : 1;
    size_t need = (size_t)ns + nc + 1;
    if (grow((void **)&sc->P1, &sc->cap_P1, need, 8)) return FM_E_NOMEM;
    if (grow((void **)&sc->P2, &sc->cap_P2, need, 8)) return FM_E_NOMEM;
    double s1 = 0.0, s2 = 0.0;
    sc->P1[0] = 0.0; sc->P2[0] = 0.0;
    for (uint32_t t = 0; t < ns + nc; ++t) {
      double d = (double)(t < ns ? src[t] : cur[t - ns]) - x0;
      s1 += d;
      s2 = fma(d, d, s2);
      sc->P1[t + 1] = s1; sc->P2[t + 1] = s2;
    }
    o->status |= 1u << (FM_S_ALG_SHIFT + 1);
    for (uint32_t j = 0; j < nc; ++j) {
      uint32_t q = ns + j, cnt = W < q ? W : q;
      double a1 = sc->P1[q] - sc->P1[q - cnt], a2 = sc->P2[q] - sc->P2[q - cnt];
      double m1 = a1 / (double)cnt, mean = x0 + m1;
      double var = fma(-m1, m1, a2 / (double)cnt);
      if (!(var > 0.0)) var = 0.0;
      double sd = sqrt(var), up, lo;
      band(mean, sd, &r, &up, &lo);
      o->sigma[1] = sd;
      if (outside((double)cur[j], up, lo, r.bound)) note(o, 1, j);
      if (want_band_alg == FM_ALG_MA_ROLLING) { bu[j] = (float)up; bl[j] = (float)lo; }
    }
  }
  /* ---- EWMA (design.md:69) */
  if (p->algorithms & FM_ALG_EWMA) {
    double s = (double)src[0], sse = 0.0;
    for (uint32_t t = 1; t < ns; ++t) {
      double e = (double)src[t] - s;
      sse = fma(e, e, sse);
      s = fma(p->alpha, e, s);
    }
    double sigma = ns >= 2 ? sqrt(sse / (double)(ns - 1)) : 0.0;
    o->sigma[2] = sigma;
    o->status |= 1u << (FM_S_ALG_SHIFT + 2);
    for (uint32_t j = 0; j < nc; ++j) {
      double x = (double)cur[j], up, lo;
      band(s, sigma, &r, &up, &lo);
      if (outside(x, up, lo, r.bound)) note(o, 2, j);
      if (want_band_alg == FM_ALG_EWMA) { bu[j] = (float)up; bl[j] = (float)lo; }
      double e = x - s;
      s = fma(p->alpha, e, s);
    }
  }
  /* ---- DES, Holt linear (design.md:70) */
  if ((p->algorithms & FM_ALG_DES) && ns >= 2) {
    double l = (double)src[0], b = (double)src[1] - (double)src[0], sse = 0.0;
    double ab = p->alpha * p->beta;
    for (uint32_t t = 1; t < ns; ++t) {
      double f = l + b, e = (double)src[t] - f;
      sse = fma(e, e, sse);
      l = fma(p->alpha, e, f);
      b = fma(ab, e, b);
    }
    double sigma = sqrt(sse / (double)(ns - 1));
    o->sigma[3] = sigma;
    o->status |= 1u << (FM_S_ALG_SHIFT + 3);
    for (uint32_t j = 0; j < nc; ++j) {
      double x = (double)cur[j], f = l + b, up, lo;
      band(f, sigma, &r, &up, &lo);
      if (outside(x, up, lo, r.bound)) note(o, 3, j);
      if (want_band_alg == FM_ALG_DES) { bu[j] = (float)up; bl[j] = (float)lo; }
      double e = x - f;
      l = fma(p->alpha, e, f);
      b = fma(ab, e, b);
    }
  }
  /* ---- Holt-Winters additive (design.md:71; 7-day history metricsquery.go:95) */
  uint32_t mper = p->period;
  if ((p->algorithms & FM_ALG_HW) && mper >= 2 && ns >= 2 * (uint64_t)mper) {
    if (grow((void **)&sc->S, &sc->cap_S, mper, 8)) return FM_E_NOMEM;
    double a0 = 0.0, a1 = 0.0;
    for (uint32_t i = 0; i < mper; ++i) a0 += (double)src[i];
    for (uint32_t i = mper; i < 2 * mper; ++i) a1 += (double)src[i];
    double L0 = a0 / (double)mper, M2 = a1 / (double)mper;
    double l = L0, b = (M2 - L0) / (double)mper, sse = 0.0;
    double ab = p->alpha * p->beta, g1 = p->gamma * (1.0 - p->alpha);
    for (uint32_t i = 0; i < mper; ++i) sc->S[i] = (double)src[i] - L0;
    uint32_t idx = 0; /* t mod m, t starts at m */
    for (uint32_t t = mper; t < ns; ++t) {
      double lb = l + b, f = lb + sc->S[idx], e = (double)src[t] - f;
      sse = fma(e, e, sse);
      l = fma(p->alpha, e, lb);
      b = fma(ab, e, b);
      sc->S[idx] = fma(g1, e, sc->S[idx]);
      if (++idx == mper) idx = 0;
    }
    double sigma = sqrt(sse / (double)(ns - mper));
    o->sigma[4] = sigma;
    o->status |= 1u << (FM_S_ALG_SHIFT + 4);
    for (uint32_t j = 0; j < nc; ++j) {
      double x = (double)cur[j], lb = l + b, f = lb + sc->S[idx], up, lo;
      band(f, sigma, &r, &up, &lo);
      /* guarded comparison (DECISIONS "HW"): excursions below 1e-9 relative do not count, so that an
       * implementation that re-associates the recurrence (the CUDA scan) flags the same points */
      double guard = 1e-9 * (fabs(f) + fabs(x));
      if (((r.bound & 1u) && x - up > guard) || ((r.bound & 2u) && lo - x > guard)) note(o, 4, j);
      if (want_band_alg == FM_ALG_HW) { bu[j] = (float)up; bl[j] = (float)lo; }
      double e = x - f;
      l = fma(p->alpha, e, lb);
      b = fma(ab, e, b);
      sc->S[idx] = fma(g1, e, sc->S[idx]);
      if (++idx == mper) idx = 0;
    }
  }
  /* ---- 5. unit status (D12) */
  for (int a = 0; a < FM_N_ALG; ++a)
    if ((o->status >> (FM_S_ALG_SHIFT + a) & 1u) && o->n_anom[a] >= p->min_anomaly_points && o->n_anom[a] > 0)
      o->status |= FM_S_ANOMALY;
  return 0;
}

int fmo_score(const fm_csr *csr, const fm_params *p, fm_series_score *scores, float *band_upper,
              float *band_lower, int threads) {
  if (!csr || !p || !scores || (csr->n_units && (!csr->values || !csr->offsets || !csr->seg_lens)))
    return FM_E_INVALID;
  int rc = 0;
#ifdef _OPENMP
  if (threads <= 0) threads = omp_get_max_threads();
#else
  threads = 1;
#endif
#pragma omp parallel num_threads(threads)
  {
    scratch_t sc;
    memset(&sc, 0, sizeof(sc));
#pragma omp for schedule(static)
    for (int64_t u = 0; u < (int64_t)csr->n_units; ++u) {
      uint32_t nh = csr->seg_lens[3 * u], nb = csr->seg_lens[3 * u + 1], nc = csr->seg_lens[3 * u + 2];
      const float *h = csr->values + csr->offsets[u];
      int64_t co = csr->offsets[u] + nh + nb;
      int e = score_unit(p, h, nh, h + nh, nb, h + nh + nb, nc,
                         csr->metric_type ? csr->metric_type[u] : 0, &scores[u],
                         band_upper ? band_upper + co : NULL, band_lower ? band_lower + co : NULL, &sc);
      if (e) {
#pragma omp atomic write
        rc = e;
      }
    }
    free(sc.S); free(sc.P1); free(sc.P2); free(sc.keys); free(sc.tmp);
  }
  return rc;
}

/* Verdict fold (D13): terminal statuses of converter.go:16-21; fail fast design.md:43. */
int fmo_verdicts(const fm_csr *csr, const fm_series_score *scores, uint32_t n_services,
                 uint64_t first_unit, fm_verdict *out) {
  if (!csr || !scores || !out) return FM_E_INVALID;
  uint64_t *worst = (uint64_t *)calloc(n_services ? n_services : 1, 8);
  if (!worst) return FM_E_NOMEM;
  for (uint32_t s = 0; s < n_services; ++s) {
    memset(&out[s], 0, sizeof(out[s]));
    out[s].service_id = s;
  }
  for (uint64_t u = 0; u < csr->n_units; ++u) {
    uint32_t s = csr->service_id ? csr->service_id[u] : 0;
    if (s >= n_services) { free(worst); return FM_E_RANGE; }
    const fm_series_score *o = &scores[u];
    out[s].n_units++;
    if (o->status & FM_S_UNKNOWN) out[s].n_unknown++;
    if (o->status & FM_S_ANOMALY) {
      uint32_t pts = 0;
      for (int a = 0; a < FM_N_ALG; ++a) pts = o->n_anom[a] > pts ? o->n_anom[a] : pts;
      out[s].n_anom_units++;
      out[s].n_anom_points += pts;
      uint64_t key = ((uint64_t)pts << 32) | (uint64_t)(0xFFFFFFFFu - (uint32_t)(first_unit + u));
      if (key > worst[s]) worst[s] = key;
    }
  }
  for (uint32_t s = 0; s < n_services; ++s) {
    if (out[s].n_anom_units) {
      out[s].verdict = FM_V_UNHEALTH;
      out[s].worst_unit = 0xFFFFFFFFu - (uint32_t)(worst[s] & 0xFFFFFFFFu);
      out[s].worst_unit_points = (uint32_t)(worst[s] >> 32);
    } else {
      out[s].verdict = (out[s].n_unknown == out[s].n_units) ? FM_V_UNKNOWN : FM_V_HEALTH;
      out[s].worst_unit = FMO_NONE;
    }
  }
  free(worst);
  return 0;
}
