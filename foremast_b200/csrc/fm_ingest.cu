// fm_ingest.cu -- host-side fast path from a Prometheus query_range response to fp32 series
// ("today's real bottleneck is fetch + string->float parse").
//
// The brain fetches the URLs that foremast-service stored in the job document
// (pkg/prometheus/prometheushelper.go:13-43) and gets back a matrix:
//   {"status":"success","data":{"resultType":"matrix","result":[
//       {"metric":{...labels...},"values":[[1550000000,"0.166"],[1550000060,"NaN"],...]}, ...]}}
// (shape as consumed by foremast-dashboard/src/services/ApiService.js:36-39 and
// src/reducers/metricReducer.js:67-69).  This is a single-pass scanner over that text: no DOM, no
// allocation, values written straight into the caller's float array, label objects returned as
// byte spans.  Sample strings go string -> double -> float (the same two roundings as Python's
// float() followed by a cast), "NaN" / "+Inf" / "-Inf" are kept as such and dropped later by
// fm_batch_add_series (a gap).
//
// Timestamps matter: query_range answers on the grid start + k*step and simply OMITS a scrape that is
// missing, while everything downstream is index-based (Wilcoxon pairs sample i with sample i, the
// Holt-Winters phase is index mod period, anomaly times are t0 + index*step).  So every timestamp is
// read (integer fast path), the step is the smallest gap between neighbours, and a series reports how
// many grid slots it lacks; with FM_PROM_GRID the samples are PLACED on the grid -- holes are NaN, or
// the previous sample with FM_PROM_FILL -- so that index and time stay the same thing.
// fm_batch_add_prometheus / _many feed response bodies straight into a batch (no Python, no DOM),
// the latter parsing on a thread pool.  Host code only: nothing here touches the GPU.
#include <atomic>
#include <charconv>
#include <cmath>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

#include "../../include/foremast_gpu.h"
#include "fm_internal.h"

namespace {

struct Scan {
  const char *p, *end, *base;
  bool ok = true;

  void ws() {
    while (p < end && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
  }
  bool eat(char c) {
    ws();
    if (p < end && *p == c) { ++p; return true; }
    return false;
  }
  // p at opening quote; returns span of the raw (still escaped) contents
  bool string(const char **s, size_t *n) {
    ws();
    if (p >= end || *p != '"') return ok = false;
    const char *q = ++p;
    while (p < end && *p != '"') {
      if (*p == '\\') ++p;
      ++p;
    }
    if (p >= end) return ok = false;
    *s = q;
    *n = (size_t)(p - q);
    ++p;
    return true;
  }
  bool number(double *out) {
    ws();
    const char *q = p;
    if (p < end && *p == '+') ++q;
    auto r = std::from_chars(q, end, *out);
    if (r.ec != std::errc()) return ok = false;
    p = r.ptr;
    return true;
  }
  // a timestamp: unix seconds, usually an integer or with a few decimals ("1550000060", "1550000060.123")
  bool timestamp(double *out) {
    ws();
    const char *q = p;
    uint64_t ip = 0;
    int nd = 0;
    while (q < end && *q >= '0' && *q <= '9' && nd < 15) { ip = ip * 10 + (uint64_t)(*q - '0'); ++q; ++nd; }
    if (nd > 0 && q < end && (*q == ',' || *q == ' ')) { *out = (double)ip; p = q; return true; }
    return number(out);
  }
  // a sample value: a quoted number ("0.166", "NaN"); no escapes expected, so the closing quote is a memchr away
  bool quoted(const char **s, size_t *n) {
    ws();
    if (p >= end || *p != '"') return ok = false;
    const char *q = p + 1;
    const char *e = static_cast<const char *>(memchr(q, '"', (size_t)(end - q)));
    if (!e) return ok = false;
    if (memchr(q, '\\', (size_t)(e - q))) return string(s, n);
    *s = q;
    *n = (size_t)(e - q);
    p = e + 1;
    return true;
  }
  bool skip_quoted() { const char *s; size_t n; return quoted(&s, &n); }
  bool skip_number() {  // a JSON number, not converted
    ws();
    const char *q = p;
    while (p < end && ((*p >= '0' && *p <= '9') || *p == '.' || *p == '-' || *p == '+' || *p == 'e' || *p == 'E')) ++p;
    if (p == q) return ok = false;
    return true;
  }
  void skip_value() {  // any JSON value
    ws();
    if (p >= end) { ok = false; return; }
    if (*p == '"') { const char *s; size_t n; string(&s, &n); return; }
    if (*p == '{' || *p == '[') {
      const char open = *p, close = open == '{' ? '}' : ']';
      int depth = 0;
      while (p < end) {
        if (*p == '"') { const char *s; size_t n; if (!string(&s, &n)) return; continue; }
        if (*p == open) ++depth;
        else if (*p == close && --depth == 0) { ++p; return; }
        ++p;
      }
      ok = false;
      return;
    }
    while (p < end && *p != ',' && *p != '}' && *p != ']') ++p;  // number / true / false / null
  }
};

bool key_is(const char *s, size_t n, const char *lit) { return n == strlen(lit) && memcmp(s, lit, n) == 0; }

// Exact fast path for the short decimals most samples are ("1", "0.25", "153.7"): up to 15
// significant digits and a power of ten within 10^22 are both exact doubles, so one multiplication
// or division rounds once -- the same value from_chars returns (Clinger's observation).  Anything
// else (17-digit shortest-round-trip floats, exponents, NaN/Inf) goes to from_chars below.
bool small_decimal(const char *s, size_t n, double *out) {
  static const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                    1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
  size_t i = 0;
  bool neg = false;
  if (i < n && (s[i] == '-' || s[i] == '+')) neg = s[i++] == '-';
  uint64_t mant = 0;
  int digits = 0, frac = 0;
  bool seen_point = false, any = false;
  for (; i < n; ++i) {
    const char c = s[i];
    if (c >= '0' && c <= '9') {
      any = true;
      if (mant || c != '0') {
        if (++digits > 15) return false;
        mant = mant * 10 + (uint64_t)(c - '0');
      }
      if (seen_point) ++frac;
    } else if (c == '.' && !seen_point) {
      seen_point = true;
    } else {
      return false;  // exponent, NaN, Inf, garbage
    }
  }
  if (!any || frac > 22) return false;
  const double v = (double)mant / kPow10[frac];
  *out = neg ? -v : v;
  return true;
}

double sample_value(const char *s, size_t n) {
  while (n && (*s == ' ')) { ++s; --n; }
  double v;
  if (small_decimal(s, n, &v)) return v;
  const bool plus = n && *s == '+';
  if (plus) { ++s; --n; }
  auto r = std::from_chars(s, s + n, v);
  if (r.ec == std::errc() && r.ptr == s + n) return v;
  return std::numeric_limits<double>::quiet_NaN();  // unparsable sample = gap
}

// One pass over the response.  Series descriptors go to `series` (up to series_cap), samples to `values`
// (up to values_cap); counts keep running past the capacities so that the caller learns what to allocate.
// flags: FM_PROM_GRID places samples at round((ts - t0) / step), step = step_hint if > 0, else the smallest
// gap of the series (found by a first look at its timestamps); FM_PROM_FILL fills holes with the previous
// sample instead of NaN.
int walk_matrix(const char *json, uint64_t len, uint32_t flags, double step_hint, fm_prom_series *series, uint32_t series_cap,
                uint32_t *n_series, float *values, uint64_t values_cap, uint64_t *n_values) {
  Scan sc{json, json + len, json};
  uint32_t ns = 0;
  uint64_t nv = 0;
  const bool grid = (flags & FM_PROM_GRID) != 0, fill = (flags & FM_PROM_FILL) != 0;
  auto finish = [&](int rc) { *n_series = ns; *n_values = nv; return rc; };
  if (!sc.eat('{')) return finish(FM_E_INVALID);
  // top level: find "data"
  bool in_data = false;
  while (sc.ok) {
    const char *k; size_t kn;
    if (sc.eat('}')) break;
    if (!sc.string(&k, &kn) || !sc.eat(':')) return finish(FM_E_INVALID);
    if (key_is(k, kn, "data")) {
      in_data = true;
      if (!sc.eat('{')) return finish(FM_E_INVALID);
      while (sc.ok) {
        if (sc.eat('}')) break;
        if (!sc.string(&k, &kn) || !sc.eat(':')) return finish(FM_E_INVALID);
        if (key_is(k, kn, "result")) {
          if (!sc.eat('[')) return finish(FM_E_INVALID);
          while (sc.ok) {
            if (sc.eat(']')) break;
            if (!sc.eat('{')) return finish(FM_E_INVALID);
            fm_prom_series cur;
            memset(&cur, 0, sizeof(cur));
            cur.first_value = nv;
            while (sc.ok) {
              if (sc.eat('}')) break;
              if (!sc.string(&k, &kn) || !sc.eat(':')) return finish(FM_E_INVALID);
              if (key_is(k, kn, "metric")) {
                sc.ws();
                const char *m0 = sc.p;
                sc.skip_value();
                cur.metric_off = (uint64_t)(m0 - json);
                cur.metric_len = (uint64_t)(sc.p - m0);
              } else if (key_is(k, kn, "values")) {
                if (!sc.eat('[')) return finish(FM_E_INVALID);
                // The step of the grid is the hint (the query's own step) or, without one, the smallest gap of
                // the series -- which takes a first look at its timestamps, but only when samples are to be placed.
                double step = step_hint > 0.0 ? step_hint : 0.0;
                if (grid && !(step > 0.0)) {
                  Scan look = sc;
                  double t_prev = 0.0;
                  uint32_t seen = 0;
                  while (look.ok) {
                    if (look.eat(']')) break;
                    double ts = 0.0;
                    if (!look.eat('[') || !look.timestamp(&ts) || !look.eat(',') || !look.skip_quoted() || !look.eat(']'))
                      return finish(FM_E_INVALID);
                    const double gap = ts - t_prev;
                    if (seen && gap > 0.0 && (step == 0.0 || gap < step)) step = gap;
                    t_prev = ts;
                    ++seen;
                    look.eat(',');
                  }
                  if (!look.ok) return finish(FM_E_INVALID);
                }
                const bool place = grid && step > 0.0;
                double t0 = 0.0, t_prev = 0.0, min_gap = 0.0;
                uint32_t count = 0;
                bool ordered = true;
                uint64_t at = 0;  // next grid slot to write (place)
                float last_v = std::numeric_limits<float>::quiet_NaN();
                while (sc.ok) {
                  if (sc.eat(']')) break;
                  double ts = 0.0;
                  const char *vs; size_t vn;
                  if (!sc.eat('[') || !sc.timestamp(&ts) || !sc.eat(',') || !sc.quoted(&vs, &vn) || !sc.eat(']'))
                    return finish(FM_E_INVALID);
                  if (count == 0) t0 = ts;
                  else {
                    const double gap = ts - t_prev;
                    if (!(gap > 0.0)) ordered = false;
                    else if (min_gap == 0.0 || gap < min_gap) min_gap = gap;
                  }
                  t_prev = ts;
                  ++count;
                  const float v = (float)sample_value(vs, vn);
                  if (place && ordered) {
                    const double rel = (ts - t0) / step;
                    if (rel > 64.0 * count + 1024.0) return finish(FM_E_INVALID);  // nonsense timestamps
                    uint64_t slot = (uint64_t)llround(rel);
                    if (slot < at) slot = at;  // two samples rounding onto one slot: keep both, in order
                    for (; at < slot; ++at, ++nv)
                      if (values && nv < values_cap) values[nv] = fill ? last_v : std::numeric_limits<float>::quiet_NaN();
                  }
                  if (values && nv < values_cap) values[nv] = v;
                  if (std::isfinite(v)) last_v = v;
                  ++nv;
                  ++at;
                  sc.eat(',');
                }
                if (!(step > 0.0)) step = min_gap;
                cur.t0 = t0;
                cur.step = step;
                // grid slots up to the last sample that have no sample of their own
                cur.reserved = 0;
                if (!ordered) cur.reserved = 0xFFFFFFFFu;
                else if (count > 1 && step > 0.0) {
                  const double span = (t_prev - t0) / step;
                  if (span > 64.0 * count + 1024.0) return finish(FM_E_INVALID);
                  const uint64_t slots = (uint64_t)llround(span) + 1;
                  cur.reserved = slots > count ? (uint32_t)(slots - count) : 0u;
                }
                cur.n_values = (uint32_t)(nv - cur.first_value);
              } else {
                sc.skip_value();
              }
              sc.eat(',');
            }
            if (series && ns < series_cap) series[ns] = cur;
            ++ns;
            sc.eat(',');
          }
        } else {
          sc.skip_value();
        }
        sc.eat(',');
      }
    } else {
      sc.skip_value();
    }
    sc.eat(',');
  }
  if (!sc.ok || !in_data) return finish(FM_E_INVALID);
  if (ns > series_cap || nv > values_cap) return finish(FM_E_RANGE);  // counts tell the caller what to allocate
  return finish(FM_OK);
}

// a whole response into vectors (sized by a counting pass)
struct Parsed {
  std::vector<fm_prom_series> series;
  std::vector<float> values;
  std::vector<uint32_t> finite;  // per series: samples that are finite (the rest are gaps, dropped at ingest, D11)
  int rc = FM_OK;
};

// dst <- the finite samples of src (with -0.0 stored as +0.0, D11); returns how many
uint32_t copy_finite(float *dst, const float *src, uint32_t n, uint32_t n_finite) {
  if (n_finite == n) {
    for (uint32_t i = 0; i < n; ++i) dst[i] = src[i] + 0.0f;
    return n;
  }
  uint32_t k = 0;
  for (uint32_t i = 0; i < n; ++i)
    if (std::isfinite(src[i])) dst[k++] = src[i] + 0.0f;
  return k;
}
void parse_into(const char *json, uint64_t len, uint32_t flags, double step_hint, Parsed *out) {
  out->series.clear();
  out->values.clear();
  out->finite.clear();
  if (!json || len == 0) { out->rc = FM_OK; return; }  // a job without this query
  // ONE walk into buffers sized by what the text can hold at most -- a point is at least `[0,"0"],` (8 bytes),
  // a series at least `{"values":[]},` (14 bytes); re-gridding can add samples the text does not have, and
  // then the counting walk below sizes the buffers exactly
  uint32_t ns = 0;
  uint64_t nv = 0;
  out->series.resize(len / 14 + 1);
  out->values.resize(len / 8 + 1);
  int rc = walk_matrix(json, len, flags, step_hint, out->series.data(), (uint32_t)out->series.size(), &ns, out->values.data(),
                       out->values.size(), &nv);
  if (rc == FM_E_RANGE) {
    out->series.resize(ns);
    out->values.resize(nv);
    rc = walk_matrix(json, len, flags, step_hint, out->series.data(), ns, &ns, out->values.data(), nv, &nv);
  }
  out->rc = rc;
  if (rc != FM_OK) return;
  out->series.resize(ns);
  out->values.resize(nv);
  out->finite.assign(ns, 0u);
  for (uint32_t i = 0; i < ns; ++i) {
    const float *v = out->values.data() + out->series[i].first_value;
    uint32_t k = 0;
    for (uint32_t j = 0; j < out->series[i].n_values; ++j) k += std::isfinite(v[j]) ? 1u : 0u;
    out->finite[i] = k;
  }
}

// The worker's pairing of one job alias (foremast_b200/brain.py::build_units): one unit per CURRENT series
// (new pod); its baseline is the baseline series of the same rank (old pod i vs new pod i, the last one if
// there are fewer), its history the first historical series (the app-level query,
// foremast-barrelman/pkg/client/metrics/metricsquery.go:79-99).
int add_units(fm_batch *b, const fm_series_meta *meta, const Parsed &h, const Parsed &bs, const Parsed &c, uint32_t *n_added) {
  *n_added = 0;
  const float *hv = h.series.empty() ? nullptr : h.values.data() + h.series[0].first_value;
  const uint32_t hn = h.series.empty() ? 0u : h.series[0].n_values;
  for (size_t i = 0; i < c.series.size(); ++i) {
    const float *bv = nullptr;
    uint32_t bn = 0;
    if (!bs.series.empty()) {
      const fm_prom_series &s = bs.series[std::min(i, bs.series.size() - 1)];
      bv = bs.values.data() + s.first_value;
      bn = s.n_values;
    }
    const fm_prom_series &s = c.series[i];
    const int rc = fm_batch_add_series(b, meta, hv, hn, bv, bn, c.values.data() + s.first_value, s.n_values);
    if (rc) return rc;
    ++*n_added;
  }
  return FM_OK;
}

}  // namespace

extern "C" {

int fm_parse_prometheus_matrix(const char *json, uint64_t len, fm_prom_series *series, uint32_t series_cap,
                               uint32_t *n_series, float *values, uint64_t values_cap, uint64_t *n_values) {
  if (!json || !n_series || !n_values) return FM_E_INVALID;
  return walk_matrix(json, len, 0u, 0.0, series, series_cap, n_series, values, values_cap, n_values);
}

int fm_parse_prometheus_matrix_grid(const char *json, uint64_t len, uint32_t flags, double step_hint, fm_prom_series *series,
                                    uint32_t series_cap, uint32_t *n_series, float *values, uint64_t values_cap,
                                    uint64_t *n_values) {
  if (!json || !n_series || !n_values || (flags & ~(FM_PROM_GRID | FM_PROM_FILL))) return FM_E_INVALID;
  return walk_matrix(json, len, flags, step_hint, series, series_cap, n_series, values, values_cap, n_values);
}

int fm_batch_add_prometheus(fm_batch *b, const fm_prom_job *job, uint32_t flags, uint32_t *n_units_added) {
  uint32_t n = 0;
  if (n_units_added) *n_units_added = 0;
  if (!b || !job || (flags & ~(FM_PROM_GRID | FM_PROM_FILL))) return FM_E_INVALID;
  Parsed h, bs, c;
  parse_into(job->hist_json, job->hist_len, flags, job->step, &h);
  parse_into(job->base_json, job->base_len, flags, job->step, &bs);
  parse_into(job->cur_json, job->cur_len, flags, job->step, &c);
  if (h.rc || bs.rc || c.rc) return FM_E_INVALID;
  const int rc = add_units(b, &job->meta, h, bs, c, &n);
  if (n_units_added) *n_units_added = n;
  return rc;
}

int fm_batch_add_prometheus_many(fm_batch *b, const fm_prom_job *jobs, uint32_t n_jobs, uint32_t flags, uint32_t n_threads,
                                 uint32_t *units_per_job) {
  if (!b || (n_jobs && !jobs) || (flags & ~(FM_PROM_GRID | FM_PROM_FILL))) return FM_E_INVALID;
  struct Job3 { Parsed h, b, c; };
  std::vector<Job3> parsed(n_jobs);
  if (n_threads == 0) n_threads = std::max(1u, std::thread::hardware_concurrency());
  n_threads = std::min<uint32_t>(n_threads, std::max<uint32_t>(n_jobs, 1));
  std::atomic<uint32_t> next{0};
  auto work = [&]() {  // the scanner is reentrant: one response at a time per thread, results in job order
    for (uint32_t i = next.fetch_add(1); i < n_jobs; i = next.fetch_add(1)) {
      parse_into(jobs[i].hist_json, jobs[i].hist_len, flags, jobs[i].step, &parsed[i].h);
      parse_into(jobs[i].base_json, jobs[i].base_len, flags, jobs[i].step, &parsed[i].b);
      parse_into(jobs[i].cur_json, jobs[i].cur_len, flags, jobs[i].step, &parsed[i].c);
    }
  };
  try {
    std::vector<std::thread> pool;
    for (uint32_t t = 1; t < n_threads; ++t) pool.emplace_back(work);
    work();
    for (auto &t : pool) t.join();
  } catch (...) {
    return FM_E_NOMEM;
  }
  // Unit layout in job order (the unit numbering does not depend on the thread schedule), one reservation in
  // the batch, then the samples are copied in by the same thread pool.  A job whose responses do not parse adds
  // no units (the worker marks it preprocess_failed) and does not stop the others.
  std::vector<uint32_t> seg, sid;
  std::vector<uint8_t> mt;
  std::vector<uint64_t> first_unit(n_jobs + 1, 0);
  try {
    for (uint32_t i = 0; i < n_jobs; ++i) {
      const Job3 &j = parsed[i];
      first_unit[i + 1] = first_unit[i];
      if (j.h.rc || j.b.rc || j.c.rc) {
        if (units_per_job) units_per_job[i] = 0xFFFFFFFFu;
        continue;
      }
      const uint32_t nh = j.h.series.empty() ? 0u : j.h.finite[0];
      for (size_t c = 0; c < j.c.series.size(); ++c) {
        seg.push_back(nh);
        seg.push_back(j.b.series.empty() ? 0u : j.b.finite[std::min(c, j.b.series.size() - 1)]);
        seg.push_back(j.c.finite[c]);
        sid.push_back(jobs[i].meta.service_id);
        mt.push_back(jobs[i].meta.metric_type);
      }
      first_unit[i + 1] += j.c.series.size();
      if (units_per_job) units_per_job[i] = (uint32_t)j.c.series.size();
    }
  } catch (...) {
    return FM_E_NOMEM;
  }
  const uint64_t n_units = first_unit[n_jobs];
  if (n_units == 0) return FM_OK;
  std::vector<uint64_t> rel(n_units);
  float *dst = nullptr;
  const int rc = fm_batch_reserve_units(b, n_units, seg.data(), sid.data(), mt.data(), &dst, rel.data());
  if (rc) return rc;
  next = 0;
  auto fill = [&]() {
    for (uint32_t i = next.fetch_add(1); i < n_jobs; i = next.fetch_add(1)) {
      const Job3 &j = parsed[i];
      for (uint64_t u = first_unit[i]; u < first_unit[i + 1]; ++u) {
        const size_t c = (size_t)(u - first_unit[i]);
        float *o = dst + rel[u];
        if (!j.h.series.empty())
          o += copy_finite(o, j.h.values.data() + j.h.series[0].first_value, j.h.series[0].n_values, j.h.finite[0]);
        if (!j.b.series.empty()) {
          const size_t bi = std::min(c, j.b.series.size() - 1);
          o += copy_finite(o, j.b.values.data() + j.b.series[bi].first_value, j.b.series[bi].n_values, j.b.finite[bi]);
        }
        o += copy_finite(o, j.c.values.data() + j.c.series[c].first_value, j.c.series[c].n_values, j.c.finite[c]);
        for (uint64_t pad = (uint64_t)(o - (dst + rel[u])); pad & 3; ++pad) *o++ = 0.0f;  // unit starts are 16-byte aligned
      }
    }
  };
  try {
    std::vector<std::thread> pool;
    for (uint32_t t = 1; t < n_threads; ++t) pool.emplace_back(fill);
    fill();
    for (auto &t : pool) t.join();
  } catch (...) {
    return FM_E_NOMEM;
  }
  return FM_OK;
}

}  // extern "C"
