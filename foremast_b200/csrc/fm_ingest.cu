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
// fm_batch_add_series (a gap).  Host code only: nothing here touches the GPU.
#include <charconv>
#include <cmath>
#include <cstring>
#include <limits>

#include "../../include/foremast_gpu.h"

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

}  // namespace

extern "C" int fm_parse_prometheus_matrix(const char *json, uint64_t len, fm_prom_series *series, uint32_t series_cap,
                                          uint32_t *n_series, float *values, uint64_t values_cap, uint64_t *n_values) {
  if (!json || !n_series || !n_values) return FM_E_INVALID;
  Scan sc{json, json + len, json};
  uint32_t ns = 0;
  uint64_t nv = 0;
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
                double t_prev = 0.0;
                while (sc.ok) {
                  if (sc.eat(']')) break;
                  double ts = 0.0;
                  const char *vs; size_t vn;
                  if (!sc.eat('[')) return finish(FM_E_INVALID);
                  // only the first two timestamps carry information (t0 and the step); the rest are skipped
                  if (cur.n_values < 2 ? !sc.number(&ts) : !sc.skip_number()) return finish(FM_E_INVALID);
                  if (!sc.eat(',') || !sc.string(&vs, &vn) || !sc.eat(']')) return finish(FM_E_INVALID);
                  if (cur.n_values == 0) cur.t0 = ts;
                  else if (cur.n_values == 1) cur.step = ts - t_prev;
                  t_prev = ts;
                  if (values && nv < values_cap) values[nv] = (float)sample_value(vs, vn);
                  ++nv;
                  ++cur.n_values;
                  sc.eat(',');
                }
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
