// fm_hw_common.cuh -- what the two Holt-Winters kernels share (fm_hw.cu: seasonals in shared memory,
// any period; fm_hw_tm.cu: seasonals in tensor memory, period % 32 == 0): the per-unit gate, the warp
// scan of the chunk maps, the anomaly bookkeeping of one super-tile, and the result write.
// MODE: 0 score, 1 score + bands, 2 emit pass of the anomaly list (see fm_band.cu), 3 fit (run the
// model window only and save level, trend, sigma, phase, seasonals per unit), 4 resume (load that
// state, score the current segment from it, save the advanced state) -- 3 / 4 are the continuous-mode
// model cache (the brain's MAX_CACHE_SIZE, foremast-brain/README.md:30; jobs re-armed every cycle
// with START_TIME/END_TIME, foremast-service/cmd/manager/main.go:59-63).
#pragma once

#include "fm_common.cuh"

namespace fm_hw {

struct V2 { double l, b; };

__device__ __forceinline__ V2 matvec(const double *m, V2 v) {  // m = {a00, a01, a10, a11}
  V2 r;
  r.l = fma(m[0], v.l, m[1] * v.b);
  r.b = fma(m[2], v.l, m[3] * v.b);
  return r;
}

// One series-window as the kernels see it.
struct Unit {
  int64_t off;                 // first sample of the unit in `values`
  uint32_t nh, nb, nc;         // segment lengths
  uint32_t a0, ns;             // model window: start inside the unit, length (0 when resuming)
  uint32_t c0, n_stream;       // start of the current segment; samples streamed = ns (+ nc unless fitting)
  const float *unit;           // values + off
  double *ust;                 // model-cache record {l, b, sigma, phase | -1, S[period]} (fit / resume)
  unsigned long long emit_base;
  uint32_t st;                 // status word so far
  double kk, mlb;              // threshold (after the pairwise adjustment), lower clamp
  uint32_t bound;
  __device__ __forceinline__ const float *at(uint32_t q) const { return unit + (q < ns ? a0 + q : c0 + (q - ns)); }
};

// Decides whether unit u is scored by this launch and fills `w`; units that are not get their
// "unknown" record (resume) or "no model" mark (fit) here.  Warp-uniform.
template <int MODE>
__device__ __forceinline__ bool unit_begin(const DevBatch &b, const DevParams *__restrict__ dp, uint64_t u, unsigned lane,
                                           uint32_t m, fm_series_score *__restrict__ scores,
                                           uint32_t *__restrict__ unit_status, uint32_t *__restrict__ unit_points,
                                           const unsigned long long *__restrict__ emit_off, double *__restrict__ state,
                                           Unit &w) {
  constexpr bool EMIT = MODE == 2, FIT = MODE == 3, RESUME = MODE == 4;
  w.off = b.offsets[u];
  w.nh = b.seg_lens[3 * u]; w.nb = b.seg_lens[3 * u + 1]; w.nc = b.seg_lens[3 * u + 2];
  w.a0 = 0; w.ns = 0;
  if (w.nh >= dp->min_historical && w.nh > 0) { w.a0 = 0; w.ns = w.nh; }
  else if (w.nb >= dp->min_historical && w.nb > 0) { w.a0 = w.nh; w.ns = w.nb; }
  w.ust = (FIT || RESUME) ? state + u * (uint64_t)(m + 4) : nullptr;
  if (FIT) {
    if ((uint64_t)w.ns < 2ull * m) {
      if (lane == 0) w.ust[3] = -1.0;  // no model for this unit
      return false;
    }
  } else if (RESUME) {
    const bool ok = w.nc > 0 && w.ust[3] >= 0.0;
    if (!ok) {
      if (lane == 0) {
        fm_series_score *o = scores + u;
#pragma unroll
        for (int a = 0; a < FM_N_ALG; ++a) o->first_idx[a] = FM_NONE_IDX;
        o->status = FM_S_UNKNOWN;
        o->p_mw = 1.0; o->p_w = 1.0; o->p_k = 1.0;
        unit_status[u] = FM_S_UNKNOWN;
        unit_points[u] = 0;
      }
      return false;
    }
    w.ns = 0;
  } else if (w.ns == 0 || w.nc == 0 || (uint64_t)w.ns < 2ull * m) {
    return false;  // unknown unit, or not two seasons of model window
  }
  w.emit_base = 0;
  if (EMIT) {
    w.emit_base = emit_off[u];
    if (emit_off[u + 1] == w.emit_base) return false;
  }
  w.c0 = w.nh + w.nb;
  w.n_stream = FIT ? w.ns : w.ns + w.nc;
  w.unit = b.values + w.off;
  w.st = (FIT || RESUME) ? 0u : scores[u].status;
  const DevRule &dr = dp->rules[b.metric_type[u] < FM_MAX_METRIC_TYPES ? b.metric_type[u] : 0];
  w.kk = (w.st & FM_S_DIFFERS) ? dr.threshold * dp->differ_scale : dr.threshold;
  w.mlb = dr.min_lower_bound;
  w.bound = dr.bound;
  return true;
}

// State entering each lane's chunk: inclusive warp scan of the chunk maps v -> A^k v + u with the
// precomputed A^(k*2^i) (apow), shifted one lane, plus the carry of the previous super-tile
// through A^(k*lane) (alane).
template <bool CARRY = true>  // false: the super-tile starts from a zero state (carry and alane are not read)
__device__ __forceinline__ V2 scan_chunks(V2 uo, V2 carry, const double *apow, const double *alane, unsigned lane) {
  V2 inc = uo;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int o = 1 << i;
    V2 p;
    p.l = __shfl_up_sync(0xffffffffu, inc.l, o);
    p.b = __shfl_up_sync(0xffffffffu, inc.b, o);
    if ((int)lane >= o) {  // inc covers the last min(o, lane+1) chunks ending here: shift the earlier block through them
      const V2 t = matvec(apow + 4 * i, p);
      inc.l += t.l;
      inc.b += t.b;
    }
  }
  V2 vin;
  vin.l = __shfl_up_sync(0xffffffffu, inc.l, 1);
  vin.b = __shfl_up_sync(0xffffffffu, inc.b, 1);
  if (lane == 0) { vin.l = 0.0; vin.b = 0.0; }
  if (CARRY) {
    const V2 cv = matvec(alane, carry);
    vin.l += cv.l;
    vin.b += cv.b;
  }
  return vin;
}

// Anomaly bookkeeping of one super-tile: fmask bit i = local sample i of this lane's chunk is
// flagged; qa = stream position of the chunk; xrow = the lane's samples (emit pass).  Chunks are in
// stream order, so an exclusive prefix over lanes gives each flag its slot in the unit's list.
template <bool EMIT>
__device__ __forceinline__ void tile_flags(unsigned long long fmask, unsigned lane, uint32_t qa, uint32_t ns,
                                           unsigned long long emit_base, uint32_t *__restrict__ emit_idx,
                                           float *__restrict__ emit_val, const float *xrow, uint32_t &n_anom,
                                           uint32_t &first) {
  const uint32_t cnt = __popcll(fmask);
  uint32_t pre = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, pre, o);
    if ((int)lane >= o) pre += t;
  }
  const uint32_t total = __shfl_sync(0xffffffffu, pre, 31);
  pre -= cnt;
  if (!total) return;
  const unsigned has = __ballot_sync(0xffffffffu, cnt != 0);
  if (n_anom == 0) {
    const int src = __ffs(has) - 1;
    const uint32_t cand = qa + (uint32_t)(__ffsll((long long)fmask) - 1) - ns;
    first = __shfl_sync(0xffffffffu, cand, src);
  }
  if (EMIT) {
    unsigned long long mk = fmask;
    uint32_t r = 0;
    while (mk) {
      const int bit = __ffsll((long long)mk) - 1;
      mk &= mk - 1;
      const unsigned long long at = emit_base + n_anom + pre + r++;
      emit_idx[at] = qa + bit - ns;
      emit_val[at] = xrow[bit];
    }
  }
  n_anom += total;
}

// Scalar part of the saved model (lane 0) -- the seasonals are written by the caller.
__device__ __forceinline__ void save_state_scalars(double *ust, V2 carry, double sigma, uint32_t phase, uint32_t n_stream, uint32_t m) {
  ust[0] = carry.l;
  ust[1] = carry.b;
  ust[2] = sigma;
  ust[3] = (double)((phase + n_stream) % m);
}

// The unit's Holt-Winters result (lane 0; not in the emit and fit passes).
template <int MODE>
__device__ __forceinline__ void unit_finish(const DevParams *__restrict__ dp, uint64_t u, const Unit &w, uint32_t n_anom,
                                            uint32_t first, double sigma, fm_series_score *__restrict__ scores,
                                            uint32_t *__restrict__ unit_status, uint32_t *__restrict__ unit_points) {
  constexpr bool RESUME = MODE == 4;
  fm_series_score *o = scores + u;
  uint32_t st = w.st;
  if (RESUME) {
#pragma unroll
    for (int a = 0; a < FM_N_ALG; ++a) o->first_idx[a] = FM_NONE_IDX;
    o->p_mw = 1.0; o->p_w = 1.0; o->p_k = 1.0;
    o->k_eff = w.kk;
    unit_points[u] = 0;
  }
  st |= 1u << (FM_S_ALG_SHIFT + 4);
  if (n_anom > 0 && n_anom >= dp->min_anomaly_points) st |= FM_S_ANOMALY;
  o->status = st;
  o->n_anom[4] = n_anom;
  o->first_idx[4] = first;
  o->sigma[4] = sigma;
  unit_status[u] = st;
  unit_points[u] = max(unit_points[u], n_anom);
}

}  // namespace fm_hw
