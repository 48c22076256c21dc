// fm_hw.cu -- Holt-Winters additive seasonal scoring: "Holt-Winters" in the
// brain's model list (docs/guides/design.md:71) over the 7-day history that barrelman requests
// (foremast-barrelman/pkg/client/metrics/metricsquery.go:95-97), period 1440 = one day at 60 s.
//
// Mapping: ONE WARP PER SERIES-WINDOW.  The seasonal array (period fp64 values, 11.25 KB at 1440)
// lives in shared memory, so a thread-per-series layout is impossible; instead the warp walks the
// stream 32 samples at a time.  Given last season's S, level and trend obey a 2-state linear
// recurrence with constant matrix
//     [l'; b'] = A [l; b] + g * (x - S),  A = [[1-a, 1-a], [-ab, 1-ab]],  g = [a; ab]
// so each 32-sample tile is an inclusive warp scan of affine maps: five SHFL.UP steps with the
// precomputed powers A^1..A^16, then the carry from the previous tile enters through A^(lane+1).
// The season update S[i] += g1*e is pointwise once the one-step error e is known.  fp64
// throughout; the scan re-associates the sums, so floats match the sequential oracle to ~1e-15
// relative (tested at 1e-5), not bit for bit.
// Algorithmic bytes per unit: 4*(n_model + n_cur) + 8 + 12 + 32.
#include "fm_common.cuh"

namespace {

constexpr int kWarpsPerBlock = 4;
constexpr int kGroup = 8;  // tiles whose loads are issued together (8 x 128 B in flight per warp)

struct V2 { double l, b; };

__device__ __forceinline__ V2 matvec(const double *m, V2 v) {  // m = {a00, a01, a10, a11}
  V2 r;
  r.l = fma(m[0], v.l, m[1] * v.b);
  r.b = fma(m[2], v.l, m[3] * v.b);
  return r;
}

// MODE 0 score, 1 score + bands, 2 emit pass of the anomaly list (see fm_band.cu)
template <int MODE>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) k_hw(DevBatch b, const DevParams *__restrict__ dp,
                                                            fm_series_score *__restrict__ scores,
                                                            uint32_t *__restrict__ unit_status,
                                                            uint32_t *__restrict__ unit_points,
                                                            float *__restrict__ band_up, float *__restrict__ band_lo,
                                                            const unsigned long long *__restrict__ emit_off,
                                                            uint32_t *__restrict__ emit_idx, float *__restrict__ emit_val) {
  constexpr bool BANDS = MODE == 1, EMIT = MODE == 2;
  extern __shared__ double s_season[];  // [warps][period]
  const unsigned lane = fm_lane(), warp = threadIdx.x >> 5;
  const uint32_t m = dp->period;
  double *const S = s_season + (size_t)warp * m;
  const double alpha = dp->alpha, ab = dp->alpha * dp->beta, g1 = dp->gamma * (1.0 - dp->alpha);
  const double *const Apow = dp->hw_pow;       // A^(2^i), i = 0..4
  const double *const Alane = dp->hw_lane + 4 * lane;  // A^(lane+1)

  for (uint64_t u = (uint64_t)blockIdx.x * kWarpsPerBlock + warp; u < b.n_units;
       u += (uint64_t)gridDim.x * kWarpsPerBlock) {
    const int64_t off = b.offsets[u];
    const uint32_t nh = b.seg_lens[3 * u], nb = b.seg_lens[3 * u + 1], nc = b.seg_lens[3 * u + 2];
    uint32_t a0 = 0, ns = 0;
    if (nh >= dp->min_historical && nh > 0) { a0 = 0; ns = nh; }
    else if (nb >= dp->min_historical && nb > 0) { a0 = nh; ns = nb; }
    if (ns == 0 || nc == 0 || (uint64_t)ns < 2ull * m) continue;  // unknown unit, or not two seasons of model window
    unsigned long long emit_base = 0;
    if (EMIT) {
      emit_base = emit_off[u];
      if (emit_off[u + 1] == emit_base) continue;
    }
    const uint32_t c0 = nh + nb, n_stream = ns + nc;
    const float *unit = b.values + off;
    auto sample = [&](uint32_t q) -> double { return (double)unit[q < ns ? a0 + q : c0 + (q - ns)]; };

    uint32_t st = scores[u].status;
    const DevRule &dr = dp->rules[b.metric_type[u] < FM_MAX_METRIC_TYPES ? b.metric_type[u] : 0];
    const double k = (st & FM_S_DIFFERS) ? dr.threshold * dp->differ_scale : dr.threshold;
    const double mlb = dr.min_lower_bound;
    const uint32_t bound = dr.bound;

    // ---- initialisation from the first two seasons (D5)
    double s0 = 0.0, s1 = 0.0;
    for (uint32_t i = lane; i < m; i += 32) { s0 += sample(i); s1 += sample(m + i); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      s0 += __shfl_xor_sync(0xffffffffu, s0, o);
      s1 += __shfl_xor_sync(0xffffffffu, s1, o);
    }
    const double L0 = s0 / (double)m, M2 = s1 / (double)m;
    V2 carry;
    carry.l = L0;
    carry.b = (M2 - L0) / (double)m;
    __syncwarp();
    for (uint32_t i = lane; i < m; i += 32) S[i] = sample(i) - L0;
    __syncwarp();

    double sse = 0.0, sigma = 0.0;
    bool sigma_ready = false;
    uint32_t n_anom = 0, first = FM_NONE_IDX;
    uint32_t idx = lane % m;  // (m + tile*32 + lane) mod m, kept incrementally
    const uint32_t n_tiles = (n_stream - m + 31) / 32;

    for (uint32_t tg = 0; tg < n_tiles; tg += kGroup) {
      double xs[kGroup];
#pragma unroll
      for (int g = 0; g < kGroup; ++g) {
        const uint32_t q = m + (tg + g) * 32 + lane;
        xs[g] = q < n_stream ? sample(q) : 0.0;
      }
#pragma unroll
      for (int g = 0; g < kGroup; ++g) {
        const uint32_t tile = tg + g;
        if (tile >= n_tiles) break;
        const uint32_t q = m + tile * 32 + lane;
        const bool valid = q < n_stream;
        const double x = xs[g];
        const double seas = S[idx];
        const double y = valid ? x - seas : 0.0;
        // inclusive scan of u_t = g*y_t under v -> A v
        V2 uv;
        uv.l = alpha * y;
        uv.b = ab * y;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
          const int o = 1 << i;
          V2 p;
          p.l = __shfl_up_sync(0xffffffffu, uv.l, o);
          p.b = __shfl_up_sync(0xffffffffu, uv.b, o);
          if ((int)lane >= o) {
            const V2 t = matvec(Apow + 4 * i, p);
            uv.l += t.l;
            uv.b += t.b;
          }
        }
        const V2 cv = matvec(Alane, carry);
        V2 v;  // state after sample q
        v.l = uv.l + cv.l;
        v.b = uv.b + cv.b;
        V2 pv;  // state before sample q
        pv.l = __shfl_up_sync(0xffffffffu, v.l, 1);
        pv.b = __shfl_up_sync(0xffffffffu, v.b, 1);
        if (lane == 0) pv = carry;
        const double f = (pv.l + pv.b) + seas;
        const double e = x - f;
        if (valid) S[idx] = fma(g1, e, seas);
        const bool in_src = valid && q < ns, in_cur = valid && q >= ns;
        if (in_src) sse = fma(e, e, sse);
        const unsigned cur_mask = __ballot_sync(0xffffffffu, in_cur);
        if (cur_mask) {
          if (!sigma_ready) {  // the model window ends in this tile: close the residual sum first
            double t = sse;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
            sigma = sqrt(t / (double)(ns - m));
            sigma_ready = true;
          }
          const double up = fma(k, sigma, f);
          double lo = fma(-k, sigma, f);
          lo = lo < mlb ? mlb : lo;
          const bool flag = in_cur && (((bound & 1u) && x > up) || ((bound & 2u) && x < lo));
          const unsigned fm = __ballot_sync(0xffffffffu, flag);
          if (fm) {
            if (n_anom == 0) first = (m + tile * 32 + (__ffs(fm) - 1)) - ns;
            if (EMIT && flag) {
              const unsigned long long at = emit_base + n_anom + __popc(fm & ((1u << lane) - 1u));
              emit_idx[at] = q - ns;
              emit_val[at] = (float)x;
            }
            n_anom += __popc(fm);
          }
          if (BANDS && in_cur) {
            band_up[off + c0 + (q - ns)] = (float)up;
            band_lo[off + c0 + (q - ns)] = (float)lo;
          }
        }
        // carry = state after the last valid sample of the tile
        const uint32_t nvalid = min(32u, n_stream - (m + tile * 32));
        carry.l = __shfl_sync(0xffffffffu, v.l, nvalid - 1);
        carry.b = __shfl_sync(0xffffffffu, v.b, nvalid - 1);
        idx += 32;
        if (idx >= m) idx -= m;
        __syncwarp();  // S[] written by this tile may be read by the next one when period < 64
      }
    }
    if (lane == 0 && !EMIT) {
      fm_series_score *o = scores + u;
      st |= 1u << (FM_S_ALG_SHIFT + 4);
      if (n_anom > 0 && n_anom >= dp->min_anomaly_points) st |= FM_S_ANOMALY;
      o->status = st;
      o->n_anom[4] = n_anom;
      o->first_idx[4] = first;
      o->sigma[4] = sigma;
      unit_status[u] = st;
      unit_points[u] = max(unit_points[u], n_anom);
    }
  }
}

}  // namespace

template <int MODE>
static cudaError_t launch_hw(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                             uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                             const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, int sm_count,
                             cudaStream_t st) {
  uint64_t blocks = (b.n_units + kWarpsPerBlock - 1) / kWarpsPerBlock;
  const uint64_t cap = (uint64_t)sm_count * 32;
  if (blocks > cap) blocks = cap;
  if (blocks == 0) return cudaSuccess;
  const int smem = kWarpsPerBlock * (int)period * (int)sizeof(double);
  cudaError_t e = cudaFuncSetAttribute(k_hw<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) return e;
  k_hw<MODE><<<(unsigned)blocks, kWarpsPerBlock * 32, smem, st>>>(b, dp, scores, unit_status, unit_points, band_up, band_lo,
                                                               emit_off, emit_idx, emit_val);
  return cudaGetLastError();
}

// Runs after fm_launch_band (it merges into the statuses that kernel wrote).  period must be in
// [32, 2048]; the caller checks.
cudaError_t fm_launch_hw(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                         uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                         int sm_count, cudaStream_t st) {
  if (band_up && band_lo)
    return launch_hw<1>(b, dp, period, scores, unit_status, unit_points, band_up, band_lo, nullptr, nullptr, nullptr, sm_count, st);
  return launch_hw<0>(b, dp, period, scores, unit_status, unit_points, nullptr, nullptr, nullptr, nullptr, nullptr, sm_count, st);
}

cudaError_t fm_launch_hw_emit(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                              const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, int sm_count,
                              cudaStream_t st) {
  return launch_hw<2>(b, dp, period, scores, nullptr, nullptr, nullptr, nullptr, emit_off, emit_idx, emit_val, sm_count, st);
}
