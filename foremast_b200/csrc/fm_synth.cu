// fm_synth.cu -- on-device synthetic workload generator and batch statistics.
//
// The generator is the counter-based recipe of oracle/DECISIONS.md "Synthetic workload":
// Philox4x32-10 keyed by (seed, 0x466F7265), one call per sample, every float operation a single
// IEEE fp32 step (the library is built with --fmad=false), so the CPU twin in oracle/fm_synth.c
// yields the same bits and benchmarks never pay PCIe for their inputs.
#include "fm_common.cuh"

namespace {

struct U4 { uint32_t v[4]; };

__device__ __forceinline__ U4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    uint32_t h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
    uint32_t h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
    uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
    c0 = n0; c1 = l1; c2 = n2; c3 = l0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  U4 o;
  o.v[0] = c0; o.v[1] = c1; o.v[2] = c2; o.v[3] = c3;
  return o;
}

__device__ __forceinline__ float u01(uint32_t w) { return (float)(w >> 8) * 5.9604644775390625e-8f; }

__constant__ float kBase[9] = {1.0f, 0.2f, 1.0f, 180.0f, 0.45f, 6.0e8f, 50.0f, 0.02f, 40.0f};
__constant__ float kSigma[9] = {0.1f, 0.15f, 0.5f, 40.0f, 0.08f, 3.0e7f, 12.0f, 0.01f, 6.0f};

__device__ __forceinline__ void unit_lens(const fm_synth_config &cfg, uint64_t gu, uint32_t lens[3]) {
  lens[0] = cfg.n_hist; lens[1] = cfg.n_base; lens[2] = cfg.n_cur;
  if (cfg.flags & FM_SYNTH_RAGGED) {
    U4 s = philox4x32_10((uint32_t)gu, (uint32_t)(gu >> 32), 1u, 0xAu, cfg.seed, 0x466F7265u);
    if (s.v[0] % 50u == 0u) {
#pragma unroll
      for (int k = 0; k < 3; ++k)
        if (lens[k] > 0) {
          uint32_t lose = (s.v[1] >> (8 * k)) % 16u;
          lens[k] = lens[k] > lose ? lens[k] - lose : 1u;
        }
    }
    if (s.v[0] % 1000u == 7u) lens[2] = 0;
    // a whole service whose new pods have not reported yet: no current window in any of its series
    const uint64_t svc = gu / cfg.units_per_service;
    U4 w = philox4x32_10((uint32_t)svc, (uint32_t)(svc >> 32), 2u, 0xAu, cfg.seed, 0x466F7265u);
    if (w.v[0] % 500u == 3u) lens[2] = 0;
  }
}

__device__ __forceinline__ float irwin_hall(const U4 &q) {
  return (((u01(q.v[0]) + u01(q.v[1])) + (u01(q.v[2]) + u01(q.v[3]))) - 2.0f) * 1.7320508f;
}

// t = unit time of the sample; baseline and current windows are CONCURRENT (old and new pods of one
// deployment over the same minutes): a current sample shares t -- season, drift, the load-driven noise
// term -- with the baseline sample beside it and adds a smaller pod-private term.
__device__ __forceinline__ float sample(const fm_synth_config &cfg, uint64_t gu, uint32_t mt, const U4 &r,
                                        uint32_t t, bool in_cur, uint32_t cur_idx) {
  float base_m = kBase[mt], sigma_m = kSigma[mt];
  float phi = u01(r.v[0]);
  float drift = (u01(r.v[1]) - 0.5f) * base_m * 1.0e-5f;
  U4 q = philox4x32_10((uint32_t)gu, (uint32_t)(gu >> 32), t, 0xBu, cfg.seed, 0x466F7265u);
  float noise = irwin_hall(q);
  if (in_cur) {
    U4 w = philox4x32_10((uint32_t)gu, (uint32_t)(gu >> 32), t, 0xCu, cfg.seed, 0x466F7265u);
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
  if (in_cur && (cfg.flags & FM_SYNTH_ANOMALIES)) {
    if (r.v[2] % 400u == 0u && cfg.n_cur > 0 && cur_idx == r.v[3] % cfg.n_cur) x = x * 20.0f + 5.0f * base_m;
    if (r.v[2] % 800u == 1u) x = x * 1.5f + sigma_m;
    if (mt == FM_MT_LATENCY && r.v[2] % 8u == 5u) x = x * 1.04f;  /* benign shift: differs, still inside the k = 10 band */
  }
  if (!(x > 0.0f)) x = 0.0f;
  if (mt == FM_MT_ERROR5XX || mt == FM_MT_ERROR4XX) x = rintf(x * 1000.0f) * 0.001f;
  return x;
}

// One thread per 4 consecutive elements of a unit's stride (one 16 B store).
__global__ void __launch_bounds__(256) k_synth(fm_synth_config cfg, uint64_t stride, float *values, int64_t *offsets,
                                               uint32_t *seg_lens, uint32_t *service_id, uint8_t *metric_type) {
  const uint64_t q4 = stride >> 2;
  const uint64_t total = cfg.n_units * q4;
  for (uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total + 8;
       idx += (uint64_t)gridDim.x * blockDim.x) {
    if (idx >= total) {  // 32 zero elements after the last unit
      reinterpret_cast<float4 *>(values)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (idx == total) offsets[cfg.n_units] = (int64_t)(cfg.n_units * stride);
      continue;
    }
    const uint64_t u = idx / q4;
    const uint32_t q = (uint32_t)(idx - u * q4);
    const uint64_t gu = cfg.first_unit + u;
    const uint32_t mt = 1u + (uint32_t)(gu % cfg.n_metric_types);
    uint32_t lens[3];
    unit_lens(cfg, gu, lens);
    if (q == 0) {
      offsets[u] = (int64_t)(u * stride);
      seg_lens[3 * u] = lens[0]; seg_lens[3 * u + 1] = lens[1]; seg_lens[3 * u + 2] = lens[2];
      service_id[u] = (uint32_t)(gu / cfg.units_per_service);
      metric_type[u] = (uint8_t)mt;
    }
    const U4 r = philox4x32_10((uint32_t)gu, (uint32_t)(gu >> 32), 0u, 0xAu, cfg.seed, 0x466F7265u);
    float o[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      uint32_t k = 4u * q + c;
      float x = 0.0f;
      if (k < lens[0]) x = sample(cfg, gu, mt, r, k, false, 0);
      else if (k < lens[0] + lens[1]) x = sample(cfg, gu, mt, r, cfg.n_hist + (k - lens[0]), false, 0);
      else if (k < lens[0] + lens[1] + lens[2]) {
        uint32_t i = k - lens[0] - lens[1];
        x = sample(cfg, gu, mt, r, cfg.n_hist + i, true, i);
      }
      o[c] = x;
    }
    reinterpret_cast<float4 *>(values)[idx] = make_float4(o[0], o[1], o[2], o[3]);
  }
}

// out6 = {max n_hist, max n_base, max n_cur, max service_id, #misaligned offsets, #overlapping units,
//         rank sort-size classes present by pair count, by longer segment (fm_rank_class)}
__global__ void __launch_bounds__(256) k_batch_stats(DevBatch b, uint32_t *out6) {
  uint32_t mh = 0, mb = 0, mc = 0, ms = 0, bad = 0, ovl = 0, kw = 0, km = 0;
  for (uint64_t u = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; u < b.n_units; u += (uint64_t)gridDim.x * blockDim.x) {
    uint32_t nh = b.seg_lens[3 * u], nb = b.seg_lens[3 * u + 1], nc = b.seg_lens[3 * u + 2];
    mh = max(mh, nh); mb = max(mb, nb); mc = max(mc, nc);
    kw |= 1u << fm_rank_class(min(nb, nc)); km |= 1u << fm_rank_class(max(nb, nc));
    ms = max(ms, b.service_id[u]);
    int64_t o = b.offsets[u], o1 = b.offsets[u + 1];
    bad += (o & 3) != 0 || o < 0;
    ovl += o + (int64_t)nh + nb + nc > o1;
  }
  mh = __reduce_max_sync(0xffffffffu, mh); mb = __reduce_max_sync(0xffffffffu, mb);
  mc = __reduce_max_sync(0xffffffffu, mc); ms = __reduce_max_sync(0xffffffffu, ms);
  bad = __reduce_add_sync(0xffffffffu, bad); ovl = __reduce_add_sync(0xffffffffu, ovl);
  kw = __reduce_or_sync(0xffffffffu, kw); km = __reduce_or_sync(0xffffffffu, km);
  if ((threadIdx.x & 31) == 0) {
    atomicMax(out6 + 0, mh); atomicMax(out6 + 1, mb); atomicMax(out6 + 2, mc); atomicMax(out6 + 3, ms);
    if (bad) atomicAdd(out6 + 4, bad);
    if (ovl) atomicAdd(out6 + 5, ovl);
    atomicOr(out6 + 6, kw); atomicOr(out6 + 7, km);
  }
}

}  // namespace

void fm_launch_synth(const fm_synth_config &cfg, uint64_t stride, float *values, int64_t *offsets,
                     uint32_t *seg_lens, uint32_t *service_id, uint8_t *metric_type, cudaStream_t st) {
  uint64_t total = cfg.n_units * (stride >> 2) + 8;
  uint64_t blocks = (total + 255) / 256;
  if (blocks > 148ull * 64) blocks = 148ull * 64;
  k_synth<<<(unsigned)blocks, 256, 0, st>>>(cfg, stride, values, offsets, seg_lens, service_id, metric_type);
}

void fm_launch_batch_stats(const DevBatch &b, uint32_t *out6, cudaStream_t st) {
  uint64_t blocks = (b.n_units + 255) / 256;
  if (blocks > 148ull * 8) blocks = 148ull * 8;
  if (blocks == 0) blocks = 1;
  k_batch_stats<<<(unsigned)blocks, 256, 0, st>>>(b, out6);
}
