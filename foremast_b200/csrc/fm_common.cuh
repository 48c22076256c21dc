// fm_common.cuh -- shared device-side definitions for libforemast_gpu (sm_100a only).
//
// Numerics contract (oracle/DECISIONS.md "Exact arithmetic"): the library is compiled with
// --fmad=false, every fused multiply-add below is an explicit fma(); that is what makes integer
// statistics and verdicts reproducible bit for bit against the CPU oracle.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/foremast_gpu.h"

#define FM_WARP 32
#define FM_NONE_IDX 0xFFFFFFFFu

// Device view of a sealed batch: ragged CSR in HBM.
struct DevBatch {
  uint64_t n_units;
  const float *values;        // fp32 samples, unit starts 16 B aligned
  const int64_t *offsets;     // [n_units+1]
  const uint32_t *seg_lens;   // [3*n_units] hist, base, cur
  const uint32_t *service_id; // [n_units]
  const uint8_t *metric_type; // [n_units]
  uint64_t first_unit;        // global index of unit 0 (multi-GPU shards)
  // optional (k_band only): thread t scores unit order[t] -- the units of a tile of 2048 sorted by length, so that the
  // 32 units of a warp are about equally long when the batch is ragged (fm_launch_band_order)
  const uint32_t *order = nullptr;
};

struct DevRule {
  double threshold, min_lower_bound;
  uint32_t bound, pad;
};

// fm_params resolved for the device (rows with bound == 0 already replaced by row 0, D16).
struct DevParams {
  uint32_t algorithms, pairwise;
  double pairwise_threshold, differ_scale;
  uint32_t min_mann_white, min_wilcoxon, min_kruskal, min_historical;
  uint32_t min_anomaly_points, ma_window, period, band_algorithm;
  double alpha, beta, gamma;
  DevRule rules[FM_MAX_METRIC_TYPES];
  // Holt-Winters level/trend transition A = [[1-a, 1-a], [-ab, 1-ab]], k = period/32:
  // hw_pow[i] = A^(k*2^i), hw_lane[l] = A^(k*l), row-major
  double hw_pow[5 * 4];
  double hw_lane[32 * 4];
  // The tensor-memory kernel (fm_hw_tm.cu) cuts a season into hw_tl chunks of hw_tk samples, one chunk per lane
  // (fm_hw_tm_lanes): its own powers hw_tpow[i] = A^(tk*2^i), hw_tlane[l] = A^(tk*l), and the chunk weights
  // hw_w[2i], hw_w[2i+1] = A^(tk-1-i) g, g = (a, ab): what sample i of a chunk contributes to the level / trend
  // at the end of the chunk (the chunk offsets as one dot product per component).
  uint32_t hw_tk, hw_tl;
  double hw_tpow[5 * 4];
  double hw_tlane[32 * 4];
  double hw_w[2 * 64];
};

// Sort-size class of a segment of n samples in the rank kernels: the sizes are 32 * {1,2,3,4,6,8,12,16,24,32,48,64} slots.
// Segments that are not ranked at all (empty, or longer than 2048: D17) count as class 0.
#ifdef __CUDACC__
__host__ __device__
#endif
static inline uint32_t fm_rank_class(uint32_t n) {
  if (n > 2048u) return 0u;
  return n <= 32 ? 0u : n <= 64 ? 1u : n <= 96 ? 2u : n <= 128 ? 3u : n <= 192 ? 4u : n <= 256 ? 5u : n <= 384 ? 6u
         : n <= 512 ? 7u : n <= 768 ? 8u : n <= 1024 ? 9u : n <= 1536 ? 10u : 11u;
}
constexpr uint32_t kRankClassSize[12] = {32, 64, 96, 128, 192, 256, 384, 512, 768, 1024, 1536, 2048};

// Lanes the tensor-memory Holt-Winters kernel uses for one season: the largest divisor of the period that is at
// most 32, provided it leaves at most 64 samples per lane and at least half the warp busy; 0 = use k_hw.
static inline uint32_t fm_hw_tm_lanes(uint32_t period) {
  if (period < 16 || period > 2048) return 0;
  for (uint32_t l = 32; l >= 16; --l)
    if (period % l == 0) return period / l <= 64 ? l : 0;
  return 0;
}

__device__ __forceinline__ unsigned fm_lane() { return threadIdx.x & 31u; }

__device__ __forceinline__ int64_t warp_sum_i64(int64_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ uint32_t warp_max_u32(uint32_t v) { return __reduce_max_sync(0xffffffffu, v); }

// 16-byte async global->shared copy (LDGSTS); the staging primitive of the band kernel.
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory");
}

// kernels (one translation unit each)
void fm_launch_synth(const fm_synth_config &cfg, uint64_t stride, float *values, int64_t *offsets,
                     uint32_t *seg_lens, uint32_t *service_id, uint8_t *metric_type, cudaStream_t st);
void fm_launch_batch_stats(const DevBatch &b, uint32_t *out6, cudaStream_t st);
// mask_w / mask_m: the sort-size classes (fm_rank_class) present among the units' pair counts min(n_base, n_cur) and
// longer segments max(n_base, n_cur), 0 = not known.  With more than one class, a scratch of 64 words and room for
// 2 * n_units unit indices, the units are grouped by class and every class runs in its own network; otherwise all
// run in the largest.
cudaError_t fm_launch_rank(const DevBatch &b, const DevParams *dp, uint32_t pairwise, uint32_t max_base, uint32_t max_cur,
                           fm_series_score *scores, int sm_count, cudaStream_t st, cudaEvent_t after_wilcoxon,
                           uint32_t mask_w = 0, uint32_t mask_m = 0, uint32_t *class_scratch = nullptr,
                           uint32_t *class_lists = nullptr);
cudaError_t fm_launch_band(const DevBatch &b, const DevParams *dp, uint32_t algorithms, bool rank_ran,
                           fm_series_score *scores, uint32_t *unit_status, uint32_t *unit_points,
                           float *band_up, float *band_lo, cudaStream_t st);
cudaError_t fm_launch_band_order(const DevBatch &b, uint32_t *order, cudaStream_t st);
cudaError_t fm_launch_hw(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                         uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                         int sm_count, cudaStream_t st);
cudaError_t fm_launch_band_emit(const DevBatch &b, const DevParams *dp, uint32_t algorithm, fm_series_score *scores,
                                const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, cudaStream_t st);
cudaError_t fm_launch_hw_emit(const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                              const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, int sm_count,
                              cudaStream_t st);
cudaError_t fm_launch_hw_fit(const DevBatch &b, const DevParams *dp, uint32_t period, double *state, int sm_count,
                             cudaStream_t st);
cudaError_t fm_launch_hw_resume(const DevBatch &b, const DevParams *dp, uint32_t period, double *state,
                                fm_series_score *scores, uint32_t *unit_status, uint32_t *unit_points, int sm_count,
                                cudaStream_t st);
// fm_hw_tm.cu: the same five modes with the seasonals in tensor memory (period % 32 == 0)
bool fm_hw_tm_supported(uint32_t period);
cudaError_t fm_launch_hw_tm(int mode, const DevBatch &b, const DevParams *dp, uint32_t period, fm_series_score *scores,
                            uint32_t *unit_status, uint32_t *unit_points, float *band_up, float *band_lo,
                            const unsigned long long *emit_off, uint32_t *emit_idx, float *emit_val, double *state,
                            int sm_count, cudaStream_t st);
uint64_t fm_scan_scratch_bytes(uint64_t n);
cudaError_t fm_launch_scan_counts(const void *in, size_t stride, uint64_t n, unsigned long long *out,
                                  unsigned long long *scratch, cudaStream_t st);
void fm_launch_verdict_reduce(const DevBatch &b, const uint32_t *unit_status, const uint32_t *unit_points,
                              uint32_t n_services, uint32_t *counts, unsigned long long *worst, cudaStream_t st);
void fm_launch_verdict_finalize(uint32_t n_services, const uint32_t *counts, const unsigned long long *worst,
                                fm_verdict *out, uint32_t first_service, cudaStream_t st);
void fm_launch_verdict_small(const DevBatch &b, const uint32_t *unit_status, const uint32_t *unit_points,
                             uint32_t n_services, uint32_t *counts, unsigned long long *worst, fm_verdict *out, cudaStream_t st);
