// fm_verdict.cu -- fold per-unit flags into per-service verdicts.
//
// Replaces the brain's final write of `status` into the job document: completed_health /
// completed_unhealth / completed_unknown (foremast-service/pkg/converter/converter.go:16-21;
// fail fast docs/guides/design.md:43; unknown when the current metric is missing design.md:25).
//
// Segmented reduction keyed by service_id into a dense per-service vector.  Units of one service
// are normally adjacent (pods x metrics of one app), so each warp first merges lanes that share a
// service (match.any) and issues one set of reductions per distinct service; the dense vector is
// also exactly what the cross-GPU merge sums/maxes (fm_result_service_buffers).
// Algorithmic bytes: 8 per unit in (+4 service id), 24 per service out.
#include "fm_common.cuh"

namespace {

__global__ void __launch_bounds__(256) k_verdict_reduce(DevBatch b, const uint32_t *__restrict__ unit_status,
                                                        const uint32_t *__restrict__ unit_points, uint32_t n_services,
                                                        uint32_t *__restrict__ counts,
                                                        unsigned long long *__restrict__ worst) {
  const uint64_t n_round = (b.n_units + 31) & ~31ull;  // whole warps stay converged for match/ballot
  for (uint64_t u = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; u < n_round; u += (uint64_t)gridDim.x * blockDim.x) {
    uint32_t sid = 0xFFFFFFFFu, st = 0, pts = 0;
    if (u < b.n_units) { sid = b.service_id[u]; st = unit_status[u]; pts = unit_points[u]; }
    if (sid >= n_services) sid = 0xFFFFFFFFu;  // padding lanes (ids were range-checked at ingest)
    const bool unk = (st & FM_S_UNKNOWN) != 0, anom = (st & FM_S_ANOMALY) != 0;
    const unsigned peers = __match_any_sync(0xffffffffu, sid);
    const unsigned unk_mask = __ballot_sync(0xffffffffu, unk);
    if (sid != 0xFFFFFFFFu) {
      if (fm_lane() == (unsigned)(__ffs(peers) - 1)) {  // one reduction per distinct service in the warp
        atomicAdd(counts + 4ull * sid + 0, (uint32_t)__popc(peers));
        const uint32_t c_unk = __popc(unk_mask & peers);
        if (c_unk) atomicAdd(counts + 4ull * sid + 1, c_unk);
      }
      if (anom) {  // rare: a per-lane reduction is fine
        atomicAdd(counts + 4ull * sid + 2, 1u);
        atomicAdd(counts + 4ull * sid + 3, pts);
        atomicMax(worst + sid, ((unsigned long long)pts << 32) |
                                   (unsigned long long)(0xFFFFFFFFu - (uint32_t)(b.first_unit + u)));
      }
    }
  }
}

// counts / worst / out point at the first service of the range, whose id is `first_service`
__global__ void __launch_bounds__(256) k_verdict_finalize(uint32_t n_services, const uint32_t *__restrict__ counts,
                                                          const unsigned long long *__restrict__ worst,
                                                          fm_verdict *__restrict__ out, uint32_t first_service) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_services) return;
  fm_verdict v;
  v.service_id = first_service + s;
  v.n_units = counts[4ull * s + 0];
  v.n_unknown = counts[4ull * s + 1];
  v.n_anom_units = counts[4ull * s + 2];
  v.n_anom_points = counts[4ull * s + 3];
  const unsigned long long w = worst[s];
  if (v.n_anom_units) {
    v.verdict = FM_V_UNHEALTH;
    v.worst_unit = 0xFFFFFFFFu - (uint32_t)(w & 0xFFFFFFFFull);
    v.worst_unit_points = (uint32_t)(w >> 32);
  } else {
    v.verdict = (v.n_unknown == v.n_units) ? FM_V_UNKNOWN : FM_V_HEALTH;
    v.worst_unit = FM_NONE_IDX;
    v.worst_unit_points = 0;
  }
  out[s] = v;
}

// One block does both steps for a small job (at most a few thousand units): reduce, block barrier, finalize --
// one launch less on the latency path of fm_score_host's small-job arena.
__device__ __forceinline__ void verdict_reduce_unit(const DevBatch &b, uint64_t u, const uint32_t *__restrict__ unit_status,
                                                    const uint32_t *__restrict__ unit_points, uint32_t n_services,
                                                    uint32_t *__restrict__ counts, unsigned long long *__restrict__ worst) {
  uint32_t sid = 0xFFFFFFFFu, st = 0, pts = 0;
  if (u < b.n_units) { sid = b.service_id[u]; st = unit_status[u]; pts = unit_points[u]; }
  if (sid >= n_services) sid = 0xFFFFFFFFu;
  const bool unk = (st & FM_S_UNKNOWN) != 0, anom = (st & FM_S_ANOMALY) != 0;
  const unsigned peers = __match_any_sync(0xffffffffu, sid);
  const unsigned unk_mask = __ballot_sync(0xffffffffu, unk);
  if (sid != 0xFFFFFFFFu) {
    if (fm_lane() == (unsigned)(__ffs(peers) - 1)) {
      atomicAdd(counts + 4ull * sid + 0, (uint32_t)__popc(peers));
      const uint32_t c_unk = __popc(unk_mask & peers);
      if (c_unk) atomicAdd(counts + 4ull * sid + 1, c_unk);
    }
    if (anom) {
      atomicAdd(counts + 4ull * sid + 2, 1u);
      atomicAdd(counts + 4ull * sid + 3, pts);
      atomicMax(worst + sid, ((unsigned long long)pts << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)(b.first_unit + u)));
    }
  }
}
__global__ void __launch_bounds__(256) k_verdict_small(DevBatch b, const uint32_t *__restrict__ unit_status,
                                                       const uint32_t *__restrict__ unit_points, uint32_t n_services,
                                                       uint32_t *counts, unsigned long long *worst, fm_verdict *__restrict__ out) {
  const uint64_t n_round = (b.n_units + 31) & ~31ull;
  for (uint64_t u = threadIdx.x; u < n_round; u += blockDim.x) verdict_reduce_unit(b, u, unit_status, unit_points, n_services, counts, worst);
  __threadfence();
  __syncthreads();
  for (uint32_t s = threadIdx.x; s < n_services; s += blockDim.x) {
    fm_verdict v;
    v.service_id = s;
    v.n_units = __ldcg(counts + 4ull * s + 0);  // the atomics above went to L2
    v.n_unknown = __ldcg(counts + 4ull * s + 1);
    v.n_anom_units = __ldcg(counts + 4ull * s + 2);
    v.n_anom_points = __ldcg(counts + 4ull * s + 3);
    const unsigned long long w = __ldcg(worst + s);
    if (v.n_anom_units) {
      v.verdict = FM_V_UNHEALTH;
      v.worst_unit = 0xFFFFFFFFu - (uint32_t)(w & 0xFFFFFFFFull);
      v.worst_unit_points = (uint32_t)(w >> 32);
    } else {
      v.verdict = (v.n_unknown == v.n_units) ? FM_V_UNKNOWN : FM_V_HEALTH;
      v.worst_unit = FM_NONE_IDX;
      v.worst_unit_points = 0;
    }
    out[s] = v;
  }
}

}  // namespace

void fm_launch_verdict_small(const DevBatch &b, const uint32_t *unit_status, const uint32_t *unit_points,
                             uint32_t n_services, uint32_t *counts, unsigned long long *worst, fm_verdict *out, cudaStream_t st) {
  k_verdict_small<<<1, 256, 0, st>>>(b, unit_status, unit_points, n_services, counts, worst, out);
}

void fm_launch_verdict_reduce(const DevBatch &b, const uint32_t *unit_status, const uint32_t *unit_points,
                              uint32_t n_services, uint32_t *counts, unsigned long long *worst, cudaStream_t st) {
  uint64_t blocks = (b.n_units + 255) / 256;
  if (blocks > 148ull * 16) blocks = 148ull * 16;
  if (blocks == 0) return;
  k_verdict_reduce<<<(unsigned)blocks, 256, 0, st>>>(b, unit_status, unit_points, n_services, counts, worst);
}

void fm_launch_verdict_finalize(uint32_t n_services, const uint32_t *counts, const unsigned long long *worst,
                                fm_verdict *out, uint32_t first_service, cudaStream_t st) {
  if (n_services == 0) return;
  k_verdict_finalize<<<(n_services + 255) / 256, 256, 0, st>>>(n_services, counts, worst, out, first_service);
}
