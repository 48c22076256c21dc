// fm_internal.h -- what the translation units of libforemast_gpu share about the opaque handles of
// include/foremast_gpu.h (host side only; fm_api.cu owns the definitions).
#pragma once
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "../../include/foremast_gpu.h"

struct fm_comm_state {
  void *comm = nullptr;  // ncclComm_t
  int rank = 0, world = 1;
};

std::mutex &fm_ctx_mutex(fm_ctx *ctx);
fm_comm_state *fm_ctx_comm(fm_ctx *ctx);
int fm_ctx_device(const fm_ctx *ctx);
cudaStream_t fm_ctx_stream(const fm_ctx *ctx);
// records the message under the context's error lock and returns `code`
int fm_set_error(fm_ctx *ctx, int code, const std::string &msg);
int fm_comm_merge_counters(fm_ctx *ctx, uint32_t *d_counts, unsigned long long *d_worst, uint32_t n_services);

// Bulk form of fm_batch_add_series for fm_ingest.cu: appends n units whose (already filtered) segment lengths
// are seg_lens[3n], returns where each unit's samples go (rel_off[k] elements past *values_out; unit starts are
// padded to 4) so that several threads can fill them side by side.  The region is NOT initialised: the caller
// writes every sample and zeroes the up to three padding elements after each unit.
int fm_batch_reserve_units(fm_batch *b, uint64_t n, const uint32_t *seg_lens, const uint32_t *service_id,
                           const uint8_t *metric_type, float **values_out, uint64_t *rel_off);
