/*
 * fm_oracle.h -- CPU restatement of the foremast-brain scoring pass.  TEST INFRASTRUCTURE ONLY.
 *
 * PARITY UNPINNED: the reference's scorer source is absent from the mount and no reference test
 * pins a score (oracle/DECISIONS.md).  This oracle follows the documented
 * pipeline, names and defaults; the only numeric pins are the scipy cross-checks on the
 * reference's demo fixture (tests/golden/).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this; the product library never links or calls it.
 */
#ifndef FM_ORACLE_H
#define FM_ORACLE_H

#include "../include/foremast_gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

/* defaults of deploy/foremast/3_brain/foremast-brain.yaml:24-79 + DECISIONS.md */
void fmo_params_default(fm_params *p);

/* Score every unit of a host CSR.  band_upper / band_lower (optional, may be NULL) are indexed
 * like csr->values and written at current-segment positions for params->band_algorithm.
 * threads <= 0 -> all cores (OpenMP).  Returns 0 or FM_E_INVALID. */
int fmo_score(const fm_csr *csr, const fm_params *p, fm_series_score *scores,
              float *band_upper, float *band_lower, int threads);

/* Fold unit scores into per-service verdicts; out has n_services entries. */
int fmo_verdicts(const fm_csr *csr, const fm_series_score *scores, uint32_t n_services,
                 uint64_t first_unit, fm_verdict *out);

/* Synthetic workload, bit-identical to the device generator. */
uint64_t fmo_synth_stride(const fm_synth_config *cfg);
int fmo_synth(const fm_synth_config *cfg, float *values /* n_units*stride + 32 */,
              int64_t *offsets /* n_units+1 */, uint32_t *seg_lens, uint32_t *service_id,
              uint8_t *metric_type, int threads);

int fmo_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
