// +build gpu

// Package gpu is the cgo binding of libforemast_gpu (include/foremast_gpu.h) for foremast-service.
//
// NOT COMPILED in the repository that ships it: the build image has no Go toolchain (`go version`
// fails there), so this file is the binding a maintainer drops into foremast-service/pkg/gpu/ and
// builds with `CGO_ENABLED=1 go build -tags gpu ./cmd/manager`.  The same C ABI is exercised end to
// end from Python (foremast_b200/, ctypes) by the repository's tests and bench.
//
// cgo rules honoured here: Go memory is only lent to C for the duration of one call (the library
// copies into its own staging); handles are opaque C pointers; no callbacks; every library entry
// point sets its CUDA device itself, so goroutines may migrate between OS threads.
package gpu

/*
#cgo CFLAGS: -I${SRCDIR}/../../../../include
#cgo LDFLAGS: -lforemast_gpu
#include <stdlib.h>
#include "foremast_gpu.h"
*/
import "C"

import (
	"errors"
	"unsafe"
)

// Ctx is one GPU (fm_ctx).  Safe for concurrent use: the library serialises GPU submission.
type Ctx struct{ h *C.fm_ctx }

// Params mirrors the brain's environment block (deploy/foremast/3_brain/foremast-brain.yaml:21-81).
type Params = C.fm_params

// Init takes the place of the brain process start-up.  device = -1: the current CUDA device.
func Init(device int) (*Ctx, error) {
	cfg := C.fm_config{device: C.int32_t(device)}
	var h *C.fm_ctx
	if rc := C.fm_init(&cfg, &h); rc != 0 {
		return nil, errors.New(C.GoString(C.fm_strerror(rc)))
	}
	return &Ctx{h}, nil
}

func (c *Ctx) Close() { C.fm_shutdown(c.h); c.h = nil }

// DefaultParams returns the documented defaults (foremast-brain.yaml:24-79).
func DefaultParams() Params {
	var p Params
	C.fm_params_default(&p)
	return p
}

func (c *Ctx) err(rc C.int) error {
	if rc == 0 {
		return nil
	}
	buf := make([]byte, 512)
	C.fm_last_error_copy(c.h, (*C.char)(unsafe.Pointer(&buf[0])), C.uint32_t(len(buf)))
	n := 0
	for n < len(buf) && buf[n] != 0 {
		n++
	}
	if n == 0 {
		return errors.New(C.GoString(C.fm_strerror(rc)))
	}
	return errors.New(string(buf[:n]))
}

// Job is one alias of one job document: the three query_range response bodies (any may be nil) that
// historicalConfig / baselineConfig / currentConfig name for it (cmd/manager/main.go:50-81).
type Job struct {
	ServiceID  uint32  // index of the job document in this sweep: the verdict key
	MetricType uint8   // FM_MT_*: error5xx, error4xx, latency, cpu, memory (foremast-brain.yaml:34-73)
	Step       float64 // `step` of the query URL in seconds (prometheushelper.go:13-43)
	Hist, Base, Cur []byte
}

// Verdict is fm_verdict: what the worker turns into `status` (converter.go:16-21) and `reason`.
type Verdict struct {
	ServiceID, Verdict, NUnits, NUnknown, NAnomUnits, NAnomPoints, WorstUnit, WorstUnitPoints uint32
}

const (
	Health   = C.FM_V_HEALTH   // completed_health
	Unknown  = C.FM_V_UNKNOWN  // completed_unknown
	Unhealth = C.FM_V_UNHEALTH // completed_unhealth
)

func cbytes(b []byte) (*C.char, C.uint64_t) {
	if len(b) == 0 {
		return nil, 0
	}
	return (*C.char)(unsafe.Pointer(&b[0])), C.uint64_t(len(b))
}

// ScoreJobs runs ONE scoring pass over all jobs of a sweep: response bodies go straight into the batch
// (fm_batch_add_prometheus_many parses them on a thread pool and re-grids missing scrapes), one
// fm_score, one verdict per ServiceID.  unitsPerJob[i] < 0 means job i's bodies did not parse.
func (c *Ctx) ScoreJobs(jobs []Job, p *Params, nServices int) (verdicts []Verdict, unitsPerJob []int, err error) {
	var b *C.fm_batch
	if err = c.err(C.fm_batch_begin(c.h, C.uint64_t(len(jobs)), 0, &b)); err != nil {
		return
	}
	defer C.fm_batch_free(b)
	// the job array lives in C memory: it holds pointers into Go byte slices, which cgo only allows for the
	// duration of the call when the array itself is not Go memory
	cj := (*[1 << 28]C.fm_prom_job)(C.calloc(C.size_t(len(jobs)+1), C.size_t(unsafe.Sizeof(C.fm_prom_job{}))))
	defer C.free(unsafe.Pointer(cj))
	pins := make([][]byte, 0, 3*len(jobs)) // keeps the bodies reachable until the call returns
	for i := range jobs {
		j := &jobs[i]
		cj[i].meta.service_id = C.uint32_t(j.ServiceID)
		cj[i].meta.metric_type = C.uint8_t(j.MetricType)
		cj[i].step = C.double(j.Step)
		cj[i].hist_json, cj[i].hist_len = cbytes(j.Hist)
		cj[i].base_json, cj[i].base_len = cbytes(j.Base)
		cj[i].cur_json, cj[i].cur_len = cbytes(j.Cur)
		pins = append(pins, j.Hist, j.Base, j.Cur)
	}
	per := make([]C.uint32_t, len(jobs)+1)
	rc := C.fm_batch_add_prometheus_many(b, &cj[0], C.uint32_t(len(jobs)), C.FM_PROM_GRID|C.FM_PROM_FILL, 0, &per[0])
	_ = pins
	if err = c.err(rc); err != nil {
		return
	}
	unitsPerJob = make([]int, len(jobs))
	for i := range jobs {
		if per[i] == 0xFFFFFFFF {
			unitsPerJob[i] = -1
		} else {
			unitsPerJob[i] = int(per[i])
		}
	}
	if err = c.err(C.fm_batch_seal(b)); err != nil {
		return
	}
	if err = c.err(C.fm_batch_set_shard(b, 0, C.uint32_t(nServices))); err != nil {
		return
	}
	var r *C.fm_result
	if err = c.err(C.fm_score(c.h, b, p, &r)); err != nil {
		return
	}
	defer C.fm_result_free(r)
	verdicts = make([]Verdict, nServices+1)
	var n C.uint64_t
	rc = C.fm_result_verdicts(r, (*C.fm_verdict)(unsafe.Pointer(&verdicts[0])), C.uint64_t(nServices), &n)
	return verdicts[:n], unitsPerJob, c.err(rc)
}

// CommInit joins this context to a multi-GPU sweep (one context per GPU): rank 0 makes the id with
// CommUniqueID and hands its 128 bytes to the others over the host's own channel.
func CommUniqueID() ([]byte, error) {
	id := make([]byte, C.FM_COMM_ID_BYTES)
	if rc := C.fm_comm_unique_id(unsafe.Pointer(&id[0]), C.uint32_t(len(id))); rc != 0 {
		return nil, errors.New(C.GoString(C.fm_strerror(rc)))
	}
	return id, nil
}

func (c *Ctx) CommInit(id []byte, rank, world int) error {
	return c.err(C.fm_comm_init(c.h, unsafe.Pointer(&id[0]), C.int32_t(rank), C.int32_t(world)))
}
