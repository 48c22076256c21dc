#!/usr/bin/env python
"""bench.py -- series-windows scored / s on the BASELINE.json headline workload.

    python bench.py --gpus N --steps K --warmup W            # our arm (libforemast_gpu)
    python bench.py --impl reference --steps K --warmup W    # CPU arm (oracle port, all host threads)

Workload (config.workload): per GPU 1,000,000 series-windows x 1440 points (720-point baseline +
720-point current, BASELINE.json configs[2] shape at the metric's 1440-pt window), all scorers
enabled: Mann-Whitney + Wilcoxon + Kruskal (ALL), MA band + EWMA + double-exponential bands,
per-service verdict fold (80 series per service).  Synthetic Philox data.
A step = one scoring pass over the whole batch.  Weak scaling: every rank scores its own 1M units
(global unit ids rank*1M...), the only exchange is the per-service counter merge (NCCL all-reduce).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from foremast_b200 import _abi as abi  # noqa: E402

METRIC = "series-windows scored/sec (1440-pt)"
UNIT = "series-windows/s"
N_BASE, N_CUR, N_HIST = 720, 720, 0
UNITS_PER_SERVICE = 80
ALGS = abi.ALG_MA_ALL | abi.ALG_MA_ROLLING | abi.ALG_EWMA | abi.ALG_DES
ALGS_R1 = abi.ALG_MA_ALL | abi.ALG_EWMA | abi.ALG_DES  # round 1's three-band step, kept under "kernels" for continuity


def workload_name(n_units, reference_sample=None):
    s = (f"{n_units} series-windows/GPU x 1440 pt (720 baseline + 720 current, concurrent windows), all scorers: "
         "MWU+Wilcoxon+Kruskal (ALL), MA_ALL+MA_ROLLING(W=30)+EWMA+DES bands, verdict fold (80 series/service); "
         "rules: documented table for 5xx/4xx/latency/cpu/memory, (5, upper) for tps/gc/threads")
    if reference_sample:
        s += f"; reference arm: {reference_sample}-unit sample of it per step"
    return s


def bench_params(defaults, algorithms=ALGS):
    """The deployment's rule table: the five documented rows (foremast-brain.yaml:34-73) plus one row for
    each of the three further metric types the 8-metric synthetic services carry."""
    p = defaults
    p.algorithms = algorithms
    p.pairwise = abi.PW_ALL
    for mt in (abi.MT_TPS, abi.MT_GC, abi.MT_THREADS):
        p.rules[mt].threshold, p.rules[mt].min_lower_bound, p.rules[mt].bound = 5.0, 0.0, abi.BOUND_UPPER
    return p


def algorithmic_bytes(n_units, n_services):
    """SURVEY 8(d) per-unit figures x units (DESIGN.md 'Algorithmic bytes')."""
    n = N_HIST + N_BASE + N_CUR
    return {
        "rank": n_units * (4 * (N_BASE + N_CUR) + 8 + 12 + 64),
        "rank_wilcoxon": n_units * (4 * (N_BASE + N_CUR) + 8 + 12 + 28),
        "rank_mw": n_units * (4 * (N_BASE + N_CUR) + 8 + 12 + 24 + 36),
        "band": n_units * (4 * n + 8 + 12 + 16),
        "verdict": n_units * (8 + 4) + n_services * 24,
    }


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                pass
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 8 for i in range(4) if r[4 + i].lower() == "active"})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def host_threads():
    """All host cores this process may use (torchrun exports OMP_NUM_THREADS=1; the oracle takes
    the thread count explicitly)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_port_rate(oracle, csr, params, threads):
    t0 = time.perf_counter()
    oracle.score(csr, params, threads=threads)
    sc_t = time.perf_counter() - t0
    return csr.n_units / sc_t, sc_t


INT_FIELDS = ["status", "n_anom", "first_idx", "w_n", "mw_2u1", "mw_2r1", "mw_ties", "w_2wplus", "w_2wminus", "w_ties"]
FLT_FIELDS = ["p_mw", "p_w", "p_k", "kw_h", "k_eff", "ma_mean", "ma_std", "sigma"]
REL_TOL = 1e-5  # north_star: smoothed values, bands and p-values within 1e-5 relative


def compare_scores(got, ref):
    """(#integer-field mismatches, worst relative error over the float fields) of two SERIES_SCORE arrays."""
    int_mis = sum(int(np.count_nonzero(got[f] != ref[f])) for f in INT_FIELDS)
    worst = 0.0
    for f in FLT_FIELDS:
        a, b = got[f].astype(np.float64), ref[f].astype(np.float64)
        err = np.abs(a - b) / np.maximum(np.abs(b), 1e-300)
        err[a == b] = 0.0
        if err.size:
            worst = max(worst, float(np.nanmax(err)))
    return int_mis, worst


def run_reference(args):
    """Reference arm: the CPU restatement of the brain's pass (oracle 'port'; the reference's own
    scorer source is not in the mount) on all host threads, same workload shape, bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle
    threads = host_threads()
    sample = args.ref_units
    csr = oracle.synth(sample, N_HIST, N_BASE, N_CUR, units_per_service=UNITS_PER_SERVICE, n_metric_types=8,
                       flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES)
    p = bench_params(oracle.params_default())
    for _ in range(args.warmup):
        oracle.score(csr, p, threads=threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sc, _, _ = oracle.score(csr, p, threads=threads)
        oracle.verdicts(csr, sc, csr.n_services)
    dt = (time.perf_counter() - t0) / args.steps
    value = sample / dt
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(1_000_000, reference_sample=sample), "sample_units_per_step": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"{sample} units of the same workload per step, OpenMP over units"},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    OUT.emit(json.dumps(line))


def run_ours(args):
    import torch
    import torch.distributed as dist
    from foremast_b200 import dist as fdist
    from foremast_b200.api import Context, default_params

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    n = args.units
    ctx = Context(local)
    fdist.init_library_comm(ctx, rank, world)  # the cross-GPU merge runs inside the library (fm_result_merge)
    total_services = (n * world + UNITS_PER_SERVICE - 1) // UNITS_PER_SERVICE
    batch = ctx.batch_synth(n, N_HIST, N_BASE, N_CUR, units_per_service=UNITS_PER_SERVICE, n_metric_types=8,
                            flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, first_unit=rank * n)
    batch.set_shard(rank * n, total_services)
    p = bench_params(default_params())

    stage_ms = {"setup": [], "rank": [], "rank_wilcoxon": [], "rank_mw": [], "band": [], "verdict": []}
    # the per-service verdicts land in pinned host memory (32 B x services: 3.2 MB per rank at 8 GPUs)
    vbuf, vbuf_ptr = ctx.host_alloc(total_services * abi.VERDICT.itemsize)
    vout = vbuf.view(abi.VERDICT)

    own_lo, own_hi = -(-(rank * n) // UNITS_PER_SERVICE), -(-((rank + 1) * n) // UNITS_PER_SERVICE)  # services that START in this shard

    def step(record, keep_series=False):
        res = ctx.score(batch, p)
        res.merge()  # NCCL all-reduce pair on the library's stream (no-op at N = 1), no host sync
        if keep_series or world == 1:
            v = res.verdicts(out=vout)  # finalize kernel + D2H of all per-service verdicts
        else:  # a rank of a sharded sweep writes back the services of its own shard: together the ranks download
            # every verdict exactly once (the untimed parity pass below downloads and checks all of them on every rank)
            v = res.verdicts_range(own_lo, own_hi - own_lo, out=vout)
        if record:
            for k, t in res.timings_ms().items():
                stage_ms[k].append(t)
        series = res.series() if keep_series else None
        res.free()
        return (v, series) if keep_series else v

    for _ in range(args.warmup):
        step(False)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    ctx.timer_start()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        verdicts = step(True)
    dev_ms = ctx.timer_stop()
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    # one more pass, untimed, whose per-unit scores come back for the full-size parity check below
    verdicts, gpu_series = step(False, keep_series=True)
    verdicts = verdicts.copy()  # out of the pinned buffer the later legs reuse
    differs_units = int(((gpu_series["status"] & abi.S_DIFFERS) != 0).sum())
    anom_units = int(((gpu_series["status"] & abi.S_ANOMALY) != 0).sum())

    # ---- per-scorer passes (north_star: "each scoring kernel" against the HBM roofline): the band
    #      kernel with ONE algorithm enabled and no rank tests, same batch, timed by its own events
    single_ms = {}
    for name, alg in (("band_ma_all", abi.ALG_MA_ALL), ("band_ma_rolling", abi.ALG_MA_ROLLING), ("band_ewma", abi.ALG_EWMA),
                      ("band_des", abi.ALG_DES), ("band_r1_three", ALGS_R1)):
        q = bench_params(default_params())
        q.algorithms, q.pairwise = alg, abi.PW_NONE
        ts = []
        for i in range(3 + 5):
            res = ctx.score(batch, q)
            t = res.timings_ms()["band"]
            res.free()
            if i >= 3:
                ts.append(t)
        single_ms[name] = sum(ts) / len(ts)

    # ---- the other BASELINE.json configs, device-resident, rank 0 only at N = 1 (parity for these
    #      shapes is in tests/; here only their kernel times and roofline fractions)
    other = {}
    if world == 1 and not args.skip_configs:
        def timed(make_batch, q, key, passes=4):
            bb = make_batch()
            ts = []
            for i in range(2 + passes):
                res = ctx.score(bb, q)
                t = res.timings_ms()
                res.free()
                if i >= 2:
                    ts.append(t[key] if key else t["setup"] + t["rank"] + t["band"] + t["verdict"])
            info = bb.info()
            bb.free()
            return sum(ts) / len(ts), info
        peak_gbs, _ = measured_peak()
        # cfg2: 10k pods x 8 metrics x 1440-pt EWMA + double-exponential smoothing bands
        q = bench_params(default_params()); q.algorithms = abi.ALG_EWMA | abi.ALG_DES; q.pairwise = abi.PW_NONE
        ms, info = timed(lambda: ctx.batch_synth(80_000, 1380, 0, 60, units_per_service=8, n_metric_types=8,
                                                 flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES), q, "band", 8)
        nbytes = 80_000 * (4 * 1440 + 36)
        other["cfg2_80k_x_1440_ewma_des"] = {"ms": ms, "series_windows_per_s": 80_000 / (ms * 1e-3), "achieved_gbs": nbytes / (ms * 1e-3) / 1e9,
                                              "frac": nbytes / (ms * 1e-3) / 1e9 / peak_gbs, "shape": "1380 model + 60 current points"}
        # cfg3: 1M series x 720-pt baseline-vs-canary rank tests = the rank stage of the headline step
        # cfg4: 1M series Holt-Winters seasonal (period 1440, 7-day history)
        q = bench_params(default_params()); q.algorithms = abi.ALG_HW; q.pairwise = abi.PW_NONE; q.period = 1440
        n4 = args.hw_units
        ms, info = timed(lambda: ctx.batch_synth(n4, 10_080, 0, 60, units_per_service=UNITS_PER_SERVICE, n_metric_types=8,
                                                 flags=abi.SYNTH_ANOMALIES), q, "band", 3)
        nbytes = n4 * (4 * 10_140 + 52)
        other["cfg4_holt_winters_10080pt_period1440"] = {"units": n4, "ms": ms, "series_per_s": n4 / (ms * 1e-3),
                                                         "achieved_gbs": nbytes / (ms * 1e-3) / 1e9, "frac": nbytes / (ms * 1e-3) / 1e9 / peak_gbs}
        # cfg4 in continuous mode (the brain re-arms a job every cycle): the Holt-Winters state of every series is fitted
        # once and kept in HBM (fm_model_fit), each cycle scores only the new window from it (fm_model_score) and
        # advances it -- the traffic of a cycle is the state, read and written (period + 4 doubles per series)
        nm = min(n4, 500_000)
        hb = ctx.batch_synth(nm, 10_080, 0, 0, units_per_service=UNITS_PER_SERVICE, n_metric_types=8, flags=abi.SYNTH_ANOMALIES)
        cb = ctx.batch_synth(nm, 0, 0, 60, units_per_service=UNITS_PER_SERVICE, n_metric_types=8, flags=abi.SYNTH_ANOMALIES)
        model = ctx.model_fit(hb, q)
        ts = []
        for i in range(5):
            res = model.score(cb)
            res.sync()
            t = res.timings_ms()["band"]
            res.free()
            if i >= 2:
                ts.append(t)
        cyc = sum(ts) / len(ts)
        cbytes = nm * (2 * (1440 + 4) * 8 + 60 * 4 + 52)
        other["cfg4_continuous_mode_model_cache"] = {"units": nm, "new_points_per_cycle": 60, "cycle_ms": cyc, "series_per_s": nm / (cyc * 1e-3),
                                                     "state_gb": nm * (1440 + 4) * 8 / 1e9, "achieved_gbs": cbytes / (cyc * 1e-3) / 1e9,
                                                     "frac": cbytes / (cyc * 1e-3) / 1e9 / peak_gbs,
                                                     "vs_rescoring_the_history": (ms * nm / n4) / cyc}
        for o in (model, hb, cb):
            o.free()
        # cfg0: the reference's own CPU-runnable case -- one canary job, 5 series, 60 vs 30 points, MA band + MWU:
        # a latency figure (launch-bound), through the one-shot host call
        rng = np.random.default_rng(0)
        units0 = [([], rng.normal(100, 5, 60).astype(np.float32), rng.normal(100, 5, 30).astype(np.float32)) for _ in range(5)]
        h0 = abi.HostCsr.from_units(units0, service_id=[0] * 5, metric_type=[1, 3, 6, 4, 5])
        q = bench_params(default_params()); q.algorithms = abi.ALG_MA_ALL; q.pairwise = abi.PW_MANN_WHITE
        for _ in range(20):
            ctx.score_host(h0, q, 1)
        t0 = time.perf_counter()
        for _ in range(200):
            ctx.score_host(h0, q, 1)
        other["cfg0_single_canary_job_5_series"] = {"us_per_call": (time.perf_counter() - t0) / 200 * 1e6,
                                                    "note": "fm_score_host small-job path: one upload, four launches, one download (the clear rides behind the previous call); no allocation"}
        # cfg0's job shape at scale: what batching jobs buys (BrainWorker.process_many puts every open job document
        # of a cycle into one batch).  200 000 canary jobs x 5 series x (60 + 30) points, MA band + Mann-Whitney U.
        ms, info = timed(lambda: ctx.batch_synth(1_000_000, 0, 60, 30, units_per_service=5, n_metric_types=5,
                                                 flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES), q, None, 4)
        other["cfg0_shape_batched_200k_jobs"] = {"ms": ms, "jobs_per_s": 200_000 / (ms * 1e-3), "series_windows_per_s": 1_000_000 / (ms * 1e-3),
                                                 "note": "one fm_score over 1 M units of 90 points; a job alone takes us_per_call above"}
        # ingest: Prometheus query_range response bodies -> sealed batch -> verdicts, cfg2's shape
        # (80 k series x 1440 points: 10 000 aliases with one 1380-point history series and eight 60-point current
        # series each), through fm_batch_add_prometheus_many on all host threads.  The same two bodies are lent
        # to every job (the scanner reads them 10 000 times; generating 450 MB of distinct text would take longer
        # than everything else in this bench).
        def prom_body(n_series, n_points, seed):
            r = np.random.default_rng(seed)
            return json.dumps({"status": "success", "data": {"resultType": "matrix", "result": [
                {"metric": {"__name__": "namespace_app_pod_cpu_usage_seconds_total", "pod": f"pod-{i}", "namespace": "dev"},
                 "values": [[1550000000 + 60 * k, f"{v:.4f}"] for k, v in enumerate(r.normal(0.45, 0.08, n_points))]}
                for i in range(n_series)]}}).encode()
        hb, cb = prom_body(1, 1380, 1), prom_body(8, 60, 2)
        n_jobs = 10_000
        jobs = ctx.prom_jobs([(j // 10, abi.MT_CPU, 60.0, hb, None, cb) for j in range(n_jobs)])  # the C array, made once
        json_mb = n_jobs * (len(hb) + len(cb)) / 1e6
        q = bench_params(default_params()); q.algorithms = abi.ALG_EWMA | abi.ALG_DES; q.pairwise = abi.PW_NONE
        ts_pack, ts_total = [], []
        for i in range(3):
            t0 = time.perf_counter()
            bb, per = ctx.batch_from_prometheus(jobs, threads=0)
            t1 = time.perf_counter()
            res = ctx.score(bb, q)
            res.verdicts()
            t2 = time.perf_counter()
            res.free()
            bb.free()
            if i:
                ts_pack.append(t1 - t0)
                ts_total.append(t2 - t0)
        t0 = time.perf_counter()
        bb, _ = ctx.batch_from_prometheus((jobs[0], n_jobs // 16, jobs[2]), threads=1)
        one_thread_s = time.perf_counter() - t0
        bb.free()
        other["ingest_prometheus_json_cfg2_shape"] = {
            "json_mb": json_mb, "units": int(per.sum()), "host_threads": host_threads(),
            "parse_pack_upload_ms": min(ts_pack) * 1e3, "json_mb_per_s_all_threads": json_mb / min(ts_pack),
            "json_mb_per_s_one_thread": json_mb / 16 / one_thread_s,
            "json_in_to_verdicts_out_ms": min(ts_total) * 1e3,
            "note": "fm_batch_add_prometheus_many (timestamps read, missing scrapes re-gridded) + fm_batch_seal + EWMA/DES scoring + verdicts"}

    # ---- end to end through the public one-shot call: pinned host CSR -> H2D -> score -> verdicts D2H
    host = batch.download()
    nbytes = host.values.nbytes
    pinned, pin_ptr = ctx.host_alloc(nbytes)
    pv = pinned.view(np.float32)
    pv[:] = host.values
    host_p = abi.HostCsr.__new__(abi.HostCsr)
    pins = [(pinned, pin_ptr)]

    def pin(a):  # the per-unit arrays live in pinned memory too, like the samples
        buf, ptr = ctx.host_alloc(a.nbytes)
        pins.append((buf, ptr))
        out = buf.view(a.dtype).reshape(a.shape)
        out[...] = a
        return out
    host_p.values, host_p.offsets, host_p.seg_lens = pv, pin(host.offsets), pin(host.seg_lens)
    host_p.service_id, host_p.metric_type = pin(host.service_id), pin(host.metric_type)
    h2d = pv.nbytes + host.offsets.nbytes + host.seg_lens.nbytes + host.service_id.nbytes + host.metric_type.nbytes
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    ctx.score_host(host_p, p, total_services, out=vout)  # warm-up
    barrier()
    t1 = time.perf_counter()
    for _ in range(e2e_steps):
        v2 = ctx.score_host(host_p, p, total_services, out=vout)
    barrier()
    e2e_s = (time.perf_counter() - t1) / e2e_steps
    # the ceiling of that number: the same pinned bytes, bare cudaMemcpyAsync host -> device, all ranks at once
    copy_s = None
    try:
        import ctypes as C
        rt = C.CDLL("libcudart.so.12")
        rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        dst = torch.empty(pv.nbytes, dtype=torch.uint8, device="cuda")
        for i in range(4):
            if i == 1:
                barrier()
                t2 = time.perf_counter()
            assert rt.cudaMemcpyAsync(dst.data_ptr(), pin_ptr, pv.nbytes, 1, None) == 0
            torch.cuda.synchronize()
        barrier()
        copy_s = (time.perf_counter() - t2) / 3
        del dst
    except Exception as e:  # the ceiling is an annotation, not part of the result
        sys.stderr.write(f"bare-copy ceiling not measured: {e}\n")
    clocks = sampler.stop() if rank == 0 else None
    # the services that lie wholly inside this rank's shard: there the merged verdicts, the one-shot call's
    # (un-merged) verdicts and the oracle fold of the shard must all be the same bytes
    s_lo, s_hi = -(-(rank * n) // UNITS_PER_SERVICE), ((rank + 1) * n) // UNITS_PER_SERVICE
    v2 = v2.copy()
    hit = v2["worst_unit"] != abi.NONE_IDX
    v2["worst_unit"][hit] += rank * n  # the one-shot call numbers units from 0, the sharded batch from rank * n
    e2e_equal = v2[s_lo:s_hi].tobytes() == verdicts[s_lo:s_hi].tobytes()

    # ---- full-size parity (north_star Target): EVERY unit of this rank's shard against the CPU oracle --
    #      integers and verdict records bit for bit, floats within 1e-5 -- on every rank.  At N = 1 the same
    #      oracle pass is the cpu_baseline timing.
    from oracle import oracle
    threads = max(1, host_threads() // world)
    n_par = min(n, args.parity_units)
    stride = int(host.offsets[1])
    sub = host if n_par == n else abi.HostCsr(host.values[:n_par * stride + 32], host.offsets[:n_par + 1], host.seg_lens[:n_par],
                                              host.service_id[:n_par], host.metric_type[:n_par])
    t_cpu = time.perf_counter()
    ref, _, _ = oracle.score(sub, p, threads=threads)
    cpu_s = time.perf_counter() - t_cpu
    cpu_rate = n_par / cpu_s
    int_mis, worst_rel = compare_scores(gpu_series[:n_par], ref)
    if n_par == n:
        v_ref = oracle.verdicts(host, ref, total_services, first_unit=rank * n)
        verdict_equal = v_ref[s_lo:s_hi].tobytes() == verdicts[s_lo:s_hi].tobytes()
        services_checked = s_hi - s_lo
    else:
        verdict_equal, services_checked = True, 0
    del ref, gpu_series

    t = torch.tensor([dev_ms, wall_ms, e2e_s * 1e3, worst_rel, float(not verdict_equal), float(not e2e_equal),
                      (copy_s or 0.0) * 1e3], dtype=torch.float64, device="cuda")
    cnt = torch.tensor([int_mis, n_par, services_checked], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    dev_ms, wall_ms, e2e_ms, worst_rel, verdict_bad, e2e_bad, copy_ms = (float(x) for x in t.tolist())
    int_mis, units_checked, services_checked = (int(x) for x in cnt.tolist())
    parity = {"units_checked": units_checked, "int_mismatches": int_mis, "max_float_rel_err": worst_rel,
              "float_tolerance": REL_TOL, "verdict_bytes_equal": verdict_bad == 0.0, "services_checked": services_checked,
              "e2e_verdicts_equal_device_run": e2e_bad == 0.0,
              "how": "every rank: oracle port over its whole shard vs fm_result_series / merged fm_result_verdicts"}
    parity_ok = int_mis == 0 and worst_rel <= REL_TOL and verdict_bad == 0.0 and e2e_bad == 0.0

    line = None
    if rank == 0:
        ms_per_step = dev_ms / args.steps
        value = n * world / (ms_per_step * 1e-3)
        peak, peak_src = measured_peak()
        ab = algorithmic_bytes(n, total_services)
        avg = {k: (sum(v) / len(v) if v else 0.0) for k, v in stage_ms.items()}
        kernels = {}
        for k in ("rank", "rank_wilcoxon", "rank_mw", "band", "verdict"):
            gbs = ab[k] / (avg[k] * 1e-3) / 1e9 if avg[k] > 0 else 0.0
            kernels[k] = {"ms": avg[k], "algorithmic_bytes": ab[k], "achieved_gbs": gbs, "frac": gbs / peak}
        for k, ms in single_ms.items():
            gbs = ab["band"] / (ms * 1e-3) / 1e9
            kernels[k] = {"ms": ms, "algorithmic_bytes": ab["band"], "achieved_gbs": gbs, "frac": gbs / peak,
                          "note": "k_band with this one algorithm, rank tests off"}
        dom = max(("rank_wilcoxon", "rank_mw", "band", "verdict"), key=lambda k: avg[k])
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            try:
                per_unit = json.load(open(tpath))["bytes_per_unit"].get(dom)
                traffic = per_unit * n if per_unit else None  # DRAM bytes per launch, from the committed ncu capture
            except Exception:
                traffic = None
        # CPU baseline = the parity pass above (the oracle port on this box's host cores over rank 0's shard);
        # a single-GPU-run figure (rank 0, N = 1): at N > 1 the ranks share the host cores
        sample = n_par
        if world > 1:
            cpu_rate = None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(n), "units_per_gpu": n, "n_services": total_services,
                       "l2": "inputs (5.76 GB/GPU) larger than L2; no flush needed",
                       "timing": "CUDA events on the library stream around the K steps, max over ranks",
                       "verdict_download": ("all services" if world == 1 else
                                            "each rank the merged verdicts of its own shard's services (every verdict reaches a host once)"),
                       "wall_ms_per_step": wall_ms / args.steps},
            "roofline": {"bound": "hbm", "kernel": {"rank_wilcoxon": "k_rank<R,0> (Wilcoxon pass)", "rank_mw": "k_mw<R,W> (Mann-Whitney/Kruskal pass) + k_rank_finalize",
                                               "band": "k_band", "verdict": "k_verdict_reduce"}[dom],
                         "achieved": kernels[dom]["achieved_gbs"], "peak": peak, "unit": "GB/s",
                         "frac": kernels[dom]["frac"], "traffic": traffic, "peak_source": peak_src,
                         "frac_of_spec_8000": kernels[dom]["achieved_gbs"] / 8000.0,
                         "note": ("the rank kernels are bound by instruction issue (a register sorting network of 1504 integer "
                                  "min/max per lane per sort, then shared-memory searches; profiles/r2_rank_mw.md), not by HBM; "
                                  "the HBM-bound kernels of this path are under \"kernels\" (band_*)") if dom.startswith("rank") else None,
                         # the bound that does apply to the rank kernels: warp instructions executed per unit (ncu, constant for
                         # this shape: profiles/r2_rank_mw.md, profiles/r2_step_1M.md) against the issue slots of the launch
                         "issue": ({"warp_instructions_per_unit": {"rank_mw": 5670, "rank_wilcoxon": 3710}[dom],
                                    "frac": {"rank_mw": 5670, "rank_wilcoxon": 3710}[dom] * n /
                                            (148 * 4 * (clocks.get("sm_mhz") or 1965) * 1e6 * kernels[dom]["ms"] * 1e-3),
                                    "note": "share of the issue slots (148 SMs x 4 schedulers x SM clock) the kernel's instructions fill"}
                                   if dom in ("rank_mw", "rank_wilcoxon") else None)},
            "kernels": kernels,
            "other_configs": other,
            "parity": parity,
            "cpu_baseline": ({"value": cpu_rate, "unit": UNIT, "cores": threads, "kind": "port",
                              "sample": f"first {sample} units of rank 0's batch (the parity pass), one pass, {cpu_s:.2f} s"}
                             if cpu_rate is not None else None),
            "e2e": {"value": n * world / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": int(total_services * 32), "ms_per_step": e2e_ms, "steps": e2e_steps,
                    # bare concurrent pinned host -> device copy of the same sample bytes on all N ranks: what the link
                    # and the host memory system allow; the call's own time as a fraction of it
                    "bare_h2d_copy_ms": copy_ms or None,
                    "frac_of_bare_copy": (copy_ms / e2e_ms) if copy_ms else None},
            # inside the timed region, per step: k_wilcoxon, k_mw, k_rank_finalize, k_band_order (launches of 131 072
            # units and more), k_band, k_verdict_reduce, k_verdict_finalize (+ NCCL's own all-reduce kernel at N > 1,
            # not counted: it is not ours)
            "gpu_launches": (7 if n >= 131072 and not os.environ.get("FOREMAST_GPU_NO_ORDER") else 6) * args.steps,
            # the untimed legs: the parity pass, the per-scorer passes, the e2e calls (the five scoring kernels once per
            # ~256 MiB upload chunk + one finalize), cfg2 / cfg4 / cfg0 / ingest, their generators
            "gpu_launches_other_legs": 6 + 5 * 8 * 2 + (e2e_steps + 1) * (5 * (-(-pv.nbytes // (256 << 20))) + 1) + 1
            + (10 * 3 + 5 * 4 + 2 + 220 * 5 + 4 * 5 if other else 0),
            "clocks": clocks,
            "verdict_summary": {"services": int(verdicts.size), "unhealth": int((verdicts["verdict"] == abi.V_UNHEALTH).sum()),
                                "unknown": int((verdicts["verdict"] == abi.V_UNKNOWN).sum()),
                                "health": int((verdicts["verdict"] == abi.V_HEALTH).sum()),
                                "units_differs_rank0": differs_units, "units_anomalous_rank0": anom_units},
        }
    del host_p
    for _, ptr in pins:
        ctx.host_free(ptr)
    batch.free()

    # ---- cfg5 (BASELINE.json configs[4]): a 10 M-series sweep, 125 000 services of 80, sharded over the N
    #      GPUs by service (strong scaling: 57.6 GB of samples in total, 1/N of it per GPU), all scorers,
    #      per-service verdict merge through fm_result_merge.  Timed like the headline: CUDA events on the
    #      library stream around K passes, barrier on both sides, max over ranks.
    cfg5 = None
    if args.cfg5_units > 0:
        n5 = args.cfg5_units
        lo, hi = fdist.shard_range(n5, rank, world, UNITS_PER_SERVICE)
        ns5 = (n5 + UNITS_PER_SERVICE - 1) // UNITS_PER_SERVICE
        b5 = ctx.batch_synth(hi - lo, N_HIST, N_BASE, N_CUR, units_per_service=UNITS_PER_SERVICE, n_metric_types=8,
                             flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, first_unit=lo)
        b5.set_shard(lo, ns5)
        v5buf, v5ptr = ctx.host_alloc(ns5 * abi.VERDICT.itemsize)
        v5 = v5buf.view(abi.VERDICT)

        s5_lo, s5_hi = -(-lo // UNITS_PER_SERVICE), -(-hi // UNITS_PER_SERVICE)

        def step5(everything=False):
            r5 = ctx.score(b5, p)
            r5.merge()
            out = r5.verdicts(out=v5) if everything or world == 1 else r5.verdicts_range(s5_lo, s5_hi - s5_lo, out=v5)
            r5.free()
            return out
        for _ in range(2):
            step5()
        barrier()
        ctx.timer_start()
        for _ in range(args.cfg5_steps):
            step5()
        ms5 = ctx.timer_stop() / args.cfg5_steps
        barrier()
        vv = step5(everything=True).copy()
        # the merge alone: K all-reduce pairs on a scratch result (its counters are garbage afterwards)
        r5 = ctx.score(b5, p)
        r5.sync()
        barrier()
        ctx.timer_start()
        for _ in range(20):
            r5.merge()
        merge_us = ctx.timer_stop() / 20 * 1e3
        r5.free()
        t5 = torch.tensor([ms5, merge_us], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t5, op=dist.ReduceOp.MAX)
        ms5, merge_us = (float(x) for x in t5.tolist())
        cfg5 = {"units_total": n5, "units_this_rank": hi - lo, "n_services": ns5, "ms_per_step": ms5,
                "series_windows_per_s": n5 / (ms5 * 1e-3), "verdict_merge_us": merge_us, "steps": args.cfg5_steps,
                "scaling": "strong", "unhealth": int((vv["verdict"] == abi.V_UNHEALTH).sum()),
                "unknown": int((vv["verdict"] == abi.V_UNKNOWN).sum()),
                "note": "sharded by service with fm_batch_set_shard; merge = fm_result_merge (NCCL all-reduce pair, one group); "
                        "every rank downloads the merged verdicts of its own services (fm_result_verdicts_range)"}
        b5.free()
        ctx.host_free(v5ptr)
    ctx.host_free(vbuf_ptr)
    ctx.close()
    if rank == 0 and line is not None:
        if cfg5 is not None:
            line["other_configs"]["cfg5_10M_x_1440_sharded"] = cfg5
        OUT.emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    if not parity_ok:
        sys.stderr.write(f"PARITY FAILED: {parity}\n")
        sys.exit(1)


class QuietStdout:
    """Everything that is not the result line goes to stderr: libraries (NCCL prints its version
    on stdout) must not break the one-JSON-line contract.  emit() writes to the real stdout."""

    def __init__(self):
        sys.stdout.flush()
        self._saved = os.dup(1)
        os.dup2(2, 1)

    def emit(self, line):
        sys.stdout.flush()
        os.write(self._saved, (line + "\n").encode())


OUT = None


def main():
    global OUT
    OUT = QuietStdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--units", type=int, default=1_000_000, help="series-windows per GPU")
    ap.add_argument("--parity-units", type=int, default=1_000_000_000,
                    help="units per rank checked against the oracle (default: the whole shard; it is also the cpu_baseline sample)")
    ap.add_argument("--ref-units", type=int, default=65536, help="--impl reference: units per step")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--hw-units", type=int, default=1_000_000, help="cfg4 Holt-Winters units (10080 points each)")
    ap.add_argument("--skip-configs", action="store_true", help="skip the extra cfg2 / cfg4 timings")
    ap.add_argument("--cfg5-units", type=int, default=10_000_000, help="cfg5: series of the sharded sweep in total (0 = skip)")
    ap.add_argument("--cfg5-steps", type=int, default=3)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
