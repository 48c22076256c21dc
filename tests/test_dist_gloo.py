"""Multi-process (world_size 2, gloo, CPU) test of the N>1 path's host logic: contiguous sharding
snapped to service boundaries and the one exchange step of the path, the merge of the dense
per-service counters (sum) and worst-unit keys (max).  Per-shard scores come from the oracle here
(no GPU in this container); the GPU run uses the same shard_range / merge_service_counters code
with NCCL (bench.py --gpus N)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from foremast_b200 import _abi as abi
from foremast_b200 import dist as fdist

N_UNITS, UPS = 1000, 40


def verdicts_from_counters(counts, worst):
    """Host twin of k_verdict_finalize, for this test only."""
    c = counts.reshape(-1, 4)
    v = np.zeros(c.shape[0], dtype=abi.VERDICT)
    v["service_id"] = np.arange(c.shape[0])
    v["n_units"], v["n_unknown"], v["n_anom_units"], v["n_anom_points"] = c[:, 0], c[:, 1], c[:, 2], c[:, 3]
    anom = c[:, 2] > 0
    v["verdict"] = np.where(anom, abi.V_UNHEALTH, np.where(c[:, 1] == c[:, 0], abi.V_UNKNOWN, abi.V_HEALTH))
    v["worst_unit"] = np.where(anom, 0xFFFFFFFF - (worst & 0xFFFFFFFF), 0xFFFFFFFF)
    v["worst_unit_points"] = np.where(anom, worst >> 32, 0)
    return v


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle
    n_services = (N_UNITS + UPS - 1) // UPS
    lo, hi = fdist.shard_range(N_UNITS, rank, world, UPS)
    csr = oracle.synth(hi - lo, 120, 30, 30, units_per_service=UPS, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES,
                       first_unit=lo, threads=2)
    p = oracle.params_default()
    p.algorithms = abi.ALG_MA_ALL | abi.ALG_EWMA
    sc, _, _ = oracle.score(csr, p, threads=2)
    v = oracle.verdicts(csr, sc, n_services, first_unit=lo)
    counts, worst = fdist.counters_from_verdicts(v)
    tc, tw = torch.from_numpy(counts), torch.from_numpy(worst)
    fdist.merge_service_counters(tc, tw)
    merged = verdicts_from_counters(tc.numpy(), tw.numpy())
    np.save(os.path.join(out_dir, f"v{rank}.npy"), merged)
    dist.destroy_process_group()


def test_two_rank_verdict_merge_matches_single_process(tmp_path):
    from oracle import oracle
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    whole = oracle.synth(N_UNITS, 120, 30, 30, units_per_service=UPS, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES)
    p = oracle.params_default()
    p.algorithms = abi.ALG_MA_ALL | abi.ALG_EWMA
    sc, _, _ = oracle.score(whole, p)
    ref = oracle.verdicts(whole, sc, whole.n_services)
    for r in range(2):
        got = np.load(os.path.join(tmp_path, f"v{r}.npy"))
        assert got.tobytes() == ref.tobytes()
    assert (ref["verdict"] == abi.V_UNHEALTH).any()


def test_shard_range_covers_and_respects_services():
    for n, world, ups in [(1000, 2, 40), (1000, 8, 80), (7, 4, 3), (10_000_000, 8, 80), (5, 8, 1)]:
        edges = [fdist.shard_range(n, r, world, ups) for r in range(world)]
        assert edges[0][0] == 0 and edges[-1][1] == n
        for (a0, a1), (b0, b1) in zip(edges, edges[1:]):
            assert a1 == b0 and a0 <= a1
        for lo, hi in edges:
            assert lo % ups == 0 or lo == n
