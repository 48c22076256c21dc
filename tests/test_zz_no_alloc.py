"""fm_scan_counts_static, the entry point for callers that may not allocate GPU memory (a worker
on a fixed device-memory budget, or a scoring loop that cannot afford the allocator's stalls).

It must give the same offsets as the allocating path (fm_result_anomalies_build) -- exactly, these
are integers -- and it must do so twice in a row, because its scratch is one fixed array that keeps
the previous call's tile states and has to be cleared by the call itself.  The file name sorts last
on purpose: if this addition breaks, every other test has already run."""
import numpy as np
import pytest

from foremast_b200 import _abi as abi

pytestmark = pytest.mark.gpu


def test_static_scan_matches_allocating_path_twice():
    import torch
    from foremast_b200.api import Context, default_params, scan_counts_static
    from foremast_b200._lib import ForemastGpuError

    torch.cuda.set_device(0)
    with Context(0) as ctx:
        # > 4 tiles of 2048 counts so that the look-back across tiles is exercised
        n = 9000
        b = ctx.batch_synth(n, 120, 0, 60, flags=abi.SYNTH_RAGGED | abi.SYNTH_ANOMALIES, seed=99)
        p = default_params()
        p.algorithms = abi.ALG_MA_ALL | abi.ALG_EWMA
        p.pairwise = abi.PW_NONE
        res = ctx.score(b, p)
        scores = res.series()
        for call, (alg, slot) in enumerate(((abi.ALG_EWMA, 2), (abi.ALG_MA_ALL, 0))):  # two calls in a row
            ref_offsets, _, _ = res.anomalies(alg)                      # existing, allocating entry point
            counts = np.ascontiguousarray(scores["n_anom"][:, slot], dtype=np.uint32)
            d_counts = torch.from_numpy(counts.view(np.int32)).cuda()   # buffers are the caller's
            d_offsets = torch.full((n + 1,), -1, dtype=torch.int64, device="cuda")
            torch.cuda.synchronize()
            scan_counts_static(d_counts.data_ptr(), n, d_offsets.data_ptr(), 0)
            torch.cuda.synchronize()
            got = d_offsets.cpu().numpy().astype(np.uint64)
            assert np.array_equal(got, ref_offsets), f"call {call}: offsets differ from fm_result_anomalies_build"
            assert int(got[-1]) == int(counts.sum()) > 0
        # empty input and the capacity limit
        d_one = torch.full((1,), -1, dtype=torch.int64, device="cuda")
        scan_counts_static(0, 0, d_one.data_ptr(), 0)
        torch.cuda.synchronize()
        assert int(d_one.item()) == 0
        with pytest.raises(ForemastGpuError) as ei:
            scan_counts_static(d_counts.data_ptr(), abi.SCAN_STATIC_MAX_COUNTS + 1, d_offsets.data_ptr(), 0)
        assert ei.value.code == abi.E_TOO_LARGE
        res.free()
        b.free()


def test_static_scan_against_host_cumsum_sizes_and_free_memory():
    """The same entry point against an independent answer (numpy's cumsum on the host), on sizes that reach
    the multi-round look-back (more than 32 tiles of 2048 counts), with a large call followed by a small
    one (the small call must not see the large call's leftover tile states), and with the free device
    memory read before and after: a caller on a fixed budget relies on the call taking none."""
    import torch
    from foremast_b200.api import scan_counts_static

    torch.cuda.set_device(0)
    rng = np.random.default_rng(5)
    sizes = [150_000, 3, 70_001, 2048, 2049, 1, 1_000_000, 5000]  # large -> small transitions on purpose
    bufs = []
    for n in sizes:  # every buffer exists before the first measured call
        counts = rng.integers(0, 7, n, dtype=np.uint32)
        counts[rng.integers(0, n, max(1, n // 50))] = rng.integers(0, 100_000, max(1, n // 50), dtype=np.uint32)
        bufs.append((n, counts, torch.from_numpy(counts.view(np.int32)).cuda(),
                     torch.full((n + 1,), -1, dtype=torch.int64, device="cuda")))
    scan_counts_static(bufs[0][2].data_ptr(), 8, bufs[0][3].data_ptr(), 0)  # the first call on a device resolves the symbol
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for n, counts, d_counts, d_offsets in bufs:
        scan_counts_static(d_counts.data_ptr(), n, d_offsets.data_ptr(), 0)
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, "fm_scan_counts_static changed the amount of free device memory"
    for n, counts, d_counts, d_offsets in bufs:
        want = np.concatenate([[0], np.cumsum(counts.astype(np.uint64))]).astype(np.uint64)
        got = d_offsets.cpu().numpy().astype(np.uint64)
        assert np.array_equal(got, want), f"n={n}: offsets differ from numpy.cumsum"
