"""Multi-GPU plumbing: one process per GPU, series sharded by contiguous unit ranges snapped to
service boundaries, and the single exchange step of the path -- the merge of the dense
per-service counters.  On GPUs the merge lives behind the C ABI (fm_comm_init /
fm_result_merge: NCCL on the library's own stream); this module only carries the 128-byte NCCL id to
the ranks (`init_library_comm`) and keeps the host-side twin of the merge -- two torch.distributed
all-reduces -- that the CPU tests run over gloo.  The reference scales the same way, with
shared-nothing brain pods (docs/guides/design.md:37-41); it has no collective at all.
"""
import numpy as np


def shard_range(n_units, rank, world, units_per_service=1):
    """Contiguous [first, last) unit range of `rank`; boundaries fall on service boundaries."""
    n_services = (n_units + units_per_service - 1) // units_per_service
    per = (n_services + world - 1) // world
    s0 = min(rank * per, n_services)
    s1 = min(s0 + per, n_services)
    return min(s0 * units_per_service, n_units), min(s1 * units_per_service, n_units)


class _DevArray:
    """Minimal __cuda_array_interface__ wrapper so torch can alias a library-owned device buffer."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 2}


def alias_service_buffers(result):
    """torch tensors aliasing the result's dense per-service counters (no copy)."""
    import torch
    counts_ptr, worst_ptr, ns = result.service_buffers()
    counts = torch.as_tensor(_DevArray(counts_ptr, (4 * ns,), "<i4"), device="cuda")
    worst = torch.as_tensor(_DevArray(worst_ptr, (ns,), "<i8"), device="cuda")
    return counts, worst


def init_library_comm(ctx, rank, world, group=None):
    """Creates the library's NCCL communicator on `ctx` (one context per rank): rank 0 makes the id,
    torch.distributed (any backend) carries its 128 bytes to the other ranks."""
    import torch.distributed as dist
    from .api import Context
    if world == 1:
        return
    box = [Context.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    ctx.comm_init(box[0], rank, world)


def merge_service_counters(counts, worst, group=None):
    """In-place merge across ranks: counters add, worst-unit keys max.  Keys are
    (points << 32 | ~unit) < 2^63, so a signed 64-bit max is exact.  Host-side twin of
    fm_result_merge.  With tensors that alias a result's device buffers the caller owns the ordering:
    result.sync() before this call (the library writes them on its private stream) and a device
    synchronise after it, before result.verdicts()."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return counts, worst
    dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(worst, op=dist.ReduceOp.MAX, group=group)
    return counts, worst


def counters_from_verdicts(v):
    """Dense counters equivalent to a VERDICT array (used to feed the merge in CPU tests)."""
    counts = np.stack([v["n_units"], v["n_unknown"], v["n_anom_units"], v["n_anom_points"]], axis=1).astype(np.int32).ravel()
    worst = np.where(v["n_anom_units"] > 0,
                     (v["worst_unit_points"].astype(np.int64) << 32) | (0xFFFFFFFF - v["worst_unit"].astype(np.int64)), 0)
    return counts, worst.astype(np.int64)
