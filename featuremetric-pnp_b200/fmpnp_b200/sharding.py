"""Query sharding across the GPUs of one box.

Queries are independent: each rank owns a contiguous block of queries together with their
feature maps, refines them with no per-iteration communication, and the fixed-size result
records are gathered once at the end (NCCL all-gather over NVLink; gloo in the CPU tests).
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_range(n_queries, rank, world_size):
    """Contiguous block [lo, hi) of ceil(B / G) queries owned by `rank` (last ranks may be short)."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank / world size")
    per = -(-n_queries // world_size)
    lo = min(n_queries, rank * per)
    hi = min(n_queries, lo + per)
    return lo, hi


_gather_buffers = {}


def shared_result_records(n_queries, group=None, ctx=None, owner_rank=0):
    """The multi-GPU result gather without a collective: rank `owner_rank` allocates one buffer for the records of
    ALL queries (fmpnp_shared_buffer_create) and broadcasts its 64-byte handle; every other rank maps it.  A rank then
    passes `records.block(lo, hi - lo)` (its shard_range) to refine_batch(out=...): the kernel's own stores of the
    176-byte records travel over NVLink peer access.  The owner reads them (`records.read()`) after every rank has
    synchronised its stream and a barrier.  NCCL / CUDA only (gloo ranks use gather_result_records)."""
    from .binding import SharedRecords
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return SharedRecords(n_queries, ctx=ctx)
    box = [None]
    rec = None
    if rank == owner_rank:
        rec = SharedRecords(n_queries, ctx=ctx)
        box[0] = rec.handle
    dist.broadcast_object_list(box, src=owner_rank, group=group)
    if rank != owner_rank:
        rec = SharedRecords(n_queries, handle=box[0], ctx=ctx)
    return rec


def gather_result_records(local_records, n_queries, group=None, async_op=False):
    """All-gathers per-rank result records.

    local_records: uint8 tensor [n_local, record_bytes] (device tensor under NCCL, CPU under gloo)
    holding this rank's block as laid out by shard_range.  Returns [n_queries, record_bytes] on
    every rank, in global query order (a view of a buffer that the next call with the same shapes overwrites).
    With `async_op=True` returns (records, work): the collective runs
    on the backend's own stream, so the caller's next kernels do not wait for the other ranks; call
    `work.wait()` (if not None) before reading `records`."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return (local_records[:n_queries], None) if async_op else local_records[:n_queries]
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    per = -(-n_queries // world)
    rec = local_records.shape[1]
    lo, hi = shard_range(n_queries, rank, world)
    if local_records.shape[0] != hi - lo:
        raise ValueError("rank %d holds %d records, expected %d" % (rank, local_records.shape[0], hi - lo))
    key = (str(local_records.device), world, per, rec)  # send / receive buffers are allocated once per shape
    if key not in _gather_buffers:
        _gather_buffers.clear()
        _gather_buffers[key] = (torch.zeros((per, rec), dtype=torch.uint8, device=local_records.device),
                                torch.empty((world * per, rec), dtype=torch.uint8, device=local_records.device))
    pad, out = _gather_buffers[key]
    pad[: hi - lo] = local_records
    work = dist.all_gather_into_tensor(out, pad, group=group, async_op=async_op)
    return (out[:n_queries], work) if async_op else out[:n_queries]


def records_to_numpy(records, dtype):
    return np.ascontiguousarray(records.cpu().numpy()).view(dtype).reshape(-1)


class host_near_gpu:
    """Context manager: runs the calling process on the CPU cores next to `device` (NVML's ideal affinity), so
    that pinned host buffers allocated inside are on the GPU's NUMA node and H2D copies of several ranks do not
    cross the socket interconnect.  Restores the previous affinity on exit; does nothing when NVML or the
    affinity calls are unavailable."""

    def __init__(self, device_index):
        self.device_index, self.saved, self.cores = device_index, None, None

    def __enter__(self):
        try:
            import os
            import pynvml
            pynvml.nvmlInit()
            uuid = torch.cuda.get_device_properties(self.device_index).uuid
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + str(uuid)).encode())
            n_words = (os.cpu_count() + 63) // 64
            words = pynvml.nvmlDeviceGetCpuAffinity(h, n_words)
            cores = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
            allowed = os.sched_getaffinity(0)
            cores &= allowed
            if cores:
                self.saved = allowed
                os.sched_setaffinity(0, cores)
                self.cores = len(cores)
        except Exception:
            self.saved = None
        return self

    def __exit__(self, *a):
        if self.saved is not None:
            import os
            try:
                os.sched_setaffinity(0, self.saved)
            except Exception:
                pass
        return False
