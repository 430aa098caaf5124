"""Sample-split multi-GPU rendering: one process per GPU, scene replicated,
one exchange step that sums the float films.

The data path is in libptb.so (include/ptb.h "groups": NCCL communicators, the
flattened-scene broadcast, ncclReduce / reduce-scatter / the fused peer-memory
reduce+resolve kernel).  This module is the launcher-side plumbing around it for
torchrun-style jobs: it reads RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*, hands
rank 0's 128-byte group id to the other ranks through torch.distributed, and
offers the scalar reductions bench.py needs for its timing (max / sum over
ranks).  The reference itself is single-device (src/core/vulkan.cpp:1485-1506).
"""
import os

import torch
import torch.distributed as dist


def init_process_group(backend=None):
    """Reads RANK / WORLD_SIZE / LOCAL_RANK / MASTER_* from the environment (torchrun)."""
    if dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    if world == 1:
        return 0, 1
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29512")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    kw = {}
    if backend == "nccl":
        kw["device_id"] = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, world


def frame_schedule(rank, world, frames_per_rank, base=0):
    """First FrameIndex (= RNG seed, basic.cpp:315,323) of rank `rank` when every rank
    renders `frames_per_rank` frames: rank r owns seeds base + r*(F+1) + [0, F] (the reset
    dispatch uses the first one).  Blocks are disjoint and gap-free over ranks.
    Same rule as ptb_group_first_frame in the C ABI."""
    stride = frames_per_rank + 1
    return base + rank * stride


def broadcast_bytes(payload, src=0, device=None):
    """Hands a small byte string from `src` to every rank (the group id, the camera records)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return payload
    box = [payload if dist.get_rank() == src else None]
    dist.broadcast_object_list(box, src=src, device=device)
    return box[0]


def create_group(ctx, device=None):
    """The C-ABI communicator of this job (None for a single process): rank 0 draws the id, everybody joins."""
    from . import integrator as I
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return None
    rank, world = dist.get_rank(), dist.get_world_size()
    gid = broadcast_bytes(I.Group.unique_id() if rank == 0 else None, 0, device)
    return I.Group(ctx, gid, rank, world)


def reduce_film(accum, dst=0, all_ranks=False):
    """torch.distributed sum of RGBA32F accumulator tensors (kept for hosts that own the film as a tensor;
    the product path is Group.reduce / Group.reduce_resolve on library-owned films)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return accum
    if all_ranks:
        dist.all_reduce(accum, op=dist.ReduceOp.SUM)
    else:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum


def barrier():
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def max_over_ranks(value, device):
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(value, device):
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def shutdown():
    """Tears the process group down (no-op for a single process)."""
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
