"""Sample-split multi-GPU rendering: one process per GPU, scene replicated,
one sum-reduce of the float accumulation buffers.

The reference is single-device (src/core/vulkan.cpp:1485-1506).  Samples are
independent given distinct RNG streams and the film is additive (rgb = sum of
XYZ, a = count; resolve divides by a, resolve.glsl:117-118), so partial films
simply add.  There is no per-round exchange: the only collective is the final
reduce of 16*W*H bytes over NCCL (NVLink 5 / NVSwitch), issued through
torch.distributed, which is plumbing here, not the product.
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
    renders `frames_per_rank` frames: rank r owns seeds base + r*F + [0, F] (the reset
    dispatch uses the first one).  Blocks are disjoint and gap-free over ranks."""
    stride = frames_per_rank + 1
    return base + rank * stride


def reduce_film(accum, dst=0, all_ranks=False):
    """Sums the RGBA32F accumulators of all ranks in place (on `dst`, or everywhere)."""
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
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
