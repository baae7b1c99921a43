"""Multi-GPU driver of the render path: one process per GPU, sample-range shards, ONE reduce.

The reference's only parallelism is a shared pixel queue over std::threads
(Raytracing_n.cpp:815-825, :932-941).  Here every rank renders Sobol indices
[ns*rank/world, ns*(rank+1)/world) of EVERY pixel (the reference reuses the same ns points for
every pixel, :834-835, so the ranges are disjoint and exhaustive) into its own accumulation buffer;
the buffers are summed onto rank 0 with a single reduce (NCCL over NVLink/NVSwitch on GPUs), and rank 0
resolves (mean -> sqrt -> 8 bit, :848-862).  DESIGN.md §6.

Two buffer forms: float32 (`render_sharded`: srt_render + reduce + srt_resolve; the sum depends on the
reduce order in the last float bits) and int64 fixed point (`render_sharded_fixed`: srt_render_fixed +
an exact integer reduce + srt_resolve_fixed — the form bench.py times: no clear pass, no conversion
kernel, and the N-GPU image is bit-identical to the 1-GPU image).

The shard arithmetic and the reduce are device-agnostic (tests/test_multigpu_cpu.py runs them over
gloo with CPU tensors); rendering and resolving need the CUDA backend.
"""
from __future__ import annotations

from typing import Optional, Tuple


def sample_shard(ns: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Sobol index range [s_begin, s_end) of `rank`: contiguous, disjoint, exhaustive, sizes differ by <= 1."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    if ns < 0:
        raise ValueError("negative sample count")
    return (ns * rank) // world_size, (ns * (rank + 1)) // world_size


def reduce_accum(accum, dst: int = 0, group=None):
    """The path's one exchange step: sum the per-rank accumulation buffers onto `dst` (in place).
    No-op without an initialised process group (single GPU)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum


def render_sharded(scene, nx: int, ny: int, ns: int, accum, rgb8=None, max_depth: int = 50, seed: int = 0x5EED,
                   rank: Optional[int] = None, world_size: Optional[int] = None, group=None):
    """One frame on all ranks: zero `accum` (device float32 [ny*nx*3]), render this rank's shard
    into it, reduce onto rank 0, resolve there into `rgb8` (device or host uint8 [ny*nx*3]).
    Returns this rank's (s_begin, s_end)."""
    import torch.distributed as dist
    from . import render
    if rank is None or world_size is None:
        if dist.is_available() and dist.is_initialized():
            rank, world_size = dist.get_rank(group), dist.get_world_size(group)
        else:
            rank, world_size = 0, 1
    s0, s1 = sample_shard(ns, rank, world_size)
    accum.zero_()
    scene.render(render.render_params(nx, ny, ns, s0, s1, max_depth=max_depth, seed=seed), accum)
    reduce_accum(accum, 0, group)
    if rank == 0 and rgb8 is not None:
        scene.ctx.resolve(accum, nx * ny, ns, rgb8)
    return s0, s1


def render_sharded_fixed(scene, nx: int, ny: int, ns: int, fixed, rgb8=None, max_depth: int = 50, seed: int = 0x5EED,
                         rank: Optional[int] = None, world_size: Optional[int] = None, group=None, flags=None):
    """One frame on all ranks in the fixed-point form.  `fixed` (device int64 [ny*nx*3]) must be ZERO on
    entry and is zero again on return: the kernels add this rank's shard into it (srt_render_fixed), the
    buffers are summed onto rank 0 with one integer reduce, rank 0 resolves into `rgb8` (device or host uint8
    [ny*nx*3]) and clears in the same kernel; the other ranks clear theirs.  Returns (s_begin, s_end)."""
    import torch.distributed as dist
    from . import render
    if rank is None or world_size is None:
        if dist.is_available() and dist.is_initialized():
            rank, world_size = dist.get_rank(group), dist.get_world_size(group)
        else:
            rank, world_size = 0, 1
    s0, s1 = sample_shard(ns, rank, world_size)
    scene.render_fixed(render.render_params(nx, ny, ns, s0, s1, max_depth=max_depth, seed=seed, flags=flags), fixed)
    reduce_accum(fixed, 0, group)
    if rank == 0:
        scene.ctx.resolve_fixed(fixed, nx * ny, ns, rgb8, None, clear=True)
    else:
        fixed.zero_()
    return s0, s1
