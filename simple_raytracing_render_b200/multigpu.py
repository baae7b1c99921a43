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

The exchange step bench.py times is the TILED form of the same sum (`TileExchange`, `exchange_tiles`): one
reduce-scatter of the int64 buffers (rank g ends up with the exact sums of pixel tile g: every link carries 1/N of
the buffer instead of rank 0 receiving all of it), the resolve of that tile on its rank, and one all-gather of the
8-bit tiles — 0.35 -> ~0.1 ms of a 3.7 ms C2 frame on 8 GPUs, and every rank holds the finished image.

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


class TileExchange:
    """Buffers of the tiled exchange step for images of `n_pixels` pixels on `world_size` ranks: the pixel range is
    padded to a multiple of world_size and cut into equal tiles, tile g belongs to rank g.
      fixed    int64 [padded_pixels * 3]  this rank's fixed-point sums (srt_render_fixed adds into its first n_pixels * 3)
      tile     int64 [tile_pixels * 3]    the summed tile after the reduce-scatter
      rgb_tile uint8 [tile_pixels * 3]    the resolved tile
      rgb      uint8 [padded_pixels * 3]  the gathered image (its first n_pixels * 3 values)"""

    def __init__(self, n_pixels: int, world_size: int, device):
        import torch
        self.n_pixels, self.world_size = int(n_pixels), int(world_size)
        self.tile_pixels = -(-self.n_pixels // self.world_size)
        self.padded_pixels = self.tile_pixels * self.world_size
        self.fixed = torch.zeros(self.padded_pixels * 3, dtype=torch.int64, device=device)
        self.tile = torch.empty(self.tile_pixels * 3, dtype=torch.int64, device=device)
        self.rgb_tile = torch.empty(self.tile_pixels * 3, dtype=torch.uint8, device=device)
        self.rgb = torch.empty(self.padded_pixels * 3, dtype=torch.uint8, device=device)

    def image(self):
        return self.rgb[: self.n_pixels * 3]


def exchange_tiles(ex: TileExchange, n_samples: int, resolve_tile, group=None):
    """The exchange step in tiles.  `ex.fixed` holds this rank's sums on entry and is ZERO on return;
    `resolve_tile(tile_int64, n_tile_pixels, n_samples, rgb_tile_uint8)` is the backend's mean -> sqrt -> 8-bit
    (Context.resolve_fixed on the GPU; the CPU test passes a numpy restatement).  Every rank returns with the whole
    image in ex.image()."""
    import torch.distributed as dist
    dist.reduce_scatter_tensor(ex.tile, ex.fixed, op=dist.ReduceOp.SUM, group=group)
    resolve_tile(ex.tile, ex.tile_pixels, n_samples, ex.rgb_tile)
    ex.fixed.zero_()
    dist.all_gather_into_tensor(ex.rgb, ex.rgb_tile, group=group)
    return ex.image()


class PeerExchange:
    """The exchange step as ONE kernel over NVLink peer memory (include/srt_ffi.h: srt_exchange_*): every
    rank's accumulation buffers and image live in a block all ranks of the box have mapped (CUDA IPC); rank g waits
    inside the kernel until every rank has rendered, sums pixel tile g with peer loads, resolves it and stores the 8-bit
    tile into every rank's image.  No NCCL call per frame: torch.distributed is used once, to pass the IPC handles round.

        ex = PeerExchange(ctx, n_pixels)          # collective: every rank of the group
        scene.render_fixed(params, ex.fixed())    # this rank's shard into this frame's buffer
        ex.run(ns, host_rgb)                      # collective; host_rgb (optional): the finished image, synchronises
    """

    def __init__(self, ctx, n_pixels: int, group=None):
        import ctypes as C
        import torch
        import torch.distributed as dist
        from . import ffi
        self.ctx, self.lib = ctx, ctx.lib
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.n_pixels = int(n_pixels)
        dev = torch.device("cuda", torch.cuda.current_device())
        self.h = None
        err = None
        # every step that can fail locally (allocation, IPC export, peer mapping) is followed by a COLLECTIVE verdict:
        # either every rank has an exchange or every rank raises — a rank must not be left waiting for the others
        try:
            h = C.c_void_p()
            ffi.check(self.lib, self.lib.srt_exchange_create(ctx.h, self.n_pixels, self.world, self.rank, C.byref(h)))
            self.h = h
            mine = (C.c_uint8 * ffi.EXCHANGE_HANDLE_BYTES)()
            ffi.check(self.lib, self.lib.srt_exchange_handle(self.h, mine))
            t = torch.tensor(list(mine), dtype=torch.uint8, device=dev)
        except Exception as e:  # noqa: BLE001
            err = e
            t = torch.zeros(ffi.EXCHANGE_HANDLE_BYTES, dtype=torch.uint8, device=dev)
        allh = torch.empty(self.world * ffi.EXCHANGE_HANDLE_BYTES, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(allh, t, group=group)
        if err is None:
            try:
                ffi.check(self.lib, self.lib.srt_exchange_connect(self.h, bytes(allh.cpu().numpy().tobytes())))
            except Exception as e:  # noqa: BLE001
                err = e
        ok = torch.tensor([0 if err is not None else 1], dtype=torch.int32, device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)  # also: nobody starts a frame before every rank has mapped every block
        if int(ok.item()) == 0:
            self.close()
            raise RuntimeError(f"peer-memory exchange unavailable on this box ({err if err is not None else 'another rank failed'})")

    def fixed(self) -> int:
        """Device address of the int64 buffer the NEXT run() will exchange (zero on entry): pass it to render_fixed."""
        return int(self.lib.srt_exchange_fixed(self.h))

    def image(self) -> int:
        """Device address of the finished 8-bit image (valid once run()'s work has executed on the stream)."""
        return int(self.lib.srt_exchange_image(self.h))

    def run(self, n_samples: int, rgb8_host=None):
        from . import ffi
        from .render import _ptr
        ffi.check(self.lib, self.lib.srt_exchange_run(self.h, int(n_samples), _ptr(rgb8_host)))

    def close(self):
        if getattr(self, "h", None):
            self.lib.srt_exchange_destroy(self.h)
            self.h = None
