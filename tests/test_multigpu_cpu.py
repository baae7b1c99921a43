"""The N > 1 path on CPU: world_size-2 gloo process group, sample-range shards + the one reduce
(DESIGN.md §6).  Rendering itself needs a GPU; here every rank stands in for its render with a
deterministic per-sample contribution, so that what is checked is exactly the host logic of the
multi-GPU path: shards partition [0, ns), the reduce lands the whole sum on rank 0 only."""
import os
import socket

import numpy as np
import pytest

from simple_raytracing_render_b200 import multigpu


def test_sample_shards_partition_the_range():
    for ns in (0, 1, 7, 16, 64, 1000, 1024):
        for world in (1, 2, 3, 4, 8):
            cuts = [multigpu.sample_shard(ns, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == ns
            for (a0, a1), (b0, b1) in zip(cuts, cuts[1:]):
                assert a1 == b0 and a0 <= a1
            sizes = [b - a for a, b in cuts]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        multigpu.sample_shard(8, 2, 2)


def _contribution(s, n):
    """Stand-in for the radiance sums of Sobol index s over n values (exact in float32)."""
    i = np.arange(n, dtype=np.int64)
    return (((i * 7 + s * 13) % 32) / 8.0).astype(np.float32)


def _resolve_numpy(fixed, n_samples):
    """Raytracing_n.cpp:848-862 on 2^-31 fixed-point sums: mean, sqrt, int(255.99 c)."""
    c = np.sqrt(fixed.astype(np.float64) / 2.0 ** 31 / n_samples)
    return np.clip((255.99 * c).astype(np.int64), 0, 255).astype(np.uint8)


def _resolve_tile_numpy(tile, n_tile_pixels, n_samples, rgb_tile):
    import torch
    rgb_tile.copy_(torch.from_numpy(_resolve_numpy(tile.numpy(), n_samples)))


def _worker(rank, world, port, ns, n, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        s0, s1 = multigpu.sample_shard(ns, rank, world)
        acc = np.zeros(n, np.float32)
        for s in range(s0, s1):
            acc += _contribution(s, n)
        t = torch.from_numpy(acc.copy())
        multigpu.reduce_accum(t, dst=0)
        np.save(os.path.join(out_dir, f"rank{rank}.npy"), t.numpy())
        # the fixed-point form (srt_render_fixed): int64 sums with 31 fractional bits, reduced exactly
        fx = torch.from_numpy(np.round(acc.astype(np.float64) * 2.0 ** 31).astype(np.int64))
        multigpu.reduce_accum(fx, dst=0)
        np.save(os.path.join(out_dir, f"fixed{rank}.npy"), fx.numpy())
        np.save(os.path.join(out_dir, f"own{rank}.npy"), acc)
        # the tiled exchange bench.py times: reduce-scatter, resolve per tile, all-gather of the 8-bit tiles
        ex = multigpu.TileExchange(n // 3, world, "cpu")
        ex.fixed[:n] = torch.from_numpy(np.round(acc.astype(np.float64) * 2.0 ** 31).astype(np.int64))
        img = multigpu.exchange_tiles(ex, ns, _resolve_tile_numpy)
        np.save(os.path.join(out_dir, f"tiles{rank}.npy"), img.numpy())
        assert int(ex.fixed.abs().sum()) == 0
    finally:
        dist.destroy_process_group()


def test_two_rank_reduce_equals_unsharded(tmp_path):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ns, n, world = 13, 3 * 37 * 21, 2
    mp.spawn(_worker, args=(world, port, ns, n, str(tmp_path)), nprocs=world, join=True)
    whole = np.zeros(n, np.float32)
    for s in range(ns):
        whole += _contribution(s, n)
    got = np.load(tmp_path / "rank0.npy")
    assert np.array_equal(got, whole)  # contributions are exact in float32: order cannot matter
    fixed = np.load(tmp_path / "fixed0.npy")
    assert fixed.dtype == np.int64 and np.array_equal(fixed, np.round(whole.astype(np.float64) * 2.0 ** 31).astype(np.int64))
    own = [np.load(tmp_path / f"own{r}.npy") for r in range(world)]
    assert np.array_equal(own[0] + own[1], whole)
    # tiled exchange: every rank ends with the image the unsharded sums resolve to (37 * 21 pixels do not divide by 2:
    # the last tile is padded)
    want = _resolve_numpy(np.round(whole.astype(np.float64) * 2.0 ** 31).astype(np.int64), ns)
    for r in range(world):
        assert np.array_equal(np.load(tmp_path / f"tiles{r}.npy"), want)


def test_reduce_is_a_noop_without_a_process_group():
    import torch
    t = torch.arange(6, dtype=torch.float32)
    assert multigpu.reduce_accum(t) is t and t.tolist() == [0, 1, 2, 3, 4, 5]
