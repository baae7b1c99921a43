"""torchrun --nproc-per-node N tools/peer_exchange_check.py : the peer-memory exchange (multigpu.PeerExchange) against the
NCCL tiled exchange and against the unsharded render, bit for bit, over several frames; then their times."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from simple_raytracing_render_b200 import multigpu, render, scenes, scene as S
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
stream = torch.cuda.Stream(device=dev)
ctx = render.Context(local, stream.cuda_stream)
nx, ny, ns = 301, 199, 16  # pixels do not divide by the ranks: the last tile is padded
sc = render.Scene(ctx, S.flatten(*scenes.by_name("spheres")))
n = nx * ny
s0, s1 = multigpu.sample_shard(ns, rank, world)
with torch.cuda.stream(stream):
    peer = multigpu.PeerExchange(ctx, n)
    ex = multigpu.TileExchange(n, world, dev)
    whole = torch.zeros(n * 3, dtype=torch.int64, device=dev)
    ref8 = torch.zeros(n * 3, dtype=torch.uint8, device=dev)
    ok = True
    for frame in range(5):
        seed = 0x5EED + frame
        p = render.render_params(nx, ny, ns, s0, s1, seed=seed)
        sc.render_fixed(p, peer.fixed())
        host = np.zeros(n * 3, np.uint8)
        peer.run(ns, host)
        sc.render_fixed(p, ex.fixed)
        img_nccl = multigpu.exchange_tiles(ex, ns, lambda t, m, k, o: ctx.resolve_fixed(t, m, k, o, None, clear=False)).cpu().numpy()
        sc.render_fixed(render.render_params(nx, ny, ns, seed=seed), whole)
        ctx.resolve_fixed(whole, n, ns, ref8, None, clear=True)
        torch.cuda.synchronize()
        want = ref8.cpu().numpy()
        a, b = np.array_equal(host, want), np.array_equal(img_nccl, want)
        ok = ok and a and b
        print(f"rank {rank} frame {frame}: peer == unsharded {a}, nccl tiles == unsharded {b}, mean {want.mean():.3f}", flush=True)
    # times (device, events): the exchange step alone on zero buffers
    for kind in ("peer", "nccl"):
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 50
        e0.record(stream)
        for _ in range(reps):
            if kind == "peer":
                peer.run(ns)
            else:
                multigpu.exchange_tiles(ex, ns, lambda t, m, k, o: ctx.resolve_fixed(t, m, k, o, None, clear=False))
        e1.record(stream)
        torch.cuda.synchronize()
        if rank == 0:
            print(f"{kind} exchange of a {nx}x{ny} image on {world} GPUs: {e0.elapsed_time(e1) / reps * 1e3:.1f} us per frame", flush=True)
peer.close(); sc.close(); ctx.close()
dist.destroy_process_group()
print("rank", rank, "OK" if ok else "MISMATCH", flush=True)
sys.exit(0 if ok else 1)
