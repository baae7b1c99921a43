# multi-GPU check: tools/bench_multigpu.sh <N> [extra bench flags]; the JSON line lands in gpurun_out/r02_bench_n<N>.json
N=$1; shift
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 "$@" > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err
tail -4 gpurun_out/r02_bench_n$N.err
python - $N <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r02_bench_n{n}.json"))
    print(f"N={n} {d['config']['workload']}: {d['ms_per_step']:.3f} ms/step value {d['value']/1e9:.3f} G paths/s e2e {d['e2e']['value']/1e9:.3f} kernel_ms {d['kernel_ms']:.3f}")
    for p in d.get('per_config',[]):
        print(f"   {p['config']}: {p['ms_per_step']:.2f} ms/step value {p['value']/1e9:.3f} G paths/s e2e {p['e2e']['value']/1e9:.3f} kernel_ms {p['kernel_ms']:.2f}")
except Exception as e:
    print("FAILED", e)
PY
