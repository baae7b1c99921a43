#!/bin/bash
# A/B harness for kernel experiments on the GPU box: tools/ab_bench.sh <tag> [ENV=VAL ...]
# prints one line per config with the kernel time bench.py measured.
tag=$1; shift
for cfg in ${SRT_AB_CONFIGS:-spheres cornell}; do
  env "$@" python bench.py --config $cfg --steps 5 --warmup 3 --no-cpu-baseline --no-per-config > gpurun_out/ab_${tag}_${cfg}.json 2> gpurun_out/ab_${tag}_${cfg}.err
  python - "$tag" "$cfg" <<'PY'
import json,sys
tag,cfg=sys.argv[1:3]
try:
    d=json.load(open(f"gpurun_out/ab_{tag}_{cfg}.json"))
    print(f"{tag:28s} {cfg:8s} kernel_ms {d['kernel_ms']:9.3f}  Mpaths/s {d['value']/1e6:9.1f}  e2e_ms {d['e2e']['ms_per_step']:8.2f} parts {d['e2e'].get('parts_ms')}")
except Exception as e:
    print(tag,cfg,"FAILED",e)
PY
done
