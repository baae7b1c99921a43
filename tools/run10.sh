mkdir -p gpurun_out
export SRT_PARITY_JSON=$PWD/gpurun_out/parity.json
rm -f gpurun_out/parity.json
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r10_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r10_pytest.log
tail -5 gpurun_out/r10_pytest.log
for c in teapot tessellated_10m; do SRT_SCENE_TIMING=1 SRT_DEVICE_BVH=1 python tools/bvh_ab.py $c 2>&1 | grep "device_bvh=1\|lbvh" | tail -3; done
( time python bench.py > gpurun_out/r10_bench.json 2> gpurun_out/r10_bench.err ) 2>&1 | tail -4
cat gpurun_out/r10_bench.json | head -c 6000
