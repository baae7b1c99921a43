mkdir -p gpurun_out
for c in teapot tessellated tessellated_10m; do SRT_SCENE_TIMING=1 python tools/bvh_ab.py $c 2>&1 | grep -v "^$"; done > gpurun_out/r8_bvh_ab.log 2>&1
cat gpurun_out/r8_bvh_ab.log | tail -80
timeout 1200 python -m pytest tests -m gpu -q -x -k "hit or bvh or 10m or tess or soldier or teapot or secondary" > gpurun_out/r8_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r8_pytest.log
tail -5 gpurun_out/r8_pytest.log
