mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_features.py -m gpu -q -x -k "device_reference or device_scene" > gpurun_out/r9_pytest_a.log 2>&1; echo "pytest rc $?" >> gpurun_out/r9_pytest_a.log
tail -15 gpurun_out/r9_pytest_a.log
for c in teapot tessellated tessellated_10m; do SRT_SCENE_TIMING=1 python tools/bvh_ab.py $c 2>&1 | grep -v "^$"; done > gpurun_out/r9_bvh_ab.log 2>&1
grep -v "item boxes\|leaf arrays\|upload complete" gpurun_out/r9_bvh_ab.log | tail -70
timeout 1200 python -m pytest tests -m gpu -q -x -k "hit or bvh or 10m or tess or soldier or teapot or secondary" > gpurun_out/r9_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r9_pytest.log
tail -5 gpurun_out/r9_pytest.log
