mkdir -p gpurun_out
export SRT_PARITY_JSON=$PWD/gpurun_out/parity.json
rm -f gpurun_out/parity.json
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r7_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r7_pytest.log
tail -5 gpurun_out/r7_pytest.log
for c in teapot soldier tessellated tessellated_10m; do python tools/bvh_ab.py $c 2>&1 | grep -v "^$"; done
