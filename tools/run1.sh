set -x
mkdir -p gpurun_out
export SRT_PARITY_JSON=$PWD/gpurun_out/parity.json
rm -f gpurun_out/parity.json
timeout 900 python -m pytest tests -m gpu -x -q -s > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r1_pytest.log
tail -5 gpurun_out/r1_pytest.log
export SRT_AB_CONFIGS="spheres teapot soldier"
timeout 600 bash tools/ab_bench.sh persist1 SRT_L2_PERSIST=1 > gpurun_out/r1_ab.log 2>&1
timeout 600 bash tools/ab_bench.sh persist0 SRT_L2_PERSIST=0 >> gpurun_out/r1_ab.log 2>&1
cat gpurun_out/r1_ab.log
