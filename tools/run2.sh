set -x
mkdir -p gpurun_out
export SRT_PARITY_JSON=$PWD/gpurun_out/parity.json
rm -f gpurun_out/parity.json
timeout 1200 python -m pytest tests -m gpu -q -s > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc $?" >> gpurun_out/r2_pytest.log
tail -8 gpurun_out/r2_pytest.log
( time timeout 1500 python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err ) 2> gpurun_out/r2_bench.time
tail -3 gpurun_out/r2_bench.time; tail -30 gpurun_out/r2_bench.err
M=dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
for cfg in spheres teapot soldier; do
  for P in 1 0; do
    SRT_L2_PERSIST=$P timeout 600 ncu --metrics $M --clock-control none -k regex:k_wf_sorted --launch-skip 1 -c 1 --csv --log-file gpurun_out/r2_dram_${cfg}_p$P.csv python bench.py --config $cfg --steps 1 --warmup 3 --no-per-config --no-cpu-baseline > /dev/null 2>&1
    tail -3 gpurun_out/r2_dram_${cfg}_p$P.csv
  done
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_wf_sorted --launch-skip 1 -c 1 -o gpurun_out/r2_sorted_spheres python bench.py --config spheres --steps 1 --warmup 3 --no-per-config --no-cpu-baseline > /dev/null 2>&1
ls -la gpurun_out/*.ncu-rep
