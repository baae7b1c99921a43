mkdir -p gpurun_out
cd _r01
for cfg in tessellated teapot; do
python bench.py --config $cfg --steps 3 --warmup 3 --no-cpu-baseline > ../gpurun_out/r01_$cfg.json 2>/dev/null
python -c "
import json;d=json.load(open('../gpurun_out/r01_$cfg.json'));print('r01 $cfg', d['kernel_ms'], d['clocks'])"
done
cd ..
for cfg in tessellated teapot; do
python bench.py --config $cfg --steps 3 --warmup 3 --no-cpu-baseline --no-per-config > gpurun_out/r02_$cfg.json 2>/dev/null
python -c "
import json;d=json.load(open('gpurun_out/r02_$cfg.json'));print('r02 $cfg', d['kernel_ms'], d['clocks'])"
done
