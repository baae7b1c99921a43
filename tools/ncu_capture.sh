# ncu --set full of k_wf_sorted on the 10 M-triangle configuration (32 spp: the replays stay short, the frame is
# long enough that ramp-up and drain do not dominate) and on C3
mkdir -p gpurun_out
AB_NS=32 AB_REPS=2 python tools/ab_render.py tessellated_10m fixed 2>&1 | tail -2
AB_NS=32 AB_REPS=1 timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_wf_sorted -c 1 -f -o gpurun_out/r2_sorted_tess10m python tools/ab_render.py tessellated_10m fixed > gpurun_out/r11_ncu.log 2>&1
tail -3 gpurun_out/r11_ncu.log
AB_NS=64 AB_REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_sorted -c 1 -f -o gpurun_out/r2_sorted_teapot python tools/ab_render.py teapot fixed > gpurun_out/r11_ncu2.log 2>&1
tail -3 gpurun_out/r11_ncu2.log
AB_NS=64 AB_REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_sorted -c 1 -f -o gpurun_out/r2_sorted_soldier python tools/ab_render.py soldier fixed > gpurun_out/r11_ncu3.log 2>&1
tail -3 gpurun_out/r11_ncu3.log
ls -la gpurun_out/*.ncu-rep
