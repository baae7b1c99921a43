# ncu --set full of k_wf_sorted on the HBM-resident configuration (8 spp so that the replays stay short) and on C2
mkdir -p gpurun_out
AB_NS=8 AB_REPS=2 python tools/ab_render.py tessellated_10m fixed 2>&1 | tail -2
AB_NS=8 AB_REPS=1 timeout 1500 ncu --set full --clock-control none --import-source on -k regex:k_wf_sorted -c 1 -f -o gpurun_out/r2_sorted_tess10m python tools/ab_render.py tessellated_10m fixed > gpurun_out/r11_ncu.log 2>&1
tail -3 gpurun_out/r11_ncu.log
AB_REPS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_wf_sorted -c 1 -f -o gpurun_out/r2_sorted_spheres python tools/ab_render.py spheres fixed > gpurun_out/r11_ncu2.log 2>&1
tail -3 gpurun_out/r11_ncu2.log
ls -la gpurun_out/*.ncu-rep
