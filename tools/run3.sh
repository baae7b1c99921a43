mkdir -p gpurun_out
export SRT_AB_CONFIGS="tessellated"
timeout 900 bash tools/ab_bench.sh t_p1 SRT_L2_PERSIST=1 
timeout 900 bash tools/ab_bench.sh t_p0 SRT_L2_PERSIST=0 
timeout 900 bash tools/ab_bench.sh t_p1_s2048 SRT_L2_PERSIST=1 SRT_WF_SLOTS_PER_SM=2048
timeout 900 bash tools/ab_bench.sh t_p0_s2048 SRT_L2_PERSIST=0 SRT_WF_SLOTS_PER_SM=2048
export SRT_AB_CONFIGS="tessellated_10m"
timeout 900 bash tools/ab_bench.sh t10_p0 SRT_L2_PERSIST=0
timeout 900 bash tools/ab_bench.sh t10_p1 SRT_L2_PERSIST=1
