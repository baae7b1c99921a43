# DRAM traffic of k_wf_sorted on C2 against the size of the slot pool and the persisting-L2 window (one ncu pass each)
python - <<'PY'
import torch
p = torch.cuda.get_device_properties(0)
print("L2", p.L2_cache_size / 2**20, "MiB")
PY
for n in 4096 5120 6144; do
  for persist in 1 0; do
    echo -n "slots_per_sm $n persist $persist: "
    SRT_L2_PERSIST=$persist SRT_WF_SLOTS_PER_SM=$n AB_REPS=1 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_wf_sorted -c 1 python tools/ab_render.py spheres fixed 2>&1 | grep -E "dram__bytes|gpu__time" | awk '{printf "%s %s %s; ", $1, $3, $2}'
    echo
  done
done
