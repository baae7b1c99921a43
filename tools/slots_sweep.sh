# slot-pool size sweep of the shipped library (SRT_WF_SLOTS_PER_SM), kernel ms
for n in 4096 5120 6144 7168 8192; do
  echo "slots_per_sm $n"
  SRT_WF_SLOTS_PER_SM=$n bash tools/ab_libs.sh "default" "${1:-spheres teapot soldier tessellated}" 2>&1 | grep -v "^$" | awk '{print $1, $NF}' | sort | uniq -c | awk '{print}' | sort -k2,2 -k3,3n | awk '!seen[$2]++'
done
