# small frames (what one rank of an 8-GPU run renders): pool size sweep, kernel ms
for cfg in spheres:8 spheres:16 soldier:128 tessellated_10m:32; do
  c=${cfg%%:*}; ns=${cfg##*:}
  for n in 1280 1920 2560 3840 5120 6144; do
    echo -n "$c @ $ns spp, slots_per_sm $n: "
    AB_NS=$ns AB_REPS=4 SRT_WF_SLOTS_PER_SM=$n python tools/ab_render.py $c fixed 2>&1 | grep " ms " | awk '{print $NF}' | sort -n | head -1
  done
done
