# tools/ab_env.sh "<lib[:ctas]> ..." "<configs>" [reps] : kernel ms + checksum of library builds (csrc/build/libsrt_<lib>.so,
# "default" = the shipped one) with SRT_WF_CTAS_PER_SM=<ctas>
mkdir -p gpurun_out
B=$PWD/simple_raytracing_render_b200/csrc/build
for cfg in $2; do
  for combo in $1; do
    lib=${combo%%:*}; unset SRT_WF_CTAS_PER_SM; n="variant's own"; [ "$combo" != "$lib" ] && { n=${combo##*:}; export SRT_WF_CTAS_PER_SM=$n; }
    if [ $lib = default ]; then unset SRT_B200_LIB; else export SRT_B200_LIB=$B/libsrt_$lib.so; fi
    echo "== $cfg $lib ctas=$n"
    AB_REPS=${3:-3} timeout 400 python tools/ab_render.py $cfg fixed 2>&1 | grep -E " ms |checksum" | tail -3
  done
done 2>&1 | tee -a gpurun_out/ab_env.log
