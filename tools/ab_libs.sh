# A/B of library builds: tools/ab_libs.sh "<libs>" "<configs>"   (default = the shipped library)
mkdir -p gpurun_out
B=simple_raytracing_render_b200/csrc/build
for cfg in ${2:-spheres teapot soldier tessellated}; do
  for lib in ${1:-default base}; do
    if [ $lib = default ]; then unset SRT_B200_LIB; else export SRT_B200_LIB=$PWD/$B/libsrt_$lib.so; fi
    AB_REPS=3 python tools/ab_render.py $cfg fixed 2>&1 | grep -E " ms |phases" | tail -3
  done
done 2>&1 | tee -a gpurun_out/r12_ab.log
