#!/bin/bash
# tools/build_alt.sh <name> [extra nvcc flags...] : builds csrc/build/libsrt_<name>.so with extra
# compile flags (kernel experiments; the shipped library is built by __graft_entry__.build()).
set -e
name=$1; shift
cd "$(dirname "$0")/.."
C=simple_raytracing_render_b200/csrc
B=$C/build/alt_$name
mkdir -p $B
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -prec-div=false -prec-sqrt=false -Xcompiler -fPIC -I include -I $C"
pids=()
nvcc $FLAGS "$@" -c $C/srt_api.cu -o $B/srt_api.o & pids+=($!)
nvcc $FLAGS "$@" -c $C/srt_bvh_build.cpp -o $B/srt_bvh_build.o & pids+=($!)
for v in 0 1 2 3 4 5 6 7; do
  nvcc $FLAGS "$@" -DSRT_VARIANT=$v -c $C/srt_render_variant.cu -o $B/v$v.o & pids+=($!)
done
for v in 0 1 2 3 4 5 6 7; do
  nvcc $FLAGS "$@" -DSRT_WF_VARIANT=$v -c $C/srt_wavefront_variant.cu -o $B/w$v.o & pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -gencode arch=compute_100a,code=sm_100a $B/*.o -o $C/build/libsrt_$name.so
echo built $C/build/libsrt_$name.so
