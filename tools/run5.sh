python tools/ab_render.py tessellated float
python tools/ab_render.py tessellated fixed
python tools/ab_render.py tessellated float peaks
SRT_B200_LIB=$PWD/simple_raytracing_render_b200/csrc/build/libsrt_noovf.so python tools/ab_render.py tessellated float
