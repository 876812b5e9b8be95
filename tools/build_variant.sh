#!/bin/bash
# Builds librtc.so + librtc_driver.so with extra nvcc flags into gpurun_variants/NAME/ (shipped to the GPU box; select it with
# RTC_LIB_DIR=gpurun_variants/NAME).  Usage: tools/build_variant.sh NAME -DRTC_TRACE_MIN_BLOCKS_TRANSPARENT=5 ...
set -e
NAME=$1; shift
OUT=gpurun_variants/$NAME; mkdir -p $OUT build/obj_$NAME
C=raytracer_challenge_b200/csrc
ARCH="-gencode arch=compute_100a,code=sm_100a -Wno-deprecated-gpu-targets"
nvcc -std=c++17 -O3 -lineinfo -fmad=false -Xcompiler -fPIC $ARCH "$@" -Xptxas -v -x cu -c $C/render_kernels.cu -o build/obj_$NAME/render_kernels.o 2> build/obj_$NAME/ptxas.log
nvcc -shared -o $OUT/librtc.so build/obj_$NAME/render_kernels.o build/obj/bvh_build.cu.o build/obj/rtc_api.cpp.o build/obj/scene_build.cpp.o build/obj/microbench.cu.o build/obj/noalloc.cu.o $ARCH -cudart shared
g++ -std=c++17 -O2 -fPIC -shared -o $OUT/librtc_driver.so driver/driver.cpp -L$OUT -lrtc -Wl,-rpath,'$ORIGIN'
grep -E "Compiling entry|Used|spill" build/obj_$NAME/ptxas.log | paste - - - | sed -e 's/ptxas info    : //g' -e 's/Compiling entry function //' -e "s/for 'sm_100a'//" | grep -E "trace_spawnILi[17]ELb0" | cut -c1-260
