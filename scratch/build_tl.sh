#!/bin/bash
# debug build with in-kernel clock stamps (-DSPS_TIMELINE) -> build/libsps_tl.so
set -e
cd /root/repo
mkdir -p build/tl
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -DSPS_TIMELINE"
for s in scaleprotoseg_b200/csrc/*.cu; do
  o=build/tl/$(basename ${s%.cu}).o
  if [ ! -f $o ] || [ $s -nt $o ] || [ -n "$(find scaleprotoseg_b200/csrc -name '*.cuh' -newer $o)" ]; then
    /usr/local/cuda/bin/nvcc $F -c $s -o $o &
  fi
done
wait
/usr/local/cuda/bin/nvcc -shared -o build/libsps_tl.so build/tl/*.o -lcudart 2>&1 | grep -v "sm_75" || true
