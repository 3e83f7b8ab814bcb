#!/bin/bash
# Development: builds diagnostic variants of the library (wfn_ws_kernel.cuh, C2_WS_DIAG bits) next to the shipped one as
# critic2_b200/libc2gpu_diag<bits>.so.  They are git-ignored, never loaded by the product (C2GPU_LIB selects them for
# tools/probe.py) and their field values are meaningless -- they remove work to see what the tensor pipe waits for.
set -e
cd "$(dirname "$0")/../critic2_b200/csrc"
ARCH="-gencode arch=compute_100a,code=sm_100a"
for bits in "$@"; do
  d=_build/diag$bits
  mkdir -p $d
  for f in c2gpu wfn_launch promol_launch cube_format; do
    if [ $f = wfn_launch ] || [ ! -f $d/$f.o ]; then
      nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -DC2_WS_DIAG=$bits -c $f.cu -o $d/$f.o &
    fi
  done
  wait
  nvcc $ARCH -shared -o ../libc2gpu_diag$bits.so $d/c2gpu.o $d/wfn_launch.o $d/promol_launch.o $d/cube_format.o -cudart static
  echo built libc2gpu_diag$bits.so
done
