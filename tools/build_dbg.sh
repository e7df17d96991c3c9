#!/bin/sh
# debug build with -DLGBCU_STAGE_TIMERS (device-side stage timers / stamps) -> tools/_dbg/lib_dbg.so
set -e
cd "$(dirname "$0")/.."
make lib >/dev/null
mkdir -p tools/_dbg build
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 --fmad=false -std=c++17 -Xcompiler -fPIC -Xcompiler -fopenmp \
  -Iinclude -ccbin /usr/bin/g++ -DLGBCU_STAGE_TIMERS -c lightgbmv1_b200/csrc/cuda/lgbcu.cu -o build/lgbcu_dbg.o
nvcc -gencode arch=compute_100a,code=sm_100a -shared -ccbin /usr/bin/g++ -Xcompiler -fopenmp -o tools/_dbg/lib_dbg.so \
  build/lgbcu_dbg.o build/host_*.o -ldl
