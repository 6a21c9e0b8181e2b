#!/bin/bash
# Diagnostic builds of the GEMM: tools/_diag/libdiag_<name>.so = libmaestro_b200.so with gemm_tc.cu compiled with extra
# -D flags (MB_DIAG_EPI=n: timing only, wrong results; MB_EPI_TMA_STORE=1: the TMA-store epilogue).
#   tools/_diag/build_diag.sh epi1:-DMB_DIAG_EPI=1 tma:-DMB_EPI_TMA_STORE=1
# Load with MAESTRO_B200_LIB=<path>.  Run after the normal build.
set -e
cd "$(dirname "$0")/../.."
B=maestro_b200/build
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr"
OTHERS=$(ls $B/*.o | grep -v gemm_tc.o)
for spec in "$@"; do
  n=${spec%%:*}; d=${spec#*:}
  nvcc $FLAGS ${d//,/ } -c maestro_b200/csrc/gemm_tc.cu -o tools/_diag/gemm_tc_$n.o &
done
wait
for spec in "$@"; do
  n=${spec%%:*}
  nvcc -shared -Wno-deprecated-gpu-targets -o tools/_diag/libdiag_$n.so tools/_diag/gemm_tc_$n.o $OTHERS -lcudart
done
