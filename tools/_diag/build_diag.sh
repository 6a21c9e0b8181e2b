#!/bin/bash
# Diagnostic builds: tools/_diag/libdiag_<name>.so = libmaestro_b200.so with ONE source (default gemm_tc) compiled with
# extra -D flags (MB_DIAG_EPI=n: timing only, wrong results; MB_EPI_TMA_STORE=0: coalesced-store epilogue; ...).
#   tools/_diag/build_diag.sh epi1:-DMB_DIAG_EPI=1 bc64:-DMB_ATTN_BC_SMALL=64:attn_bwd_tc
# Load with MAESTRO_B200_LIB=<path>.  Run after the normal build.
set -e
cd "$(dirname "$0")/../.."
B=maestro_b200/build
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr"
for spec in "$@"; do
  IFS=: read -r n d src <<< "$spec"; src=${src:-gemm_tc}
  nvcc $FLAGS ${d//,/ } -c maestro_b200/csrc/$src.cu -o tools/_diag/${src}_$n.o &
done
wait
for spec in "$@"; do
  IFS=: read -r n d src <<< "$spec"; src=${src:-gemm_tc}
  OTHERS=$(ls $B/*.o | grep -v "/$src.o")
  nvcc -shared -Wno-deprecated-gpu-targets -o tools/_diag/libdiag_$n.so tools/_diag/${src}_$n.o $OTHERS -lcudart
done
