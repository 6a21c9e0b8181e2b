#!/bin/bash
# Where does the short-K GEMM time go?  default / stripped-epilogue / no-TMA-store / direct-store builds, single-CTA and pair tiles.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
OUT=gpurun_out/epi_probe.txt
: > $OUT
for lib in default epi1 epi2 epi3; do
  for pair in 2 1; do
    echo "=== lib=$lib MB_GEMM_PAIR=$pair" >> $OUT
    if [ $lib = default ]; then unset MAESTRO_B200_LIB; else export MAESTRO_B200_LIB=$PWD/tools/_diag/libdiag_$lib.so; fi
    MB_GEMM_PAIR=$pair timeout 300 python tools/gemm_bench.py --only inter/,dec_aerial/ --iters 10 >> $OUT 2>&1
  done
done
unset MAESTRO_B200_LIB
echo "=== grouped default" >> $OUT
timeout 300 python tools/gemm_bench.py --grouped --iters 10 >> $OUT 2>&1
