#!/bin/bash
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
OUT=gpurun_out/epi_probe4.txt
: > $OUT
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k gemm > gpurun_out/epi_probe4_tests.txt 2>&1
tail -3 gpurun_out/epi_probe4_tests.txt >> $OUT
for lib in default tma epi1; do
for pair in 2 1; do
echo "=== lib=$lib MB_GEMM_PAIR=$pair" >> $OUT
if [ $lib = default ]; then unset MAESTRO_B200_LIB; else export MAESTRO_B200_LIB=$PWD/tools/_diag/libdiag_$lib.so; fi
MB_GEMM_PAIR=$pair timeout 300 python tools/gemm_bench.py --only inter/,dec_aerial/ --iters 10 >> $OUT 2>&1
done
done
