#!/bin/bash
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
OUT=gpurun_out/epi_probe5.txt
: > $OUT
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k gemm 2>&1 | tail -2 >> $OUT
for lib in default buf1; do
echo "=== lib=$lib" >> $OUT
if [ $lib = default ]; then unset MAESTRO_B200_LIB; else export MAESTRO_B200_LIB=$PWD/tools/_diag/libdiag_$lib.so; fi
timeout 300 python tools/gemm_bench.py --grouped --iters 10 --only fwd,dgrad >> $OUT 2>&1
done
