#!/bin/bash
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
OUT=gpurun_out/epi_probe6.txt
: > $OUT
for lib in default pf0 pipe1 pf0pipe1; do
echo "=== lib=$lib" >> $OUT
if [ $lib = default ]; then unset MAESTRO_B200_LIB; else export MAESTRO_B200_LIB=$PWD/tools/_diag/libdiag_$lib.so; fi
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k gemm 2>&1 | tail -1 >> $OUT
timeout 300 python tools/gemm_bench.py --grouped --iters 10 --only fwd_out,fwd_down,dgrad_down >> $OUT 2>&1
done
