#!/bin/bash
# usage: attn_probe.sh lib1 lib2 ... ("default" = the product library)
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
OUT=gpurun_out/attn_probe.txt
: > $OUT
for lib in "$@"; do
echo "=== lib=$lib" >> $OUT
if [ $lib = default ]; then unset MAESTRO_B200_LIB; else export MAESTRO_B200_LIB=$PWD/tools/_diag/libdiag_$lib.so; fi
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k attn 2>&1 | tail -2 >> $OUT
timeout 300 python tools/kernel_bench.py --only attn --iters 10 >> $OUT 2>&1
done
