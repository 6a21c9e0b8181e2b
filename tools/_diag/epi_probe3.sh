#!/bin/bash
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
OUT=gpurun_out/epi_probe3.txt
: > $OUT
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k gemm > gpurun_out/epi_probe3_tests.txt 2>&1
tail -3 gpurun_out/epi_probe3_tests.txt >> $OUT
for pair in 2 1; do
echo "=== MB_GEMM_PAIR=$pair" >> $OUT
MB_GEMM_PAIR=$pair timeout 300 python tools/gemm_bench.py --only inter/,dec_aerial/ --iters 10 >> $OUT 2>&1
done
echo "=== grouped MINK=4000" >> $OUT
MB_GEMM_PAIR_MINK=4000 timeout 300 python tools/gemm_bench.py --grouped --iters 10 >> $OUT 2>&1
