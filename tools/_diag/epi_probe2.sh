#!/bin/bash
# After the coalesced-store epilogue / tile-width / split-K changes: correctness of the GEMM paths, then timings.
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
OUT=gpurun_out/epi_probe2.txt
: > $OUT
timeout 600 python -m pytest tests/test_kernels_gpu.py -x -q -m gpu -k gemm > gpurun_out/epi_probe2_tests.txt 2>&1
tail -3 gpurun_out/epi_probe2_tests.txt >> $OUT
echo "=== default" >> $OUT
timeout 300 python tools/gemm_bench.py --only inter/,dec_aerial/ --iters 10 >> $OUT 2>&1
echo "=== grouped default" >> $OUT
timeout 300 python tools/gemm_bench.py --grouped --iters 10 >> $OUT 2>&1
echo "=== grouped MINK=4000" >> $OUT
MB_GEMM_PAIR_MINK=4000 timeout 300 python tools/gemm_bench.py --grouped --iters 10 --only wgrad >> $OUT 2>&1
echo "=== single MINK=4000" >> $OUT
MB_GEMM_PAIR_MINK=4000 timeout 300 python tools/gemm_bench.py --iters 10 --only enc_aerial/wgrad,enc_s2/wgrad >> $OUT 2>&1
