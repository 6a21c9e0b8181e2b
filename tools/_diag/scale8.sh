#!/bin/bash
run() {
  b=$1; shift
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29515 $b --gpus 8 --steps 12 --warmup 4 --no-cpu-baseline --no-hbm-kernels "$@" 2> gpurun_out/r2_8gpu.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$b $*', round(d['value'],1), round(d['ms_per_step'],2), round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],2))"
  grep "rank 0\]" gpurun_out/r2_8gpu.err | tail -2 | cut -c1-200
}
run tools/_diag/bench_r1.py
run bench.py
