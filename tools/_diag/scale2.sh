#!/bin/bash
run() {
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 15 --warmup 4 --no-cpu-baseline --no-hbm-kernels "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2))"
}
run
run --e2e-ablate h2d
run --e2e-ablate readback
run --h2d-gate 3
