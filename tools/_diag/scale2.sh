#!/bin/bash
run() {
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline --no-hbm-kernels 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2))"
}
run A=0
run MB_DIAG_REDUCE=none
run MB_DIAG_REDUCE=serial
