#!/bin/bash
run() {
  python bench.py --steps 15 --warmup 4 --no-cpu-baseline --no-hbm-kernels "$@" 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$*', round(d['ms_per_step'],2), round(d['e2e']['ms_per_step'],2))"
}
run
run --h2d-gate 0
run --h2d-gate 4
run --h2d-gate 5
run --h2d-gate 6
run --e2e-ablate h2d
run --e2e-ablate readback
