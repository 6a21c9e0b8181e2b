"""Print the headline fields of a bench.py JSON line: python tools/print_bench.py FILE [FILE ...]"""
import json
import sys

for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    r = d.get("roofline", {})
    print(f"{f}: value {d['value']:.1f} {d['unit']} ({d['ms_per_step']:.2f} ms/step, {d['n_gpus']} GPU) "
          f"e2e {d['e2e']['value']:.1f}  clocks {d.get('clocks')}  "
          f"gemm {r.get('achieved', 0):.0f}/{r.get('peak', 0):.0f} {r.get('unit')} share {r.get('share_of_step', 0):.2f} "
          f"launches {d.get('gpu_launches')}")
    for k in d.get("hbm_kernels") or []:
        print(f"    {k['kernel'][:60]:60s} {k['us']:9.1f} us {k['achieved']:8.1f} {k['unit']}  frac {k['frac']:.2f}")
