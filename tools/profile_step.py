"""Run a few eager (non-graph) pretraining steps for profilers: `ncu ... python tools/profile_step.py --steps 2`.

Prints the number of kernel launches of ours per step so that a launch list can be cut to the last step.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

import bench  # noqa: E402
from maestro_b200.conf import DATASET_SPECS, MODEL_SIZES, MaskSpec  # noqa: E402
from maestro_b200.engine import PretrainEngine  # noqa: E402
from maestro_b200.ops import Kernels  # noqa: E402
from maestro_b200.optim import FusedAdamW  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--dataset", default="flair")
    ap.add_argument("--size", default="medium")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    spec, hp = DATASET_SPECS[args.dataset](), MODEL_SIZES[args.size]
    K = Kernels()
    eng = PretrainEngine(spec, hp, MaskSpec(), K, dev, torch.bfloat16, inter_depth=3, loss="l1_norm")
    eng.init_random(0)
    opt = FusedAdamW(eng, lr=1e-4, weight_decay=0.01)
    batch = {k: v.to(dev) for k, v in bench.synthetic_batch(spec, args.batch, seed=1).items()}
    torch.manual_seed(0)
    for i in range(args.steps):
        l0 = K.launches
        loss = eng.step(dict(batch))
        opt.step()
        torch.cuda.synchronize()
        print(f"step {i}: loss {float(loss):.5f}, {K.launches - l0} launches of ours", flush=True)


if __name__ == "__main__":
    main()
