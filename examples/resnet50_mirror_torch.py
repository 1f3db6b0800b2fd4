"""imagenet-resnet50-mirror.py, torch flavour: ONE process drives every GPU of the box (the
MirroredStrategy shape, reference line :21) through a *local* b200dp context.

Per step: every replica runs forward/backward on its own GPU and batch shard, then
``plan.pack_step`` sums the 214 gradients across replicas (rank-order, bit-identical to TF's
ReductionToOneDevice), applies Adam on each replica's shard and all-gathers the weights -- the
work `batch_reduce` + 214 ResourceApplyAdam launches do under `model.fit` (:76, :81).

    python examples/resnet50_mirror_torch.py --steps 3 --batch 32          # all visible GPUs
"""
import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200dp  # noqa: E402
from b200dp import harness  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--batch", type=int, default=32, help="per replica (the script uses 32*num_replicas, :54)")
    ap.add_argument("--gpus", type=int, default=0, help="0 = all")
    args = ap.parse_args()
    n = args.gpus or torch.cuda.device_count()
    devices = list(range(n))
    torch.backends.cudnn.allow_tf32 = True
    numels = b200dp.resnet50_numels()
    ctx = b200dp.Context.local(devices)                       # strategy = MirroredStrategy()
    plan = b200dp.Plan(ctx, b200dp.Layout(numels, nranks=n))
    steppers, data = [], []
    for r in devices:
        with torch.cuda.device(r):
            torch.manual_seed(42)                              # with strategy.scope(): identical mirrored variables
            model = harness.KerasResNet50().cuda(r).to(memory_format=torch.channels_last).train()
            steppers.append(harness.DataParallelStep(model, plan, dev=r, optimizer="adam", lr=1e-3))
            g = torch.Generator(device="cuda:%d" % r).manual_seed(7 + r)
            x = (torch.rand(args.batch, 3, 224, 224, generator=g, device="cuda:%d" % r) * 255).contiguous(
                memory_format=torch.channels_last)
            y = torch.randint(0, 1000, (args.batch,), generator=g, device="cuda:%d" % r)
            data.append((x, y))
    for step in range(args.steps):
        t0 = time.time()
        losses = []
        for r in devices:                                      # replicas run concurrently: launches are async
            with torch.cuda.device(r):
                losses.append(steppers[r].forward_backward(*data[r]))
        with b200dp.issue_group():                             # nothing may block the host between the replicas' calls
            for r in devices:                                  # one collective call per replica, same order
                with torch.cuda.device(r):
                    steppers[r].hot_path()
        for r in devices:
            torch.cuda.synchronize(r)
        print("step %d  loss %.4f  %.1f images/s" % (step + 1, sum(float(l) for l in losses),
                                                     n * args.batch / (time.time() - t0)))
    w0 = plan.tensor(0, b200dp.BUF_WEIGHTS)
    same = all(torch.equal(w0.cpu(), plan.tensor(r, b200dp.BUF_WEIGHTS).cpu()) for r in devices[1:])
    for r in devices:
        with torch.cuda.device(r):
            plan.check(r)
    print("replicas hold identical weights:", same)
    return 0 if same else 1


if __name__ == "__main__":
    sys.exit(main())
