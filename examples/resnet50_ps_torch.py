"""imagenet-resnet50-ps.py, torch flavour: the ParameterServerStrategy push / pull mapped onto
reduce-scatter / all-gather shards (BASELINE.json configs[4]).

In the script every variable lives on a parameter server (sharded by
``MinSizePartitioner(min_shard_bytes=256 << 10, max_shards=NUM_PS)``, :75-78); a worker PUSHES its
gradients there, the update runs on the server, and the worker PULLS fresh weights (:80-82, fit
:142).  Here rank r is the parameter server of slice r of every bucket:

    push  = b2dp_reduce_scatter_apply  (gradients of every replica -> owner of the slice; Adam there)
    pull  = b2dp_all_gather            (updated slice -> every replica)

Two launches per step instead of the fused one, same result bit for bit.  Unlike the
reference's PS training this is SYNCHRONOUS (every step uses every replica's gradient; no stale
reads): the deviation is stated in DESIGN.md.  The script's `model.save(...)` (:145-148) becomes
``checkpoint.save`` with the same partitioner, which names the axis-0 pieces part_0..part_k.

    python examples/resnet50_ps_torch.py --steps 2 --batch 8 [--gpus N] [--save /tmp/ck.npz]
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200dp  # noqa: E402
from b200dp import checkpoint, harness  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--batch", type=int, default=32)          # :118
    ap.add_argument("--gpus", type=int, default=0, help="0 = all (NUM_WORKERS = NUM_PS = number of GPUs)")
    ap.add_argument("--save", default="")
    args = ap.parse_args()
    n = args.gpus or torch.cuda.device_count()
    devices = list(range(n))
    torch.backends.cudnn.allow_tf32 = True
    numels = b200dp.resnet50_numels()
    ctx = b200dp.Context.local(devices)                        # create_in_process_cluster (:31-65): one process
    plan = b200dp.Plan(ctx, b200dp.Layout(numels, nranks=n))
    steppers, data = [], []
    for r in devices:
        with torch.cuda.device(r):
            torch.manual_seed(42)
            model = harness.KerasResNet50().cuda(r).to(memory_format=torch.channels_last).train()
            steppers.append(harness.DataParallelStep(model, plan, dev=r, optimizer="adam", lr=1e-3))
            g = torch.Generator(device="cuda:%d" % r).manual_seed(7 + r)
            x = (torch.rand(args.batch, 3, 224, 224, generator=g, device="cuda:%d" % r) * 255).contiguous(
                memory_format=torch.channels_last)
            data.append((x, torch.randint(0, 1000, (args.batch,), generator=g, device="cuda:%d" % r)))
    for step in range(1, args.steps + 1):
        losses = []
        for r in devices:
            with torch.cuda.device(r):
                losses.append(steppers[r].forward_backward(*data[r]))
        h = b200dp.adam_hyper(lr=1e-3, step=step)
        for r in devices:                                      # workers hand their gradients over
            with torch.cuda.device(r):
                plan.pack(r, [p.grad for p in steppers[r].params])
        with b200dp.issue_group():
            for r in devices:                                  # PUSH: reduce on the slice owner + apply there
                with torch.cuda.device(r):
                    plan.reduce_scatter_apply(r, "adam", h)
            for r in devices:                                  # PULL: every replica fetches the fresh slices
                with torch.cuda.device(r):
                    plan.all_gather(r)
        for r in devices:
            torch.cuda.synchronize(r)
        print("step %d  loss %.4f" % (step, sum(float(l) for l in losses)))
    w0 = plan.tensor(0, b200dp.BUF_WEIGHTS)
    same = all(torch.equal(w0.cpu(), plan.tensor(r, b200dp.BUF_WEIGHTS).cpu()) for r in devices[1:])
    for r in devices:
        with torch.cuda.device(r):
            plan.check(r)
    print("workers hold identical weights:", same)
    if args.save:
        part = checkpoint.MinSizePartitioner(min_shard_bytes=256 << 10, max_shards=n)   # :75-78
        written = checkpoint.save(plan, args.save, step=args.steps, partitioner=part)
        print("Saving model to ", args.save, "(%d arrays, %d partitioned pieces)"
              % (len(written), sum("/part_" in k for k in written)))
    return 0 if same else 1


if __name__ == "__main__":
    sys.exit(main())
