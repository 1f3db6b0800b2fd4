"""Runs the n=8 collective kernels as EMULATED ranks on one GPU (every rank's block on cuda:0,
flag waits off), so that ncu -- which must never wrap a multi-rank command -- can still capture
step_kernel<8,...>, the all-reduce variant, gather_kernel and unpack.  On one GPU the peer
loads are local, so these captures show the kernels' HBM behaviour, not NVLink."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200dp  # noqa: E402


def main():
    n = 8
    numels = b200dp.resnet50_numels()
    ctx = b200dp.Context.local([0] * n)
    lay = b200dp.Layout(numels, nranks=n)
    plan = b200dp.Plan(ctx, lay)
    g = torch.randn(sum(numels), device="cuda") * 1e-3
    views, o = [], 0
    for k in numels:
        views.append(g[o:o + k])
        o += k
    outs = [torch.empty(k, device="cuda") for k in numels]
    for it in range(3):
        for r in range(n):
            plan.pack(r, views)
        for r in range(n):
            plan.step(r, "adam", b200dp.adam_hyper(step=it + 1, want_stats=True))
        for r in range(n):
            plan.allreduce(r, postscale=1.0 / n)
        for r in range(n):
            plan.reduce_scatter_apply(r, "sgdm", b200dp.sgdm_hyper())
        for r in range(n):
            plan.all_gather(r)
        plan.unpack(0, outs)
    torch.cuda.synchronize()
    print("ok")


if __name__ == "__main__":
    main()
