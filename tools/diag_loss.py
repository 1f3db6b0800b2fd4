"""Developer probe: does the synthetic training run of bench.py stay finite?  Prints the loss
per step for a few learning rates (Keras-shaped ResNet-50, BN in inference mode as the scripts
have it, random images/labels)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200dp  # noqa: E402
from b200dp import harness  # noqa: E402

torch.backends.cudnn.benchmark = True
torch.backends.cudnn.allow_tf32 = True
numels = b200dp.resnet50_numels()
x = (torch.rand(64, 3, 224, 224, device="cuda") * 255).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 1000, (64,), device="cuda")
for lr in (1e-3, 1e-4, 1e-5):
    ctx = b200dp.Context.local([0])
    plan = b200dp.Plan(ctx, b200dp.Layout(numels, nranks=1))
    model = harness.KerasResNet50().cuda().to(memory_format=torch.channels_last).train()
    st = harness.DataParallelStep(model, plan, lr=lr)
    losses = [float(st(x, y)) for _ in range(60)]
    print("lr %g:" % lr, " ".join("%.3g" % l for l in losses[:8]), "...", " ".join("%.3g" % l for l in losses[-4:]))
    del st, model, plan, ctx
