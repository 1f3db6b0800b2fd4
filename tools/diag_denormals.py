"""Developer probe: after 25 real training steps (the state bench.py's roofline timing sees), how
many operands of the Adam kernel's sqrt/div are denormal or zero, and what does the kernel take?"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200dp  # noqa: E402
from b200dp import harness  # noqa: E402
from investigate_n1 import time_us  # noqa: E402

torch.backends.cudnn.benchmark = True
torch.backends.cudnn.allow_tf32 = True
numels = b200dp.resnet50_numels()
ctx = b200dp.Context.local([0])
plan = b200dp.Plan(ctx, b200dp.Layout(numels, nranks=1))
model = harness.KerasResNet50().cuda().to(memory_format=torch.channels_last).train()
st = harness.DataParallelStep(model, plan)
x = (torch.rand(256, 3, 224, 224, device="cuda") * 255).contiguous(memory_format=torch.channels_last)
y = torch.randint(0, 1000, (256,), device="cuda")
TINY = 1.1754944e-38


def census(tag):
    g = torch.cat([p.grad.flatten() for p in st.params])
    m = plan.tensor(0, b200dp.BUF_STATE0)
    v = plan.tensor(0, b200dp.BUF_STATE1)
    den = lambda t: int(((t != 0) & (t.abs() < TINY)).sum())
    g2 = g * g
    print("%s: loss %.4g | g: zero %d denormal %d | g*g: zero %d denormal %d | m: zero %d denormal %d | v: zero %d denormal %d | N=%d" % (
        tag, float(st.last), int((g == 0).sum()), den(g), int((g2 == 0).sum()), den(g2), int((m == 0).sum()), den(m),
        int((v == 0).sum()), den(v), g.numel()))


t = [0]
for steps in (5, 20, 25):
    for _ in range(steps):
        st.last = st(x, y).detach()
    torch.cuda.synchronize()
    census("after +%d steps" % steps)
    ptrs = b200dp.ptr_array([p.grad for p in st.params])
    # time on a COPY of the state so the real training state is not disturbed
    w, m, v = (plan.tensor(0, b).clone() for b in (b200dp.BUF_WEIGHTS, b200dp.BUF_STATE0, b200dp.BUF_STATE1))

    def call():
        plan.pack_step(0, ptrs, "adam", b200dp.adam_hyper(step=st.t + 1))
    us = time_us(call)
    for b, saved in ((b200dp.BUF_WEIGHTS, w), (b200dp.BUF_STATE0, m), (b200dp.BUF_STATE1, v)):
        plan.tensor(0, b).copy_(saved)
    print("   fused kernel on this state: %.1f us" % us)
