"""Developer probe: why is the fused n=1 kernel ~6% slower inside the training loop (117.6 us)
than in the hot-path-only bench (110.7 us)?  Times pack_step on (a) the autograd gradients where
they lie, (b) the same values copied into one contiguous buffer, before and after a burst of
conv work, with the SM clock sampled."""
import os
import subprocess
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200dp  # noqa: E402
from b200dp import harness  # noqa: E402


def clock():
    try:
        return subprocess.check_output(["nvidia-smi", "--query-gpu=clocks.sm,clocks.mem,power.draw", "--format=csv,noheader"],
                                       text=True).strip()
    except Exception:
        return "?"


def time_us(fn, reps=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    torch.cuda._sleep(int(2e7))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


_copy_src = None


def copy_gbs():
    """torch copy of 512 MiB -> 512 MiB (the same kind of figure as MEASURED_PEAKS.json's hbm_gbs)."""
    global _copy_src
    if _copy_src is None:
        _copy_src = (torch.empty(1 << 27, device="cuda"), torch.empty(1 << 27, device="cuda"))
    a, b = _copy_src
    us = time_us(lambda: b.copy_(a), reps=20)
    return 2 * a.numel() * 4 / us / 1e3


def main():
    torch.backends.cudnn.benchmark = True
    torch.backends.cudnn.allow_tf32 = True
    numels = b200dp.resnet50_numels()
    ctx = b200dp.Context.local([0])
    plan = b200dp.Plan(ctx, b200dp.Layout(numels, nranks=1))
    model = harness.KerasResNet50().cuda().to(memory_format=torch.channels_last).train()
    st = harness.DataParallelStep(model, plan)
    x = (torch.rand(256, 3, 224, 224, device="cuda") * 255).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 1000, (256,), device="cuda")
    t = [100]

    def mk(ptrs):
        def f():
            t[0] += 1
            plan.pack_step(0, ptrs, "adam", b200dp.adam_hyper(step=t[0]))
        return f

    flat = torch.randn(sum(numels), device="cuda") * 1e-3
    views, o = [], 0
    for k in numels:
        views.append(flat[o:o + k])
        o += k
    p_flat = b200dp.ptr_array(views)
    print("cold start, contiguous grads : %.1f us   copy %.0f GB/s   [%s]" % (time_us(mk(p_flat)), copy_gbs(), clock()))
    for _ in range(8):
        st(x, y)
    torch.cuda.synchronize()
    grads = [p.grad for p in st.params]
    p_auto = b200dp.ptr_array(grads)
    print("after 8 train steps, autograd grads in place : %.1f us   [%s]" % (time_us(mk(p_auto)), clock()))
    print("after 8 train steps, contiguous grads        : %.1f us   [%s]" % (time_us(mk(p_flat)), clock()))
    al = sorted(set(g.data_ptr() % 256 for g in grads))
    print("autograd grad pointer alignment mod 256:", al[:8], "count of distinct 2MB pages:",
          len(set(g.data_ptr() >> 21 for g in grads)))
    for rounds in range(5):
        for _ in range(8):
            st.last_loss = st(x, y).detach()
        torch.cuda.synchronize()
        gr = [p.grad for p in st.params]
        wbuf = plan.tensor(0, b200dp.BUF_WEIGHTS)
        vbuf = plan.tensor(0, b200dp.BUF_STATE1)
        gcat = torch.cat([g.flatten() for g in gr])
        tiny = 1.1754944e-38
        print("   loss %s  non-finite w: %d  g: %d   denormal v: %d  g: %d  zero g: %d" % (
            float(st.last_loss) if hasattr(st, "last_loss") else "?", int((~torch.isfinite(wbuf)).sum()),
            int((~torch.isfinite(gcat)).sum()), int(((vbuf != 0) & (vbuf.abs() < tiny)).sum()),
            int(((gcat != 0) & (gcat.abs() < tiny)).sum()), int((gcat == 0).sum())))
        print("after %d steps, autograd grads : %.1f us   contiguous: %.1f us   copy %.0f GB/s   [%s]" % (
            16 + 8 * rounds, time_us(mk(b200dp.ptr_array(gr))), time_us(mk(p_flat)), copy_gbs(), clock()))
    w_ptr, _ = plan.buffer_ptr(0, b200dp.BUF_WEIGHTS)
    print("W at", hex(w_ptr), " S0 at", hex(plan.buffer_ptr(0, b200dp.BUF_STATE0)[0]), " S1 at", hex(plan.buffer_ptr(0, b200dp.BUF_STATE1)[0]))


if __name__ == "__main__":
    main()
