"""Single-process sweep of the collective kernels' launch geometry over 2/4/8 GPUs of one box
(local context = MirroredStrategy's shape).  Developer tool; prints one line per config.
Times are CUDA-event device times, max over ranks, every launch queued behind a device-side
sleep so host launch overhead is excluded."""
import itertools
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200dp  # noqa: E402

S = 4 * b200dp.NUM_ELEMENTS


def views(flat, numels):
    out, o = [], 0
    for k in numels:
        out.append(flat[o:o + k])
        o += k
    return out


def time_all(n, fn, reps=20):
    for _ in range(3):
        for r in range(n):
            with torch.cuda.device(r):
                fn(r)
    for r in range(n):
        torch.cuda.synchronize(r)
    ev = []
    for r in range(n):
        with torch.cuda.device(r):
            torch.cuda._sleep(int(4e7))
            e0 = torch.cuda.Event(enable_timing=True)
            e0.record()
            ev.append([e0, None])
    for _ in range(reps):
        for r in range(n):
            with torch.cuda.device(r):
                fn(r)
    for r in range(n):
        with torch.cuda.device(r):
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            ev[r][1] = e1
    for r in range(n):
        torch.cuda.synchronize(r)
    return max(e0.elapsed_time(e1) for e0, e1 in ev) / reps * 1e3   # us


def main():
    ndev = torch.cuda.device_count()
    numels = b200dp.resnet50_numels()
    results = []
    for n in [x for x in (2, 4, 8) if x <= ndev]:
        ctx = b200dp.Context.local(list(range(n)))
        ctx.set_timeout(5.0)
        lay = b200dp.Layout(numels, nranks=n)
        plan = b200dp.Plan(ctx, lay)
        plan.enable_multicast()
        for r in range(n):
            with torch.cuda.device(r):
                g = torch.randn(sum(numels), device="cuda:%d" % r) * 1e-3
                plan.pack(r, views(g, numels))
                plan.tensor(r, b200dp.BUF_WEIGHTS).uniform_(-0.05, 0.05)
        t = [0]

        def step(r):
            if r == 0:
                t[0] += 1
            plan.step(r, "adam", b200dp.adam_hyper(step=t[0] + 1))

        def allreduce(r):
            plan.allreduce(r, postscale=1.0 / n)

        def rs_only(r):
            plan.reduce_scatter_apply(r, "adam", b200dp.adam_hyper(step=t[0] + 1))

        def ag_only(r):
            plan.all_gather(r)

        if os.environ.get("TUNE_FULL"):
            configs = [("multicast", dict(mc_unroll=u, mc_grid_cap=c)) for u, c in itertools.product((1, 2, 4), (37, 74, 148, 296))]
            configs += [("peer", dict(peer_unroll=u, grid_cap=c)) for u, c in itertools.product((1, 2), (74, 148, 296, 592))]
        else:   # the chosen geometry only, plus the two halves of the step on their own
            configs = [("multicast", dict(mc_unroll=1, mc_grid_cap=74)), ("peer", dict(peer_unroll=0, grid_cap=296, peer_vec=4))]
            configs += [("peer", dict(peer_unroll=0, grid_cap=c, peer_vec=8)) for c in (148, 296)]
        for path, knobs in configs:
            plan.set_path(path)
            plan.set_tuning(**knobs)
            us_step = time_all(n, step)
            us_ar = time_all(n, allreduce)
            for r in range(n):
                with torch.cuda.device(r):
                    plan.check(r)
            us_rs = time_all(n, rs_only)
            us_ag = time_all(n, ag_only)
            row = {"n": n, "path": path, **knobs, "step_us": round(us_step, 1), "allreduce_us": round(us_ar, 1),
                   "reduce_scatter_apply_us": round(us_rs, 1), "all_gather_us": round(us_ag, 1),
                   "step_busbw": round(S / us_step * 2 * (n - 1) / n / 1e3, 1)}
            results.append(row)
            print(json.dumps(row), flush=True)
        del plan, ctx
    # Can the two directions overlap better than the fused kernel manages?  Run the two halves
    # CONCURRENTLY (two plans, so each has its own flag words; two streams per device).
    if os.environ.get("TUNE_CONCURRENT", "0") == "1":
        for n in [x for x in (8,) if x <= ndev]:
            ctx = b200dp.Context.local(list(range(n)))
            ctx.set_timeout(5.0)
            lay = b200dp.Layout(numels, nranks=n)
            plans = [b200dp.Plan(ctx, lay), b200dp.Plan(ctx, lay)]
            for pl in plans:
                pl.enable_multicast()
            s1 = [torch.cuda.Stream(device=r) for r in range(n)]
            s2 = [torch.cuda.Stream(device=r) for r in range(n)]
            for r in range(n):
                with torch.cuda.device(r):
                    g = torch.randn(sum(numels), device="cuda:%d" % r) * 1e-3
                    for pl in plans:
                        pl.pack(r, views(g, numels))
                        pl.tensor(r, b200dp.BUF_WEIGHTS).uniform_(-0.05, 0.05)
            for path, knobs in (("multicast", dict(mc_unroll=1, mc_grid_cap=74)), ("multicast", dict(mc_unroll=1, mc_grid_cap=37)),
                                ("peer", dict(peer_unroll=0, grid_cap=148)), ("peer", dict(peer_unroll=0, grid_cap=296))):
                for pl in plans:
                    pl.set_path(path)
                    pl.set_tuning(**knobs)

                def both(r):
                    cur = torch.cuda.current_stream()
                    ev = torch.cuda.Event()
                    ev.record(cur)
                    s1[r].wait_event(ev)
                    s2[r].wait_event(ev)
                    plans[0].reduce_scatter_apply(r, "adam", b200dp.adam_hyper(step=5), stream=s1[r])
                    plans[1].all_gather(r, stream=s2[r])
                    cur.wait_stream(s1[r])
                    cur.wait_stream(s2[r])

                us = time_all(n, both)
                for r in range(n):
                    with torch.cuda.device(r):
                        plans[0].check(r)
                        plans[1].check(r)
                print(json.dumps({"n": n, "path": path, **knobs, "concurrent_rs_and_ag_us": round(us, 1)}), flush=True)
            del plans, ctx
    best = {}
    for row in results:
        k = (row["n"], row["path"])
        if k not in best or row["step_us"] < best[k]["step_us"]:
            best[k] = row
    print("BEST")
    for k in sorted(best):
        print(json.dumps(best[k]))


if __name__ == "__main__":
    main()
