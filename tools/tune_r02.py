"""Round-2 sweep on the GPUs of one box (local context, one process): the per-step call
(pack_step) with 1 / 2 / 4 buckets (multi-bucket calls pack bucket b+1 under bucket b's step),
the peer-path geometry (grid cap, 256-bit unroll), and the bf16 wire on both paths.
Developer tool; prints one JSON line per configuration.  TUNE_N=2,4,8 selects rank counts, TUNE_BUCKETS=1,2,4 the bucket counts (default 1)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200dp  # noqa: E402
from tune_multi import time_all, views  # noqa: E402

S = 4 * b200dp.NUM_ELEMENTS


def main():
    ndev = torch.cuda.device_count()
    want = [int(x) for x in os.environ.get("TUNE_N", "2,4,8").split(",")]
    numels = b200dp.resnet50_numels()
    for n in [x for x in want if x <= ndev]:
        for wire in ("f32", "bf16"):
            for nb in [int(x) for x in os.environ.get("TUNE_BUCKETS", "1").split(",")]:
                bucket_bytes = 0 if nb == 1 else int(S / nb * (1.02 if wire == "f32" else 0.51))
                ctx = b200dp.Context.local(list(range(n)))
                ctx.set_timeout(5.0)
                lay = b200dp.Layout(numels, nranks=n, wire=wire, bucket_bytes=bucket_bytes)
                plan = b200dp.Plan(ctx, lay)
                have_mc = True
                try:
                    plan.enable_multicast()
                except b200dp.B200dpError:
                    have_mc = False
                grads, keep = [], []
                for r in range(n):
                    with torch.cuda.device(r):
                        g = torch.randn(sum(numels), device="cuda:%d" % r) * 1e-3
                        keep.append(g)
                        grads.append(b200dp.ptr_array(views(g, numels)))
                        plan.pack(r, grads[r])
                        plan.tensor(r, b200dp.BUF_WEIGHTS).uniform_(-0.05, 0.05)
                t = [0]

                def call(r):
                    if r == 0:
                        t[0] += 1
                    plan.pack_step(r, grads[r], "adam", b200dp.adam_hyper(step=t[0] + 1))

                def step(r):
                    plan.step(r, "adam", b200dp.adam_hyper(step=t[0] + 1))

                def pack(r):
                    plan.pack(r, grads[r])

                configs = []
                if wire == "f32":
                    configs += [("peer", dict(peer_vec=8, peer_unroll=0, grid_cap=c)) for c in (148, 222, 296)]
                    configs += [("peer", dict(peer_vec=8, peer_unroll=0, grid_cap=296, skip_own=1))]   # own slice read in place
                    if n == 2:
                        configs += [("peer", dict(peer_vec=8, peer_unroll=1, grid_cap=296))]
                    if have_mc:
                        configs += [("multicast", dict(mc_unroll=1, mc_grid_cap=c)) for c in (74, 148)]
                else:
                    configs += [("peer", dict(peer_unroll=0, grid_cap=c)) for c in (148, 296)]
                    if have_mc:
                        configs += [("multicast", dict(mc_bf16_grid_cap=c)) for c in (74, 148, 296)]
                if nb > 1:   # bucket pipelining: the geometry question was answered with one bucket
                    configs = [c for c in configs if c[1].get("grid_cap", 296) == 296 and c[1].get("peer_unroll", 0) == 0
                               and c[1].get("mc_grid_cap", 74) == 74 and c[1].get("mc_bf16_grid_cap", 148) == 148]
                for path, knobs in configs:
                    plan.set_path(path)
                    plan.set_tuning(skip_own=0)
                    plan.set_tuning(**knobs)
                    us_call = time_all(n, call)
                    us_step = time_all(n, step)
                    us_pack = time_all(n, pack)
                    for r in range(n):
                        with torch.cuda.device(r):
                            plan.check(r)
                    print(json.dumps({"n": n, "wire": wire, "buckets": lay.num_buckets, "path": path, **knobs,
                                      "call_us": round(us_call, 1), "step_us": round(us_step, 1), "pack_us": round(us_pack, 1),
                                      "call_minus_step_us": round(us_call - us_step, 1),
                                      "step_busbw": round(S / us_step * 2 * (n - 1) / n / 1e3, 1)}), flush=True)
                del plan, ctx, grads, keep


if __name__ == "__main__":
    main()
