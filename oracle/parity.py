"""Drives a b200dp plan through 3 optimizer steps on the seeded gradients of SURVEY.md 8d and
compares the result with the CPU oracle (TEST INFRASTRUCTURE ONLY: the checker, used by
tests/, __graft_entry__.smoke() and bench.py's untimed parity leg).

  peer path (rank-order sum):  weights on every rank and this rank's m / v shard BIT-EXACT
                               against oracle.addn + oracle.adam_ (f32 and bf16 wire);
  multicast path:              every element inside its own interval bound (oracle/bounds.py);
  NCCL (when a process group is up): all_reduce of the same gradients inside the order-free
                               bound (n-1)*2^-24*sum|g_i| around the oracle's sum.

lr schedule [1e-3, 1e-3, 1e-4] (lr is a per-call argument), Adam steps 1..3.
"""
import numpy as np
import torch

import b200dp

from . import adam_, adam_alpha, addn, scale, set_threads
from . import bounds, synth

LR_SCHEDULE = (1e-3, 1e-3, 1e-4)


def _views(flat, numels):
    out, o = [], 0
    for k in numels:
        out.append(flat[o:o + k])
        o += k
    return out


def flat_index(lay, numels):
    """Padded flat offset of every element of the unpadded tensor list."""
    parts = [np.arange(lay.tensor_offset(k), lay.tensor_offset(k) + m, dtype=np.int64) for k, m in enumerate(numels) if m]
    return np.concatenate(parts) if parts else np.zeros(0, np.int64)


def run(plan, locals_, dist_name="D1", average=False, nccl=False, threads=None, steps=3):
    """locals_: list of (dev_index_in_plan, global_rank, cuda_device) for the ranks this process drives.
    Every process of the job must call this with the same arguments (the step is collective)."""
    lay, ctx = plan.layout, plan.ctx
    numels, n = lay.numels, ctx.nranks
    N = sum(numels)
    wire_bf16 = lay.wire == b200dp.BF16
    path = plan.path if n > 1 else "single"
    if threads:
        set_threads(threads)
    post = 1.0 / n if average else 1.0
    idx = flat_index(lay, numels)
    gen_dev = torch.device("cuda", locals_[0][2])

    w0 = (torch.rand(N, generator=torch.Generator().manual_seed(42)) - 0.5) * 0.1
    for dev, _rank, cdev in locals_:
        with torch.cuda.device(cdev):
            plan.tensor(dev, b200dp.BUF_STATE0).zero_()
            plan.tensor(dev, b200dp.BUF_STATE1).zero_()
            plan.load_weights(dev, _views(w0.cuda(cdev), numels))
    w_ref, m_ref, v_ref = w0.numpy().copy(), np.zeros(N, np.float32), np.zeros(N, np.float32)
    track = bounds.AdamBound(N) if path == "multicast" else None
    nccl_ok, nccl_ratio = None, None

    for s in range(steps):
        lr = LR_SCHEDULE[s % len(LR_SCHEDULE)]
        dg = synth.all_ranks(dist_name, n, s, numels, device=gen_dev)
        hg = [g.cpu().numpy() for g in dg]
        h = b200dp.adam_hyper(lr=lr, step=s + 1, postscale=post)
        emulated = len(set(ctx.devices)) < len(ctx.devices)   # ranks share one GPU: no kernel may wait on another
        # every local rank's gradient on its own GPU BEFORE the first collective is issued: a
        # cross-device copy queued behind a kernel that waits for its peers would never run
        gl = [dg[rank] if cdev == gen_dev.index else dg[rank].to("cuda:%d" % cdev) for _dev, rank, cdev in locals_]
        for _dev, _rank, cdev in locals_:
            torch.cuda.synchronize(cdev)
        with b200dp.issue_group():
            for (dev, rank, cdev), g in zip(locals_, gl):
                with torch.cuda.device(cdev):
                    if emulated:
                        plan.pack(dev, _views(g, numels))
                    else:
                        plan.pack_step(dev, _views(g, numels), "adam", h)
        if emulated:
            for dev, rank, cdev in locals_:
                plan.step(dev, "adam", h)
        if nccl and s == 0:
            import torch.distributed as tdist
            t = dg[locals_[0][1]].clone()
            tdist.all_reduce(t)
            ref32 = addn(hg)
            ref = ref32.astype(np.float64)
            B = bounds.sum_bound_f32(hg, ref32)
            ratio = np.abs(t.cpu().numpy().astype(np.float64) - ref) / B
            nccl_ok, nccl_ratio = bool(np.all(ratio <= 1.0)), float(ratio.max())
        gsum = addn(hg, bf16_wire=wire_bf16)
        if track is not None:
            if wire_bf16:
                from .oracle_np import bf16_round
                B = bounds.sum_bound_bf16_switch([bf16_round(g) for g in hg], gsum)
            else:
                B = bounds.sum_bound_f32(hg, gsum)
            B = B * post
        if average:
            scale(gsum, post)
        m_old, v_old = (m_ref.copy(), v_ref.copy()) if track is not None else (None, None)
        adam_(w_ref, m_ref, v_ref, gsum, lr=lr, step=s + 1)
        if track is not None:
            track.step(gsum, B, m_old, v_old, m_ref, v_ref, w_ref, adam_alpha(lr, 0.9, 0.999, s + 1))
        del dg, gl

    for _dev, _rank, cdev in locals_:
        torch.cuda.synchronize(cdev)
    pad = lambda x: _padded(x, idx, lay.total_elems)          # noqa: E731
    m_pad, v_pad = pad(m_ref), pad(v_ref)
    Em_pad = pad(track.Em) if track is not None else None
    Ev_pad = pad(track.Ev) if track is not None else None
    out = {"checked": True, "path": path, "wire": "bf16" if wire_bf16 else "f32", "dist": dist_name, "steps": steps,
           "lr_schedule": list(LR_SCHEDULE[:steps]), "ranks": n, "ok": True, "max_abs_err": 0.0}
    bit_exact = True
    checks = []
    for dev, rank, cdev in locals_:
        plan.check(dev)
        got_w = plan.tensor(dev, b200dp.BUF_WEIGHTS).cpu().numpy()[idx]
        shard_m = plan.tensor(dev, b200dp.BUF_STATE0).cpu().numpy()
        shard_v = plan.tensor(dev, b200dp.BUF_STATE1).cpu().numpy()
        got_m, ref_m, got_v, ref_v, bm, bv = [], [], [], [], [], []
        for b in range(lay.num_buckets):
            info = lay.bucket_info(b)
            lo = info["offset"] + rank * info["slice"]
            sl, sh = slice(lo, lo + info["slice"]), slice(info["shard_off"], info["shard_off"] + info["slice"])
            got_m.append(shard_m[sh]); ref_m.append(m_pad[sl]); got_v.append(shard_v[sh]); ref_v.append(v_pad[sl])
            if track is not None:
                bm.append(Em_pad[sl]); bv.append(Ev_pad[sl])
        got_m, ref_m, got_v, ref_v = (np.concatenate(x) for x in (got_m, ref_m, got_v, ref_v))
        same = all(np.array_equal(a.view(np.uint32), b_.view(np.uint32))
                   for a, b_ in ((got_w, w_ref), (got_m, ref_m), (got_v, ref_v)))
        bit_exact &= same
        out["max_abs_err"] = max(out["max_abs_err"], float(np.max(np.abs(got_w.astype(np.float64) - w_ref))))
        if track is not None:
            floor = 1e-38
            checks += [bounds.check(got_w, w_ref, track.Ew, "w rank %d" % rank),
                       bounds.check(got_m, ref_m, np.concatenate(bm) + floor, "m shard rank %d" % rank),
                       bounds.check(got_v, ref_v, np.concatenate(bv) + floor, "v shard rank %d" % rank)]
    out["bit_exact"] = bool(bit_exact)
    if track is None:
        out["bound"] = "bit-exact (rank-order fp32 sum, one rounding per operation)"
        out["ok"] = bool(bit_exact)
    else:
        out["bound"] = ("interval bound propagated per element from |g^-g| <= %s through 3 Adam steps (oracle/bounds.py)"
                        % ("2^-7|sum b|+(n-1)2^-24 sum|b_i|" if wire_bf16 else "(n-1)2^-24 sum|g_i|"))
        out["ok"] = all(c["ok"] for c in checks)
        out["violations"] = sum(c["violations"] for c in checks)
        out["max_err_over_bound"] = max(c["max_err_over_bound"] for c in checks)
        wchecks = [c for c in checks if c["what"].startswith("w ")]
        out["w_median_bound"] = max(c["median_bound"] for c in wchecks)
        out["w_max_bound"] = max(c["max_bound"] for c in wchecks)
        out["w_frac_bit_equal"] = min(c["frac_bit_equal"] for c in wchecks)
    if nccl_ok is not None:
        out["nccl_within_bound"] = nccl_ok
        out["nccl_max_err_over_bound"] = nccl_ratio
        out["ok"] = out["ok"] and nccl_ok
    return out


def _padded(x, idx, total):
    p = np.zeros(total, x.dtype)
    p[idx] = x
    return p
