"""Real multi-GPU parity (skipped on a one-GPU box; run with `gpurun --gpus 2|4|8`).

 - local context over distinct GPUs (MirroredStrategy's shape: one process, peer access,
   in-kernel flag barriers) vs the oracle, bit-exact;
 - rank contexts, one process per GPU (Horovod / MWMS shape: fd exchange over unix sockets,
   VMM import) vs the oracle bit-exact, and vs NCCL's all_reduce -- "the reference's
   reduction as actually executed"  -- under the order-free tolerance
   |err| <= (n-1) * 2^-24 * sum_i |g_i| + |oracle - exact sum| (oracle/bounds.py::sum_bound_f32:
   SURVEY's figure bounds one fp32 summation against the exact sum; the oracle's own, computed,
   error is added because two fp32 summations are being compared)."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import b200dp  # noqa: E402
import oracle  # noqa: E402

F = np.float32
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0
needs2 = pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")


def bits(a):
    return np.ascontiguousarray(a, dtype=F).view(np.uint32)


def host_grads(n, numels, step):
    """Every rank can rebuild every rank's gradients (seeded per rank, tensor, step)."""
    out = []
    for r in range(n):
        g = torch.Generator().manual_seed(1234 + 1000 * r + 100000 * step)
        x = torch.randn(sum(numels), generator=g) * 1e-3
        x[::5] += (-1.0) ** r * 50.0
        out.append(x.numpy())
    return out


def views(flat, numels):
    out, o = [], 0
    for k in numels:
        out.append(flat[o:o + k])
        o += k
    return out


@needs2
@pytest.mark.parametrize("wire", ["f32", "bf16"])
def test_local_context_distinct_gpus_bit_exact(wire):
    n = min(NGPU, 8)
    numels = b200dp.resnet50_numels()
    N = sum(numels)
    ctx = b200dp.Context.local(list(range(n)))
    ctx.set_timeout(5.0)
    lay = b200dp.Layout(numels, nranks=n, wire=wire)
    plan = b200dp.Plan(ctx, lay)
    w0 = (torch.rand(N, generator=torch.Generator().manual_seed(42)) - 0.5) * 0.1
    for r in range(n):
        with torch.cuda.device(r):
            plan.load_weights(r, views(w0.cuda(r), numels))
    w_ref, m_ref, v_ref = w0.numpy().copy(), np.zeros(N, F), np.zeros(N, F)
    for s in range(2):
        hg = host_grads(n, numels, s)
        dg = [torch.from_numpy(hg[r]).cuda(r) for r in range(n)]
        for r in range(n):
            with torch.cuda.device(r):
                plan.pack(r, views(dg[r], numels))
        with b200dp.issue_group():
            for r in range(n):      # all ranks' kernels are in flight together; each waits for its peers
                with torch.cuda.device(r):
                    plan.step(r, "adam", b200dp.adam_hyper(step=s + 1))
        for r in range(n):
            torch.cuda.synchronize(r)
        oracle.adam_(w_ref, m_ref, v_ref, oracle.addn(hg, bf16_wire=(wire == "bf16")), step=s + 1)
    for r in range(n):
        with torch.cuda.device(r):
            plan.check(r)
            got = plan.tensor(r, b200dp.BUF_WEIGHTS).cpu().numpy()
        o = 0
        for k, m in enumerate(numels):
            off = lay.tensor_offset(k)
            assert np.array_equal(bits(got[off:off + m]), bits(w_ref[o:o + m])), "rank %d tensor %d" % (r, k)
            o += m


def _rank_worker(rank, n, port, wire, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(n))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=n, device_id=torch.device("cuda", rank))
    try:
        numels = b200dp.resnet50_numels()
        N = sum(numels)
        ctx = b200dp.Context.rank(rank, n, rank)
        ctx.set_timeout(5.0)
        lay = b200dp.Layout(numels, nranks=n, wire=wire)
        plan = b200dp.Plan(ctx, lay)
        plan.exchange("t-%d" % port)
        assert plan.ready
        dist.barrier()
        w0 = (torch.rand(N, generator=torch.Generator().manual_seed(42)) - 0.5) * 0.1
        plan.load_weights(0, views(w0.cuda(), numels))
        w_ref, m_ref, v_ref = w0.numpy().copy(), np.zeros(N, F), np.zeros(N, F)
        ok_w = ok_ar = ok_nccl = True
        for s in range(2):
            hg = host_grads(n, numels, s)
            mine = torch.from_numpy(hg[rank]).cuda()
            gsum = oracle.addn(hg, bf16_wire=(wire == "bf16"))
            # hvd-shaped op: all-reduce only
            plan.pack(0, views(mine, numels))
            plan.allreduce(0)
            torch.cuda.synchronize()
            red = plan.tensor(0, b200dp.BUF_REDUCED).cpu().numpy()
            flat_red = np.concatenate([red[lay.tensor_offset(k):lay.tensor_offset(k) + m] for k, m in enumerate(numels)])
            ok_ar &= bool(np.array_equal(bits(flat_red), bits(gsum)))
            if wire == "f32":   # NCCL on the same inputs: order-free tolerance
                t = mine.clone()
                dist.all_reduce(t)
                from oracle import bounds
                ok_nccl &= bool(np.all(np.abs(t.cpu().numpy().astype(np.float64) - gsum) <= bounds.sum_bound_f32(hg, gsum)))
            # fused step (the all-reduce above ran in place, so pack again)
            plan.pack(0, views(mine, numels))
            plan.step(0, "adam", b200dp.adam_hyper(step=s + 1))
            oracle.adam_(w_ref, m_ref, v_ref, gsum, step=s + 1)
        torch.cuda.synchronize()
        plan.check(0)
        got = plan.tensor(0, b200dp.BUF_WEIGHTS).cpu().numpy()
        flat_w = np.concatenate([got[lay.tensor_offset(k):lay.tensor_offset(k) + m] for k, m in enumerate(numels)])
        ok_w = bool(np.array_equal(bits(flat_w), bits(w_ref)))
        # step-0 / per-epoch helpers: broadcast from the last rank, small all-reduce (Average)
        plan.tensor(0, b200dp.BUF_WEIGHTS).fill_(float(rank))
        torch.cuda.synchronize()
        dist.barrier()
        plan.broadcast(0, root=n - 1)
        vals = [np.array([r + 0.25, -r, 1e-3 * r], F) for r in range(n)]
        out = torch.zeros(3, device="cuda")
        plan.allreduce_small(0, torch.from_numpy(vals[rank]).cuda(), out, postscale=1.0 / n)
        torch.cuda.synchronize()
        plan.check(0)
        ok_w &= bool(torch.all(plan.tensor(0, b200dp.BUF_WEIGHTS) == float(n - 1)))
        ok_w &= bool(np.array_equal(bits(out.cpu().numpy()), bits(oracle.scale(oracle.addn(vals), 1.0 / n))))
        dist.barrier()
        q.put((rank, ok_w, ok_ar, ok_nccl, ""))
    except Exception as e:  # report instead of hanging the parent
        import traceback
        q.put((rank, False, False, False, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@needs2
@pytest.mark.parametrize("wire", ["f32", "bf16"])
def test_rank_contexts_one_process_per_gpu(wire):
    import torch.multiprocessing as mp
    n = min(NGPU, 8)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_worker, args=(r, n, port, wire, q)) for r in range(n)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(n)]
    for p in procs:
        p.join(60)
    for rank, ok_w, ok_ar, ok_nccl, err in sorted(results):
        assert not err, err
        assert ok_ar, "rank %d: all-reduce differs from the oracle" % rank
        assert ok_w, "rank %d: weights after 2 fused steps differ from the oracle" % rank
        assert ok_nccl, "rank %d: NCCL all_reduce outside the order-free bound" % rank


# ------------------------------------------------------------------ Path B (multicast)
# The switch adds the n copies in its own order, so Path B is held to the order-free bound
#   |sum_hat - sum| <= (n-1) * 2^-24 * sum_i |g_i| + |oracle - exact|   ("Path B / NCCL-ordered",
#                                                                        plus the oracle's own computed error)
# for the reduced gradient, and to tolerances derived from it for the updated weights:
#   SGD (linear in g):   |dw| <= lr * bound + 2 ulp(w)
#   Adam (3 steps, lr schedule, D1 and D3 gradients of SURVEY 8d -- any signs, zeros, 2^-24..2^8):
#   w on every rank and the m / v shards inside the interval bound that oracle/bounds.py
#   propagates ELEMENT BY ELEMENT from the bound on the sum through TF's update formula.  No
#   element is excluded; where sqrt(v) ~ eps or the sum sits near zero the bound is simply wide.
#   The check also demands that the bound is informative (median a few ulp of |w| <= 0.05).


def _flat(plan, lay, numels, which, dev=0):
    buf = plan.tensor(dev, which).cpu().numpy()
    return np.concatenate([buf[lay.tensor_offset(k):lay.tensor_offset(k) + m] for k, m in enumerate(numels)])


def _path_b_checks(plan, lay, n, numels, local_devs, ops, sync, wire="f32"):
    """Shared by the local and the rank-context tests.  ops.put(list_of_host_arrays) packs
    every local rank's gradient, ops.allreduce()/load(w)/step(opt, h) issue the call for every
    local rank; sync() waits for the device(s)."""
    N = sum(numels)
    assert plan.path == "multicast"
    # (1) all-reduce on cancellation-stress data: order-free bound, every element
    hg = host_grads(n, numels, 0)
    ops.put(hg)
    ops.allreduce()
    sync()
    from oracle import bounds, oracle_np
    if wire == "bf16":
        # the switch adds the bf16 copies in fp32 and rounds the sum once more to bf16
        ref32 = oracle.addn(hg, bf16_wire=True)
        rb = [oracle_np.bf16_round(g) for g in hg]
        bound = bounds.sum_bound_bf16_switch(rb, ref32)
        if os.environ.get("B2DP_REPORT_SWITCH_ROUNDING"):       # how the switch rounds its fp32 sum to bf16, measured
            for dev in local_devs:
                got = _flat(plan, lay, numels, b200dp.BUF_REDUCED, dev).astype(np.float64)
                nz = np.abs(ref32) > 0
                rel = np.abs(got - ref32)[nz] / np.abs(ref32[nz].astype(np.float64))
                print("SWITCH_BF16_ROUNDING n=%d max_rel_err=%.6g (2^-8=%.6g 2^-7=%.6g) frac_over_2^-8=%.3g frac_on_bf16_grid=%.4f"
                      % (n, rel.max(), 2.0 ** -8, 2.0 ** -7, float((rel > 2.0 ** -8 * 1.0001).mean()),
                         float(np.mean(oracle_np.bf16_round(got.astype(F)) == got.astype(F)))), flush=True)
        # vs the fp32 oracle: every g_i moves by <= 2^-8|g_i| when cast, the switch's result by <= 2^-7|sum|
        bound_vs_fp32 = 3 * 2.0 ** -8 * np.sum(np.abs(np.stack(hg)), axis=0) + 1e-30
        ref_fp32 = oracle.addn(hg).astype(np.float64)
    else:
        ref32 = oracle.addn(hg)
        bound = bounds.sum_bound_f32(hg, ref32)     # any-order fp32 sum vs the exact sum + the oracle's own (computed) error
    ref = ref32.astype(np.float64)
    for dev in local_devs:
        got = _flat(plan, lay, numels, b200dp.BUF_REDUCED, dev).astype(np.float64)
        assert np.all(np.abs(got - ref) <= bound), "all-reduce outside the order-free bound"
        if wire == "bf16":
            assert np.all(np.abs(got - ref_fp32) <= bound_vs_fp32), "bf16 gradient mode outside 3 * 2^-8 * sum|g_i| of the fp32 oracle"
    # (2) fused SGD step on the same data: linear error bound
    w0 = ((torch.rand(N, generator=torch.Generator().manual_seed(42)) - 0.5) * 0.1).numpy()
    ops.load(w0)
    ops.put(hg)
    lr = 0.01
    ops.step("sgdm", b200dp.sgdm_hyper(lr=lr, momentum=0.9))
    sync()
    w_ref, a_ref = w0.copy(), np.zeros(N, F)
    oracle.sgdm_(w_ref, a_ref, ref32.copy(), lr=lr, mu=0.9)
    for dev in local_devs:
        got = _flat(plan, lay, numels, b200dp.BUF_WEIGHTS, dev)
        assert np.all(np.abs(got.astype(np.float64) - w_ref) <= lr * bound + 2 * np.spacing(np.abs(w_ref).astype(F)))
    # (2b) the two-launch form (PS push / pull) through the multicast mapping: same bound
    for dev in local_devs:
        plan.tensor(dev, b200dp.BUF_STATE0).zero_()
    sync()
    ops.load(w0)
    ops.put(hg)
    ops.two_launch("sgdm", b200dp.sgdm_hyper(lr=lr, momentum=0.9))
    sync()
    for dev in local_devs:
        got = _flat(plan, lay, numels, b200dp.BUF_WEIGHTS, dev)
        assert np.all(np.abs(got.astype(np.float64) - w_ref) <= lr * bound + 2 * np.spacing(np.abs(w_ref).astype(F)))
    # (3) fused Adam, three steps with the lr schedule, D1 and D3 data: per-element interval bound
    from oracle import parity
    for dist_name in ("D1", "D3"):
        res = parity.run(plan, ops.locals, dist_name=dist_name)
        assert res["ok"], res
        assert res["violations"] == 0 and res["max_err_over_bound"] <= 1.0
        if wire == "f32":
            assert res["w_median_bound"] <= 1e-7, res  # the bound says something: a few ulp of |w|
        else:                                          # a bf16 wire moves the gradient by up to 2^-7 of itself
            assert res["w_median_bound"] <= 1e-4, res


class _Ops:
    pass


@needs2
@pytest.mark.parametrize("wire", ["f32", "bf16"])
def test_path_b_multicast_local_context(wire):
    n = min(NGPU, 8)
    numels = b200dp.resnet50_numels()
    ctx = b200dp.Context.local(list(range(n)))
    ctx.set_timeout(5.0)
    lay = b200dp.Layout(numels, nranks=n, wire=wire)
    plan = b200dp.Plan(ctx, lay)
    plan.enable_multicast()

    def each(fn):
        with b200dp.issue_group():
            for r in range(n):
                with torch.cuda.device(r):
                    fn(r)

    ops = _Ops()
    ops.put = lambda hg: each(lambda r: plan.pack(r, views(torch.from_numpy(hg[r]).cuda(r), numels)))
    ops.allreduce = lambda: each(lambda r: plan.allreduce(r))
    ops.load = lambda w: each(lambda r: plan.load_weights(r, views(torch.from_numpy(w).cuda(r), numels)))
    ops.step = lambda opt, h: each(lambda r: plan.step(r, opt, h))
    ops.two_launch = lambda opt, h: (each(lambda r: plan.reduce_scatter_apply(r, opt, h)), each(lambda r: plan.all_gather(r)))
    ops.locals = [(r, r, r) for r in range(n)]

    def sync():
        for r in range(n):
            torch.cuda.synchronize(r)

    _path_b_checks(plan, lay, n, numels, list(range(n)), ops, sync, wire)
    each(lambda r: plan.check(r))


def _rank_worker_mc(rank, n, port, q, wire="f32"):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(n))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=n, device_id=torch.device("cuda", rank))
    try:
        numels = b200dp.resnet50_numels()
        ctx = b200dp.Context.rank(rank, n, rank)
        ctx.set_timeout(5.0)
        lay = b200dp.Layout(numels, nranks=n, wire=wire)
        plan = b200dp.Plan(ctx, lay)
        plan.exchange("m-%d" % port)
        plan.enable_multicast("m-%d" % port)
        dist.barrier()
        ops = _Ops()
        ops.put = lambda hg: plan.pack(0, views(torch.from_numpy(hg[rank]).cuda(), numels))
        ops.allreduce = lambda: plan.allreduce(0)
        ops.load = lambda w: plan.load_weights(0, views(torch.from_numpy(w).cuda(), numels))
        ops.step = lambda opt, h: plan.step(0, opt, h)
        ops.two_launch = lambda opt, h: (plan.reduce_scatter_apply(0, opt, h), plan.all_gather(0))
        ops.locals = [(0, rank, rank)]
        _path_b_checks(plan, lay, n, numels, [0], ops, torch.cuda.synchronize, wire)
        plan.check(0)
        dist.barrier()
        q.put((rank, ""))
    except Exception:
        import traceback
        q.put((rank, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@needs2
@pytest.mark.parametrize("wire", ["f32", "bf16"])
def test_path_b_multicast_rank_contexts(wire):
    import torch.multiprocessing as mp
    n = min(NGPU, 8)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_rank_worker_mc, args=(r, n, port, q, wire)) for r in range(n)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(n)]
    for p in procs:
        p.join(60)
    for rank, err in sorted(results):
        assert not err, "rank %d:\n%s" % (rank, err)


# ------------------------------------------------------------------ hvd-shaped module, n ranks
def _hvd_worker(rank, n, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(n),
                      LOCAL_RANK=str(rank))
    try:
        from b200dp import hvd
        hvd.init()
        assert (hvd.rank(), hvd.size(), hvd.local_rank()) == (rank, n, rank)
        shapes = [(64, 3, 7, 7), (64,), (1000, 2048), (1000,), (3,)]
        numels = [int(np.prod(s)) for s in shapes]
        hg = host_grads(n, numels, 3)
        mine = [t.view(s) for t, s in zip(views(torch.from_numpy(hg[rank]).cuda(), numels), shapes)]
        # hvd.allreduce(op=Average) on a fused group == oracle sum * float(1/n)
        outs = hvd.grouped_allreduce(mine, op=hvd.Average)
        torch.cuda.synchronize()
        ref = oracle.scale(oracle.addn(hg), 1.0 / n)
        got = np.concatenate([o.cpu().numpy().ravel() for o in outs])
        ok_avg = bool(np.array_equal(bits(got), bits(ref)))
        # fused Adam with Average
        rng = np.random.default_rng(5)
        w0 = [rng.standard_normal(s).astype(F) * 0.05 for s in shapes]
        params = [torch.nn.Parameter(torch.from_numpy(x).cuda()) for x in w0]
        # hvd.broadcast_parameters goes through b2dp_broadcast: ranks that start different end equal
        # to the root, bit for bit (non-zero root)
        scratch = [torch.nn.Parameter(torch.from_numpy(x).cuda() + float(rank)) for x in w0]
        hvd.broadcast_parameters(scratch, root_rank=n - 1)
        torch.cuda.synchronize()
        ok_bcast = all(np.array_equal(bits(p.detach().cpu().numpy()), bits(x + F(n - 1))) for p, x in zip(scratch, w0))
        opt = hvd.FusedDistributedAdam(params, lr=1e-3)
        # the three callbacks of imagenet-resnet50-hvd.py:111-115, driven the way Keras drives them
        cbs = [hvd.callbacks.BroadcastGlobalVariablesCallback(0), hvd.callbacks.MetricAverageCallback(),
               hvd.callbacks.LearningRateWarmupCallback(1e-3 * n, warmup_epochs=1, steps_per_epoch=4)]
        for cb in cbs:
            cb.set_optimizer(opt)
            cb.on_train_begin()
        cbs[2].on_epoch_begin(0)
        cbs[2].on_batch_begin(2)
        ok_bcast &= abs(opt.lr - 1e-3 * n / n * (0.5 * (n - 1) + 1)) < 1e-12
        logs = {"loss": 1.0 + rank, "accuracy": 0.125 * rank, "note": "text"}
        cbs[1].on_epoch_end(0, logs)
        want = oracle.scale(oracle.addn([np.array([0.125 * r, 1.0 + r], F) for r in range(n)]), 1.0 / n)
        ok_bcast &= bool(np.array_equal(bits(np.array([logs["accuracy"], logs["loss"]], F)), bits(want))) and logs["note"] == "text"
        opt.lr = 1e-3
        w = np.concatenate([x.ravel() for x in w0])
        m, v = np.zeros_like(w), np.zeros_like(w)
        for step in (1, 2):
            hg = host_grads(n, numels, 10 + step)
            for p, g, s in zip(params, views(torch.from_numpy(hg[rank]).cuda(), numels), shapes):
                p.grad = g.view(s)
            opt.step()
            oracle.adam_(w, m, v, oracle.scale(oracle.addn(hg), 1.0 / n), lr=1e-3, step=step)
        torch.cuda.synchronize()
        got = np.concatenate([p.detach().cpu().numpy().ravel() for p in params])
        ok_adam = bool(np.array_equal(bits(got), bits(w)))
        import torch.distributed as dist
        dist.barrier()
        q.put((rank, ok_avg, ok_adam and ok_bcast, ""))
        dist.destroy_process_group()
    except Exception:
        import traceback
        q.put((rank, False, False, traceback.format_exc()))


@needs2
def test_hvd_module_one_process_per_gpu():
    import torch.multiprocessing as mp
    n = min(NGPU, 8)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_hvd_worker, args=(r, n, port, q)) for r in range(n)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(n)]
    for p in procs:
        p.join(60)
    for rank, ok_avg, ok_adam, err in sorted(results):
        assert not err, err
        assert ok_avg, "rank %d: hvd Average differs from the oracle" % rank
        assert ok_adam, "rank %d: fused Adam differs from the oracle" % rank


# ------------------------------------------------------------------ MWMS-shaped CrossDeviceOps, n ranks
def _cdo_worker(rank, n, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(n))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=n, device_id=torch.device("cuda", rank))
    try:
        from b200dp import cross_device_ops as cdo
        ops = cdo.B200CrossDeviceOps(rank=rank, nranks=n, device=rank)     # one worker of MultiWorkerMirroredStrategy
        shapes = [s for _, s in b200dp.resnet50_trainable_variables()]
        numels = [int(np.prod(s)) for s in shapes]
        hg = host_grads(n, numels, 5)
        mine = [t.view(s) for t, s in zip(views(torch.from_numpy(hg[rank]).cuda(), numels), shapes)]
        opts = cdo.CommunicationOptions(bytes_per_pack=32 << 20, implementation="NCCL")
        ok = True
        for op, ref in ((cdo.ReduceOp.SUM, oracle.addn(hg)), (cdo.ReduceOp.MEAN, oracle.div_n(oracle.addn(hg), n))):
            out = ops.batch_reduce(op, [([t], None) for t in mine], opts)
            torch.cuda.synchronize()
            got = np.concatenate([o[0].cpu().numpy().ravel() for o in out])
            ok &= bool(np.array_equal(bits(got), bits(ref)))
        # replica-context hook of newer TF, tensor and list forms
        one = ops._all_reduce(cdo.ReduceOp.SUM, mine[0], 0, None)
        two = ops._all_reduce(cdo.ReduceOp.SUM, [mine[1], mine[2]], 0, None)
        torch.cuda.synchronize()
        ref = oracle.addn(hg)
        ok &= bool(np.array_equal(bits(one.cpu().numpy().ravel()), bits(ref[:numels[0]])))
        ok &= bool(np.array_equal(bits(two[1].cpu().numpy().ravel()), bits(ref[numels[0] + numels[1]:numels[0] + numels[1] + numels[2]])))
        gathered = ops.gather([torch.full((3, 2), float(rank), device="cuda")], axis=1)[0]
        torch.cuda.synchronize()
        ok &= gathered.shape == (3, 2 * n) and gathered[0].tolist() == [float(r) for r in range(n) for _ in range(2)]
        dist.barrier()
        q.put((rank, ok, ""))
    except Exception:
        import traceback
        q.put((rank, False, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@needs2
def test_rank_context_cross_device_ops_one_process_per_gpu():
    """B200CrossDeviceOps(rank=, nranks=, device=): the drop-in for MWMS' _cross_device_ops
    (imagenet-resnet50-multiworkers.py:20-25, fit :95), bit-exact against AddN / AddN/n."""
    import torch.multiprocessing as mp
    n = min(NGPU, 8)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_cdo_worker, args=(r, n, port, q)) for r in range(n)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(n)]
    for p in procs:
        p.join(60)
    for rank, ok, err in sorted(results):
        assert not err, err
        assert ok, "rank %d: CrossDeviceOps result differs from the oracle" % rank


# ------------------------------------------------------------------ 8f-1: bucket overlap, n ranks
def _overlap_worker(rank, n, port, q):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(n))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=n, device_id=torch.device("cuda", rank))
    try:
        from b200dp import harness
        from test_gpu_glue import _TinyNet
        torch.backends.cuda.matmul.allow_tf32 = False
        results = []
        for tag, overlap, bucket_bytes in (("serial", False, 0), ("overlap", True, 600_000)):
            torch.manual_seed(0)                      # same initial weights on every rank
            model = _TinyNet().cuda()
            numels = [p.numel() for p in model.keras_order_params()]
            ctx = b200dp.Context.rank(rank, n, rank)
            ctx.set_timeout(5.0)
            lay = b200dp.Layout(numels, nranks=n, bucket_bytes=bucket_bytes)
            plan = b200dp.Plan(ctx, lay)
            plan.exchange("ov-%d-%s" % (port, tag))
            dist.barrier()
            stepper = harness.DataParallelStep(model, plan, overlap=overlap)
            g = torch.Generator(device="cuda").manual_seed(100 + rank)     # different data per rank
            for _ in range(3):
                x = torch.randn(64, 256, generator=g, device="cuda")
                y = torch.randint(0, 10, (64,), generator=g, device="cuda")
                stepper(x, y)
            torch.cuda.synchronize()
            plan.check(0)
            results.append(torch.cat([p.detach().flatten() for p in model.keras_order_params()]).clone())
            dist.barrier()
        same_as_serial = bool(torch.equal(results[0].view(torch.int32), results[1].view(torch.int32)))
        gathered = [torch.empty_like(results[1]) for _ in range(n)]
        dist.all_gather(gathered, results[1])
        same_on_all = all(bool(torch.equal(gathered[0], t)) for t in gathered)
        q.put((rank, same_as_serial, same_on_all, ""))
    except Exception:
        import traceback
        q.put((rank, False, False, traceback.format_exc()))
    finally:
        dist.destroy_process_group()


@needs2
def test_bucket_overlap_with_backward_rank_contexts():
    import torch.multiprocessing as mp
    n = min(NGPU, 8)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_overlap_worker, args=(r, n, port, q)) for r in range(n)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(n)]
    for p in procs:
        p.join(60)
    for rank, same_as_serial, same_on_all, err in sorted(results):
        assert not err, err
        assert same_as_serial, "rank %d: overlapped bucket pipeline differs from the serial step" % rank
        assert same_on_all, "ranks ended with different weights"


# ------------------------------------------------------------------ failure detection
@needs2
def test_missing_peer_times_out_instead_of_hanging():
    """If a rank never reaches a collective, the waiting kernel gives up after the context's
    timeout and b2dp_plan_check reports B2DP_ERR_TIMEOUT -- the GPU is never left spinning."""
    import time
    ctx = b200dp.Context.local([0, 1])
    ctx.set_timeout(0.5)
    lay = b200dp.Layout([4096], nranks=2)
    plan = b200dp.Plan(ctx, lay)
    with torch.cuda.device(0):
        plan.pack(0, [torch.ones(4096, device="cuda:0")])
        t0 = time.time()
        plan.step(0, "adam", b200dp.adam_hyper(step=1))      # rank 1 never calls step
        with pytest.raises(b200dp.B200dpError) as e:
            plan.check(0)
        assert e.value.code == b200dp._lib.ERR_TIMEOUT
        assert time.time() - t0 < 20
        plan.check(0)                                        # the status word was cleared


# ------------------------------------------------------------------ own slice read in place
@needs2
@pytest.mark.parametrize("bucket_bytes", [0, 300_000])
def test_pack_step_reads_the_own_slice_in_place_bit_identically(bucket_bytes):
    """Peer path, f32 wire: b2dp_pack_step leaves rank r's own slice of every bucket unpacked and the
    step kernel reads it from the gradient tensors through the tile table.  Ragged, empty and
    misaligned tensors, a prescale != 1, one and several buckets: same bits as packing everything
    (skip_own = 0) and as the oracle."""
    n = 8 if NGPU >= 8 else (4 if NGPU >= 4 else 2)
    numels = [40000, 7, 0, 1027, 64, 333, 147456, 5, 2048, 70001]
    N = sum(numels)
    rng = np.random.default_rng(17)
    w0 = (rng.standard_normal(N) * 0.05).astype(F)
    pre = 0.5
    results = []
    for skip in (1, 0):
        ctx = b200dp.Context.local(list(range(n)))
        ctx.set_timeout(10.0)
        lay = b200dp.Layout(numels, nranks=n, bucket_bytes=bucket_bytes)
        plan = b200dp.Plan(ctx, lay)
        plan.set_tuning(skip_own=skip)
        for r in range(n):
            with torch.cuda.device(r):
                plan.load_weights(r, views(torch.from_numpy(w0).cuda(r), numels))
        w_ref, m_ref, v_ref = w0.copy(), np.zeros(N, F), np.zeros(N, F)
        g_rng = np.random.default_rng(23)
        for s in range(3):
            hg = [(g_rng.standard_normal(N) * 1e-2).astype(F) for _ in range(n)]
            # one float of slack in front: every tensor pointer is only 4-byte aligned
            dg = [torch.from_numpy(np.concatenate([[0.0], g]).astype(F)).cuda(r)[1:] for r, g in enumerate(hg)]
            for r in range(n):
                torch.cuda.synchronize(r)
            h = b200dp.adam_hyper(step=s + 1, want_stats=True)
            with b200dp.issue_group():
                for r in range(n):
                    with torch.cuda.device(r):
                        plan.pack_step(r, views(dg[r], numels), "adam", h, prescale=pre)
            oracle.adam_(w_ref, m_ref, v_ref, oracle.addn(hg, prescale=pre), step=s + 1)
            for r in range(n):
                torch.cuda.synchronize(r)
        got = []
        for r in range(n):
            with torch.cuda.device(r):
                plan.check(r)
            got.append(_flat(plan, lay, numels, b200dp.BUF_WEIGHTS, r))
            assert np.array_equal(bits(got[-1]), bits(w_ref)), "skip_own=%d rank %d differs from the oracle" % (skip, r)
        stats = sum(float(plan.tensor(r, b200dp.BUF_STATS).cpu()[0]) for r in range(n))
        want = oracle.sumsq(oracle.addn(hg, prescale=pre))
        assert abs(stats - want) <= 1e-6 * want
        results.append(got[0])
    assert np.array_equal(bits(results[0]), bits(results[1]))
