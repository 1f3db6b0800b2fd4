"""Soak / stress tests of the cross-GPU flag protocol (>= 2 real GPUs; skipped otherwise).

compute-sanitizer is closed on this GPU pool, so the evidence that the in-kernel barriers hold
is behavioural: hundreds of consecutive collectives with random per-rank launch skew, mixed
entry points, several buckets and both paths, compared with the CPU oracle along the way.

The gradients are "exactly summable" (multiples of 2^-10 with |g| < 1): every partial sum of
<= 8 of them is exact in fp32 whatever the order, so the NVSwitch path -- which adds in an
order of its own -- must match the oracle BIT FOR BIT here too, and any torn read or
premature overwrite shows up as a bit difference.
"""
import threading

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import b200dp  # noqa: E402
import oracle  # noqa: E402

F = np.float32
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0
needs2 = pytest.mark.skipif(NGPU < 2, reason="needs >= 2 GPUs")

NUMELS = [262144, 64, 1000, 589824, 7, 2048, 147456, 1048576, 333]


def bits(a):
    return np.ascontiguousarray(a, dtype=F).view(np.uint32)


def views(flat, numels):
    out, o = [], 0
    for k in numels:
        out.append(flat[o:o + k])
        o += k
    return out


def summable_grad(rank, step, N, device):
    g = torch.Generator(device=device).manual_seed(99 + 1000 * rank + 7919 * step)
    return torch.randint(-1023, 1024, (N,), generator=g, device=device).to(torch.float32) * (2.0 ** -10)


def unpadded(plan, lay, dev, which, numels):
    buf = plan.tensor(dev, which).cpu().numpy()
    return np.concatenate([buf[lay.tensor_offset(k):lay.tensor_offset(k) + m] for k, m in enumerate(numels)])


@needs2
@pytest.mark.parametrize("path,bucket_bytes", [("peer", 0), ("peer", 1 << 20), ("multicast", 0), ("multicast", 1 << 20)])
def test_300_steps_with_random_rank_skew(path, bucket_bytes):
    n = 8 if NGPU >= 8 else (4 if NGPU >= 4 else 2)
    N = sum(NUMELS)
    ctx = b200dp.Context.local(list(range(n)))
    ctx.set_timeout(20.0)
    lay = b200dp.Layout(NUMELS, nranks=n, bucket_bytes=bucket_bytes)
    plan = b200dp.Plan(ctx, lay)
    if path == "multicast":
        try:
            plan.enable_multicast()
        except b200dp.B200dpError as e:
            pytest.skip("no NVSwitch multicast on this box: %s" % e)
    assert (lay.num_buckets > 1) == (bucket_bytes > 0)
    rng = np.random.default_rng(5)
    w0 = ((torch.rand(N, generator=torch.Generator().manual_seed(42)) - 0.5) * 0.1)
    for r in range(n):
        with torch.cuda.device(r):
            plan.load_weights(r, views(w0.cuda(r), NUMELS))
    w_ref, m_ref, v_ref = w0.numpy().copy(), np.zeros(N, F), np.zeros(N, F)
    small_in = [torch.zeros(5, device="cuda:%d" % r) for r in range(n)]
    small_out = [torch.zeros(5, device="cuda:%d" % r) for r in range(n)]
    checked = 0
    for step in range(1, 301):
        dg = [summable_grad(r, step, N, "cuda:%d" % r) for r in range(n)]
        hg = [g.cpu().numpy() for g in dg]
        slow = int(rng.integers(n))                       # one rank arrives late, by up to ~2 ms
        delay = int(rng.integers(0, 4_000_000))
        h = b200dp.adam_hyper(lr=1e-3, step=step)
        order = list(rng.permutation(n))                  # and the host issues the ranks in a random order
        with b200dp.issue_group():                        # nothing may block the host between the ranks' launches
            for r in order:
                with torch.cuda.device(int(r)):
                    if r == slow:
                        torch.cuda._sleep(delay)
                    try:
                        plan.pack_step(int(r), views(dg[r], NUMELS), "adam", h)
                    except b200dp.B200dpError as e:
                        raise AssertionError("step %d, rank %d (issue order %s, slow rank %d): %s" % (step, r, order, slow, e))
        oracle.adam_(w_ref, m_ref, v_ref, oracle.addn(hg), lr=1e-3, step=step)
        if step % 10 == 0:                                # metric averaging between steps
            vals = [np.array([r + step, -r, 0.5 * r, step, 1.0], F) for r in range(n)]
            for r in range(n):
                small_in[r].copy_(torch.from_numpy(vals[r]))      # a blocking copy: before the group, not inside it
            with b200dp.issue_group():
                for r in order:
                    with torch.cuda.device(int(r)):
                        plan.allreduce_small(int(r), small_in[r], small_out[r], postscale=1.0 / n)
            want = oracle.scale(oracle.addn(vals), 1.0 / n)
            for r in range(n):
                assert np.array_equal(bits(small_out[r].cpu().numpy()), bits(want)), "small all-reduce, step %d rank %d" % (step, r)
        if step % 50 == 0:                                # a re-broadcast of the (identical) weights from a rotating root
            root = (step // 50) % n
            with b200dp.issue_group():
                for r in order:
                    with torch.cuda.device(int(r)):
                        plan.broadcast(int(r), root=root)
        if step % 25 == 0:
            r = (step // 25) % n
            torch.cuda.synchronize(r)
            got = unpadded(plan, lay, r, b200dp.BUF_WEIGHTS, NUMELS)
            assert np.array_equal(bits(got), bits(w_ref)), "weights differ from the oracle at step %d on rank %d (%s)" % (step, r, path)
            checked += 1
    for r in range(n):
        torch.cuda.synchronize(r)
        plan.check(r)
        assert np.array_equal(bits(unpadded(plan, lay, r, b200dp.BUF_WEIGHTS, NUMELS)), bits(w_ref)), "final weights, rank %d" % r
    assert checked == 12


@needs2
def test_one_host_thread_per_device():
    """include/b200dp.h: a local context may be driven by one host thread per device concurrently
    (MirroredStrategy runs one Python thread per replica).  n threads, each issuing its own rank's
    pack_step for 40 steps with no ordering between them but a per-step barrier on the host."""
    n = 8 if NGPU >= 8 else (4 if NGPU >= 4 else 2)
    N = sum(NUMELS)
    ctx = b200dp.Context.local(list(range(n)))
    ctx.set_timeout(20.0)
    lay = b200dp.Layout(NUMELS, nranks=n)
    plan = b200dp.Plan(ctx, lay)
    w0 = ((torch.rand(N, generator=torch.Generator().manual_seed(42)) - 0.5) * 0.1)
    for r in range(n):
        with torch.cuda.device(r):
            plan.load_weights(r, views(w0.cuda(r), NUMELS))
        torch.cuda.synchronize(r)
    steps = 40
    errors = []
    barrier = threading.Barrier(n)

    def worker(r):
        try:
            torch.cuda.set_device(r)
            for step in range(1, steps + 1):
                g = summable_grad(r, step, N, "cuda:%d" % r)
                plan.pack_step(r, views(g, NUMELS), "adam", b200dp.adam_hyper(lr=1e-3, step=step))
                if step % 8 == 0:
                    barrier.wait(timeout=60)      # let the host threads drift apart in between
            torch.cuda.synchronize(r)
            plan.check(r)
        except Exception as e:   # noqa: BLE001
            errors.append((r, repr(e)))
            barrier.abort()

    ts = [threading.Thread(target=worker, args=(r,)) for r in range(n)]
    for t in ts:
        t.start()
    for t in ts:
        t.join(300)
    assert not errors, errors
    w_ref, m_ref, v_ref = w0.numpy().copy(), np.zeros(N, F), np.zeros(N, F)
    for step in range(1, steps + 1):
        hg = [summable_grad(r, step, N, "cuda:%d" % r).cpu().numpy() for r in range(n)]
        oracle.adam_(w_ref, m_ref, v_ref, oracle.addn(hg), lr=1e-3, step=step)
    for r in range(n):
        assert np.array_equal(bits(unpadded(plan, lay, r, b200dp.BUF_WEIGHTS, NUMELS)), bits(w_ref)), "rank %d" % r


@needs2
def test_timed_out_rank_skips_its_slice_and_later_calls_fail_fast():
    """A rank whose peer never arrives must not apply anything (its peers' buckets may not be
    packed), and the NEXT collective call must fail on the host without launching -- training
    cannot silently run on with partly updated weights."""
    ctx = b200dp.Context.local([0, 1])
    ctx.set_timeout(0.5)
    numels = [40000, 777]
    lay = b200dp.Layout(numels, nranks=2)
    plan = b200dp.Plan(ctx, lay)
    w0 = torch.full((sum(numels),), 0.25)
    with torch.cuda.device(0):
        plan.load_weights(0, views(w0.cuda(0), numels))
        plan.pack(0, views(torch.ones(sum(numels), device="cuda:0"), numels))
        plan.step(0, "adam", b200dp.adam_hyper(step=1))          # rank 1 never calls step
        torch.cuda.synchronize(0)
        with pytest.raises(b200dp.B200dpError) as e:              # refused on the host, no sync needed
            plan.step(0, "adam", b200dp.adam_hyper(step=2))
        assert e.value.code == b200dp._lib.ERR_TIMEOUT
        with pytest.raises(b200dp.B200dpError) as e:
            plan.allreduce(0)
        assert e.value.code == b200dp._lib.ERR_TIMEOUT
        with pytest.raises(b200dp.B200dpError) as e:
            plan.check(0)
        assert e.value.code == b200dp._lib.ERR_TIMEOUT
        plan.check(0)                                             # reported once, then cleared
        got = unpadded(plan, lay, 0, b200dp.BUF_WEIGHTS, numels)
        assert np.all(got == F(0.25)), "the timed-out rank touched its slice"
        assert float(plan.tensor(0, b200dp.BUF_STATE0).abs().sum()) == 0.0
