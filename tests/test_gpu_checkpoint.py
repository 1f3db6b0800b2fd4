"""Checkpoint round trip (model.save at the end of every script,
imagenet-resnet50-mirror.py:83-86, and the PS script's MinSizePartitioner, -ps.py:75-78):
train 2 steps on 4 ranks -> save (sharded Adam state gathered, variables partitioned and named)
-> load into plans cut for 2 ranks and for 1 rank -> the third step is bit-identical to the
uninterrupted 4-rank run and to the oracle.  Emulated ranks, so it runs on a one-GPU box."""
import os
import tempfile

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import b200dp  # noqa: E402
import oracle  # noqa: E402
from b200dp import checkpoint as ck  # noqa: E402

F = np.float32
VARS = [("a/kernel", (3, 3, 16, 32)), ("a/bias", (32,)), ("b/kernel", (700, 130)), ("b/bias", (130,)),
        ("c/gamma", (7,)), ("d/kernel", (1, 1, 64, 256))]
NUMELS = [int(np.prod(s)) for _, s in VARS]
N = sum(NUMELS)


def bits(a):
    return np.ascontiguousarray(a, dtype=F).view(np.uint32)


def split(flat):
    out, o = [], 0
    for k in NUMELS:
        out.append(torch.from_numpy(np.ascontiguousarray(flat[o:o + k])).cuda())
        o += k
    return out


def plan_for(n, bucket_bytes=0):
    ctx = b200dp.Context.local([0] * n)
    lay = b200dp.Layout(NUMELS, nranks=n, bucket_bytes=bucket_bytes)
    return ctx, lay, b200dp.Plan(ctx, lay)


def run_step(plan, n, grads, step, lr):
    for r in range(n):
        plan.pack(r, split(grads[r]))
    for r in range(n):
        plan.step(r, "adam", b200dp.adam_hyper(lr=lr, step=step))


def weights(plan, lay, r=0):
    buf = plan.tensor(r, b200dp.BUF_WEIGHTS).cpu().numpy()
    return np.concatenate([buf[lay.tensor_offset(k):lay.tensor_offset(k) + m] for k, m in enumerate(NUMELS)])


@pytest.mark.parametrize("partitioned", [False, True])
def test_save_at_4_ranks_continue_at_2_and_1(partitioned):
    rng = np.random.default_rng(3)
    w0 = (rng.standard_normal(N) * 0.05).astype(F)
    # steps 1-2: ordinary gradients; step 3: multiples of 2^-10, so any grouping of the 4 ranks'
    # gradients into 2 or 1 sums to exactly the same fp32 values
    g12 = [[(rng.standard_normal(N) * 1e-2).astype(F) for _ in range(4)] for _ in range(2)]
    g3 = [(rng.integers(-1023, 1024, N) * 2.0 ** -10).astype(F) for _ in range(4)]
    lrs = (1e-3, 1e-3, 1e-4)
    w_ref, m_ref, v_ref = w0.copy(), np.zeros(N, F), np.zeros(N, F)
    for s in range(2):
        oracle.adam_(w_ref, m_ref, v_ref, oracle.addn(g12[s]), lr=lrs[s], step=s + 1)
    m2, v2 = m_ref.copy(), v_ref.copy()
    oracle.adam_(w_ref, m_ref, v_ref, oracle.addn(g3), lr=lrs[2], step=3)

    ctx4, lay4, plan4 = plan_for(4, bucket_bytes=200_000)
    for r in range(4):
        plan4.load_weights(r, split(w0))
    for s in range(2):
        run_step(plan4, 4, g12[s], s + 1, lrs[s])
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "ImageNet-b200dp-reuse.npz")
        part = ck.MinSizePartitioner(min_shard_bytes=256 << 10, max_shards=3) if partitioned else None
        written = ck.save(plan4, path, step=2, variables=VARS, partitioner=part)
        names = set(written)
        if partitioned:      # b/kernel is 364,000 B: one full 256 KiB shard + a remainder -> ceil(700/505) = 2 parts
            assert {"b/kernel/part_0", "b/kernel/part_1", "optimizer/b/kernel/m/part_0", "optimizer/b/kernel/v/part_1"} <= names
            assert "b/kernel" not in names and "a/bias" in names
        else:
            assert {"a/kernel", "optimizer/a/kernel/m", "optimizer/a/kernel/v", "optimizer/iterations"} <= names
        saved = ck.read(path)
        assert saved["a/kernel"].shape == (3, 3, 16, 32)
        # the gathered slots are the oracle's m, v after two steps, per variable
        o = 0
        for (name, shape), k in zip(VARS, NUMELS):
            assert np.array_equal(bits(saved["optimizer/%s/m" % name].ravel()), bits(m2[o:o + k])), name
            assert np.array_equal(bits(saved["optimizer/%s/v" % name].ravel()), bits(v2[o:o + k])), name
            o += k
        # uninterrupted run: third step on the 4-rank plan
        run_step(plan4, 4, g3, 3, lrs[2])
        torch.cuda.synchronize()
        assert np.array_equal(bits(weights(plan4, lay4)), bits(w_ref))
        for n, groups in ((2, [(0, 1), (2, 3)]), (1, [(0, 1, 2, 3)])):
            ctx, lay, plan = plan_for(n)
            assert ck.load(plan, path, variables=VARS) == 2
            gn = [np.sum([g3[i] for i in grp], axis=0, dtype=F) for grp in groups]
            run_step(plan, n, gn, 3, lrs[2])
            torch.cuda.synchronize()
            for r in range(n):
                assert np.array_equal(bits(weights(plan, lay, r)), bits(w_ref)), "resumed at n=%d, rank %d" % (n, r)
                plan.check(r)


def test_state_tensors_gathers_before_unpacking():
    """Plan.state_tensors(which) = gather_state + unpack (it used to skip the gather)."""
    rng = np.random.default_rng(4)
    ctx, lay, plan = plan_for(1)
    w0 = (rng.standard_normal(N) * 0.05).astype(F)
    g = (rng.standard_normal(N) * 1e-2).astype(F)
    plan.load_weights(0, split(w0))
    run_step(plan, 1, [g], 1, 1e-3)
    w, m, v = w0.copy(), np.zeros(N, F), np.zeros(N, F)
    oracle.adam_(w, m, v, g, lr=1e-3, step=1)
    like = split(w0)
    got_m = torch.cat([t.flatten() for t in plan.state_tensors(0, b200dp.BUF_STATE0, like)]).cpu().numpy()
    got_v = torch.cat([t.flatten() for t in plan.state_tensors(0, b200dp.BUF_STATE1, like)]).cpu().numpy()
    assert np.array_equal(bits(got_m), bits(m)) and np.array_equal(bits(got_v), bits(v))
