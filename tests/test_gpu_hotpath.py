"""GPU parity tests: the CUDA hot path, called through the C ABI (ctypes), against the CPU
oracle and the committed golden vectors.  Bar: BIT-EXACT for the fp32 rank-order sum and
for Adam / SGD-momentum state and weights (the kernels use one rounding per operation, no
FMA contraction, like the oracle); the only tolerance in this file is the gradient-norm
statistic (double accumulation in a different order: rel 1e-6).

n > 1 on a one-GPU box runs as *emulated ranks* (every rank's buffers on cuda:0, the
collective kernels launched rank after rank with the flag waits off), which exercises
exactly the peer-pointer arithmetic the multi-GPU run uses."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import b200dp  # noqa: E402
import oracle  # noqa: E402

F = np.float32


def bits(a):
    return np.ascontiguousarray(a, dtype=F).view(np.uint32)


def split(flat, numels):
    out, o = [], 0
    for n in numels:
        out.append(np.ascontiguousarray(flat[o:o + n]))
        o += n
    return out


def dev(arrs):
    return [torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in arrs]


def read_tensors(plan, r, numels, which=None):
    outs = [torch.empty(n, device="cuda", dtype=torch.float32) for n in numels]
    plan.unpack(r, outs, which=b200dp.BUF_WEIGHTS if which is None else which)
    torch.cuda.synchronize()
    return np.concatenate([o.cpu().numpy() for o in outs]) if outs else np.zeros(0, F)


def gather_state(plan, lay, n, which, numels):
    """Reassembles a sharded state (m / v / a) into tensor order from every rank's shard."""
    flat = np.zeros(lay.total_elems, F)
    for r in range(n):
        shard = plan.tensor(r, which).cpu().numpy()
        for b in range(lay.num_buckets):
            info = lay.bucket_info(b)
            lo = info["offset"] + r * info["slice"]
            flat[lo:lo + info["slice"]] = shard[info["shard_off"]:info["shard_off"] + info["slice"]]
    return np.concatenate([flat[lay.tensor_offset(k):lay.tensor_offset(k) + m] for k, m in enumerate(numels)]) \
        if numels else np.zeros(0, F)


def make_plan(numels, n, wire="f32", bucket_bytes=0):
    ctx = b200dp.Context.local([0] * n)
    lay = b200dp.Layout(numels, nranks=n, wire=wire, bucket_bytes=bucket_bytes)
    return ctx, lay, b200dp.Plan(ctx, lay)


# ------------------------------------------------------------------ golden vectors
@pytest.mark.parametrize("wire", ["f32", "bf16"])
@pytest.mark.parametrize("bucket_bytes", [0, 4096])
def test_golden_adam_three_steps(golden, wire, bucket_bytes):
    numels = [int(x) for x in golden["numels"]]
    n = golden["grads"].shape[1]
    ctx, lay, plan = make_plan(numels, n, wire, bucket_bytes)
    w0 = dev(split(golden["w0"], numels))
    for r in range(n):
        plan.load_weights(r, w0)
    for s in range(3):
        gts = [dev(split(golden["grads"][s, r], numels)) for r in range(n)]
        for r in range(n):
            plan.pack(r, gts[r])
        for r in range(n):
            plan.step(r, "adam", b200dp.adam_hyper(lr=float(golden["lrs"][s]), step=s + 1))
    for r in range(n):
        assert np.array_equal(bits(read_tensors(plan, r, numels)), bits(golden["adam_w_" + wire])), "rank %d" % r
        plan.check(r)
    assert np.array_equal(bits(gather_state(plan, lay, n, b200dp.BUF_STATE0, numels)), bits(golden["adam_m_" + wire]))
    assert np.array_equal(bits(gather_state(plan, lay, n, b200dp.BUF_STATE1, numels)), bits(golden["adam_v_" + wire]))


@pytest.mark.parametrize("wire", ["f32", "bf16"])
def test_golden_allreduce_sum(golden, wire):
    numels = [int(x) for x in golden["numels"]]
    n = golden["grads"].shape[1]
    ctx, lay, plan = make_plan(numels, n, wire)
    for s in range(3):
        gts = [dev(split(golden["grads"][s, r], numels)) for r in range(n)]
        for r in range(n):
            plan.pack(r, gts[r])
        for r in range(n):
            plan.allreduce(r)
        for r in range(n):
            got = read_tensors(plan, r, numels, b200dp.BUF_REDUCED)
            assert np.array_equal(bits(got), bits(golden["sum_" + wire][s])), "step %d rank %d" % (s, r)


def test_golden_hvd_average_then_adam(golden):
    numels = [int(x) for x in golden["numels"]]
    n = golden["grads"].shape[1]
    ctx, lay, plan = make_plan(numels, n)
    w0 = dev(split(golden["w0"], numels))
    for r in range(n):
        plan.load_weights(r, w0)
    for s in range(3):
        gts = [dev(split(golden["grads"][s, r], numels)) for r in range(n)]
        for r in range(n):
            plan.pack(r, gts[r])
        for r in range(n):
            plan.step(r, "adam", b200dp.adam_hyper(lr=float(golden["lrs"][s]), step=s + 1, postscale=1.0 / n))
    assert np.array_equal(bits(read_tensors(plan, 0, numels)), bits(golden["hvdavg_adam_w_f32"]))


@pytest.mark.parametrize("nest", [False, True])
@pytest.mark.parametrize("wire", ["f32", "bf16"])
def test_golden_sgdm(golden, nest, wire):
    numels = [int(x) for x in golden["numels"]]
    n = golden["grads"].shape[1]
    ctx, lay, plan = make_plan(numels, n, wire)
    w0 = dev(split(golden["w0"], numels))
    for r in range(n):
        plan.load_weights(r, w0)
    for s in range(3):
        gts = [dev(split(golden["grads"][s, r], numels)) for r in range(n)]
        for r in range(n):
            plan.pack(r, gts[r])
        for r in range(n):
            plan.step(r, "sgdm", b200dp.sgdm_hyper(lr=float(golden["sgd_lrs"][s]), momentum=0.9, nesterov=nest,
                                                   weight_decay=1e-4))
    tag = "sgdm_nest" if nest else "sgdm"
    for r in range(n):
        assert np.array_equal(bits(read_tensors(plan, r, numels)), bits(golden[tag + "_w_" + wire]))
    assert np.array_equal(bits(gather_state(plan, lay, n, b200dp.BUF_STATE0, numels)), bits(golden[tag + "_a_" + wire]))


# ------------------------------------------------------- full-size ResNet-50 vs oracle
def resnet_inputs(n, seed):
    """Synthetic per-rank gradients (SURVEY 8d D1 + D2 mix) generated on the GPU, mirrored to host."""
    numels = b200dp.resnet50_numels()
    g = torch.Generator(device="cuda").manual_seed(seed)
    flat = []
    for r in range(n):
        x = torch.randn(sum(numels), generator=g, device="cuda") * 1e-3
        x[::7] += (-1.0) ** r * 1e2          # cancellation stress on every 7th element
        flat.append(x)
    return numels, flat


def views(flat, numels):
    out, o = [], 0
    for n in numels:
        out.append(flat[o:o + n])
        o += n
    return out


@pytest.mark.parametrize("n,wire", [(1, "f32"), (2, "f32"), (8, "f32"), (8, "bf16"), (3, "f32")])
def test_resnet50_adam_steps_bit_exact(n, wire):
    numels, _ = resnet_inputs(0, 0)
    N = sum(numels)
    ctx, lay, plan = make_plan(numels, n, wire)
    gen = torch.Generator(device="cuda").manual_seed(42)
    w_flat = (torch.rand(N, generator=gen, device="cuda") - 0.5) * 0.1
    for r in range(n):
        plan.load_weights(r, views(w_flat, numels))
    w_ref = w_flat.cpu().numpy().copy()
    m_ref, v_ref = np.zeros(N, F), np.zeros(N, F)
    lrs = [1e-3, 1e-3, 1e-4]
    for s in range(3):
        _, flats = resnet_inputs(n, 1234 + s)
        for r in range(n):
            plan.pack(r, views(flats[r], numels))
        for r in range(n):
            plan.step(r, "adam", b200dp.adam_hyper(lr=lrs[s], step=s + 1, want_stats=True))
        host = [f.cpu().numpy() for f in flats]
        gsum = oracle.addn(host, bf16_wire=(wire == "bf16"))
        oracle.adam_(w_ref, m_ref, v_ref, gsum, lr=lrs[s], step=s + 1)
        # gradient-norm statistic: per-rank partial sums of squares add up to the oracle's
        torch.cuda.synchronize()
        tot = sum(float(plan.tensor(r, b200dp.BUF_STATS)[0]) for r in range(n))
        bad = sum(float(plan.tensor(r, b200dp.BUF_STATS)[1]) for r in range(n))
        assert abs(tot - oracle.sumsq(gsum)) <= 1e-6 * oracle.sumsq(gsum)   # tolerance: rel 1e-6
        assert bad == 0
    for r in range(n):
        assert np.array_equal(bits(read_tensors(plan, r, numels)), bits(w_ref)), "weights, rank %d" % r
        plan.check(r)
    assert np.array_equal(bits(gather_state(plan, lay, n, b200dp.BUF_STATE0, numels)), bits(m_ref))
    assert np.array_equal(bits(gather_state(plan, lay, n, b200dp.BUF_STATE1, numels)), bits(v_ref))


def test_resnet50_weights_alias_flat_buffer():
    """A harness may make its variables views of B2DP_BUF_WEIGHTS (no unpack needed)."""
    numels = b200dp.resnet50_numels()
    ctx, lay, plan = make_plan(numels, 1)
    wflat = plan.tensor(0, b200dp.BUF_WEIGHTS)
    assert wflat.numel() == lay.total_elems and wflat.dtype == torch.float32
    wflat.fill_(0.25)
    g = torch.full((sum(numels),), 0.5, device="cuda")
    plan.pack(0, views(g, numels))
    plan.step(0, "sgdm", b200dp.sgdm_hyper(lr=0.5, momentum=0.0))
    torch.cuda.synchronize()
    k = 100
    off = lay.tensor_offset(k)
    assert torch.all(wflat[off:off + numels[k]] == 0.0)      # 0.25 - 0.5*0.5, exact


# --------------------------------------------------------- properties at full size
def test_pack_unpack_roundtrip_and_unaligned_sources():
    """pack then unpack(GRADS==REDUCED for an f32 wire) is the identity, including tensors
    whose device pointers are only 4-byte aligned (scalar path) and ragged sizes."""
    rng = np.random.default_rng(5)
    numels = [1, 3, 4099, 64, 0, 12345, 4096, 8191, 77]
    big = torch.randn(sum(numels) + len(numels) + 8, device="cuda")
    srcs, o = [], 1            # start at element 1: every tensor is deliberately misaligned
    for n in numels:
        srcs.append(big[o:o + n])
        o += n + 1
    ctx, lay, plan = make_plan(numels, 1)
    plan.pack(0, srcs, prescale=1.0)
    got = read_tensors(plan, 0, numels, b200dp.BUF_REDUCED)
    want = np.concatenate([s.cpu().numpy() for s in srcs])
    assert np.array_equal(bits(got), bits(want))
    # padding between tensors stays zero
    flat = plan.tensor(0, b200dp.BUF_GRADS).cpu().numpy()
    mask = np.ones(lay.total_elems, bool)
    for k, n in enumerate(numels):
        mask[lay.tensor_offset(k):lay.tensor_offset(k) + n] = False
    assert not flat[mask].any()


def test_prescale_and_bf16_pack_match_oracle():
    numels = [1000, 4096, 33]
    x = torch.randn(sum(numels), device="cuda") * 3
    ctx, lay, plan = make_plan(numels, 1, "bf16")
    plan.pack(0, views(x, numels), prescale=0.125)
    torch.cuda.synchronize()
    flat = plan.tensor(0, b200dp.BUF_GRADS).view(torch.int16).cpu().numpy().view(np.uint16)
    want = oracle.cast_bf16((x.cpu().numpy() * F(0.125)).astype(F))
    o = 0
    for k, n in enumerate(numels):
        off = lay.tensor_offset(k)
        assert np.array_equal(flat[off:off + n], want[o:o + n])
        o += n


def test_allreduce_of_identical_ranks_is_exact_multiple():
    """Linearity: n equal inputs sum to n*g exactly when n is a power of two."""
    numels = b200dp.resnet50_numels()
    n = 4
    ctx, lay, plan = make_plan(numels, n)
    g = torch.randn(sum(numels), device="cuda")
    for r in range(n):
        plan.pack(r, views(g, numels))
    for r in range(n):
        plan.allreduce(r, postscale=1.0 / n)
    torch.cuda.synchronize()
    for r in range(n):
        red = plan.tensor(r, b200dp.BUF_REDUCED)
        for k in (0, 57, 213):
            off = lay.tensor_offset(k)
            assert torch.equal(red[off:off + numels[k]], views(g, numels)[k])


def test_two_launch_form_equals_fused_step():
    """reduce_scatter_apply + all_gather (PS push / pull) == the fused step, bit for bit."""
    numels = [5000, 64, 70001]
    n = 4
    rng = np.random.default_rng(9)
    grads = [[(rng.standard_normal(k) * 1e-2).astype(F) for k in numels] for _ in range(n)]
    w0 = [rng.standard_normal(k).astype(F) for k in numels]
    results = []
    for fused in (True, False):
        ctx, lay, plan = make_plan(numels, n)
        for r in range(n):
            plan.load_weights(r, dev(w0))
            plan.pack(r, dev(grads[r]))
        h = b200dp.adam_hyper(step=1)
        if fused:
            for r in range(n):
                plan.step(r, "adam", h)
        else:
            for r in range(n):
                plan.reduce_scatter_apply(r, "adam", h)
            for r in range(n):
                plan.all_gather(r)
        results.append([read_tensors(plan, r, numels) for r in range(n)])
    for r in range(n):
        assert np.array_equal(bits(results[0][r]), bits(results[1][r]))
        assert np.array_equal(bits(results[0][r]), bits(results[0][0]))


def test_nonfinite_gradients_propagate_and_are_counted():
    numels = [4096]
    ctx, lay, plan = make_plan(numels, 2)
    g0 = torch.ones(4096, device="cuda")
    g1 = torch.ones(4096, device="cuda")
    g0[5] = float("inf")
    g1[5] = float("-inf")
    g1[4000] = float("nan")
    for r, g in enumerate((g0, g1)):
        plan.load_weights(r, [torch.zeros(4096, device="cuda")])
        plan.pack(r, [g])
    for r in range(2):
        plan.step(r, "adam", b200dp.adam_hyper(step=1, want_stats=True))
    torch.cuda.synchronize()
    bad = sum(float(plan.tensor(r, b200dp.BUF_STATS)[1]) for r in range(2))
    assert bad == 2
    w = read_tensors(plan, 0, numels)
    assert np.isnan(w[5]) and np.isnan(w[4000]) and np.isfinite(w[6])


def test_empty_and_all_empty_plans():
    for numels in ([], [0, 0]):
        ctx, lay, plan = make_plan(numels, 2)
        ts = [torch.empty(0, device="cuda") for _ in numels]
        for r in range(2):
            plan.pack(r, ts)
        for r in range(2):
            plan.step(r, "adam", b200dp.adam_hyper(step=1))
            plan.allreduce(r)
            plan.check(r)


def test_argument_errors():
    ctx, lay, plan = make_plan([128, 128], 1)
    with pytest.raises(ValueError):
        plan.pack(0, [torch.zeros(128, device="cuda")])          # wrong tensor count
    with pytest.raises(b200dp.B200dpError):
        plan.step(0, "none", b200dp.adam_hyper())                 # NONE is allreduce's job
    with pytest.raises(b200dp.B200dpError):
        plan.step(0, "adam", b200dp.adam_hyper(step=0))           # Adam step is 1-based
    with pytest.raises(b200dp.B200dpError):
        plan.step(1, "adam", b200dp.adam_hyper())                 # no such local rank
    with pytest.raises(b200dp.B200dpError):
        plan.step(0, "adam", b200dp.adam_hyper(), bucket=3)       # no such bucket
    lay8 = b200dp.Layout([128], nranks=8)
    with pytest.raises(b200dp.B200dpError):
        b200dp.Plan(ctx, lay8)                                    # layout cut for another rank count


def test_lr_is_a_per_call_argument():
    """The scripts' callbacks change lr while training (ReduceLROnPlateau, warm-up)."""
    numels = [2048]
    ctx, lay, plan = make_plan(numels, 1)
    w = np.ones(2048, F)
    plan.load_weights(0, dev([w]))
    w_ref, m, v = w.copy(), np.zeros(2048, F), np.zeros(2048, F)
    g = (np.arange(2048, dtype=F) - 1000) * F(1e-3)
    for step, lr in enumerate([0.8, 0.08, 1e-5], start=1):
        plan.pack(0, dev([g]))
        plan.step(0, "adam", b200dp.adam_hyper(lr=lr, step=step))
        oracle.adam_(w_ref, m, v, g, lr=lr, step=step)
    assert np.array_equal(bits(read_tensors(plan, 0, numels)), bits(w_ref))


# ------------------------------------------------------------- pack_step (one call per step)
def test_pack_step_single_rank_reads_gradients_in_place_bit_exact():
    """n = 1: one kernel, no bucket copy; ResNet-50, 3 steps, prescale != 1, statistics on."""
    numels, _ = resnet_inputs(0, 0)
    N = sum(numels)
    ctx, lay, plan = make_plan(numels, 1)
    gen = torch.Generator(device="cuda").manual_seed(42)
    w_flat = (torch.rand(N, generator=gen, device="cuda") - 0.5) * 0.1
    plan.load_weights(0, views(w_flat, numels))
    w_ref = w_flat.cpu().numpy().copy()
    m_ref, v_ref = np.zeros(N, F), np.zeros(N, F)
    for s in range(3):
        _, flats = resnet_inputs(1, 77 + s)
        plan.pack_step(0, views(flats[0], numels), "adam", b200dp.adam_hyper(lr=1e-3, step=s + 1, want_stats=True),
                       prescale=0.5)
        g = oracle.addn([flats[0].cpu().numpy()], prescale=0.5)
        oracle.adam_(w_ref, m_ref, v_ref, g, lr=1e-3, step=s + 1)
        torch.cuda.synchronize()
        st = plan.tensor(0, b200dp.BUF_STATS).cpu().numpy()
        assert abs(st[0] - oracle.sumsq(g)) <= 1e-6 * oracle.sumsq(g) and st[1] == 0      # tolerance: rel 1e-6
    assert np.array_equal(bits(read_tensors(plan, 0, numels)), bits(w_ref))
    assert np.array_equal(bits(gather_state(plan, lay, 1, b200dp.BUF_STATE0, numels)), bits(m_ref))
    assert np.array_equal(bits(gather_state(plan, lay, 1, b200dp.BUF_STATE1, numels)), bits(v_ref))
    # the gradient bucket was never written
    assert not plan.tensor(0, b200dp.BUF_GRADS).any()


@pytest.mark.parametrize("opt", ["adam", "sgdm"])
@pytest.mark.parametrize("bucket_bytes", [0, 20000])
def test_pack_step_ragged_and_misaligned_sources(opt, bucket_bytes):
    rng = np.random.default_rng(11)
    numels = [1, 3, 8195, 64, 0, 12345, 8192, 16383, 77, 9]
    big = torch.randn(sum(numels) + 4 * len(numels) + 8, device="cuda") * 1e-2
    for shift in (0, 1, 4):            # 32-byte aligned, 4-byte aligned, 16-byte aligned sources
        srcs, o = [], shift
        for n in numels:
            srcs.append(big[o:o + n])
            o += (n + 7) // 8 * 8
        ctx, lay, plan = make_plan(numels, 1, bucket_bytes=bucket_bytes)
        w0 = [rng.standard_normal(n).astype(F) for n in numels]
        plan.load_weights(0, dev(w0))
        w_ref = np.concatenate(w0)
        s0, s1 = np.zeros_like(w_ref), np.zeros_like(w_ref)
        g = np.concatenate([s.cpu().numpy() for s in srcs])
        for step in (1, 2):
            if opt == "adam":
                plan.pack_step(0, srcs, "adam", b200dp.adam_hyper(step=step))
                oracle.adam_(w_ref, s0, s1, g, step=step)
            else:
                plan.pack_step(0, srcs, "sgdm", b200dp.sgdm_hyper(lr=0.1, momentum=0.9, nesterov=True, weight_decay=1e-3))
                oracle.sgdm_(w_ref, s0, g, lr=0.1, mu=0.9, nesterov=True, wd=1e-3)
        assert np.array_equal(bits(read_tensors(plan, 0, numels)), bits(w_ref)), "shift %d" % shift


def test_pack_step_with_ranks_equals_pack_then_step(golden):
    numels = [int(x) for x in golden["numels"]]
    n = golden["grads"].shape[1]
    ctx, lay, plan = make_plan(numels, n)
    w0 = dev(split(golden["w0"], numels))
    for r in range(n):
        plan.load_weights(r, w0)
    for s in range(3):
        gts = [dev(split(golden["grads"][s, r], numels)) for r in range(n)]
        # emulated ranks: every rank's pack must precede any rank's step, so pack_step is issued
        # for ranks 0..n-2 as pack only and the collective part afterwards (what real ranks do
        # concurrently); rank n-1 can use the one-call form
        for r in range(n - 1):
            plan.pack(r, gts[r])
        h = b200dp.adam_hyper(lr=float(golden["lrs"][s]), step=s + 1)
        plan.pack(n - 1, gts[n - 1])
        for r in range(n):
            plan.step(r, "adam", h)
    assert np.array_equal(bits(read_tensors(plan, 0, numels)), bits(golden["adam_w_f32"]))
    with pytest.raises(b200dp.B200dpError):
        plan.pack_step(0, dev(split(golden["grads"][0, 0], numels)), "none", b200dp.adam_hyper())


# ------------------------------------------------- step-0 / per-epoch helpers
def test_broadcast_from_root_emulated_ranks():
    numels = [5000, 64, 70001, 3]
    n = 4
    ctx, lay, plan = make_plan(numels, n)
    rng = np.random.default_rng(21)
    per_rank = [[rng.standard_normal(k).astype(F) for k in numels] for _ in range(n)]
    for r in range(n):
        plan.load_weights(r, dev(per_rank[r]))
    for r in range(n):
        plan.broadcast(r, root=2)
    want = np.concatenate(per_rank[2])
    for r in range(n):
        assert np.array_equal(bits(read_tensors(plan, r, numels)), bits(want))
    with pytest.raises(b200dp.B200dpError):
        plan.broadcast(0, root=7)


@pytest.mark.parametrize("count", [1, 7, 1024])
def test_allreduce_small_matches_rank_order_sum(count):
    n = 4
    ctx, lay, plan = make_plan([256], n)
    rng = np.random.default_rng(count)
    for call in range(3):          # consecutive calls alternate the slot halves
        vals = [(rng.standard_normal(count) * 10.0 ** rng.integers(-3, 3)).astype(F) for _ in range(n)]
        ins = dev(vals)
        outs = [torch.zeros(count, device="cuda") for _ in range(n)]
        for r in range(n):
            plan.allreduce_small(r, ins[r], outs[r], postscale=1.0 / n)
        torch.cuda.synchronize()
        ref = oracle.scale(oracle.addn(vals), 1.0 / n)
        for r in range(n):
            assert np.array_equal(bits(outs[r].cpu().numpy()), bits(ref)), "call %d rank %d" % (call, r)


def test_allreduce_small_rejects_large_messages():
    ctx, lay, plan = make_plan([256], 1)
    x = torch.zeros(1025, device="cuda")
    with pytest.raises(b200dp.B200dpError) as e:
        plan.allreduce_small(0, x, x.clone())
    assert e.value.code == b200dp._lib.ERR_TOO_LARGE
    y = torch.arange(10, device="cuda", dtype=torch.float32)
    out = torch.empty_like(y)
    plan.allreduce_small(0, y, out, postscale=0.5)      # n = 1: value * postscale
    assert torch.equal(out, y * 0.5)


@pytest.mark.parametrize("n,bucket_bytes", [(1, 0), (4, 0), (4, 4096)])
def test_gather_state_reassembles_sharded_adam_moments(golden, n, bucket_bytes):
    """After training, m and v live sharded across ranks; gather_state + unpack returns them
    per variable on every rank (what a checkpoint needs), equal to the oracle's."""
    numels = [int(x) for x in golden["numels"]]
    ng = golden["grads"].shape[1]
    ctx, lay, plan = make_plan(numels, n, bucket_bytes=bucket_bytes)
    w0 = dev(split(golden["w0"], numels))
    for r in range(n):
        plan.load_weights(r, w0)
    w_ref = golden["w0"].copy()
    m_ref, v_ref = np.zeros_like(w_ref), np.zeros_like(w_ref)
    for s in range(2):
        hg = [golden["grads"][s, r % ng] for r in range(n)]
        for r in range(n):
            plan.pack(r, dev(split(hg[r], numels)))
        for r in range(n):
            plan.step(r, "adam", b200dp.adam_hyper(step=s + 1))
        oracle.adam_(w_ref, m_ref, v_ref, oracle.addn(hg), step=s + 1)
    for which, ref in ((b200dp.BUF_STATE0, m_ref), (b200dp.BUF_STATE1, v_ref)):
        for r in range(n):
            plan.gather_state(r, which)
        for r in range(n):
            got = read_tensors(plan, r, numels, b200dp.BUF_REDUCED)
            assert np.array_equal(bits(got), bits(ref)), "rank %d" % r
    with pytest.raises(b200dp.B200dpError):
        plan.gather_state(0, b200dp.BUF_WEIGHTS)


def test_large_plan_offsets_beyond_32_bits():
    """Maximum-size case: one 2^29+12,345-element tensor (2.1 GB per buffer), so byte offsets
    pass 2^31; three 1M-element windows (start, middle, ragged end) are checked against the
    oracle, bit-exact, through both the in-place fused kernel and pack -> step."""
    n_el = (1 << 29) + 12345
    numels = [4096, n_el, 777]
    free, _ = torch.cuda.mem_get_info()
    if free < 16 * (1 << 30):
        pytest.skip("needs ~12 GB of free device memory")
    ctx, lay, plan = make_plan(numels, 1)
    gen = torch.Generator(device="cuda").manual_seed(9)
    g = [torch.randn(k, generator=gen, device="cuda") * 1e-2 for k in numels]
    w = [torch.rand(k, generator=gen, device="cuda") - 0.5 for k in numels]
    plan.load_weights(0, w)
    plan.pack_step(0, g, "adam", b200dp.adam_hyper(step=1))          # fused in-place kernel
    plan.pack(0, g)
    plan.step(0, "adam", b200dp.adam_hyper(step=2))                  # pack -> apply1 kernel
    torch.cuda.synchronize()
    flat_w = plan.tensor(0, b200dp.BUF_WEIGHTS)
    off = lay.tensor_offset(1)
    for lo in (0, n_el // 2 + 3, n_el - 1_000_000):
        hi = min(n_el, lo + 1_000_000)
        ws = w[1][lo:hi].cpu().numpy().copy()
        gs = g[1][lo:hi].cpu().numpy()
        ms, vs = np.zeros_like(ws), np.zeros_like(ws)
        oracle.adam_(ws, ms, vs, gs, step=1)
        oracle.adam_(ws, ms, vs, gs, step=2)
        got = flat_w[off + lo:off + hi].cpu().numpy()
        assert np.array_equal(bits(got), bits(ws)), "window at %d" % lo
    # the small tensor after the big one sits past 2^31 bytes too
    off2 = lay.tensor_offset(2)
    ws = w[2].cpu().numpy().copy()
    ms, vs = np.zeros_like(ws), np.zeros_like(ws)
    oracle.adam_(ws, ms, vs, g[2].cpu().numpy(), step=1)
    oracle.adam_(ws, ms, vs, g[2].cpu().numpy(), step=2)
    assert np.array_equal(bits(flat_w[off2:off2 + 777].cpu().numpy()), bits(ws))


def test_bf16_gradient_mode_within_stated_bound_of_fp32_oracle():
    """bf16 wire vs the FP32 oracle: |err| <= 2^-8 * sum_i |g_i|, every element;
    (vs the bf16-wire oracle it is bit-exact, tested above)."""
    numels = [100_000, 4096, 33]
    n = 8
    ctx, lay, plan = make_plan(numels, n, "bf16")
    rng = np.random.default_rng(123)
    hg = [np.concatenate([(rng.standard_normal(k) * 10.0 ** rng.integers(-4, 2)).astype(F) for k in numels])
          for _ in range(n)]
    for r in range(n):
        plan.pack(r, dev(split(hg[r], numels)))
    for r in range(n):
        plan.allreduce(r)
    got = read_tensors(plan, 0, numels, b200dp.BUF_REDUCED).astype(np.float64)
    ref = oracle.addn(hg).astype(np.float64)
    bound = 2.0 ** -8 * np.sum(np.abs(np.stack(hg)), axis=0)
    assert np.all(np.abs(got - ref) <= bound)
    assert np.array_equal(bits(got.astype(F)), bits(oracle.addn(hg, bf16_wire=True)))


@pytest.mark.parametrize("n", [2, 4, 8])
def test_256_bit_peer_kernel_is_bit_identical(golden, n):
    """B2DP_TUNE_PEER_VEC=8 selects the LDG.256/STG.256 form of the peer-path step: same sums, same
    optimizer arithmetic, so the golden answers (n=4) and the 128-bit kernel's results (any n) hold."""
    numels = [int(x) for x in golden["numels"]]
    ng = golden["grads"].shape[1]
    res = []
    for vec in (4, 8):
        ctx, lay, plan = make_plan(numels, n, bucket_bytes=4096)
        plan.set_tuning(peer_vec=vec)
        w0 = dev(split(golden["w0"], numels))
        for r in range(n):
            plan.load_weights(r, w0)
        for s in range(3):
            for r in range(n):
                plan.pack(r, dev(split(golden["grads"][s, r % ng], numels)))
            for r in range(n):
                plan.step(r, "adam", b200dp.adam_hyper(lr=float(golden["lrs"][s]), step=s + 1, want_stats=True))
        for r in range(n):
            plan.pack(r, dev(split(golden["grads"][0, r % ng], numels)))
        for r in range(n):
            plan.allreduce(r)
        res.append((read_tensors(plan, n - 1, numels), read_tensors(plan, 0, numels, b200dp.BUF_REDUCED)))
    assert np.array_equal(bits(res[0][0]), bits(res[1][0])) and np.array_equal(bits(res[0][1]), bits(res[1][1]))
    if n == ng:
        assert np.array_equal(bits(res[1][0]), bits(golden["adam_w_f32"]))
        assert np.array_equal(bits(res[1][1]), bits(golden["sum_f32"][0]))


# ------------------------------------------------------------------ statistics across buckets
@pytest.mark.parametrize("n", [1, 2, 4])
@pytest.mark.parametrize("call", ["step", "pack_step", "reduce_scatter_apply"])
def test_stats_cover_every_bucket_of_the_call(n, call):
    """want_stats with a multi-bucket plan: one call launches one kernel per bucket, and
    B2DP_BUF_STATS must hold the sum over ALL of them (summed over ranks = the oracle's sum of
    squares of the whole reduced gradient), including a non-finite element planted in bucket 0."""
    numels = [5000, 64, 3000, 9001, 128, 7000, 2048, 4100]
    N = sum(numels)
    ctx, lay, plan = make_plan(numels, n, bucket_bytes=20000)
    assert lay.num_buckets >= 4
    rng = np.random.default_rng(11)
    for plant_inf in (False, True):
        hg = [(rng.standard_normal(N) * 1e-2).astype(F) for _ in range(n)]
        if plant_inf:
            hg[0][17] = np.inf                      # tensor 0 -> bucket 0
        gts = [dev(split(g, numels)) for g in hg]
        h = b200dp.adam_hyper(step=1, want_stats=True)
        if call == "pack_step" and n == 1:
            plan.pack_step(0, gts[0], "adam", h)
        else:
            for r in range(n):
                plan.pack(r, gts[r])
            for r in range(n):
                getattr(plan, "step" if call == "pack_step" else call)(r, "adam", h)
        torch.cuda.synchronize()
        stats = np.stack([plan.tensor(r, b200dp.BUF_STATS).cpu().numpy() for r in range(n)])
        gsum = oracle.addn(hg)
        if plant_inf:
            assert stats[:, 1].sum() == 1, stats
            assert np.isinf(stats[:, 0].sum())
        else:
            want = oracle.sumsq(gsum)
            assert abs(stats[:, 0].sum() - want) <= 1e-6 * want, (stats, want)   # tolerance: rel 1e-6
            assert stats[:, 1].sum() == 0


def test_more_tensors_than_fit_in_launch_parameters():
    """Up to 448 tensor pointers ride in the resolve kernel's launch parameters; a longer list goes
    through a device-resident pointer table.  600 tensors: pack -> fused step -> unpack vs the
    oracle, then the same with every gradient at a NEW address (the table must be re-resolved)."""
    rng = np.random.default_rng(21)
    numels = [int(x) for x in rng.integers(1, 300, 600)]
    N = sum(numels)
    for n in (1, 2):
        ctx, lay, plan = make_plan(numels, n)
        w0 = (rng.standard_normal(N) * 0.05).astype(F)
        for r in range(n):
            plan.load_weights(r, dev(split(w0, numels)))
        w, m, v = w0.copy(), np.zeros(N, F), np.zeros(N, F)
        for step in (1, 2):
            hg = [(rng.standard_normal(N) * 1e-2).astype(F) for _ in range(n)]
            gts = [dev(split(g, numels)) for g in hg]          # fresh tensors: new addresses every step
            if n == 1:
                plan.pack_step(0, gts[0], "adam", b200dp.adam_hyper(step=step))
            else:
                for r in range(n):
                    plan.pack(r, gts[r])
                for r in range(n):
                    plan.step(r, "adam", b200dp.adam_hyper(step=step))
            oracle.adam_(w, m, v, oracle.addn(hg), step=step)
        for r in range(n):
            assert np.array_equal(bits(read_tensors(plan, r, numels)), bits(w)), "n=%d rank %d" % (n, r)
