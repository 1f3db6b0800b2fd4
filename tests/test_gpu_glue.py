"""The reference-shaped glue on a GPU: B200CrossDeviceOps.batch_reduce (MirroredStrategy's
hook) and the Horovod-shaped module, against the oracle (bit-exact: peer path)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import b200dp  # noqa: E402
import oracle  # noqa: E402
from b200dp import cross_device_ops as cdo  # noqa: E402

F = np.float32


def test_batch_reduce_sum_and_mean_two_replicas_on_one_gpu():
    """214 ResNet-50-shaped gradients, 2 replicas (emulated on cuda:0), SUM and MEAN."""
    shapes = [s for _, s in b200dp.resnet50_trainable_variables()]
    g = torch.Generator(device="cuda").manual_seed(1)
    vals = [[torch.randn(s, generator=g, device="cuda") * 1e-2 for _ in range(2)] for s in shapes]
    ops = cdo.B200CrossDeviceOps()
    for op in (cdo.ReduceOp.SUM, cdo.ReduceOp.MEAN):
        out = ops.batch_reduce(op, [(v, None) for v in vals], cdo.CommunicationOptions(bytes_per_pack=8 << 20))
        torch.cuda.synchronize()
        assert len(out) == len(vals) and all(len(o) == 2 for o in out)
        for k in (0, 1, 57, 100, 212, 213):
            ref = oracle.addn([vals[k][0].cpu().numpy().ravel(), vals[k][1].cpu().numpy().ravel()])
            if op == cdo.ReduceOp.MEAN:
                ref = oracle.div_n(ref, 2)
            for r in range(2):
                assert out[k][r].shape == vals[k][r].shape
                assert np.array_equal(out[k][r].cpu().numpy().ravel().view(np.uint32), ref.view(np.uint32))


def test_reduce_single_value():
    v = [torch.full((1000,), 1.5, device="cuda"), torch.full((1000,), 2.25, device="cuda")]
    out = cdo.B200CrossDeviceOps().reduce(cdo.ReduceOp.SUM, v)
    assert torch.all(out[0] == 3.75) and torch.all(out[1] == 3.75)


def test_hvd_single_process_surface():
    from b200dp import hvd
    hvd.shutdown()
    hvd.init()
    assert (hvd.rank(), hvd.size(), hvd.local_rank()) == (0, 1, 0)
    x = torch.randn(12345, device="cuda")
    assert torch.equal(hvd.allreduce(x), x)                       # Average over one rank
    assert torch.equal(hvd.allreduce(x, op=hvd.Sum, postscale_factor=0.5), x * 0.5)
    with pytest.raises(ValueError):
        hvd.allreduce(x, average=True, op=hvd.Sum)
    # DistributedOptimizer(size 1) == the wrapped optimizer
    p1 = torch.nn.Parameter(torch.ones(100, device="cuda"))
    p2 = torch.nn.Parameter(torch.ones(100, device="cuda"))
    o1 = hvd.DistributedOptimizer(torch.optim.SGD([p1], lr=0.1))
    o2 = torch.optim.SGD([p2], lr=0.1)
    p1.grad = torch.full((100,), 0.5, device="cuda")
    p2.grad = torch.full((100,), 0.5, device="cuda")
    o1.step()
    o2.step()
    assert torch.equal(p1, p2)


def test_hvd_fused_adam_matches_oracle():
    from b200dp import hvd
    hvd.shutdown()
    hvd.init()
    rng = np.random.default_rng(3)
    shapes = [(64, 3, 7, 7), (64,), (1000, 2048), (1000,)]
    params = [torch.nn.Parameter(torch.from_numpy(rng.standard_normal(s).astype(F) * 0.05).cuda()) for s in shapes]
    w = np.concatenate([p.detach().cpu().numpy().ravel() for p in params])
    m, v = np.zeros_like(w), np.zeros_like(w)
    opt = hvd.FusedDistributedAdam(params, lr=1e-3)
    for step in (1, 2, 3):
        gs = [rng.standard_normal(s).astype(F) * 1e-2 for s in shapes]
        for p, g in zip(params, gs):
            p.grad = torch.from_numpy(g).cuda()
        opt.step()
        oracle.adam_(w, m, v, np.concatenate([g.ravel() for g in gs]), lr=1e-3, step=step)
    torch.cuda.synchronize()
    got = np.concatenate([p.detach().cpu().numpy().ravel() for p in params])
    assert np.array_equal(got.view(np.uint32), w.view(np.uint32))


class _TinyNet(torch.nn.Module):
    """Small deterministic model (matmuls only) for the overlap test."""

    def __init__(self):
        super().__init__()
        self.l = torch.nn.ModuleList([torch.nn.Linear(256, 512), torch.nn.Linear(512, 512), torch.nn.Linear(512, 300),
                                      torch.nn.Linear(300, 10)])

    def forward(self, x):
        for i, m in enumerate(self.l):
            x = m(x)
            if i < len(self.l) - 1:
                x = torch.relu(x)
        return x

    def keras_order_params(self):
        out = []
        for m in self.l:
            out += [m.weight, m.bias]
        return out


def _train(overlap, bucket_bytes, steps=3):
    from b200dp import harness
    torch.manual_seed(0)
    torch.backends.cuda.matmul.allow_tf32 = False
    model = _TinyNet().cuda()
    numels = [p.numel() for p in model.keras_order_params()]
    ctx = b200dp.Context.local([0])
    lay = b200dp.Layout(numels, nranks=1, bucket_bytes=bucket_bytes)
    plan = b200dp.Plan(ctx, lay)
    stepper = harness.DataParallelStep(model, plan, overlap=overlap)
    g = torch.Generator(device="cuda").manual_seed(5)
    for _ in range(steps):
        x = torch.randn(64, 256, generator=g, device="cuda")
        y = torch.randint(0, 10, (64,), generator=g, device="cuda")
        stepper(x, y)
    torch.cuda.synchronize()
    return lay.num_buckets, torch.cat([p.detach().flatten() for p in model.keras_order_params()]).cpu()


def test_bucket_pipeline_overlapped_with_backward_equals_serial_step():
    """issuing each bucket's pack+step from autograd hooks on a side stream, while
    backward still runs, gives bit-identical weights to the serial step."""
    nb0, w_serial = _train(False, 0)
    nb1, w_overlap = _train(True, 600_000)
    assert nb0 == 1 and nb1 >= 3
    assert torch.equal(w_serial.view(torch.int32), w_overlap.view(torch.int32))


def test_batch_reduce_mean_three_replicas_divides_like_tf():
    """n = 3 is not a power of two: MEAN must be the AddN sum DIVIDED by 3 (TF), not times float(1/3)."""
    g = torch.Generator(device="cuda").manual_seed(3)
    vals = [[torch.randn(5001, generator=g, device="cuda") for _ in range(3)] for _ in range(4)]
    out = cdo.B200CrossDeviceOps().batch_reduce(cdo.ReduceOp.MEAN, [(v, None) for v in vals])
    torch.cuda.synchronize()
    for k in range(4):
        ref = oracle.div_n(oracle.addn([t.cpu().numpy() for t in vals[k]]), 3)
        for r in range(3):
            assert np.array_equal(out[k][r].cpu().numpy().view(np.uint32), ref.view(np.uint32))


def test_gather_concatenates_replicas_along_any_axis():
    """CrossDeviceOps._gather_implementation (newer TF): exact copies, rank order, any axis (3
    emulated replicas on cuda:0; -0.0 and NaN payloads survive because nothing is added)."""
    ops = cdo.B200CrossDeviceOps()
    g = torch.Generator(device="cuda").manual_seed(9)
    vals = [torch.randn(5, 7, generator=g, device="cuda") for _ in range(3)]
    vals[1][0, 0] = -0.0
    for axis in (0, 1):
        outs = ops.gather(vals, axis=axis)
        torch.cuda.synchronize()
        want = torch.cat(vals, dim=axis)
        assert len(outs) == 3
        for o in outs:
            assert torch.equal(o.view(torch.int32), want.view(torch.int32))
    scal = ops.gather([torch.tensor(float(r), device="cuda") for r in range(3)], axis=0)
    assert scal[0].tolist() == [0.0, 1.0, 2.0]


def test_rank_context_cross_device_ops_single_worker():
    """The MWMS-shaped constructor with one worker: batch_reduce, _all_reduce (tensor and list) and
    gather all pass through the library and return the input (SUM over one replica)."""
    ops = cdo.B200CrossDeviceOps(rank=0, nranks=1, device=0)
    a, b = torch.randn(1000, device="cuda"), torch.randn(3, 5, device="cuda")
    out = ops.batch_reduce(cdo.ReduceOp.SUM, [([a], None), ([b], None)])
    assert torch.equal(out[0][0], a) and torch.equal(out[1][0], b)
    assert torch.equal(ops._all_reduce(cdo.ReduceOp.MEAN, a, 0, None), a)
    red = ops._all_reduce("SUM", [a, b])
    assert torch.equal(red[0], a) and torch.equal(red[1], b)
    assert torch.equal(ops.gather([b], axis=0)[0], b)
    with pytest.raises(ValueError):
        ops.batch_reduce(cdo.ReduceOp.SUM, [([a, a], None)])     # one process per GPU: one local replica


def test_two_fused_adam_optimizers_over_equal_shapes_do_not_alias():
    """Two FusedDistributedAdam instances over models with the same parameter sizes (an EMA or
    teacher copy) own separate plans: separate weights, separate Adam moments."""
    from b200dp import hvd
    hvd.shutdown()
    hvd.init()
    rng = np.random.default_rng(8)
    shapes = [(300, 40), (40,)]
    init = [[rng.standard_normal(s).astype(F) * 0.05 for s in shapes] for _ in range(2)]
    models = [[torch.nn.Parameter(torch.from_numpy(x).cuda()) for x in ws] for ws in init]
    opts = [hvd.FusedDistributedAdam(ps, lr=1e-3) for ps in models]
    assert opts[0].plan is not opts[1].plan
    refs = [[np.concatenate([x.ravel() for x in ws]), None, None] for ws in init]
    for ref in refs:
        ref[1], ref[2] = np.zeros_like(ref[0]), np.zeros_like(ref[0])
    for step in (1, 2):
        for i in (0, 1):
            gs = [rng.standard_normal(s).astype(F) * 1e-2 for s in shapes]
            for p, g in zip(models[i], gs):
                p.grad = torch.from_numpy(g).cuda()
            opts[i].step()
            oracle.adam_(refs[i][0], refs[i][1], refs[i][2], np.concatenate([g.ravel() for g in gs]), lr=1e-3, step=step)
    torch.cuda.synchronize()
    for i in (0, 1):
        got = np.concatenate([p.detach().cpu().numpy().ravel() for p in models[i]])
        assert np.array_equal(got.view(np.uint32), refs[i][0].view(np.uint32)), "model %d" % i


def test_hvd_broadcast_parameters_single_rank_is_a_noop():
    from b200dp import hvd
    hvd.shutdown()
    hvd.init()
    p = torch.nn.Parameter(torch.arange(10.0, device="cuda"))
    hvd.broadcast_parameters([p], root_rank=0)
    assert p.tolist() == list(range(10))
