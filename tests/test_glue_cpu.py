"""Host-side behaviour of the reference-shaped glue (no GPU): error behaviour of the
CrossDeviceOps mirror, the TF-facing modules' import guards, and hvd's rank/size bootstrap
under a 2-process gloo group."""
import multiprocessing as mp
import os
import socket

import pytest

import b200dp
from b200dp import cross_device_ops as cdo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tf_facing_modules_say_why_they_cannot_load():
    with pytest.raises(ImportError, match="TensorFlow"):
        import b200dp.hvd_keras  # noqa: F401
    with pytest.raises(ImportError, match="TensorFlow"):
        cdo.make_tf_cross_device_ops()


def test_cross_device_ops_rejects_bad_values_like_upstream():
    import torch
    ops = cdo.B200CrossDeviceOps()
    a, b = torch.zeros(4), torch.zeros(5)
    with pytest.raises(ValueError):
        ops.batch_reduce(cdo.ReduceOp.SUM, [([a, b], None)])          # replicas disagree on shape
    with pytest.raises(ValueError):
        ops.batch_reduce(cdo.ReduceOp.SUM, [([a, a], None)])          # CPU tensors: no CPU path
    with pytest.raises(ValueError):
        ops.batch_reduce(cdo.ReduceOp.SUM, [([a, a], None), ([a], None)])   # replica counts differ
    with pytest.raises(ValueError):
        ops.batch_reduce(cdo.ReduceOp.SUM, [(a, None)])               # not a per-replica value
    with pytest.raises(ValueError):
        cdo.CommunicationOptions(bytes_per_pack=-1)
    assert ops.batch_reduce(cdo.ReduceOp.SUM, []) == []


def test_hvd_requires_init():
    from b200dp import hvd
    hvd.shutdown()
    with pytest.raises(ValueError, match="hvd.init"):
        hvd.rank()


def _hvd_worker(rank, world, port, q):
    import sys
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    from b200dp import hvd
    hvd.init()
    import torch.distributed as dist
    ok = (hvd.rank(), hvd.size(), hvd.local_rank()) == (rank, world, rank)
    sessions = [None] * world
    dist.all_gather_object(sessions, hvd._state["session"])
    q.put((rank, ok, len(set(sessions)) == 1))
    dist.destroy_process_group()


def test_hvd_bootstrap_two_processes_gloo():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_hvd_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(30)
    for rank, ok, same_session in res:
        assert ok and same_session


def test_learning_rate_warmup_follows_horovods_formula(monkeypatch):
    """hvd.callbacks.LearningRateWarmupCallback(0.1 * size, warmup_epochs=3) (imagenet-resnet50-hvd.py:115):
    lr(epoch) = initial_lr / size * (epoch * (size - 1) / warmup_epochs + 1), fractional epochs,
    set before every batch; exactly initial_lr once warm-up is over."""
    from b200dp import hvd

    class Opt:
        lr = -1.0

    for size in (1, 2, 8):
        monkeypatch.setitem(hvd._state, "size", size)
        monkeypatch.setitem(hvd._state, "init", True)
        init = 0.1 * size
        cb = hvd.callbacks.LearningRateWarmupCallback(init, warmup_epochs=3, steps_per_epoch=10, verbose=0)
        opt = Opt()
        cb.set_optimizer(opt)
        seen = []
        for epoch in range(5):
            cb.on_epoch_begin(epoch)
            for batch in range(10):
                cb.on_batch_begin(batch)
                seen.append(opt.lr)
            cb.on_epoch_end(epoch)
        for i, lr in enumerate(seen[:30]):
            e = i / 10.0
            assert lr == pytest.approx(init / size * (e * (size - 1) / 3 + 1), rel=1e-12)
        assert seen[0] == pytest.approx(0.1)                 # starts at the single-worker rate
        assert all(lr == init for lr in seen[30:])           # and stays at initial_lr afterwards
        assert all(b >= a for a, b in zip(seen, seen[1:]))
    with pytest.raises(ValueError, match="steps_per_epoch"):
        cb2 = hvd.callbacks.LearningRateWarmupCallback(0.1, warmup_epochs=3).set_optimizer(Opt())
        cb2.on_batch_begin(0)

    # a torch-style optimizer: param_groups carry the rate
    class TorchLike:
        param_groups = [{"lr": 0.0}, {"lr": 0.0}]

    monkeypatch.setitem(hvd._state, "size", 4)
    t = TorchLike()
    cb = hvd.callbacks.LearningRateWarmupCallback(0.4, warmup_epochs=2, steps_per_epoch=4).set_optimizer(t)
    cb.on_epoch_begin(1)
    cb.on_batch_begin(2)
    assert [g["lr"] for g in t.param_groups] == [pytest.approx(0.4 / 4 * (1.5 * 3 / 2 + 1))] * 2


def test_metric_average_and_broadcast_callbacks_are_noops_on_one_rank(monkeypatch):
    from b200dp import hvd
    monkeypatch.setitem(hvd._state, "size", 1)
    monkeypatch.setitem(hvd._state, "init", True)
    logs = {"loss": 1.5, "accuracy": 0.25, "name": "x"}
    hvd.callbacks.MetricAverageCallback().on_epoch_end(0, logs)
    assert logs == {"loss": 1.5, "accuracy": 0.25, "name": "x"}
    cb = hvd.callbacks.BroadcastGlobalVariablesCallback(0).set_optimizer(object(), params=[])
    cb.on_train_begin()
    assert cb.broadcast_done
    with pytest.raises(ValueError, match="set_optimizer"):
        hvd.callbacks.BroadcastGlobalVariablesCallback(0).on_batch_end(0)


def test_rank_context_cross_device_ops_argument_errors():
    with pytest.raises(ValueError, match="both rank and nranks"):
        cdo.B200CrossDeviceOps(rank=0)
    with pytest.raises(ValueError, match="outside"):
        cdo.B200CrossDeviceOps(rank=3, nranks=2)
    ops = cdo.B200CrossDeviceOps()
    with pytest.raises(ValueError, match="replica context"):
        ops._all_reduce(cdo.ReduceOp.SUM, None)

    class FakeExtended:
        _cross_device_ops = None

    class FakeStrategy:
        extended = FakeExtended()

    s = FakeStrategy()
    assert ops.swap_into(s) is s and s.extended._cross_device_ops is ops


def _session_worker(rank, world, port, q):
    import sys
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from b200dp import cross_device_ops as cdo
    eng = cdo.B200CrossDeviceOps(rank=rank, nranks=world, device=0)._engine
    q.put((rank, eng._session(), eng._session()))
    dist.destroy_process_group()


def test_rank_context_cross_device_ops_agree_on_a_session_two_processes_gloo():
    """One process per worker (the MWMS shape): the job-unique string the fd exchange needs is agreed
    through the process group -- same on every rank, different for every plan built."""
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_session_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(30)
    assert res[0][1] == res[1][1] and res[0][2] == res[1][2] and res[0][1] != res[0][2]
    with pytest.raises(ValueError, match="session"):
        cdo.B200CrossDeviceOps(rank=0, nranks=2, device=0)._engine._session()    # no group, no session string
