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
