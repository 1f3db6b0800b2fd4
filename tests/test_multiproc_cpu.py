"""world_size-2 checks of the host-side multi-process logic, on CPU:
 - b2dp_fd_exchange moves real fds between processes (the transport a rank context uses
   for its peer-mappable blocks; here the fds are memfds);
 - under torch.distributed (gloo) every rank derives the same layout and the n-rank
   reduce-scatter -> apply -> all-gather slicing reproduces the oracle's one-copy update."""
import ctypes
import multiprocessing as mp
import os

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _fd_worker(rank, n, session, q):
    import sys
    sys.path.insert(0, ROOT)
    from b200dp import _lib
    L = _lib.load()
    fd = os.memfd_create("b2dp-test-%d" % rank)
    os.write(fd, b"hello from rank %d" % rank)
    fds = (ctypes.c_int * n)()
    rc = L.b2dp_fd_exchange(rank, n, session.encode(), fd, fds, 20000)
    got = {}
    if rc == 0:
        for r in range(n):
            got[r] = os.pread(fds[r], 64, 0).decode()   # pread: the open file description is shared
    q.put((rank, rc, got, L.b2dp_last_error().decode()))


@pytest.mark.parametrize("n", [2, 4])
def test_fd_exchange_between_processes(n):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    session = "t%d-%d" % (os.getpid(), n)
    procs = [ctx.Process(target=_fd_worker, args=(r, n, session, q)) for r in range(n)]
    for p in procs:
        p.start()
    results = [q.get(timeout=60) for _ in range(n)]
    for p in procs:
        p.join(30)
    for rank, rc, got, err in results:
        assert rc == 0, err
        assert got == {r: "hello from rank %d" % r for r in range(n)}


def test_fd_exchange_times_out_without_peer():
    import sys
    sys.path.insert(0, ROOT)
    from b200dp import _lib
    L = _lib.load()
    fd = os.memfd_create("lonely")
    fds = (ctypes.c_int * 2)()
    rc = L.b2dp_fd_exchange(0, 2, b"nobody-%d" % os.getpid(), fd, fds, 300)
    assert rc == _lib.ERR_TIMEOUT
    os.close(fd)


def _gloo_worker(rank, world, port, q):
    import sys
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    import b200dp
    import oracle
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        numels = [64, 1000, 7, 0, 1027, 512]
        L = b200dp.Layout(numels, nranks=world, bucket_bytes=4096)
        desc = [L.total_elems, L.shard_elems, [L.bucket_info(b) for b in range(L.num_buckets)]]
        all_desc = [None] * world
        dist.all_gather_object(all_desc, desc)
        same = all(d == all_desc[0] for d in all_desc)

        # n-rank simulation of reduce-scatter -> Adam on the owned slice -> all-gather,
        # with gloo moving the data; compared with the oracle's single logical copy.
        rng = np.random.default_rng(100 + rank)
        total = L.total_elems
        g = np.zeros(total, np.float32)
        for k, n in enumerate(numels):
            off = L.tensor_offset(k)
            g[off:off + n] = rng.standard_normal(n).astype(np.float32)
        w = np.linspace(-1, 1, total).astype(np.float32)
        m = np.zeros(total, np.float32)
        v = np.zeros(total, np.float32)
        gl = [torch.zeros(total) for _ in range(world)]
        dist.all_gather(gl, torch.from_numpy(g))
        gl = [t.numpy() for t in gl]
        w_new = w.copy()
        for b in range(L.num_buckets):
            info = L.bucket_info(b)
            lo = info["offset"] + rank * info["slice"]
            hi = lo + info["slice"]
            gs = oracle.addn([x[lo:hi] for x in gl])           # rank-order sum of my slice
            ws, ms, vs = w[lo:hi].copy(), m[lo:hi].copy(), v[lo:hi].copy()
            oracle.adam_(ws, ms, vs, gs, step=1)
            w_new[lo:hi] = ws
        # all-gather of the updated slices
        pieces = [torch.zeros(total) for _ in range(world)]
        mask = np.zeros(total, np.float32)
        for b in range(L.num_buckets):
            info = L.bucket_info(b)
            lo = info["offset"] + rank * info["slice"]
            mask[lo:lo + info["slice"]] = 1
        dist.all_gather(pieces, torch.from_numpy(w_new * mask))
        w_gathered = sum(p.numpy() for p in pieces)
        # oracle: one logical copy
        w_ref, m_ref, v_ref = w.copy(), m.copy(), v.copy()
        oracle.adam_(w_ref, m_ref, v_ref, oracle.addn(gl), step=1)
        q.put((rank, same, bool(np.array_equal(w_gathered, w_ref))))
    finally:
        dist.destroy_process_group()


def test_gloo_two_ranks_layout_agreement_and_sharded_step():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(2)]
    for p in procs:
        p.join(30)
    for rank, same, ok in results:
        assert same, "ranks derived different layouts"
        assert ok, "sharded RS->apply->AG differs from the oracle's single-copy update"
