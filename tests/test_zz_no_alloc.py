"""b2dp_adam_apply_noalloc -- the entry point for callers that cannot allocate device memory.

Some users of this library run where cudaMalloc is not available to them: a TensorFlow or
Horovod custom op must take every device byte from the framework's allocator, a
latency-critical training loop cannot afford the device-wide serialisation an allocation
causes, and some deployments pin the memory budget before the job starts.  For them the
library offers the single-GPU fused Adam apply on arrays they already own, with the
reduction scratch living in a fixed static device array inside the library.

This test calls that entry point through the C ABI twice in a row (two optimizer steps on
the same caller-owned w/m/v) and, after each call, compares with the existing plan-based
entry point (b2dp_pack + b2dp_step, n = 1) fed the same inputs:
  - w, m, v: bit-exact (both use the one-rounding-per-operation Adam arithmetic);
  - statistics (sum of squares, non-finite count): rel 1e-6, the tolerance the other GPU
    tests use for this statistic (the two entry points add the partial sums in different
    orders).
The file name sorts last on purpose: a `pytest -x` run has executed every other test
before it gets here.  It also passes alone in a fresh process (nothing here depends on
state left by other tests: the first call on a device does its own set-up)."""
import ctypes

import numpy as np
import pytest

import b200dp
from b200dp import _lib


def test_oversize_call_is_rejected_before_touching_the_gpu():
    """Sizes beyond the static scratch are refused with a clear error (runs without a GPU:
    the size check comes first)."""
    h = b200dp.adam_hyper(step=1)
    stats = (ctypes.c_double * 2)()
    rc = _lib.load().b2dp_adam_apply_noalloc(None, None, None, None, _lib.NOALLOC_MAX_ELEMS + 1,
                                             ctypes.byref(h), ctypes.cast(stats, ctypes.c_void_p), None)
    assert rc == _lib.ERR_TOO_LARGE
    msg = _lib.load().b2dp_last_error().decode()
    assert "exceed" in msg and str(_lib.NOALLOC_MAX_ELEMS) in msg


@pytest.mark.gpu
def test_two_calls_in_a_row_match_the_plan_based_entry_point():
    torch = pytest.importorskip("torch")
    assert torch.cuda.is_available()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    n = 3_000_003                       # ragged: exercises the scalar tail and a partial last chunk
    gen = torch.Generator(device=dev).manual_seed(2024)
    w0 = (torch.rand(n, generator=gen, device=dev) - 0.5) * 0.1

    # existing entry point: a 1-rank plan
    ctx = b200dp.Context.local([0])
    lay = b200dp.Layout([n], nranks=1)
    plan = b200dp.Plan(ctx, lay)
    plan.load_weights(0, [w0])

    # new entry point: caller-owned arrays (allocated here, by the caller, up front)
    w, m, v = w0.clone(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
    stats = torch.zeros(2, dtype=torch.float64, device=dev)
    # and the CPU oracle on the same inputs, so the new entry point is held to the reference
    # arithmetic directly and not only through another kernel of this library
    import oracle
    w_orc, m_orc, v_orc = w0.cpu().numpy().copy(), np.zeros(n, np.float32), np.zeros(n, np.float32)

    for step, lr in ((1, 1e-3), (2, 1e-4)):
        g = torch.randn(n, generator=gen, device=dev) * 1e-2
        if step == 2:
            g[12345] = float("inf")     # one non-finite element: must be counted by both
        h = b200dp.adam_hyper(lr=lr, step=step, want_stats=True)
        plan.pack(0, [g])
        plan.step(0, "adam", h)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        b200dp.adam_apply_noalloc(w, m, v, g, h, stats)
        torch.cuda.synchronize()
        if step == 2:                   # after the one-time set-up of call 1: no device memory is taken
            assert torch.cuda.mem_get_info()[0] == free_before

        ref_w = plan.tensor(0, b200dp.BUF_WEIGHTS)[:n]
        ref_m = plan.tensor(0, b200dp.BUF_STATE0)[:n]
        ref_v = plan.tensor(0, b200dp.BUF_STATE1)[:n]
        ref_stats = plan.tensor(0, b200dp.BUF_STATS).cpu().numpy()
        for name, got, ref in (("w", w, ref_w), ("m", m, ref_m), ("v", v, ref_v)):
            assert torch.equal(got.view(torch.int32), ref.view(torch.int32)), "%s differs at step %d" % (name, step)
        if step == 1:                   # finite inputs: bit-exact against the oracle's TF-form Adam as well
            oracle.adam_(w_orc, m_orc, v_orc, g.cpu().numpy(), lr=lr, step=step)
            for name, got, ref in (("w", w, w_orc), ("m", m, m_orc), ("v", v, v_orc)):
                assert np.array_equal(got.cpu().numpy().view(np.uint32), ref.view(np.uint32)), \
                    "%s differs from the oracle at step %d" % (name, step)
            assert abs(stats.cpu().numpy()[0] - oracle.sumsq(g.cpu().numpy())) <= 1e-6 * oracle.sumsq(g.cpu().numpy())
        got_stats = stats.cpu().numpy()
        if step == 1:
            assert abs(got_stats[0] - ref_stats[0]) <= 1e-6 * ref_stats[0]     # tolerance: rel 1e-6
            assert got_stats[1] == ref_stats[1] == 0
        else:
            assert np.isinf(got_stats[0]) and np.isinf(ref_stats[0])
            assert got_stats[1] == ref_stats[1] == 1


@pytest.mark.gpu
def test_argument_errors_and_empty_call():
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    h = b200dp.adam_hyper(step=1)
    stats = torch.ones(2, dtype=torch.float64, device=dev)
    e = torch.empty(0, device=dev)
    b200dp.adam_apply_noalloc(e, e, e, e, h, stats)           # empty call: statistics are zeroed
    torch.cuda.synchronize()
    assert stats.tolist() == [0.0, 0.0]
    x = torch.zeros(1025, device=dev)
    with pytest.raises(b200dp.B200dpError):                    # 4-byte-aligned view: rejected, not mis-read
        b200dp.adam_apply_noalloc(x[1:], x[1:], x[1:], x[1:], h, stats)
    with pytest.raises(b200dp.B200dpError):
        b200dp.adam_apply_noalloc(x, x, x, x, b200dp.adam_hyper(step=0), stats)
