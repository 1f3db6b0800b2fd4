"""Pins the CPU oracle (oracle/b2dp_oracle.c) -- the checker every GPU parity test uses.

The reference has no tests, golden vectors or runnable implementation of this path
(parity unpinned), so the oracle is pinned by: the committed golden vectors
(tests/golden, produced by the independent numpy restatement), torch's bf16 cast, and
torch.optim.Adam / SGD under a stated tolerance (their formulas differ from TF's in
epsilon placement / lr placement, so those two are tolerance checks, not bitwise)."""
import numpy as np
import pytest

import oracle
from oracle import oracle_np as onp

F = np.float32


def _split(flat, numels):
    out, o = [], 0
    for n in numels:
        out.append(flat[o:o + n])
        o += n
    return out


def test_bf16_cast_known_answers(golden):
    x = golden["bf16_in"]
    assert np.array_equal(oracle.cast_bf16(x), golden["bf16_bits"])
    assert np.array_equal(onp.f32_to_bf16_bits(x), golden["bf16_bits"])
    # hand-checked: 1.00390625 = 1 + 2^-8 is a tie -> even (1.0); 1.01171875 = 1+3*2^-8 tie -> 1.015625
    assert golden["bf16_bits"][3] == 0x3F80 and golden["bf16_bits"][5] == 0x3F82
    assert golden["bf16_bits"][13] == 0x7F80  # FLT_MAX rounds to +inf


def test_bf16_cast_matches_torch():
    import torch
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.standard_normal(100000).astype(F) * F(10.0) ** rng.integers(-30, 30, 100000).astype(F),
                        np.array([0.0, -0.0, np.inf, -np.inf, 1e-40, -1e-45], dtype=F)])
    ref = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(oracle.cast_bf16(x), ref)
    assert np.array_equal(onp.f32_to_bf16_bits(x), ref)


@pytest.mark.parametrize("wire", ["f32", "bf16"])
def test_sum_matches_golden(golden, wire):
    for s in range(golden["grads"].shape[0]):
        got = oracle.addn(list(golden["grads"][s]), bf16_wire=(wire == "bf16"))
        assert np.array_equal(got.view(np.uint32), golden["sum_" + wire][s].view(np.uint32))


def test_sum_is_order_dependent_and_rank_ordered(golden):
    """D2 tensors cancel (+1e4, -1e4, ...): a different association changes bits, so a
    bitwise match really pins the rank order ((g0+g1)+g2)+g3."""
    g = list(golden["grads"][0])
    fwd = oracle.addn(g)
    rev = oracle.addn(g[::-1])
    pair = ((g[0] + g[2]) + (g[1] + g[3])).astype(F)
    assert not np.array_equal(fwd, rev) or not np.array_equal(fwd, pair)
    manual = ((g[0] + g[1]).astype(F) + g[2]).astype(F) + g[3]
    assert np.array_equal(fwd, manual.astype(F))


@pytest.mark.parametrize("wire", ["f32", "bf16"])
def test_adam_three_steps_matches_golden(golden, wire):
    w = golden["w0"].copy()
    m, v = np.zeros_like(w), np.zeros_like(w)
    for s in range(3):
        assert oracle.adam_alpha(float(golden["lrs"][s]), 0.9, 0.999, s + 1) == golden["adam_alpha"][s]
        g = oracle.addn(list(golden["grads"][s]), bf16_wire=(wire == "bf16"))
        oracle.adam_(w, m, v, g, lr=float(golden["lrs"][s]), step=s + 1)
    for name, arr in (("w", w), ("m", m), ("v", v)):
        assert np.array_equal(arr.view(np.uint32), golden["adam_%s_%s" % (name, wire)].view(np.uint32)), name


def test_hvd_average_matches_golden(golden):
    w = golden["w0"].copy()
    m, v = np.zeros_like(w), np.zeros_like(w)
    n = golden["grads"].shape[1]
    for s in range(3):
        g = oracle.scale(oracle.addn(list(golden["grads"][s])), 1.0 / n)
        oracle.adam_(w, m, v, g, lr=float(golden["lrs"][s]), step=s + 1)
    assert np.array_equal(w.view(np.uint32), golden["hvdavg_adam_w_f32"].view(np.uint32))


def test_mean_div_equals_scale_for_power_of_two():
    rng = np.random.default_rng(1)
    x = rng.standard_normal(4096).astype(F)
    for n in (2, 4, 8):
        assert np.array_equal(oracle.div_n(x.copy(), n), oracle.scale(x.copy(), 1.0 / n))


@pytest.mark.parametrize("nest", [False, True])
@pytest.mark.parametrize("wire", ["f32", "bf16"])
def test_sgdm_matches_golden(golden, nest, wire):
    w = golden["w0"].copy()
    a = np.zeros_like(w)
    for s in range(3):
        g = oracle.addn(list(golden["grads"][s]), bf16_wire=(wire == "bf16"))
        oracle.sgdm_(w, a, g, lr=float(golden["sgd_lrs"][s]), mu=0.9, nesterov=nest, wd=1e-4)
    tag = "sgdm_nest" if nest else "sgdm"
    assert np.array_equal(w.view(np.uint32), golden[tag + "_w_" + wire].view(np.uint32))
    assert np.array_equal(a.view(np.uint32), golden[tag + "_a_" + wire].view(np.uint32))


def test_c_oracle_equals_numpy_twin_on_random_inputs():
    rng = np.random.default_rng(7)
    n = 50001
    gs = [(rng.standard_normal(n) * 10.0 ** rng.integers(-6, 3)).astype(F) for _ in range(8)]
    for bw in (False, True):
        assert np.array_equal(oracle.addn(gs, prescale=0.125, bf16_wire=bw), onp.addn(gs, prescale=0.125, bf16_wire=bw))
    w = rng.standard_normal(n).astype(F)
    m = (rng.standard_normal(n) * 1e-3).astype(F)
    v = (rng.standard_normal(n) ** 2 * 1e-6).astype(F)
    g = oracle.addn(gs)
    w2, m2, v2 = onp.adam(w, m, v, g, lr=3e-4, step=17)
    oracle.adam_(w, m, v, g, lr=3e-4, step=17)
    assert np.array_equal(w, w2) and np.array_equal(m, m2) and np.array_equal(v, v2)


def test_adam_close_to_torch_adam():
    """torch.optim.Adam puts eps outside the bias-corrected sqrt, TF inside the fused
    alpha (the "epsilon-hat" form): same update up to O(eps/sqrt(v)).  Gradients are kept
    away from zero (|g| >= 5e-3) so that term is <= 1e-3 of a ~1e-3 update: atol 5e-6."""
    import torch
    rng = np.random.default_rng(3)
    n = 10000
    w0 = rng.standard_normal(n).astype(F)
    grads = []
    for _ in range(3):
        z = rng.standard_normal(n)
        grads.append((np.sign(z) * (0.5 + np.abs(z)) * 1e-2).astype(F))
    w, m, v = w0.copy(), np.zeros(n, F), np.zeros(n, F)
    p = torch.nn.Parameter(torch.from_numpy(w0.copy()))
    opt = torch.optim.Adam([p], lr=1e-3, betas=(0.9, 0.999), eps=1e-7)
    for s, g in enumerate(grads):
        oracle.adam_(w, m, v, g, lr=1e-3, step=s + 1)
        p.grad = torch.from_numpy(g.copy())
        opt.step()
    np.testing.assert_allclose(w, p.detach().numpy(), rtol=0, atol=5e-6)


def test_sgdm_close_to_torch_sgd():
    """Keras keeps lr inside the accumulator (a = mu*a - lr*g), torch outside; identical for
    a constant lr up to rounding -- tolerance 1e-5."""
    import torch
    rng = np.random.default_rng(4)
    n = 10000
    w0 = rng.standard_normal(n).astype(F)
    grads = [(rng.standard_normal(n) * 1e-2).astype(F) for _ in range(4)]
    w, a = w0.copy(), np.zeros(n, F)
    p = torch.nn.Parameter(torch.from_numpy(w0.copy()))
    opt = torch.optim.SGD([p], lr=0.05, momentum=0.9)
    for g in grads:
        oracle.sgdm_(w, a, g, lr=0.05, mu=0.9)
        p.grad = torch.from_numpy(g.copy())
        opt.step()
    np.testing.assert_allclose(w, p.detach().numpy(), rtol=1e-5, atol=1e-6)


def test_nonfinite_propagates():
    g0 = np.array([np.inf, 1.0, np.nan, 1.0], dtype=F)
    g1 = np.array([-np.inf, np.inf, 1.0, 2.0], dtype=F)
    s = oracle.addn([g0, g1])
    assert np.isnan(s[0]) and np.isinf(s[1]) and np.isnan(s[2]) and s[3] == 3.0


def test_empty_input_is_a_noop():
    e = np.zeros(0, dtype=F)
    assert oracle.addn([e, e]).size == 0
    oracle.adam_(e, e.copy(), e.copy(), e.copy())


# ------------------------------------------------------------------ bounds for order-free paths
@pytest.mark.parametrize("dist", ["D1", "D3"])
@pytest.mark.parametrize("order", ["reversed", "pairwise", "float64"])
def test_interval_bound_holds_for_any_summation_order(dist, order):
    """oracle/bounds.py: a sum taken in another order (what the NVSwitch or NCCL may do) stays inside
    (n-1)*2^-24*sum|g_i| of the rank-order sum, and three Adam steps driven by it stay inside the
    per-element interval AdamBound propagates -- while that interval remains a few ulp of |w| wide."""
    import torch
    from oracle import bounds, synth
    numels = [64, 1000, 7, 0, 1027, 512, 40000]
    N, n = sum(numels), 8
    w = ((torch.rand(N, generator=torch.Generator().manual_seed(42)) - 0.5) * 0.1).numpy()
    m, v = np.zeros(N, F), np.zeros(N, F)
    w2, m2, v2 = w.copy(), m.copy(), v.copy()
    tr = bounds.AdamBound(N)
    for s, lr in enumerate((1e-3, 1e-3, 1e-4)):
        hg = [g.numpy() for g in synth.all_ranks(dist, n, s, numels)]
        gs = oracle.addn(hg)
        if order == "reversed":
            gh = oracle.addn(hg[::-1])
        elif order == "float64":
            gh = np.sum(np.stack(hg).astype(np.float64), axis=0).astype(F)
        else:
            a = [(hg[2 * i] + hg[2 * i + 1]).astype(F) for i in range(4)]
            gh = ((a[0] + a[1]).astype(F) + (a[2] + a[3]).astype(F)).astype(F)
        B = bounds.sum_bound_f32(hg, gs)
        assert np.all(np.abs(gh.astype(np.float64) - gs) <= B)
        mo, vo = m.copy(), v.copy()
        oracle.adam_(w, m, v, gs, lr=lr, step=s + 1)
        oracle.adam_(w2, m2, v2, gh, lr=lr, step=s + 1)
        tr.step(gs, B, mo, vo, m, v, w, oracle.adam_alpha(lr, 0.9, 0.999, s + 1))
    for got, ref, bound in ((w2, w, tr.Ew), (m2, m, tr.Em + 1e-38), (v2, v, tr.Ev + 1e-38)):
        c = bounds.check(got, ref, bound, "x")
        assert c["ok"] and c["max_err_over_bound"] <= 1.0, c
    assert np.median(tr.Ew) <= 1e-7          # informative: |w| <= 0.05 has ulp 3.7e-9
    # and it is not vacuous: a gradient error 1000 times the bound is caught
    w3, m3, v3 = w.copy(), m.copy(), v.copy()
    oracle.adam_(w3, m3, v3, (gs + 1000 * B).astype(F), lr=1e-3, step=4)
    w4, m4, v4 = w.copy(), m.copy(), v.copy()
    mo, vo = m.copy(), v.copy()
    oracle.adam_(w4, m4, v4, gs, lr=1e-3, step=4)
    tr.step(gs, B, mo, vo, m4, v4, w4, oracle.adam_alpha(1e-3, 0.9, 0.999, 4))
    assert not bounds.check(w3, w4, tr.Ew, "x")["ok"]


def test_synthetic_gradients_are_reproducible_and_shaped_as_specified():
    from oracle import synth
    numels = [64, 0, 1000, 4096]
    a = synth.all_ranks("D3", 2, 1, numels)
    b = synth.all_ranks("D3", 2, 1, numels)
    assert all(np.array_equal(x.numpy(), y.numpy()) for x, y in zip(a, b))
    d3 = a[0].numpy()
    nz = np.abs(d3[d3 != 0])
    assert 0 < (d3 == 0).mean() < 0.05 and nz.min() >= 2.0 ** -24.01 and nz.max() <= 2.0 ** 8.01
    d2 = synth.all_ranks("D2", 2, 0, numels)
    assert d2[0].mean() > 9e3 and d2[1].mean() < -9e3
    d1 = synth.rank_grad("D1", 0, 0, numels).numpy()
    assert abs(d1.std() - 1e-3) < 1e-4 and not np.array_equal(d1, synth.rank_grad("D1", 1, 0, numels).numpy())
