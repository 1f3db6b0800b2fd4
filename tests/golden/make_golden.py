"""Generates tests/golden/hotpath_golden.npz -- known-answer vectors for the hot path.

The reference (/root/reference) holds no golden vectors, no tests and no runnable
implementation of this path (TensorFlow/Horovod absent), so these vectors come from the
numpy restatement oracle/oracle_np.py of the published upstream algorithm: PARITY
UNPINNED against TF itself.  They pin (a) the C oracle, (b) the CUDA kernels and (c) any
later refactor to one fixed answer.  Inputs follow SURVEY.md section 8d: per (rank r,
tensor k) generator seeded 1234 + 1000*r + k (+ 100000*step), distributions D1/D2/D3
cycling over tensors; weights seed 42; 3 steps, lr schedule [1e-3, 1e-3, 1e-4].

Run from the repo root:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle_np as onp  # noqa: E402

F = np.float32
NUMELS = [64, 1000, 7, 0, 1027, 512]   # ragged on purpose: odd sizes, an empty tensor
NRANKS = 4
STEPS = 3
LRS = [1e-3, 1e-3, 1e-4]
SGD_LRS = [1e-2, 1e-2, 1e-3]


def synth_grad(rank, k, numel, step=0):
    """SURVEY.md 8d distributions; D = k % 3."""
    rng = np.random.default_rng(1234 + 1000 * rank + k + 100000 * step)
    d = k % 3
    if d == 0:    # D1: N(0,1)*1e-3
        g = rng.standard_normal(numel) * 1e-3
    elif d == 1:  # D2: cancellation stress, exposes summation order
        g = (-1.0) ** rank * 1e4 + rng.standard_normal(numel) * 1e-3
    else:         # D3: +-2^U(-24,8), 1% exact zeros (dynamic range, bf16 subnormal-ish)
        g = np.sign(rng.standard_normal(numel)) * np.exp2(rng.uniform(-24, 8, numel))
        g[rng.uniform(size=numel) < 0.01] = 0.0
    return g.astype(F)


def synth_weights(numels):
    rng = np.random.default_rng(42)
    return [rng.uniform(-0.05, 0.05, n).astype(F) for n in numels]


def main():
    total = sum(NUMELS)
    grads = np.zeros((STEPS, NRANKS, total), dtype=F)
    for s in range(STEPS):
        for r in range(NRANKS):
            grads[s, r] = np.concatenate([synth_grad(r, k, n, s) for k, n in enumerate(NUMELS)])
    w0 = np.concatenate(synth_weights(NUMELS))
    out = {"numels": np.array(NUMELS, dtype=np.int64), "grads": grads, "w0": w0, "lrs": np.array(LRS, dtype=F),
           "sgd_lrs": np.array(SGD_LRS, dtype=F)}

    for wire in ("f32", "bf16"):
        bw = wire == "bf16"
        # Adam, SUM reduction (tf.distribute: loss pre-scaled by 1/n, so SUM is the mean gradient)
        w, m, v = w0.copy(), np.zeros_like(w0), np.zeros_like(w0)
        sums, alphas = [], []
        for s in range(STEPS):
            g = onp.addn(list(grads[s]), bf16_wire=bw)
            sums.append(g)
            alphas.append(onp.adam_alpha(LRS[s], 0.9, 0.999, s + 1))
            w, m, v = onp.adam(w, m, v, g, lr=LRS[s], step=s + 1)
        out["sum_" + wire] = np.stack(sums)
        out["adam_w_" + wire], out["adam_m_" + wire], out["adam_v_" + wire] = w, m, v
        out["adam_alpha"] = np.array(alphas, dtype=F)
        # hvd Average: sum * float(1/n), then Adam
        w, m, v = w0.copy(), np.zeros_like(w0), np.zeros_like(w0)
        for s in range(STEPS):
            g = (onp.addn(list(grads[s]), bf16_wire=bw) * F(1.0 / NRANKS)).astype(F)
            w, m, v = onp.adam(w, m, v, g, lr=LRS[s], step=s + 1)
        out["hvdavg_adam_w_" + wire] = w
        # SGD momentum 0.9 + weight decay 1e-4, plain and nesterov
        for nest in (False, True):
            w, a = w0.copy(), np.zeros_like(w0)
            for s in range(STEPS):
                g = onp.addn(list(grads[s]), bf16_wire=bw)
                w, a = onp.sgdm(w, a, g, lr=SGD_LRS[s], mu=0.9, nesterov=nest, wd=1e-4)
            tag = "sgdm_nest" if nest else "sgdm"
            out[tag + "_w_" + wire], out[tag + "_a_" + wire] = w, a
    # bf16 cast known answers (bit patterns) incl. ties, NaN, inf, subnormals
    specials = np.array([0.0, -0.0, 1.0, 1.00390625, 1.005859375, 1.01171875, np.inf, -np.inf, np.nan,
                         3.0e-39, 1e-45, 65504.0, 3.3895314e38, 3.4028235e38, -2.5, 0.1], dtype=F)
    out["bf16_in"] = specials
    out["bf16_bits"] = onp.f32_to_bf16_bits(specials)
    path = os.path.join(ROOT, "tests", "golden", "hotpath_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
