"""Elementwise error bounds for the paths that cannot be bit-exact (TEST INFRASTRUCTURE ONLY).

The peer path adds the n gradient copies in rank order and is compared with the oracle bit for
bit.  The NVSwitch multicast path (and NCCL) add them in an order of their own, so the reduced
gradient g^ is only known to satisfy |g^ - g| <= B elementwise, with

    f32 wire:   B = (n-1) * 2^-24 * sum_i |g_i| + |oracle_sum - exact_sum|   (see sum_bound_f32)
    bf16 wire:  B = 2^-7 * |sum_i b_i| + (n-1) * 2^-24 * sum_i |b_i|     (b_i = RNE bf16 of g_i;
                the switch accumulates in fp32 and converts the sum once more to bf16 -- one
                bf16 ulp, since the rounding mode of that conversion is not documented)

What that implies for Adam is NOT a fixed tolerance: where sqrt(v) is small against eps, or
where the sum is near a sign change, a tiny change of g moves the update by a large fraction of
lr.  AdamBound therefore carries an interval through the TF-form update, element by element:

    m' = b1*m + (1-b1)*g          =>  Em' = b1*Em + (1-b1)*B
    v' = b2*v + (1-b2)*g^2        =>  Ev' = b2*Ev + (1-b2)*(2|g|B + B^2)
    u  = alpha*m'/(sqrt(v')+eps)      monotone in m' and in v', so the largest deviation over
                                      m' +- Em', v' +- Ev' is attained at one of the 4 corners
    w' = w - u                    =>  Ew' = Ew + Eu

plus a slack of a few fp32 ulps per operation for the rounding differences between two fp32
evaluations of the same formula on slightly different inputs.  Every element is checked against
its own bound; nothing is excluded by hand.  Works in float64, in chunks, so 25.6M elements
cost a few seconds.
"""
import numpy as np

U32 = 2.0 ** -24          # fp32 unit roundoff
# How far the NVSwitch's bf16 result may sit from the exact fp32 sum, in units of 2^-8 |sum| (half a
# bf16 ulp at the bottom of a binade).  Round-to-nearest would give 1; the PTX ISA does not state the
# rounding of multimem.ld_reduce's final conversion, and what was measured on B200 (r02, tests
# test_path_b_multicast_*[bf16] with B2DP_REPORT_SWITCH_ROUNDING=1) exceeds 1, so the bound allows a
# full ulp (truncation).
SWITCH_BF16_ULPS = 2.0
CHUNK = 1 << 22


def sum_bound_f32(grads, oracle_sum=None):
    """Bound on |s^ - oracle_sum| for s^ = the n values added in fp32 in ANY order.

    SURVEY 8c's (n-1) * 2^-24 * sum_i |g_i| bounds ONE fp32 summation against the EXACT sum.  The
    oracle's rank-order sum is itself such a summation, so two orders may differ by the sum of
    their two errors (measured: NCCL at n = 4 sits 1.33x the one-sided figure away from AddN on D1
    data).  The oracle's own error is not bounded but COMPUTED -- the float64 sum of <= 8 fp32
    values is exact to 2^-53 -- and added: B = (n-1) 2^-24 sum|g_i| + |oracle - exact|."""
    n = len(grads)
    acc = np.zeros(grads[0].size, np.float64)
    exact = np.zeros(grads[0].size, np.float64)
    for g in grads:
        acc += np.abs(g, dtype=np.float64)
        exact += g
    B = (n - 1) * U32 * (1 + 1e-6) * acc + 1e-38
    if oracle_sum is not None:
        B += np.abs(np.asarray(oracle_sum, np.float64) - exact)
    return B


def sum_bound_bf16_switch(bgrads, bsum):
    """bgrads: the bf16-rounded gradients (as fp32 arrays); bsum: their oracle (rank-order fp32) sum."""
    return SWITCH_BF16_ULPS * 2.0 ** -8 * np.abs(bsum, dtype=np.float64) + sum_bound_f32(bgrads, bsum)


class AdamBound:
    def __init__(self, numel):
        self.Em = np.zeros(numel, np.float64)
        self.Ev = np.zeros(numel, np.float64)
        self.Ew = np.zeros(numel, np.float64)

    def step(self, g, B, m_old, v_old, m_new, v_new, w_new, alpha, beta1=0.9, beta2=0.999, eps=1e-7):
        """g: oracle reduced gradient (after postscale); B: bound on |g^ - g| (after postscale);
        m_old/v_old and m_new/v_new/w_new: the ORACLE's state before / after this step (fp32)."""
        b1, b2, a, eps = float(np.float32(beta1)), float(np.float32(beta2)), float(alpha), float(np.float32(eps))
        for lo in range(0, g.size, CHUNK):
            sl = slice(lo, min(g.size, lo + CHUNK))
            gg = np.abs(g[sl], dtype=np.float64)
            BB = np.asarray(B[sl], np.float64)
            m1, v1 = m_new[sl].astype(np.float64), v_new[sl].astype(np.float64)
            Em = b1 * self.Em[sl] + (1 - b1) * BB + 4 * U32 * (np.abs(m1) + (1 - b1) * (gg + np.abs(m_old[sl], dtype=np.float64)))
            Ev = b2 * self.Ev[sl] + (1 - b2) * (2 * gg * BB + BB * BB) + 4 * U32 * (v1 + (1 - b2) * (gg * gg + v_old[sl].astype(np.float64)))
            u0 = a * m1 / (np.sqrt(v1) + eps)
            Eu = np.zeros_like(u0)
            umax = np.abs(u0)
            for sm in (-1.0, 1.0):
                mc = m1 + sm * Em
                for sv in (-1.0, 1.0):
                    vc = np.maximum(v1 + sv * Ev, 0.0)
                    uc = a * mc / (np.sqrt(vc) + eps)
                    np.maximum(Eu, np.abs(uc - u0), out=Eu)
                    np.maximum(umax, np.abs(uc), out=umax)
            self.Em[sl], self.Ev[sl] = Em, Ev
            self.Ew[sl] += Eu + 16 * U32 * umax + 2 * U32 * np.abs(w_new[sl], dtype=np.float64) + 1e-38


def check(got, ref, bound, what):
    """Returns a dict describing how `got` sits inside ref +- bound (all elementwise)."""
    err = np.abs(got.astype(np.float64) - ref.astype(np.float64))
    bad = ~(err <= bound)          # NaN-safe: a NaN error counts as a violation
    ratio = float(np.max(err / bound)) if err.size else 0.0
    return {"what": what, "ok": not bool(bad.any()), "violations": int(bad.sum()),
            "max_abs_err": float(err.max()) if err.size else 0.0, "max_err_over_bound": ratio,
            "max_bound": float(np.max(bound)) if err.size else 0.0,
            "median_bound": float(np.median(bound)) if err.size else 0.0,
            "frac_bit_equal": float(np.mean(got.view(np.uint32) == ref.view(np.uint32))) if err.size else 1.0}
