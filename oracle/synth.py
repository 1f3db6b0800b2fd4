"""Seeded synthetic gradients of SURVEY.md 8d (TEST INFRASTRUCTURE ONLY).

For rank r, tensor k, step s the generator is seeded 1234 + 1000*r + k (+ 1_000_003*s, so
that the three steps of a parity run see different data).  Three distributions:

  D1  N(0,1) * 1e-3                                   timing and default parity
  D2  (-1)^r * 1e4 + N(0,1) * 1e-3                    cancellation stress: exposes summation order
  D3  +-2^U(-24,8) with 1 % exact zeros               dynamic range, bf16 subnormals

Generated with torch on `device`: on a GPU box every rank can rebuild every rank's gradients in
milliseconds (same GPU model and torch build => same Philox stream), on CPU the tests use the
CPU generator.  The two devices give DIFFERENT numbers for one seed; a run uses one of them.
"""
import torch


def rank_grad(dist, rank, step, numels, device="cpu"):
    """Flat fp32 gradient of one rank (tensors laid end to end, no padding)."""
    dev = torch.device(device)
    gen = torch.Generator(device=dev)
    out = torch.empty(sum(numels), dtype=torch.float32, device=dev)
    o = 0
    for k, m in enumerate(numels):
        if m == 0:
            continue
        gen.manual_seed(1234 + 1000 * rank + k + 1_000_003 * step)
        seg = out[o:o + m]
        if dist == "D1":
            seg.normal_(0.0, 1.0, generator=gen).mul_(1e-3)
        elif dist == "D2":
            seg.normal_(0.0, 1.0, generator=gen).mul_(1e-3).add_(1e4 if rank % 2 == 0 else -1e4)
        elif dist == "D3":
            e = torch.empty(m, device=dev).uniform_(-24.0, 8.0, generator=gen)
            sgn = torch.empty(m, device=dev).uniform_(0.0, 1.0, generator=gen)
            z = torch.empty(m, device=dev).uniform_(0.0, 1.0, generator=gen)
            seg.copy_(torch.exp2(e) * torch.where(sgn < 0.5, -1.0, 1.0) * (z >= 0.01))
        else:
            raise ValueError("unknown distribution %r" % (dist,))
        o += m
    return out


def all_ranks(dist, n, step, numels, device="cpu"):
    return [rank_grad(dist, r, step, numels, device) for r in range(n)]
