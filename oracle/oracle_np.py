"""numpy float32 restatement of oracle/b2dp_oracle.c (TEST INFRASTRUCTURE ONLY).

Written independently of the C file so that the two can pin each other: numpy
float32 array arithmetic rounds to nearest-even after every ufunc and never
fuses a multiply with an add, which is the oracle's arithmetic contract.
PARITY UNPINNED against TensorFlow/Horovod themselves (absent from this image);
see the C header for what each function restates and the reference call sites
(imagenet-resnet50-mirror.py:21,76,81; imagenet-resnet50-hvd.py:99,101).
"""
import numpy as np

F = np.float32


def f32_to_bf16_bits(x):
    """RNE fp32 -> bf16, returned as uint16 bit patterns (NaN stays quiet NaN)."""
    u = np.ascontiguousarray(x, dtype=F).view(np.uint32)
    nan = (u & np.uint32(0x7FFFFFFF)) > np.uint32(0x7F800000)
    lsb = (u >> np.uint32(16)) & np.uint32(1)
    r = (u + np.uint32(0x7FFF) + lsb) >> np.uint32(16)
    r = np.where(nan, (u >> np.uint32(16)) | np.uint32(0x0040), r)
    return r.astype(np.uint16)


def bf16_bits_to_f32(h):
    return (np.asarray(h, dtype=np.uint16).astype(np.uint32) << np.uint32(16)).view(F)


def bf16_round(x):
    return bf16_bits_to_f32(f32_to_bf16_bits(x))


def addn(grads, prescale=1.0, bf16_wire=False):
    """((g0+g1)+g2)+... in fp32, replica order (TF AddN / ReductionToOneDevice)."""
    ps = F(prescale)
    acc = None
    for g in grads:
        x = np.asarray(g, dtype=F) * ps
        if bf16_wire:
            x = bf16_round(x)
        acc = x.copy() if acc is None else (acc + x).astype(F)
    return acc


def adam_alpha(lr, beta1, beta2, step):
    """lr*sqrt(1-beta2^t)/(1-beta1^t) in fp32.  The two powers are the correctly
    rounded fp32 values of the exact powers (computed in float64, rounded once), which is
    what glibc's powf returns; numpy's own float32 power is off by an ulp at some steps."""
    b1p = F(np.float64(F(beta1)) ** int(step))
    b2p = F(np.float64(F(beta2)) ** int(step))
    num = np.sqrt(F(1.0) - b2p, dtype=F)
    den = F(1.0) - b1p
    return F(F(F(lr) * num) / den)


def adam(w, m, v, g, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7, step=1, alpha=None):
    """TF fused Adam, returns (w, m, v) new arrays.  `alpha` overrides the host
    scalar (lets a test feed the C library's own alpha so powf differences between
    libm and numpy cannot matter)."""
    w, m, v, g = (np.asarray(a, dtype=F) for a in (w, m, v, g))
    a = adam_alpha(lr, beta1, beta2, step) if alpha is None else F(alpha)
    omb1 = F(1.0) - F(beta1)
    omb2 = F(1.0) - F(beta2)
    m2 = m + (g - m) * omb1
    v2 = v + (g * g - v) * omb2
    w2 = w - (a * m2) / (np.sqrt(v2) + F(eps))
    return w2.astype(F), m2.astype(F), v2.astype(F)


def sgdm(w, a, g, lr=0.01, mu=0.9, nesterov=False, wd=0.0):
    """Keras SGD(momentum) with optional decoupled weight decay; returns (w, a)."""
    w, a, g = (np.asarray(x, dtype=F) for x in (w, a, g))
    lr, mu, wd = F(lr), F(mu), F(wd)
    if wd != 0:
        w = w - (w * wd) * lr
    lg = lr * g
    if mu != 0:
        a2 = mu * a - lg
        w2 = w + (mu * a2 - lg) if nesterov else w + a2
    else:
        a2 = a
        w2 = w - lg
    return w2.astype(F), a2.astype(F)
