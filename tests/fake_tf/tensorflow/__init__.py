"""A STAND-IN for the slice of the TensorFlow API that b200dp's TF-facing glue touches
(cross_device_ops.make_tf_cross_device_ops, hvd_keras).  TEST INFRASTRUCTURE ONLY.

TensorFlow is not installable on this image, so the TF-facing modules cannot run against the
real thing.  This package lets the tests execute their control flow end to end -- "tensors" are
torch CUDA tensors, tf.py_function calls its function eagerly, DLPack capsules are torch's --
which catches wrong argument orders, missing attributes and broken index arithmetic in the glue.
It proves NOTHING about TensorFlow's own behaviour (graph tracing, stream ordering, placement):
those stay "unexecuted against TensorFlow" in DESIGN.md.  It is only ever put on sys.path by
tests/test_gpu_tf_stub.py.
"""
import contextlib
import enum
import types

import torch
from torch.utils import dlpack as _dl

float32 = torch.float32
IS_B200DP_TEST_STUB = True


class _ReduceOp(enum.Enum):
    SUM = "SUM"
    MEAN = "MEAN"


class DistributedValues:
    """PerReplica / Mirrored: `.values` is one tensor per replica."""

    def __init__(self, values):
        self.values = tuple(values)


class CrossDeviceOps:
    """tf.distribute.CrossDeviceOps: the public entry points validate and forward to the
    *_implementation hooks a subclass overrides."""

    def reduce(self, reduce_op, per_replica_value, destinations, options=None):
        return self.reduce_implementation(reduce_op, per_replica_value, destinations, options)

    def batch_reduce(self, reduce_op, value_destination_pairs, options=None):
        return self.batch_reduce_implementation(reduce_op, value_destination_pairs, options)

    def reduce_implementation(self, reduce_op, per_replica_value, destinations, options):
        raise NotImplementedError

    def batch_reduce_implementation(self, reduce_op, value_destination_pairs, options):
        raise NotImplementedError


distribute = types.SimpleNamespace(ReduceOp=_ReduceOp, CrossDeviceOps=CrossDeviceOps, DistributedValues=DistributedValues)
experimental = types.SimpleNamespace(dlpack=types.SimpleNamespace(to_dlpack=_dl.to_dlpack, from_dlpack=_dl.from_dlpack))


def py_function(func, inp, Tout):
    out = func(*inp)
    assert len(out) == len(Tout)
    return list(out)


@contextlib.contextmanager
def device(_name):
    yield


def reshape(t, shape):
    return t.reshape(tuple(shape))


def identity(t):
    return t


def constant(value, dtype=None):
    return torch.tensor(value, dtype=dtype or torch.float32, device="cuda")


class Variable:
    def __init__(self, value):
        self.value = value.clone() if isinstance(value, torch.Tensor) else torch.tensor(float(value), device="cuda")

    def assign(self, v):
        v = v if isinstance(v, torch.Tensor) else torch.tensor(float(v), device=self.value.device)
        self.value = v.to(self.value.device).reshape(self.value.shape).clone()

    def numpy(self):
        return self.value.cpu().numpy()


def convert_to_tensor(v):
    return v.value if isinstance(v, Variable) else v
