"""The TF-facing glue (make_tf_cross_device_ops, hvd_keras) executed against tests/fake_tf -- a
STAND-IN for the TensorFlow API surface those modules touch, because TensorFlow itself cannot
be installed here.  What this shows: the glue's control flow, index arithmetic and argument
handling run end to end and give oracle-exact numbers.  What it does not show: anything about
TensorFlow's tracing, placement or stream ordering (still unexecuted against the real thing)."""
import importlib
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
import b200dp  # noqa: E402
import oracle  # noqa: E402

FAKE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fake_tf")


@pytest.fixture
def fake_tf():
    sys.path.insert(0, FAKE)
    for m in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.")]:
        del sys.modules[m]
    import tensorflow as tf
    assert getattr(tf, "IS_B200DP_TEST_STUB", False)
    try:
        yield tf
    finally:
        sys.path.remove(FAKE)
        for m in [k for k in sys.modules if k == "tensorflow" or k.startswith("tensorflow.") or k == "b200dp.hvd_keras"]:
            del sys.modules[m]


def test_tf_cross_device_ops_subclass_reduces_per_replica_values(fake_tf):
    tf = fake_tf
    from b200dp import cross_device_ops as cdo
    ops = cdo.make_tf_cross_device_ops()
    assert isinstance(ops, tf.distribute.CrossDeviceOps)
    g = torch.Generator(device="cuda").manual_seed(2)
    shapes = [(7, 7, 3, 64), (64,), (2048, 1000)]
    vals = [tf.distribute.DistributedValues([torch.randn(s, generator=g, device="cuda") * 1e-2 for _ in range(2)]) for s in shapes]
    out = ops.batch_reduce(tf.distribute.ReduceOp.SUM, [(v, None) for v in vals])
    torch.cuda.synchronize()
    assert len(out) == len(vals)
    for v, o in zip(vals, out):
        ref = oracle.addn([t.cpu().numpy().ravel() for t in v.values])
        assert len(o.values) == 2
        for t, src in zip(o.values, v.values):
            assert t.shape == src.shape
            assert np.array_equal(t.cpu().numpy().ravel().view(np.uint32), ref.view(np.uint32))
    one = ops.reduce(tf.distribute.ReduceOp.MEAN, vals[1], None)
    ref = oracle.div_n(oracle.addn([t.cpu().numpy().ravel() for t in vals[1].values]), 2)
    assert np.array_equal(one.values[0].cpu().numpy().view(np.uint32), ref.view(np.uint32))


def test_hvd_keras_module_surface(fake_tf):
    tf = fake_tf
    from tensorflow import keras
    from b200dp import hvd as core
    core.shutdown()
    hvd = importlib.import_module("b200dp.hvd_keras")        # `import b200dp.hvd_keras as hvd` (-hvd.py:15)
    hvd.init()
    assert (hvd.rank(), hvd.size(), hvd.local_rank()) == (0, 1, 0)
    x = torch.randn(5000, device="cuda")
    assert torch.equal(hvd.allreduce(x), x)                   # Average over one rank, through py_function + DLPack
    # DistributedOptimizer wraps the optimizer's own class and keeps its configuration
    opt = hvd.DistributedOptimizer(keras.optimizers.SGD(learning_rate=0.5))
    assert isinstance(opt, keras.optimizers.SGD) and float(opt.learning_rate.value) == 0.5
    v = tf.Variable(torch.ones(10, device="cuda"))
    opt.apply_gradients([(torch.full((10,), 0.25, device="cuda"), v), (None, tf.Variable(torch.zeros(1, device="cuda")))])
    assert torch.all(v.value == 0.875) and opt.applied == 1
    # the three callbacks of imagenet-resnet50-hvd.py:111-115, driven in Keras' order
    model = keras.Model([v], opt)
    cbs = [hvd.callbacks.BroadcastGlobalVariablesCallback(0), hvd.callbacks.MetricAverageCallback(),
           hvd.callbacks.LearningRateWarmupCallback(0.1 * hvd.size(), warmup_epochs=3, verbose=1)]
    for cb in cbs:
        cb.set_model(model)
        cb.set_params({"steps": 10})
        if hasattr(cb, "on_train_begin"):
            cb.on_train_begin()
    logs = {"loss": 2.0}
    for epoch in range(4):
        for cb in cbs:
            if hasattr(cb, "on_epoch_begin"):
                cb.on_epoch_begin(epoch)
        for batch in range(10):
            for cb in cbs:
                if hasattr(cb, "on_batch_begin"):
                    cb.on_batch_begin(batch)
            for cb in cbs:
                if hasattr(cb, "on_batch_end"):
                    cb.on_batch_end(batch, logs)
        for cb in cbs:
            if hasattr(cb, "on_epoch_end"):
                cb.on_epoch_end(epoch, logs)
    assert logs == {"loss": 2.0}
    assert float(opt.learning_rate.value) == pytest.approx(0.1)     # size 1: the warm-up is flat at initial_lr
