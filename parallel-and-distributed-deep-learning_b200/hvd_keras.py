"""``import b200dp.hvd_keras as hvd`` -- the module imagenet-resnet50-hvd.py:15 imports, with
the names that script touches: init, rank, size, local_rank (:16,28,32,41), DistributedOptimizer
(:101), callbacks.BroadcastGlobalVariablesCallback / MetricAverageCallback /
LearningRateWarmupCallback (:111-115), allreduce.

UNEXECUTED ON THIS IMAGE: TensorFlow/Keras are not installed, so importing this
module raises ImportError here.  It is written against the upstream horovod.tensorflow.keras
API; gradients cross into libb200dp.so through DLPack inside tf.py_function, exactly like
cross_device_ops.make_tf_cross_device_ops().  The executed, tested equivalent is b200dp.hvd.
"""
try:
    import tensorflow as tf
    from tensorflow import keras
except ImportError as e:
    raise ImportError(
        "b200dp.hvd_keras needs TensorFlow/Keras, which are not installed on this image. "
        "Use b200dp.hvd (torch tensors, same function names) or bind include/b200dp.h as "
        "INTEGRATION.md section 4 shows.") from e

from . import hvd as _core  # pragma: no cover

Average, Sum = _core.Average, _core.Sum  # pragma: no cover
init, rank, size, local_rank, shutdown = _core.init, _core.rank, _core.size, _core.local_rank, _core.shutdown  # pragma: no cover


def _to_torch(t):  # pragma: no cover
    from torch.utils import dlpack
    return dlpack.from_dlpack(tf.experimental.dlpack.to_dlpack(t))


def _to_tf(t):  # pragma: no cover
    from torch.utils import dlpack
    return tf.experimental.dlpack.from_dlpack(dlpack.to_dlpack(t))


def grouped_allreduce(tensors, average=None, op=None, prescale_factor=1.0, postscale_factor=1.0, name=None):  # pragma: no cover
    def run(*ts):
        import torch
        outs = _core.grouped_allreduce([_to_torch(t) for t in ts], average, op, prescale_factor, postscale_factor)
        torch.cuda.synchronize()
        return [_to_tf(o) for o in outs]
    outs = tf.py_function(run, list(tensors), [t.dtype for t in tensors])
    return [tf.reshape(o, t.shape) for o, t in zip(outs, tensors)]


def allreduce(tensor, average=None, op=None, prescale_factor=1.0, postscale_factor=1.0, name=None):  # pragma: no cover
    return grouped_allreduce([tensor], average, op, prescale_factor, postscale_factor, name)[0]


def DistributedOptimizer(optimizer, name=None, op=Average, **kwargs):  # pragma: no cover
    """Wraps a Keras optimizer: gradients are averaged across ranks (one fused message) before
    apply_gradients, as horovod.tensorflow.keras.DistributedOptimizer does."""
    base = optimizer.__class__

    class _Distributed(base):
        def apply_gradients(self, grads_and_vars, *a, **k):
            gv = [(g, v) for g, v in grads_and_vars if g is not None]
            reduced = grouped_allreduce([g for g, _ in gv], op=op)
            return super().apply_gradients(list(zip(reduced, [v for _, v in gv])), *a, **k)

    return _Distributed.from_config(optimizer.get_config())


class callbacks:  # pragma: no cover
    """keras.callbacks.Callback subclasses over the framework-neutral logic in b200dp.hvd.callbacks
    (which is what runs and is tested on this image)."""

    class BroadcastGlobalVariablesCallback(keras.callbacks.Callback):
        def __init__(self, root_rank=0, device=""):
            super().__init__()
            self.root_rank, self.done = root_rank, False

        def on_batch_end(self, batch, logs=None):
            if self.done or size() == 1:
                return
            import torch
            vs = list(self.model.variables) + list(self.model.optimizer.variables())
            ts = [_to_torch(tf.convert_to_tensor(v)).clone() for v in vs]
            _core.broadcast_parameters(ts, self.root_rank)      # load_weights -> b2dp_broadcast -> unpack
            torch.cuda.synchronize()
            for v, t in zip(vs, ts):
                v.assign(_to_tf(t))
            self.done = True

    class MetricAverageCallback(keras.callbacks.Callback):
        def __init__(self, device=""):
            super().__init__()
            self._core = _core.callbacks.MetricAverageCallback()

        def on_epoch_end(self, epoch, logs=None):
            self._core.on_epoch_end(epoch, logs)                # one b2dp_allreduce_small over the sorted keys

    class LearningRateWarmupCallback(keras.callbacks.Callback):
        """lr ramps from initial_lr/size to initial_lr over warmup_epochs (the run-time lr change
        that makes lr a per-call argument of the apply kernel)."""

        class _KerasLr:
            def __init__(self, cb):
                self._cb = cb

            def __setattr__(self, name, value):
                if name == "lr":
                    self._cb.model.optimizer.learning_rate.assign(value)
                else:
                    object.__setattr__(self, name, value)

        def __init__(self, initial_lr, warmup_epochs=5, momentum_correction=True, steps_per_epoch=None, verbose=0):
            super().__init__()
            self._core = _core.callbacks.LearningRateWarmupCallback(initial_lr, warmup_epochs, momentum_correction,
                                                                    steps_per_epoch, verbose)
            self._core.set_optimizer(self._KerasLr(self))

        def on_train_begin(self, logs=None):
            if not self._core.steps_per_epoch:
                self._core.steps_per_epoch = self.params.get("steps")

        def on_epoch_begin(self, epoch, logs=None):
            self._core.on_epoch_begin(epoch, logs)

        def on_batch_begin(self, batch, logs=None):
            self._core.on_batch_begin(batch, logs)

        def on_epoch_end(self, epoch, logs=None):
            self._core.on_epoch_end(epoch, logs)
