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
    class BroadcastGlobalVariablesCallback(keras.callbacks.Callback):
        def __init__(self, root_rank=0):
            super().__init__()
            self.root_rank, self.done = root_rank, False

        def on_batch_end(self, batch, logs=None):
            if self.done or size() == 1:
                return
            import torch.distributed as dist
            for v in self.model.variables:
                t = _to_torch(tf.convert_to_tensor(v))
                dist.broadcast(t, src=self.root_rank)
                v.assign(_to_tf(t))
            self.done = True

    class MetricAverageCallback(keras.callbacks.Callback):
        def on_epoch_end(self, epoch, logs=None):
            if logs and size() > 1:
                for k in sorted(logs):
                    logs[k] = float(allreduce(tf.constant([logs[k]], tf.float32), op=Average)[0])

    class LearningRateWarmupCallback(keras.callbacks.Callback):
        """lr ramps from initial_lr/size to initial_lr over warmup_epochs (the run-time lr change
        that makes lr a per-call argument of the apply kernel)."""

        def __init__(self, initial_lr, warmup_epochs=5, verbose=0, steps_per_epoch=None):
            super().__init__()
            self.initial_lr, self.warmup_epochs, self.verbose, self.steps = initial_lr, warmup_epochs, verbose, steps_per_epoch
            self.epoch = 0

        def on_epoch_begin(self, epoch, logs=None):
            self.epoch = epoch

        def on_batch_begin(self, batch, logs=None):
            if self.epoch >= self.warmup_epochs:
                return
            steps = self.steps or self.params.get("steps") or 1
            e = self.epoch + batch / steps
            lr = self.initial_lr / size() * (e * (size() - 1) / self.warmup_epochs + 1)
            self.model.optimizer.learning_rate.assign(lr)
