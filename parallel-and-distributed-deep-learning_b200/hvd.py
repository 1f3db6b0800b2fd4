"""Horovod-shaped module over libb200dp.so (torch-tensor flavour, executed and tested here).

Reference boundary: ``import horovod.tensorflow.keras as hvd`` (imagenet-resnet50-hvd.py:15),
``hvd.init()`` (:16), ``hvd.rank()/size()/local_rank()`` (:28,32,41,...),
``hvd.DistributedOptimizer(opt)`` (:101).  One process per GPU; rank and size come from the
launcher's environment (torchrun's RANK/WORLD_SIZE/LOCAL_RANK, or Open MPI's
OMPI_COMM_WORLD_*, which is what horovodrun/mpirun set).  torch.distributed is used for
rendezvous only (a session string); every gradient, weight and metric byte moves through the
library's own kernels (b2dp_allreduce / b2dp_step / b2dp_broadcast / b2dp_allreduce_small).
``hvd.callbacks`` holds the three callbacks the script constructs (:111-115) in a
framework-neutral form: plain objects with Keras' hook names that act on an optimizer.

The Keras/TensorFlow flavour with the script's exact import name is hvd_keras.py
(unexecuted on this image: TensorFlow is absent).
"""
import os

from . import api
from ._lib import BUF_REDUCED

Average = "Average"
Sum = "Sum"

_state = {"init": False, "rank": 0, "size": 1, "local_rank": 0, "plans": {}, "session": None, "counter": 0}


def _env_int(*names, default=0):
    for n in names:
        if n in os.environ:
            return int(os.environ[n])
    return default


def init():
    """hvd.init(): read rank/size from the launcher, pin the GPU, join the rendezvous."""
    if _state["init"]:
        return
    import torch
    _state["rank"] = _env_int("RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK")
    _state["size"] = _env_int("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", default=1)
    _state["local_rank"] = _env_int("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", default=_state["rank"])
    if torch.cuda.is_available():
        torch.cuda.set_device(_state["local_rank"])
    if _state["size"] > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            backend = "nccl" if torch.cuda.is_available() else "gloo"
            dist.init_process_group(backend)
        obj = ["hvd-%s-%d" % (os.environ.get("MASTER_PORT", "0"), os.getpid())]
        dist.broadcast_object_list(obj, src=0)
        _state["session"] = obj[0]
    _state["init"] = True


def shutdown():
    _state["plans"].clear()
    _state["init"] = False


def _need_init():
    if not _state["init"]:
        raise ValueError("Horovod has not been initialized; use hvd.init().")


def rank():
    _need_init()
    return _state["rank"]


def size():
    _need_init()
    return _state["size"]


def local_rank():
    _need_init()
    return _state["local_rank"]


def _new_plan(numels):
    """A fresh plan over this job's ranks (collective: every rank must create it at the same point)."""
    ctx = api.Context.rank(_state["rank"], _state["size"], _state["local_rank"])
    lay = api.Layout(numels, nranks=_state["size"])
    plan = api.Plan(ctx, lay)
    if _state["size"] > 1:
        _state["counter"] += 1
        plan.exchange("%s-%d" % (_state["session"], _state["counter"]))
        import torch.distributed as dist
        dist.barrier()
    plan._keep = (ctx, lay)
    return plan


def _plan_for(numels):
    """Cached plan for the STATELESS entry points (allreduce, grouped_allreduce, broadcast_parameters):
    they only pass data through the plan's buffers.  Anything that keeps weights or optimizer state
    in a plan (FusedDistributedAdam) owns a plan of its own -- two optimizers over models with equal
    parameter sizes must not alias one weight buffer and one set of Adam moments."""
    key = tuple(numels)
    if key not in _state["plans"]:
        _state["plans"][key] = _new_plan(numels)
    return _state["plans"][key]


def _scales(average, op, prescale_factor, postscale_factor):
    if average is not None and op is not None:
        raise ValueError("The op parameter supersedes average. Please provide only one of them.")
    if op is None:
        op = Average if (average is None or average) else Sum
    if op not in (Average, Sum):
        raise ValueError("op must be hvd.Average or hvd.Sum (got %r)" % (op,))
    post = postscale_factor * (1.0 / _state["size"] if op == Average else 1.0)
    return prescale_factor, post


def grouped_allreduce(tensors, average=None, op=None, prescale_factor=1.0, postscale_factor=1.0, name=None):
    """All-reduce a list of tensors as one fused message (what Horovod's fusion buffer does
    behind per-tensor hvd.allreduce calls).  Returns new tensors."""
    _need_init()
    import torch
    pre, post = _scales(average, op, prescale_factor, postscale_factor)
    plan = _plan_for([t.numel() for t in tensors])
    plan.pack(0, [t.contiguous() for t in tensors], prescale=pre)
    plan.allreduce(0, postscale=post)
    outs = [torch.empty_like(t, memory_format=torch.contiguous_format) for t in tensors]
    plan.unpack(0, outs, which=BUF_REDUCED)
    return outs


def allreduce(tensor, average=None, op=None, prescale_factor=1.0, postscale_factor=1.0, name=None):
    """hvd.allreduce(tensor, average=None, op=None, prescale_factor=1.0, postscale_factor=1.0, name=None).
    Up to 1024 values (metric averaging, imagenet-resnet50-hvd.py:113) take the one-shot
    latency-bound kernel; anything larger goes through the fused bucket path."""
    _need_init()
    if tensor.numel() <= 1024 and prescale_factor == 1.0 and str(tensor.dtype) == "torch.float32" and tensor.is_cuda:
        import torch
        _, post = _scales(average, op, prescale_factor, postscale_factor)
        plan = _plan_for([256])     # any plan carries the small-message slots
        src = tensor.contiguous()
        out = torch.empty_like(src)
        plan.allreduce_small(0, src, out, postscale=post)
        return out
    return grouped_allreduce([tensor], average, op, prescale_factor, postscale_factor, name)[0]


def broadcast_parameters(params, root_rank=0):
    """hvd.broadcast_parameters / BroadcastGlobalVariablesCallback(0) (imagenet-resnet50-hvd.py:111):
    every rank's tensors are overwritten with the root's, through the library (load_weights ->
    b2dp_broadcast -> unpack: one kernel over the flat buffer instead of 214 NCCL broadcasts)."""
    _need_init()
    params = [p.data if hasattr(p, "data") else p for p in params]
    if _state["size"] == 1 or not params:
        return
    if not 0 <= root_rank < _state["size"]:
        raise ValueError("root_rank %d outside 0..%d" % (root_rank, _state["size"] - 1))
    from ._lib import BUF_WEIGHTS
    plan = _plan_for([p.numel() for p in params])
    src = [p.contiguous() for p in params]
    plan.load_weights(0, src)
    plan.broadcast(0, root=root_rank, which=BUF_WEIGHTS)
    plan.unpack(0, src, which=BUF_WEIGHTS)
    for p, c in zip(params, src):
        if c.data_ptr() != p.data_ptr():
            p.copy_(c)


def broadcast_variables(variables, root_rank=0):
    """hvd.broadcast_variables (the name horovod.tensorflow uses)."""
    broadcast_parameters(variables, root_rank)


class DistributedOptimizer:
    """hvd.DistributedOptimizer(optimizer, op=Average): averages the gradients of the wrapped
    torch optimizer's parameters across ranks (one fused message) before every step()."""

    def __init__(self, optimizer, named_parameters=None, op=Average, prescale_factor=1.0, postscale_factor=1.0):
        _need_init()
        self._opt, self._op = optimizer, op
        self._pre, self._post = prescale_factor, postscale_factor
        self._params = [p for g in optimizer.param_groups for p in g["params"] if p.requires_grad]

    def __getattr__(self, name):
        return getattr(self._opt, name)

    def synchronize(self):
        ps = [p for p in self._params if p.grad is not None]
        if not ps:
            return
        reduced = grouped_allreduce([p.grad for p in ps], op=self._op, prescale_factor=self._pre,
                                    postscale_factor=self._post)
        for p, g in zip(ps, reduced):
            p.grad = g.view_as(p.grad)

    def step(self, closure=None):
        self.synchronize()
        return self._opt.step(closure)

    def zero_grad(self, *a, **k):
        return self._opt.zero_grad(*a, **k)


class FusedDistributedAdam:
    """DistributedOptimizer(Adam) with the optimizer fused into the reduction kernel
    (b2dp_step): gradients are averaged, Adam runs on each rank's shard, updated weights
    are all-gathered -- one kernel instead of all-reduce + 214 apply launches.  The
    parameters become views of the plan's flat weight buffer."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-7, op=Average):
        _need_init()
        from ._lib import BUF_WEIGHTS
        self.params = [p for p in params if p.requires_grad]
        self.lr, self.betas, self.eps, self.t = lr, betas, eps, 0
        self._post = 1.0 / _state["size"] if op == Average else 1.0
        self.plan = _new_plan([p.numel() for p in self.params])   # owns weights + Adam state: never shared
        self.plan.load_weights(0, [p.data.contiguous() for p in self.params])
        flat = self.plan.tensor(0, BUF_WEIGHTS)
        import torch
        torch.cuda.current_stream().synchronize()
        for k, p in enumerate(self.params):
            off = self.plan.layout.tensor_offset(k)
            p.data = flat[off:off + p.numel()].view(p.shape)

    def zero_grad(self, set_to_none=True):
        for p in self.params:
            p.grad = None

    def broadcast_weights(self, root_rank=0):
        """hvd.callbacks.BroadcastGlobalVariablesCallback(root_rank): every rank starts from the
        root's weights (one kernel over the flat weight buffer)."""
        self.plan.broadcast(0, root=root_rank)

    def step(self):
        self.t += 1
        self.plan.pack(0, [p.grad.contiguous() for p in self.params])
        self.plan.step(0, "adam", api.adam_hyper(lr=self.lr, beta1=self.betas[0], beta2=self.betas[1], eps=self.eps,
                                                 step=self.t, postscale=self._post))


def _set_lr(optimizer, lr):
    """lr of a FusedDistributedAdam (attribute) or of a torch optimizer / DistributedOptimizer (param_groups)."""
    if hasattr(optimizer, "param_groups"):
        for g in optimizer.param_groups:
            g["lr"] = lr
    else:
        optimizer.lr = lr


class callbacks:
    """horovod.tensorflow.keras.callbacks, as imagenet-resnet50-hvd.py:111-115 constructs them:

        hvd.callbacks.BroadcastGlobalVariablesCallback(0)
        hvd.callbacks.MetricAverageCallback()
        hvd.callbacks.LearningRateWarmupCallback(0.1 * hvd.size(), warmup_epochs=3, verbose=1)

    Framework-neutral: objects with Keras' hook names (on_train_begin / on_epoch_begin /
    on_batch_begin / on_batch_end / on_epoch_end) that a training loop calls in Keras' order;
    ``set_optimizer`` plays the part of Keras' ``set_model``.  hvd_keras.py wraps the same logic
    in keras.callbacks.Callback subclasses for a TensorFlow install."""

    class _Base:
        optimizer = None
        params = None

        def set_optimizer(self, optimizer, params=None):
            self.optimizer, self.params = optimizer, params
            return self

        def on_train_begin(self, logs=None):
            pass

        def on_epoch_begin(self, epoch, logs=None):
            pass

        def on_batch_begin(self, batch, logs=None):
            pass

        def on_batch_end(self, batch, logs=None):
            pass

        def on_epoch_end(self, epoch, logs=None):
            pass

    class BroadcastGlobalVariablesCallback(_Base):
        """Every rank starts from root_rank's weights (and, upstream, optimizer slots -- which are
        all zeros at step 0 here).  Runs once, after the first batch like upstream, or at
        on_train_begin when the variables already exist (they always do in this flavour)."""

        def __init__(self, root_rank=0, device=""):
            self.root_rank, self.broadcast_done = root_rank, False

        def _broadcast(self):
            if self.broadcast_done:
                return
            if hasattr(self.optimizer, "broadcast_weights"):
                self.optimizer.broadcast_weights(self.root_rank)     # flat weight buffer, one kernel
            elif self.params is not None:
                broadcast_parameters(self.params, self.root_rank)
            else:
                raise ValueError("BroadcastGlobalVariablesCallback: set_optimizer(optimizer, params) first")
            self.broadcast_done = True

        def on_train_begin(self, logs=None):
            self._broadcast()

        def on_batch_end(self, batch, logs=None):
            self._broadcast()

    class MetricAverageCallback(_Base):
        """Averages every numeric entry of the epoch's logs over the ranks, in sorted key order,
        as ONE small-message all-reduce (b2dp_allreduce_small, <= 1024 metrics)."""

        def __init__(self, device=""):
            pass

        def on_epoch_end(self, epoch, logs=None):
            if not logs or _state["size"] == 1:
                return
            import torch
            keys = sorted(k for k, v in logs.items() if isinstance(v, (int, float)))
            if not keys:
                return
            src = torch.tensor([float(logs[k]) for k in keys], dtype=torch.float32, device="cuda")
            out = allreduce(src, op=Average)
            for k, v in zip(keys, out.tolist()):
                logs[k] = v

    class LearningRateWarmupCallback(_Base):
        """Horovod's gradual warm-up (Goyal et al.): over the first warmup_epochs the learning rate
        grows from initial_lr / size to initial_lr,

            lr(epoch) = initial_lr / size * (epoch * (size - 1) / warmup_epochs + 1)

        with a fractional epoch = epoch + batch / steps_per_epoch, adjusted before every batch.
        This is why lr is a per-call argument of the apply kernel."""

        def __init__(self, initial_lr, warmup_epochs=5, momentum_correction=True, steps_per_epoch=None, verbose=0):
            self.initial_lr, self.warmup_epochs = float(initial_lr), warmup_epochs
            self.steps_per_epoch, self.verbose = steps_per_epoch, verbose
            self.current_epoch = 0

        def multiplier(self, epoch):
            size = _state["size"]
            return 1.0 / size * (epoch * (size - 1) / self.warmup_epochs + 1)

        def on_epoch_begin(self, epoch, logs=None):
            self.current_epoch = epoch

        def on_batch_begin(self, batch, logs=None):
            if self.current_epoch >= self.warmup_epochs:
                return
            if not self.steps_per_epoch:
                raise ValueError("LearningRateWarmupCallback needs steps_per_epoch (Keras supplies params['steps'])")
            epoch = self.current_epoch + float(batch) / self.steps_per_epoch
            _set_lr(self.optimizer, self.initial_lr * self.multiplier(epoch))

        def on_epoch_end(self, epoch, logs=None):
            if epoch == self.warmup_epochs - 1:
                _set_lr(self.optimizer, self.initial_lr)
                if self.verbose and _state["rank"] == 0:
                    print("\nEpoch %d: finished gradual learning rate warmup to %g." % (epoch + 1, self.initial_lr))
