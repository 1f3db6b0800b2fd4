"""Horovod-shaped module over libb200dp.so (torch-tensor flavour, executed and tested here).

Reference boundary: ``import horovod.tensorflow.keras as hvd`` (imagenet-resnet50-hvd.py:15),
``hvd.init()`` (:16), ``hvd.rank()/size()/local_rank()`` (:28,32,41,...),
``hvd.DistributedOptimizer(opt)`` (:101).  One process per GPU; rank and size come from the
launcher's environment (torchrun's RANK/WORLD_SIZE/LOCAL_RANK, or Open MPI's
OMPI_COMM_WORLD_*, which is what horovodrun/mpirun set).  torch.distributed is used for
rendezvous only (a session string, and the step-0 parameter broadcast); every gradient byte
moves through the library's own kernels.

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


def _plan_for(numels):
    key = tuple(numels)
    if key not in _state["plans"]:
        ctx = api.Context.rank(_state["rank"], _state["size"], _state["local_rank"])
        lay = api.Layout(numels, nranks=_state["size"])
        plan = api.Plan(ctx, lay)
        if _state["size"] > 1:
            _state["counter"] += 1
            plan.exchange("%s-%d" % (_state["session"], _state["counter"]))
            import torch.distributed as dist
            dist.barrier()
        _state["plans"][key] = (ctx, lay, plan)
    return _state["plans"][key][2]


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
    a step-0 concern, done with torch.distributed."""
    _need_init()
    if _state["size"] == 1:
        return
    import torch.distributed as dist
    for p in params:
        dist.broadcast(p.data if hasattr(p, "data") else p, src=root_rank)


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
        self.plan = _plan_for([p.numel() for p in self.params])
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
