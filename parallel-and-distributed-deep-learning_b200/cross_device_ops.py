"""B200CrossDeviceOps -- the tf.distribute.CrossDeviceOps plug-in point, over libb200dp.so.

Reference boundary: ``strategy = tf.distribute.MirroredStrategy()``
(imagenet-resnet50-mirror.py:21) becomes
``MirroredStrategy(cross_device_ops=B200CrossDeviceOps())``; under ``model.fit`` (:81) Keras
then calls ``batch_reduce(ReduceOp.SUM, [(PerReplica grad_k, variable_k)] * 214)`` once per
step [upstream tensorflow/python/distribute/cross_device_ops.py].

Two process shapes, one class:
  * ``B200CrossDeviceOps()`` -- one process drives every GPU (MirroredStrategy, -mirror.py:21):
    a per-replica value is a list with one tensor per local replica;
  * ``B200CrossDeviceOps(rank=r, nranks=n, device=d)`` -- one process per GPU
    (MultiWorkerMirroredStrategy, imagenet-resnet50-multiworkers.py:20-25: each worker has ONE
    local replica and the reduction runs across workers): a per-replica value is a list with the
    single local tensor, the plan is built over a rank context and the peers are attached by fd
    exchange on first use (a session string is agreed through ``session=`` or torch.distributed).
    MWMS takes no ``cross_device_ops`` argument: ``swap_into(strategy)`` replaces
    ``strategy.extended._cross_device_ops`` after construction, as SURVEY.md 8b describes.

Two classes share one engine:
  * ``B200CrossDeviceOps``      -- executed and tested here: same method names, argument meaning
    and error behaviour (reduce / batch_reduce / broadcast and the implementation hooks
    reduce_implementation / batch_reduce_implementation / _all_reduce / _gather_implementation),
    with per-replica values given as lists of torch CUDA tensors because TensorFlow is not
    installed on this image.
  * ``make_tf_cross_device_ops()`` -- returns the real ``tf.distribute.CrossDeviceOps`` subclass
    when TensorFlow is importable (tensors cross via DLPack inside a ``tf.py_function``, since
    the hook runs while the step function is being traced).  Written against the upstream API,
    **unexecuted on this image**.
"""
import enum

from . import api
from ._lib import BUF_REDUCED


class ReduceOp(enum.Enum):
    """tf.distribute.ReduceOp"""
    SUM = "SUM"
    MEAN = "MEAN"


class CommunicationOptions:
    """tf.distribute.experimental.CommunicationOptions (imagenet-resnet50-multiworkers.py:20-21):
    bytes_per_pack maps to the bucket size of the layout; 0 = one pack, as upstream."""

    def __init__(self, bytes_per_pack=0, timeout_seconds=None, implementation=None):
        if bytes_per_pack < 0:
            raise ValueError("Argument `bytes_per_pack` must be >=0, Received %d." % bytes_per_pack)
        self.bytes_per_pack = bytes_per_pack
        self.timeout_seconds = timeout_seconds
        self.implementation = implementation


class _Engine:
    """Caches one plan per (tensor sizes, devices, bucket size) and runs pack -> allreduce -> unpack.
    Local mode drives every rank; rank mode drives this process's one rank of `nranks`."""

    def __init__(self, wire="f32", rank=None, nranks=None, device=None, session=None):
        self.wire = wire
        self._plans = {}
        self.rank, self.nranks, self.device, self.session = rank, nranks, device, session
        self._counter = 0
        if (rank is None) != (nranks is None):
            raise ValueError("give both rank and nranks (one process per GPU) or neither (one process, all GPUs)")
        if rank is not None and not 0 <= rank < nranks:
            raise ValueError("rank %d outside 0..%d" % (rank, nranks - 1))

    @property
    def rank_mode(self):
        return self.rank is not None

    def _session(self):
        """A job-unique string every rank agrees on (for the fd exchange).  Given explicitly, or
        agreed through torch.distributed when a process group is up."""
        self._counter += 1
        if self.session is None:
            import os
            import torch.distributed as dist
            if not (dist.is_available() and dist.is_initialized()):
                raise ValueError("rank-context B200CrossDeviceOps needs session=<job-unique string> or an initialised "
                                 "torch.distributed process group to agree on one")
            obj = ["cdo-%s-%d" % (os.environ.get("MASTER_PORT", "0"), os.getpid())]
            dist.broadcast_object_list(obj, src=0)
            self.session = obj[0]
        return "%s-%d" % (self.session, self._counter)

    def plan_for(self, numels, devices, bucket_bytes):
        key = (tuple(numels), tuple(devices), int(bucket_bytes))
        if key not in self._plans:
            if self.rank_mode:
                ctx = api.Context.rank(self.rank, self.nranks, devices[0])
                lay = api.Layout(numels, nranks=self.nranks, wire=self.wire, bucket_bytes=bucket_bytes)
                plan = api.Plan(ctx, lay)
                if self.nranks > 1:
                    plan.exchange(self._session())   # collective: every rank builds its plans in the same order
            else:
                ctx = api.Context.local(devices)
                lay = api.Layout(numels, nranks=len(devices), wire=self.wire, bucket_bytes=bucket_bytes)
                plan = api.Plan(ctx, lay)
            self._plans[key] = (ctx, lay, plan)
        return self._plans[key][2]

    def all_reduce(self, per_replica, reduce_op, bucket_bytes):
        """per_replica[r][k] = tensor k on LOCAL replica r.  Returns the same structure, reduced
        over all ranks of the job (= the local replicas in local mode)."""
        import torch
        nlocal = len(per_replica)
        n = self.nranks if self.rank_mode else nlocal
        if self.rank_mode and nlocal != 1:
            raise ValueError("one process per GPU: a per-replica value holds exactly one local tensor, got %d" % nlocal)
        numels = [t.numel() for t in per_replica[0]]
        devices = [ts[0].device.index if ts else (self.device or 0) for ts in per_replica]
        plan = self.plan_for(numels, devices, bucket_bytes)
        # MEAN is SUM / n in TF (ReductionToOneDevice divides the AddN result by the replica count).
        # For a power-of-two n -- every BASELINE.json config -- multiplying by float(1/n) inside
        # the kernel gives the same bits; otherwise the sum is divided afterwards, exactly as TF does.
        pow2 = (n & (n - 1)) == 0
        post = 1.0 / n if (reduce_op == ReduceOp.MEAN and pow2) else 1.0
        outs = []
        for r in range(nlocal):
            with torch.cuda.device(devices[r]):
                plan.pack(r, [t.contiguous() for t in per_replica[r]])
        for r in range(nlocal):
            with torch.cuda.device(devices[r]):
                plan.allreduce(r, postscale=post)
        for r in range(nlocal):
            with torch.cuda.device(devices[r]):
                o = [torch.empty_like(t, memory_format=torch.contiguous_format) for t in per_replica[r]]
                plan.unpack(r, o, which=BUF_REDUCED)
                if reduce_op == ReduceOp.MEAN and not pow2:
                    # a device-tensor divisor: torch turns division by a host scalar into a
                    # multiplication by its reciprocal, which is not the IEEE quotient
                    d = torch.full((), float(n), device=o[0].device) if o else None
                    for t in o:
                        t.div_(d)
                outs.append(o)
        return outs

    def all_gather(self, per_replica, axis):
        """per_replica[r] = the tensor of LOCAL replica r.  Returns, per local replica, the
        concatenation along `axis` of every rank's tensor in rank order (exact copies: each rank's
        piece is placed in its own slice of the flat weight buffer and b2dp_all_gather pushes the
        slices to every rank)."""
        import torch
        from ._lib import BUF_WEIGHTS
        nlocal = len(per_replica)
        n = self.nranks if self.rank_mode else nlocal
        shape = tuple(per_replica[0].shape)
        m = per_replica[0].numel()
        piece = (m + 255) // 256 * 256                 # one rank slice: a multiple of the slice alignment
        devices = [t.device.index for t in per_replica]
        plan = self.plan_for([piece * n], devices, 0)  # a single "tensor" whose n slices are the pieces
        ranks = [self.rank] if self.rank_mode else list(range(nlocal))
        for r in range(nlocal):
            with torch.cuda.device(devices[r]):
                flat = plan.tensor(r, BUF_WEIGHTS)
                flat[ranks[r] * piece:ranks[r] * piece + m].copy_(per_replica[r].contiguous().view(-1))
        for r in range(nlocal):
            with torch.cuda.device(devices[r]):
                plan.all_gather(r)
        outs = []
        for r in range(nlocal):
            with torch.cuda.device(devices[r]):
                flat = plan.tensor(r, BUF_WEIGHTS)
                parts = [flat[q * piece:q * piece + m].view(shape) for q in range(n)]
                outs.append(torch.cat(parts, dim=axis) if shape else torch.stack(parts))
        return outs


class B200CrossDeviceOps:
    """Drop-in for tf.distribute.CrossDeviceOps on the gradient path (torch-tensor flavour).

    A *per-replica value* is a list with one tensor per replica (``PerReplica.values``); the
    result is a list with one tensor per destination replica (``Mirrored.values``), every
    entry holding the same reduced value.
    """

    def __init__(self, wire="f32", rank=None, nranks=None, device=None, session=None):
        self._engine = _Engine(wire, rank, nranks, device, session)
        self._num_between_graph_workers = nranks or 1

    def swap_into(self, strategy):
        """MultiWorkerMirroredStrategy has no cross_device_ops argument (-multiworkers.py:24-25): the
        drop-in there is replacing the object the strategy already built."""
        strategy.extended._cross_device_ops = self
        return strategy

    # -- public API, as upstream ---------------------------------------------------------
    def reduce(self, reduce_op, per_replica_value, destinations=None, options=None):
        return self.batch_reduce(reduce_op, [(per_replica_value, destinations)], options)[0]

    def batch_reduce(self, reduce_op, value_destination_pairs, options=None):
        if not isinstance(reduce_op, ReduceOp):
            reduce_op = ReduceOp(str(reduce_op).split(".")[-1].upper())
        if options is None:
            options = CommunicationOptions()
        self._validate(value_destination_pairs)
        return self.batch_reduce_implementation(reduce_op, value_destination_pairs, options)

    def broadcast(self, tensor, destinations):
        return [tensor.to("cuda:%d" % d) for d in destinations]

    def gather(self, per_replica_value, destinations=None, axis=0, options=None):
        """CrossDeviceOps._gather: concatenates the replicas' tensors along `axis`."""
        if not isinstance(per_replica_value, (list, tuple)) or not per_replica_value:
            raise ValueError("gather: not a per-replica value (a non-empty list of tensors)")
        return self._gather_implementation(per_replica_value, destinations, axis, options)

    # -- implementation hooks, as upstream ------------------------------------------------
    def reduce_implementation(self, reduce_op, per_replica_value, destinations, options):
        return self.batch_reduce_implementation(reduce_op, [(per_replica_value, destinations)], options)[0]

    def batch_reduce_implementation(self, reduce_op, value_destination_pairs, options):
        values = [v for v, _ in value_destination_pairs]
        if not values:
            return []
        n = len(values[0])
        per_replica = [[v[r] for v in values] for r in range(n)]
        reduced = self._engine.all_reduce(per_replica, reduce_op, options.bytes_per_pack)
        return [[reduced[r][k] for r in range(n)] for k in range(len(values))]

    def _all_reduce(self, reduce_op, value, replica_id=0, options=None):
        """Newer-TF hook (CollectiveAllReduce._all_reduce): all-reduce the calling replica's `value`
        -- a tensor or a flat list of tensors -- across the replicas IN REPLICA CONTEXT.  With one
        process per GPU (the MWMS shape) that is a direct collective call; in the one-process
        shape TF's replica threads each call it and the calls meet inside the collective, which a
        single-threaded caller cannot do: use batch_reduce there."""
        if not self._engine.rank_mode:
            raise ValueError("_all_reduce runs in replica context: construct B200CrossDeviceOps(rank=, nranks=) "
                             "(one process per GPU) or use batch_reduce (one process, all GPUs)")
        if not isinstance(reduce_op, ReduceOp):
            reduce_op = ReduceOp(str(reduce_op).split(".")[-1].upper())
        single = not isinstance(value, (list, tuple))
        flat = [value] if single else list(value)
        if options is None:
            options = CommunicationOptions()
        self._validate([([t], None) for t in flat])
        reduced = self._engine.all_reduce([flat], reduce_op, options.bytes_per_pack)[0]
        return reduced[0] if single else reduced

    def _gather_implementation(self, per_replica_value, destinations, axis, options):
        outs = self._engine.all_gather(list(per_replica_value), axis)
        return outs

    @staticmethod
    def _validate(pairs):
        n = None
        for k, (value, _dest) in enumerate(pairs):
            if not isinstance(value, (list, tuple)) or not value:
                raise ValueError("value %d is not a per-replica value (a non-empty list of tensors)" % k)
            if n is None:
                n = len(value)
            if len(value) != n:
                raise ValueError("value %d has %d replicas, expected %d" % (k, len(value), n))
            shape = tuple(value[0].shape)
            for t in value:
                if tuple(t.shape) != shape:
                    raise ValueError("value %d: replicas disagree on shape (%s vs %s)" % (k, tuple(t.shape), shape))
                if not t.is_cuda:
                    raise ValueError("value %d lives on %s; this library has no CPU path" % (k, t.device))
                if str(t.dtype) != "torch.float32":
                    raise ValueError("value %d has dtype %s; gradients must be float32" % (k, t.dtype))
        devs = [t.device for t in pairs[0][0]] if pairs else []
        for k, (value, _dest) in enumerate(pairs):
            if [t.device for t in value] != devs:
                raise ValueError("value %d is placed on different devices than value 0" % k)


def make_tf_cross_device_ops(wire="f32"):
    """The tf.distribute.CrossDeviceOps subclass itself.  UNEXECUTED on this image (no
    TensorFlow); raises ImportError here.  Stream ordering between TensorFlow's compute stream
    and the py_function is the known open item (SURVEY.md 7 "hard parts"): a compiled OpKernel
    over the same C ABI is the proper fix once TF headers exist."""
    try:
        import tensorflow as tf
    except ImportError as e:  # pragma: no cover - TF absent on this image
        raise ImportError("TensorFlow is not installed: the TF-facing B200CrossDeviceOps cannot be built here. "
                          "Use b200dp.cross_device_ops.B200CrossDeviceOps (torch tensors) or bind "
                          "include/b200dp.h as INTEGRATION.md section 2 shows.") from e

    import torch
    from torch.utils import dlpack as torch_dlpack

    engine = _Engine(wire)

    class TFB200CrossDeviceOps(tf.distribute.CrossDeviceOps):  # pragma: no cover
        def reduce_implementation(self, reduce_op, per_replica_value, destinations, options):
            return self.batch_reduce_implementation(reduce_op, [(per_replica_value, destinations)], options)[0]

        def batch_reduce_implementation(self, reduce_op, value_destination_pairs, options):
            values = [v for v, _ in value_destination_pairs]
            n = len(values[0].values)
            flat_in = [t for v in values for t in v.values]           # K * n tensors, k-major
            op = ReduceOp.MEAN if reduce_op == tf.distribute.ReduceOp.MEAN else ReduceOp.SUM

            def run(*tensors):
                per_replica = [[torch_dlpack.from_dlpack(tf.experimental.dlpack.to_dlpack(tensors[k * n + r]))
                                for k in range(len(values))] for r in range(n)]
                reduced = engine.all_reduce(per_replica, op, options.bytes_per_pack if options else 0)
                torch.cuda.synchronize()
                return [tf.experimental.dlpack.from_dlpack(torch_dlpack.to_dlpack(reduced[r][k]))
                        for k in range(len(values)) for r in range(n)]

            outs = tf.py_function(run, flat_in, [t.dtype for t in flat_in])
            result = []
            for k, v in enumerate(values):
                per_dev = []
                for r, src in enumerate(v.values):
                    with tf.device(src.device):
                        per_dev.append(tf.reshape(tf.identity(outs[k * n + r]), src.shape))
                result.append(tf.distribute.DistributedValues(per_dev) if hasattr(tf.distribute, "DistributedValues")
                              else per_dev)
            return result

    return TFB200CrossDeviceOps()
