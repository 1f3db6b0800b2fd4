"""Host-side mirror of the C ABI: Layout, Context, Plan.

Thin by design -- each method is one C call.  torch is used only for what the task
allows it for: device memory handed in by the caller, streams, and (in glue modules)
torch.distributed for rendezvous.  Every compute call lands in libb200dp.so.
"""
import ctypes
import os

from . import _lib
from ._lib import (BF16, BUF_GRADS, BUF_REDUCED, BUF_STATE0, BUF_STATE1, BUF_STATS, BUF_WEIGHTS, F32, OPT_ADAM,
                   OPT_NONE, OPT_SGDM, B200dpError, Hyper, check)

_DTYPES = {"f32": F32, "float32": F32, "bf16": BF16, "bfloat16": BF16, F32: F32, BF16: BF16}
_OPTS = {"adam": OPT_ADAM, "sgdm": OPT_SGDM, "sgd": OPT_SGDM, "none": OPT_NONE,
         OPT_ADAM: OPT_ADAM, OPT_SGDM: OPT_SGDM, OPT_NONE: OPT_NONE}


def adam_hyper(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-7, step=1, postscale=1.0, want_stats=False):
    """Keras Adam defaults (compile(optimizer='adam'), imagenet-resnet50-mirror.py:76)."""
    return Hyper(lr=lr, beta1=beta1, beta2=beta2, eps=eps, momentum=0.0, weight_decay=0.0, nesterov=0,
                 want_stats=int(bool(want_stats)), step=int(step), postscale=postscale, reserved=0.0)


def sgdm_hyper(lr=0.01, momentum=0.9, nesterov=False, weight_decay=0.0, postscale=1.0, want_stats=False):
    return Hyper(lr=lr, beta1=0.0, beta2=0.0, eps=0.0, momentum=momentum, weight_decay=weight_decay,
                 nesterov=int(bool(nesterov)), want_stats=int(bool(want_stats)), step=1, postscale=postscale,
                 reserved=0.0)


def adam_alpha(lr, beta1, beta2, step):
    return float(_lib.load().b2dp_adam_alpha(lr, beta1, beta2, int(step)))


def adam_apply_noalloc(w, m, v, g, hyper, stats_out, stream=None):
    """b2dp_adam_apply_noalloc on caller-owned torch tensors (fp32, contiguous, on the current
    device): the fused Adam apply for callers that may not allocate device memory.  The
    library's only scratch is a fixed static device array, so at most 2**26 elements per
    call, and calls that overlap on the GPU race on that array (see include/b200dp.h)."""
    n = w.numel()
    if not (m.numel() == n and v.numel() == n and g.numel() == n):
        raise ValueError("adam_apply_noalloc: w, m, v, g must have the same number of elements")
    if stats_out.numel() < 2:
        raise ValueError("adam_apply_noalloc: stats_out needs 2 float64 elements")
    vp = ctypes.c_void_p
    check(_lib.load().b2dp_adam_apply_noalloc(vp(w.data_ptr()), vp(m.data_ptr()), vp(v.data_ptr()),
                                              vp(g.data_ptr()), n, ctypes.byref(hyper), vp(stats_out.data_ptr()),
                                              _stream_handle(stream, w.device.index)))


class Layout:
    """The bucketing layer: K tensors -> buckets -> per-rank slices (host arithmetic only)."""

    def __init__(self, numels, nranks=1, wire="f32", bucket_bytes=0):
        L = _lib.load()
        numels = [int(x) for x in numels]
        arr = (ctypes.c_int64 * max(1, len(numels)))(*numels)
        h = ctypes.c_void_p()
        try:
            wire_id = _DTYPES[wire]
        except KeyError:
            raise ValueError("wire dtype must be f32 or bf16, got %r" % (wire,))
        rc = L.b2dp_layout_create(len(numels), arr, int(nranks), wire_id, int(bucket_bytes), ctypes.byref(h))
        if rc == _lib.ERR_INVALID:
            raise ValueError(L.b2dp_last_error().decode())
        check(rc)
        self._h = h
        self._destroy = L.b2dp_layout_destroy   # bound now: module globals may be gone at interpreter exit
        self.numels = numels
        self.nranks = int(nranks)
        self.wire = wire_id

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and getattr(self, "_destroy", None):
            self._destroy(h)

    num_tensors = property(lambda s: _lib.load().b2dp_layout_num_tensors(s._h))
    num_buckets = property(lambda s: _lib.load().b2dp_layout_num_buckets(s._h))
    total_elems = property(lambda s: _lib.load().b2dp_layout_total_elems(s._h))
    shard_elems = property(lambda s: _lib.load().b2dp_layout_shard_elems(s._h))

    def tensor_offset(self, k):
        return check(_lib.load().b2dp_layout_tensor_offset(self._h, k))

    def tensor_bucket(self, k):
        return check(_lib.load().b2dp_layout_tensor_bucket(self._h, k))

    def bucket_info(self, b):
        info = (ctypes.c_int64 * 6)()
        rc = _lib.load().b2dp_layout_bucket_info(self._h, b, info)
        if rc == _lib.ERR_INVALID:
            raise ValueError(_lib.load().b2dp_last_error().decode())
        check(rc)
        keys = ("first", "count", "offset", "elems", "slice", "shard_off")
        return dict(zip(keys, (int(x) for x in info)))


class Context:
    def __init__(self, handle, nranks, ranks, devices, local):
        self._h = handle
        self._destroy = _lib.load().b2dp_ctx_destroy
        self.nranks = nranks
        self.ranks = list(ranks)      # global rank of each local slot
        self.devices = list(devices)  # CUDA device of each local slot
        self.is_local = local

    @classmethod
    def local(cls, devices):
        """One process drives all ranks (MirroredStrategy(), imagenet-resnet50-mirror.py:21).
        Repeating a device id gives emulated ranks on one GPU."""
        devices = [int(d) for d in devices]
        arr = (ctypes.c_int32 * len(devices))(*devices)
        h = ctypes.c_void_p()
        check(_lib.load().b2dp_ctx_create_local(len(devices), arr, ctypes.byref(h)))
        return cls(h, len(devices), range(len(devices)), devices, True)

    @classmethod
    def rank(cls, rank, nranks, device):
        """One process per GPU (hvd.init(), imagenet-resnet50-hvd.py:16; one MWMS worker,
        imagenet-resnet50-multiworkers.py:24-25)."""
        h = ctypes.c_void_p()
        check(_lib.load().b2dp_ctx_create_rank(int(rank), int(nranks), int(device), ctypes.byref(h)))
        return cls(h, int(nranks), [int(rank)], [int(device)], False)

    def set_timeout(self, seconds):
        check(_lib.load().b2dp_ctx_set_timeout(self._h, float(seconds)))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and getattr(self, "_destroy", None):
            self._destroy(h)


class _DevMem:
    """__cuda_array_interface__ carrier so torch can view a library-owned buffer."""

    def __init__(self, ptr, shape, typestr, owner):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": typestr, "data": (ptr, False), "version": 2}
        self._owner = owner


def _stream_handle(stream, device):
    if stream is None:
        import torch
        return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)
    if hasattr(stream, "cuda_stream"):
        return ctypes.c_void_p(stream.cuda_stream)
    return ctypes.c_void_p(int(stream))


class issue_group:
    """``with b200dp.issue_group():`` around the loop in which ONE host thread issues a collective
    for every rank of a local context.  Between the first and the last rank's call nothing may block
    the host on a device -- the kernels already launched wait inside the GPU for the ranks that are
    still to be launched (the reason NCCL has ncclGroupStart/End).  The calls of this library never
    block; what can is Python's garbage collector running some object's destructor at an arbitrary
    point (an old Plan's destroy synchronises its devices; torch frees memory): the context manager
    keeps the cyclic collector off for the duration of the loop."""

    def __enter__(self):
        import gc
        self._was = gc.isenabled()
        gc.disable()
        return self

    def __exit__(self, *exc):
        import gc
        if self._was:
            gc.enable()
        return False


def ptr_array(tensors_or_ptrs):
    """ctypes array of device pointers; build it once and reuse it when the tensors' addresses
    are stable (214 data_ptr() calls cost more host time than the pack kernel takes)."""
    return _ptr_array(tensors_or_ptrs)


def _ptr_array(tensors_or_ptrs):
    if isinstance(tensors_or_ptrs, ctypes.Array):
        return tensors_or_ptrs
    vals = [int(t.data_ptr()) if hasattr(t, "data_ptr") else int(t) for t in tensors_or_ptrs]
    return (ctypes.c_void_p * max(1, len(vals)))(*vals)


class Plan:
    """A layout bound to a context: owns the symmetric buffers and issues the hot path."""

    def __init__(self, ctx, layout):
        self.ctx = ctx
        self.layout = layout
        h = ctypes.c_void_p()
        check(_lib.load().b2dp_plan_create(ctx._h, layout._h, ctypes.byref(h)))
        self._h = h
        self._L = _lib.load()
        self._destroy = self._L.b2dp_plan_destroy

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and getattr(self, "_destroy", None):
            self._destroy(h)

    # -- bootstrap (rank contexts) ----------------------------------------------------
    def exchange(self, session, timeout_ms=60000):
        """Moves every rank's block fd to every other rank and maps the peers."""
        if self.ctx.is_local or self.ctx.nranks == 1:
            return
        fd = ctypes.c_int(-1)
        check(self._L.b2dp_plan_export_fd(self._h, ctypes.byref(fd)))
        n, rank = self.ctx.nranks, self.ctx.ranks[0]
        fds = (ctypes.c_int * n)()
        try:
            check(self._L.b2dp_fd_exchange(rank, n, str(session).encode(), fd.value, fds, int(timeout_ms)))
            for r in range(n):
                if r != rank:
                    check(self._L.b2dp_plan_attach_peer(self._h, r, fds[r]))
        finally:
            os.close(fd.value)
            for r in range(n):
                if fds[r] >= 0:
                    os.close(fds[r])

    def enable_multicast(self, session=None, timeout_ms=60000):
        """Path B: binds every rank's block to one NVSwitch multicast object and switches the
        collectives to multimem.ld_reduce / multimem.st.  Raises if the box has no multicast.
        For rank contexts `session` must be a job-unique string (as for exchange())."""
        L = self._L
        if self.ctx.is_local:
            check(L.b2dp_plan_multicast_create(self._h, None))
        else:
            n, rank = self.ctx.nranks, self.ctx.ranks[0]
            own = ctypes.c_int(-1)
            check(L.b2dp_plan_export_fd(self._h, ctypes.byref(own)))   # dummy payload for ranks != 0
            mcfd = ctypes.c_int(-1)
            try:
                if rank == 0:
                    check(L.b2dp_plan_multicast_create(self._h, ctypes.byref(mcfd)))
                fds = (ctypes.c_int * n)()

                def rendezvous(tag, payload):
                    check(L.b2dp_fd_exchange(rank, n, ("%s-%s" % (session, tag)).encode(), payload, fds, int(timeout_ms)))
                    got = [fds[r] for r in range(n)]
                    return got

                got = rendezvous("mc0", mcfd.value if rank == 0 else own.value)
                check(L.b2dp_plan_multicast_join(self._h, got[0]))
                for f in got:
                    os.close(f)
                for f in rendezvous("mc1", own.value):     # everyone has joined
                    os.close(f)
                check(L.b2dp_plan_multicast_bind(self._h))
                for f in rendezvous("mc2", own.value):     # everyone has bound
                    os.close(f)
            finally:
                os.close(own.value)
                if mcfd.value >= 0:
                    os.close(mcfd.value)
        check(L.b2dp_plan_set_path(self._h, _lib.PATH_MULTICAST))

    def set_path(self, path):
        check(self._L.b2dp_plan_set_path(self._h, {"peer": _lib.PATH_PEER, "multicast": _lib.PATH_MULTICAST}.get(path, path)))

    def set_tuning(self, **knobs):
        """grid_cap / mc_grid_cap / mc_unroll / peer_unroll -- same values on every rank."""
        for k, v in knobs.items():
            check(self._L.b2dp_plan_set_tuning(self._h, _lib.TUNE[k], int(v)))

    @property
    def path(self):
        return "multicast" if self._L.b2dp_plan_path(self._h) == _lib.PATH_MULTICAST else "peer"

    @property
    def ready(self):
        return bool(self._L.b2dp_plan_ready(self._h))

    # -- buffers -----------------------------------------------------------------------
    def buffer_ptr(self, dev, which):
        p, n = ctypes.c_void_p(), ctypes.c_int64()
        check(self._L.b2dp_plan_buffer(self._h, dev, which, ctypes.byref(p), ctypes.byref(n)))
        return int(p.value or 0), int(n.value)

    def tensor(self, dev, which):
        """torch view (no copy) of one of the plan's flat buffers on local rank `dev`."""
        import torch
        ptr, n = self.buffer_ptr(dev, which)
        if which == BUF_STATS:
            typestr, dt = "<f8", torch.float64
        elif which == BUF_GRADS and self.layout.wire == BF16:
            typestr, dt = "<i2", torch.bfloat16
        else:
            typestr, dt = "<f4", torch.float32
        device = "cuda:%d" % self.ctx.devices[dev]
        if n == 0:
            return torch.empty(0, dtype=dt, device=device)
        t = torch.as_tensor(_DevMem(ptr, (n,), typestr, self), device=device)
        return t.view(dt) if t.dtype != dt else t

    def _stream(self, stream, dev):
        device = self.ctx.devices[dev] if 0 <= dev < len(self.ctx.devices) else 0  # the library rejects bad dev
        return _stream_handle(stream, device)

    # -- the hot path --------------------------------------------------------------------
    def pack(self, dev, grads, prescale=1.0, bucket=-1, stream=None):
        if len(grads) != max(1, self.layout.num_tensors) if isinstance(grads, ctypes.Array) \
                else len(grads) != self.layout.num_tensors:
            raise ValueError("pack: %d tensors, plan has %d" % (len(grads), self.layout.num_tensors))
        check(self._L.b2dp_pack(self._h, dev, bucket, _ptr_array(grads), F32, prescale,
                                self._stream(stream, dev)))

    def load_weights(self, dev, weights, stream=None):
        if len(weights) != self.layout.num_tensors:
            raise ValueError("load_weights: %d tensors, plan has %d" % (len(weights), self.layout.num_tensors))
        check(self._L.b2dp_load_weights(self._h, dev, _ptr_array(weights),
                                        self._stream(stream, dev)))

    def unpack(self, dev, outs, which=BUF_WEIGHTS, bucket=-1, stream=None):
        if len(outs) != self.layout.num_tensors:
            raise ValueError("unpack: %d tensors, plan has %d" % (len(outs), self.layout.num_tensors))
        check(self._L.b2dp_unpack(self._h, dev, bucket, which, _ptr_array(outs),
                                  self._stream(stream, dev)))

    def step(self, dev, opt, hyper, bucket=-1, stream=None):
        check(self._L.b2dp_step(self._h, dev, bucket, _OPTS[opt], ctypes.byref(hyper),
                                self._stream(stream, dev)))

    def pack_step(self, dev, grads, opt, hyper, prescale=1.0, bucket=-1, stream=None):
        """pack + step in one call (one kernel when there is a single rank)."""
        if not isinstance(grads, ctypes.Array) and len(grads) != self.layout.num_tensors:
            raise ValueError("pack_step: %d tensors, plan has %d" % (len(grads), self.layout.num_tensors))
        check(self._L.b2dp_pack_step(self._h, dev, bucket, _ptr_array(grads), F32, prescale, _OPTS[opt],
                                     ctypes.byref(hyper), self._stream(stream, dev)))

    def reduce_scatter_apply(self, dev, opt, hyper, bucket=-1, stream=None):
        check(self._L.b2dp_reduce_scatter_apply(self._h, dev, bucket, _OPTS[opt], ctypes.byref(hyper),
                                                self._stream(stream, dev)))

    def all_gather(self, dev, bucket=-1, stream=None):
        check(self._L.b2dp_all_gather(self._h, dev, bucket, self._stream(stream, dev)))

    def allreduce(self, dev, postscale=1.0, bucket=-1, stream=None):
        check(self._L.b2dp_allreduce(self._h, dev, bucket, postscale,
                                     self._stream(stream, dev)))

    def broadcast(self, dev, root=0, which=BUF_WEIGHTS, stream=None):
        """Collective: the root rank's flat buffer overwrites every rank's copy."""
        check(self._L.b2dp_broadcast(self._h, dev, which, int(root), self._stream(stream, dev)))

    def gather_state(self, dev, which, stream=None):
        """Collective: assemble the sharded optimizer state `which` (BUF_STATE0 / BUF_STATE1) into
        BUF_REDUCED of every rank; unpack(which=BUF_REDUCED) then yields per-variable tensors."""
        check(self._L.b2dp_gather_state(self._h, dev, which, self._stream(stream, dev)))

    def state_tensors(self, dev, which, like):
        """gather_state(which) + unpack(BUF_REDUCED) into new tensors shaped like `like`: the sharded
        optimizer state `which` (BUF_STATE0 / BUF_STATE1) as per-variable tensors.  gather_state is
        collective: with n > 1 every rank must make this call (a local context: call it for every
        dev, from one thread per device or -- emulated ranks -- use gather_state for all ranks
        first).  Overwrites BUF_REDUCED: call it between steps."""
        import torch
        self.gather_state(dev, which)
        outs = [torch.empty_like(t, memory_format=torch.contiguous_format) for t in like]
        self.unpack(dev, outs, which=BUF_REDUCED)
        return outs

    def allreduce_small(self, dev, src, dst, postscale=1.0, stream=None):
        """Collective one-shot all-reduce of <= 1024 fp32 values (torch CUDA tensors src -> dst)."""
        if src.numel() != dst.numel():
            raise ValueError("allreduce_small: src and dst sizes differ")
        vp = ctypes.c_void_p
        check(self._L.b2dp_allreduce_small(self._h, dev, vp(src.data_ptr()), vp(dst.data_ptr()), src.numel(),
                                           postscale, self._stream(stream, dev)))

    def check(self, dev, stream=None):
        check(self._L.b2dp_plan_check(self._h, dev, self._stream(stream, dev)))
