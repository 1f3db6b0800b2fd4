"""ctypes binding of libb200dp.so (include/b200dp.h).  There is no fallback: if the
library is not built this module raises, and the library itself refuses to create a
context on anything but an sm_100 GPU."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "libb200dp.so")

OK = 0
ERR_INVALID, ERR_CUDA, ERR_NO_DEVICE, ERR_ARCH, ERR_STATE, ERR_TIMEOUT, ERR_SYS, ERR_TOO_LARGE = range(-1, -9, -1)
F32, BF16 = 0, 1
OPT_NONE, OPT_ADAM, OPT_SGDM = 0, 1, 2
BUF_GRADS, BUF_WEIGHTS, BUF_REDUCED, BUF_STATE0, BUF_STATE1, BUF_STATS = range(6)
NOALLOC_MAX_ELEMS = 1 << 26
PATH_PEER, PATH_MULTICAST = 0, 1
TUNE = {"grid_cap": 0, "mc_grid_cap": 1, "mc_unroll": 2, "peer_unroll": 3, "peer_vec": 4, "mc_bf16_grid_cap": 5, "skip_own": 6}


class Hyper(ctypes.Structure):
    _fields_ = [
        ("lr", ctypes.c_float), ("beta1", ctypes.c_float), ("beta2", ctypes.c_float), ("eps", ctypes.c_float),
        ("momentum", ctypes.c_float), ("weight_decay", ctypes.c_float),
        ("nesterov", ctypes.c_int32), ("want_stats", ctypes.c_int32),
        ("step", ctypes.c_int64), ("postscale", ctypes.c_float), ("reserved", ctypes.c_float),
    ]


class B200dpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("b200dp error %d: %s" % (code, msg))
        self.code = code


_vp = ctypes.c_void_p
_i32 = ctypes.c_int32
_i64 = ctypes.c_int64

# name -> (restype, argtypes); every symbol include/b200dp.h declares
SIGNATURES = {
    "b2dp_last_error": (ctypes.c_char_p, []),
    "b2dp_abi_version": (ctypes.c_int, []),
    "b2dp_layout_create": (ctypes.c_int, [_i32, ctypes.POINTER(_i64), _i32, _i32, _i64, ctypes.POINTER(_vp)]),
    "b2dp_layout_destroy": (None, [_vp]),
    "b2dp_layout_num_tensors": (_i32, [_vp]),
    "b2dp_layout_num_buckets": (_i32, [_vp]),
    "b2dp_layout_nranks": (_i32, [_vp]),
    "b2dp_layout_total_elems": (_i64, [_vp]),
    "b2dp_layout_shard_elems": (_i64, [_vp]),
    "b2dp_layout_tensor_offset": (_i64, [_vp, _i32]),
    "b2dp_layout_tensor_bucket": (_i32, [_vp, _i32]),
    "b2dp_layout_bucket_info": (ctypes.c_int, [_vp, _i32, ctypes.POINTER(_i64)]),
    "b2dp_ctx_create_local": (ctypes.c_int, [_i32, ctypes.POINTER(_i32), ctypes.POINTER(_vp)]),
    "b2dp_ctx_create_rank": (ctypes.c_int, [_i32, _i32, _i32, ctypes.POINTER(_vp)]),
    "b2dp_ctx_destroy": (None, [_vp]),
    "b2dp_ctx_nranks": (_i32, [_vp]),
    "b2dp_ctx_num_local": (_i32, [_vp]),
    "b2dp_ctx_set_timeout": (ctypes.c_int, [_vp, ctypes.c_double]),
    "b2dp_plan_create": (ctypes.c_int, [_vp, _vp, ctypes.POINTER(_vp)]),
    "b2dp_plan_destroy": (None, [_vp]),
    "b2dp_plan_export_fd": (ctypes.c_int, [_vp, ctypes.POINTER(ctypes.c_int)]),
    "b2dp_plan_attach_peer": (ctypes.c_int, [_vp, _i32, ctypes.c_int]),
    "b2dp_plan_ready": (_i32, [_vp]),
    "b2dp_plan_buffer": (ctypes.c_int, [_vp, _i32, _i32, ctypes.POINTER(_vp), ctypes.POINTER(_i64)]),
    "b2dp_plan_check": (ctypes.c_int, [_vp, _i32, _vp]),
    "b2dp_plan_multicast_create": (ctypes.c_int, [_vp, ctypes.POINTER(ctypes.c_int)]),
    "b2dp_plan_multicast_join": (ctypes.c_int, [_vp, ctypes.c_int]),
    "b2dp_plan_multicast_bind": (ctypes.c_int, [_vp]),
    "b2dp_plan_set_path": (ctypes.c_int, [_vp, _i32]),
    "b2dp_plan_path": (_i32, [_vp]),
    "b2dp_plan_set_tuning": (ctypes.c_int, [_vp, _i32, _i32]),
    "b2dp_pack": (ctypes.c_int, [_vp, _i32, _i32, ctypes.POINTER(_vp), _i32, ctypes.c_float, _vp]),
    "b2dp_pack_step": (ctypes.c_int, [_vp, _i32, _i32, ctypes.POINTER(_vp), _i32, ctypes.c_float, _i32,
                                      ctypes.POINTER(Hyper), _vp]),
    "b2dp_step": (ctypes.c_int, [_vp, _i32, _i32, _i32, ctypes.POINTER(Hyper), _vp]),
    "b2dp_allreduce": (ctypes.c_int, [_vp, _i32, _i32, ctypes.c_float, _vp]),
    "b2dp_reduce_scatter_apply": (ctypes.c_int, [_vp, _i32, _i32, _i32, ctypes.POINTER(Hyper), _vp]),
    "b2dp_all_gather": (ctypes.c_int, [_vp, _i32, _i32, _vp]),
    "b2dp_unpack": (ctypes.c_int, [_vp, _i32, _i32, _i32, ctypes.POINTER(_vp), _vp]),
    "b2dp_load_weights": (ctypes.c_int, [_vp, _i32, ctypes.POINTER(_vp), _vp]),
    "b2dp_broadcast": (ctypes.c_int, [_vp, _i32, _i32, _i32, _vp]),
    "b2dp_gather_state": (ctypes.c_int, [_vp, _i32, _i32, _vp]),
    "b2dp_allreduce_small": (ctypes.c_int, [_vp, _i32, _vp, _vp, _i32, ctypes.c_float, _vp]),
    "b2dp_adam_alpha": (ctypes.c_float, [ctypes.c_float, ctypes.c_float, ctypes.c_float, _i64]),
    "b2dp_adam_apply_noalloc": (ctypes.c_int, [_vp, _vp, _vp, _vp, _i64, ctypes.POINTER(Hyper), _vp, _vp]),
    "b2dp_fd_exchange": (ctypes.c_int, [_i32, _i32, ctypes.c_char_p, ctypes.c_int, ctypes.POINTER(ctypes.c_int), _i32]),
}

_lib = None


def load():
    """Returns the loaded library; raises if it was never built (python __graft_entry__.py
    or `python parallel-and-distributed-deep-learning_b200/build.py`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise ImportError(
            "libb200dp.so is not built (%s missing). Build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'`. There is no CPU or PyTorch fallback." % SO_PATH)
    lib = ctypes.CDLL(SO_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header and library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc < 0:
        raise B200dpError(rc, load().b2dp_last_error().decode("utf-8", "replace"))
    return rc
