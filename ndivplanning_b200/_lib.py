"""ctypes binding of libndivplan.so (include/ndivplan.h).  No CPU fallback: if the
library is missing or the device is not sm_100 every entry point raises."""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NDP_LIB_PATH") or os.path.join(_HERE, "libndivplan.so")   # NDP_LIB_PATH: dev builds (profiling counters)

PREC_FP32, PREC_F16X3, PREC_F16, PREC_F16X2, PREC_F16_REFINE = 0, 1, 2, 3, 4
REFINE_R = 128


class NdpError(RuntimeError):
    pass


class PlannerDesc(C.Structure):
    _fields_ = [("E", C.c_int), ("nd", C.c_int), ("A", C.c_int), ("H", C.c_int), ("T", C.c_int),
                ("precision", C.c_int)]


class GanDesc(C.Structure):
    _fields_ = [("C", C.c_int32), ("nd", C.c_int32), ("A", C.c_int32), ("S", C.c_int32), ("max_F", C.c_int32),
                ("d_steps", C.c_int32), ("margin", C.c_float), ("div_factor", C.c_float), ("lr", C.c_double),
                ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double)]


_vp, _i, _i64, _sz, _f = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float
_PROTOS = {
    "ndp_abi_version": (_i, []),
    "ndp_last_error": (C.c_char_p, []),
    "ndp_device_check": (_i, []),
    "ndp_launch_count": (C.c_uint64, []),
    "ndp_planner_create": (_i, [C.POINTER(_vp), C.POINTER(PlannerDesc)]),
    "ndp_planner_destroy": (None, [_vp]),
    "ndp_planner_load": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), _i, _vp]),
    "ndp_planner_workspace_bytes": (_i, [_vp, _i, _i, C.POINTER(_sz)]),
    "ndp_rollout_score": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ndp_generator_forward": (_i, [_vp, _vp, _i, _vp, _vp]),
    "ndp_forward_step": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "ndp_sample_uniform": (_i, [C.c_uint64, C.c_uint64, _vp, _sz, _vp]),
    "ndp_select_workspace_bytes": (_i, [_i, _i, _i, C.POINTER(_sz)]),
    "ndp_select": (_i, [_vp, _i, _i, _i, _i64, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ndp_select_static_max_bytes": (_i, [C.POINTER(_sz)]),
    "ndp_select_static": (_i, [_vp, _i, _i, _i, _i64, _vp, _vp, _vp, _vp]),
    "ndp_plan_workspace_bytes": (_i, [_vp, _i, _i, _i, C.POINTER(_sz)]),
    "ndp_plan_select": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ndp_plan_fallback": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ndp_plan_select_bulk": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ndp_plan_select_refine": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i64, _f, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ndp_plan_verify_info": (_i, [_vp, _i, _i, _i, _vp, _sz, _vp, _vp]),
    "ndp_gather_best_action": (_i, [_vp, _vp, _i, _i, _i64, _vp, _vp, _vp]),
    "ndp_shard_pack": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp]),
    "ndp_shard_merge": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ndp_shard_exchange": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ndp_ndiv_workspace_bytes": (_i, [_i, _i, _i, _i, C.POINTER(_sz)]),
    "ndp_ndiv_rowsums": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "ndp_ndiv_loss_grad": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "ndp_ndiv_forward_backward": (_i, [_vp, _vp, _i, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ndp_ndiv_forward": (_i, [_vp, _vp, _i, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ndp_ndiv_gradz_workspace_bytes": (_i, [_i, _i, _i, _i, C.POINTER(_sz)]),
    "ndp_ndiv_backward": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ndp_pairwise": (_i, [_vp, _i, _i, _i, _i, _vp, _vp]),
    "ndp_pairwise_backward": (_i, [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp]),
    "ndp_gan_create": (_i, [C.POINTER(_vp), C.POINTER(GanDesc)]),
    "ndp_gan_destroy": (None, [_vp]),
    "ndp_gan_load": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), _vp]),
    "ndp_gan_export": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), _vp]),
    "ndp_gan_train_step": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp]),
    "ndp_gan_train_step_dp": (_i, [_vp, _vp, _i, _vp, _vp, _vp, _i, _vp, _vp]),
    "ndp_gan_set_lr": (_i, [_vp, C.c_double, _vp]),
    "ndp_gan_param_counts": (_i, [_vp, C.POINTER(_sz), C.POINTER(_sz)]),
    "ndp_gan_grads": (_i, [_vp, _vp, _vp, _vp]),
    "ndp_gan_last_actions": (_i, [_vp, _vp, _i, _vp]),
    "ndp_encoder_create": (_i, [C.POINTER(_vp), _i, _i]),
    "ndp_encoder_destroy": (None, [_vp]),
    "ndp_encoder_load": (_i, [_vp, C.POINTER(_vp), C.POINTER(_vp), _vp]),
    "ndp_encoder_forward": (_i, [_vp, _vp, _i, _vp, _vp]),
    "ndp_comm_unique_id": (_i, [_vp]),
    "ndp_comm_create": (_i, [C.POINTER(_vp), _vp, _i, _i]),
    "ndp_comm_allreduce_min_u64": (_i, [_vp, _vp, _i, _vp]),
    "ndp_comm_allreduce_sum_f32": (_i, [_vp, _vp, _i, _vp]),
    "ndp_comm_allgather_f32": (_i, [_vp, _vp, _vp, _i, _vp]),
    "ndp_comm_allgather_bytes": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "ndp_comm_rank": (_i, [_vp]),
    "ndp_comm_world": (_i, [_vp]),
    "ndp_comm_peer_status": (_i, [_vp, C.POINTER(_i), _vp]),
    "ndp_comm_destroy": (None, [_vp]),
    "ndp_ndiv_sharded_workspace_bytes": (_i, [_i, _i, _i, _i, C.POINTER(_sz)]),
    "ndp_ndiv_sharded_forward_backward": (_i, [_vp, _vp, _vp, _i, _i, _i, _f, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
}
EXPORTED_SYMBOLS = tuple(_PROTOS)

_lib = None


def lib():
    """Loads libndivplan.so once. Raises NdpError when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise NdpError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
                           f"g.build()'` (ndivplanning_b200/csrc/build.sh). There is no CPU fallback.")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(l, name)
            fn.restype, fn.argtypes = res, args
        if l.ndp_abi_version() != 1:
            raise NdpError("libndivplan ABI version mismatch")
        _lib = l
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise NdpError(f"libndivplan error {rc}: {lib().ndp_last_error().decode()}")


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(*tensors):
    """The product path is CUDA-only: refuse CPU tensors loudly instead of falling back."""
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NdpError("ndivplanning_b200 runs on sm_100a only; got a CPU tensor (no CPU fallback)")
    if not torch.cuda.is_available():
        raise NdpError("CUDA is not available; ndivplanning_b200 has no CPU fallback")


def f32c(t):
    return t.detach().to(torch.float32).contiguous()


def launch_count() -> int:
    return int(lib().ndp_launch_count())
