"""ctypes binding of libraygun_b200.so (the C ABI declared in include/raygun_b200.h).

There is NO fallback: if the shared library is missing or the device is not sm_100 every
operator raises.  PyTorch is used for device memory and streams only.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import torch

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "lib", "libraygun_b200.so")

RG_F32, RG_BF16, RG_U8 = 0, 1, 2
ACT_NONE, ACT_RELU, ACT_LEAKY, ACT_TANH, ACT_SIGMOID = 0, 1, 2, 3, 4
PAD_ZEROS, PAD_REFLECT, PAD_REPLICATE = 0, 1, 2
NORM_INSTANCE, NORM_BATCH = 0, 1
LOSS_MSE_CONST, LOSS_SMOOTHL1, LOSS_L1, LOSS_MEAN = 0, 1, 2, 3
E_UNSUPPORTED = -2

# every exported symbol of include/raygun_b200.h (checked by tests/test_abi.py)
SYMBOLS = [
    "rg_version", "rg_last_error", "rg_device_check",
    "rg_conv3d_direct_fprop", "rg_conv3d_direct_dgrad", "rg_conv3d_direct_wgrad", "rg_conv3d_direct_wgrad_workspace_bytes",
    "rg_plan_create", "rg_plan_destroy", "rg_plan_workspace_bytes", "rg_pack_weights",
    "rg_igemm_run", "rg_igemm_wgrad", "rg_ring_supported", "rg_ring_config", "rg_ring_pack_weights", "rg_ring_workspace_bytes",
    "rg_ring_run", "rg_ring_run_head", "rg_smallk_wgrad_supported", "rg_smallk_wgrad_workspace_bytes", "rg_smallk_wgrad",
    "rg_stem_conv_supported", "rg_stem_conv_run", "rg_upconv2_supported", "rg_upconv2_run",
    "rg_halo_wgrad_supported", "rg_halo_wgrad_workspace_bytes", "rg_halo_wgrad",
    "rg_pw_conv_fprop", "rg_pw_conv_dgrad", "rg_pw_conv_wgrad_workspace_bytes", "rg_pw_conv_wgrad", "rg_smallcin_dgrad",
    "rg_copy5", "rg_im2col_hw", "rg_pad3d_fwd", "rg_pad3d_bwd", "rg_pad3d_fold", "rg_pad_border_fill",
    "rg_reduce_ws_floats", "rg_rows_sum", "rg_norm_stats", "rg_norm_stats_static", "rg_norm_finalize", "rg_norm_apply", "rg_norm_bwd",
    "rg_act_fwd", "rg_act_bwd", "rg_act_bwd_bias",
    "rg_maxpool3d_fwd", "rg_maxpool3d_bwd", "rg_upsample_nearest_fwd", "rg_upsample_nearest_bwd",
    "rg_resize_trilinear_fwd", "rg_resize_trilinear_bwd", "rg_loss_fwd_grad", "rg_u8_normalise", "rg_denorm_clamp_u8", "rg_h2d_block", "rg_d2h_block", "rg_adam_step", "rg_adam_flat",
]


class Tensor5(C.Structure):
    _fields_ = [(k, C.c_int64) for k in ("n", "c", "d", "h", "w", "sn", "sc", "sd", "sh", "sw")] + \
               [("dtype", C.c_int32), ("_pad", C.c_int32)]


class ConvGeom(C.Structure):
    _fields_ = [("kernel", C.c_int32 * 3), ("stride", C.c_int32 * 3), ("pad", C.c_int32 * 3),
                ("groups", C.c_int32), ("act", C.c_int32), ("slope", C.c_float)]


class IgemmDesc(C.Structure):
    _fields_ = [("mode", C.c_int32), ("n", C.c_int32), ("cin", C.c_int32), ("cout", C.c_int32),
                ("in_dhw", C.c_int32 * 3), ("out_dhw", C.c_int32 * 3),
                ("kernel", C.c_int32 * 3), ("stride", C.c_int32 * 3), ("pad", C.c_int32 * 3),
                ("x_strides", C.c_int64 * 4), ("y_strides", C.c_int64 * 4),
                ("act", C.c_int32), ("slope", C.c_float),
                ("has_bias", C.c_int32), ("want_stats", C.c_int32), ("out_f32", C.c_int32)]


class Upconv2Desc(C.Structure):
    _fields_ = [("n", C.c_int32), ("cin", C.c_int32), ("cout", C.c_int32), ("in_dhw", C.c_int32 * 3),
                ("x_strides", C.c_int64 * 4), ("y_strides", C.c_int64 * 4),
                ("act", C.c_int32), ("slope", C.c_float), ("has_bias", C.c_int32), ("_pad", C.c_int32)]


_lib = None
_lock = threading.Lock()
_checked_devices = set()
launch_count = 0          # kernels-launching C-ABI calls made (bench.py reports it)


class RaygunB200Error(RuntimeError):
    pass


def lib():
    """Load the library (once).  Raises if it has not been built: no silent fallback."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise RaygunB200Error(
                        f"{LIB_PATH} not found: build it with `python -m raygun_b200._build` "
                        "(raygun_b200 has no CPU / PyTorch fallback)")
                l = C.CDLL(LIB_PATH)
                l.rg_last_error.restype = C.c_char_p
                l.rg_plan_workspace_bytes.restype = C.c_size_t
                l.rg_halo_wgrad_workspace_bytes.restype = C.c_size_t
                l.rg_ring_workspace_bytes.restype = C.c_size_t
                l.rg_pw_conv_wgrad_workspace_bytes.restype = C.c_size_t
                l.rg_conv3d_direct_wgrad_workspace_bytes.restype = C.c_size_t
                l.rg_smallk_wgrad_workspace_bytes.restype = C.c_size_t
                l.rg_reduce_ws_floats.restype = C.c_longlong
                _lib = l
    return _lib


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = lib().rg_last_error().decode(errors="replace")
        raise RaygunB200Error(f"{what}: error {rc}: {msg}")


def _dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return RG_F32
    if t.dtype == torch.bfloat16:
        return RG_BF16
    if t.dtype == torch.uint8:
        return RG_U8
    raise RaygunB200Error(f"unsupported dtype {t.dtype}")


def t5(t: torch.Tensor) -> Tensor5:
    """Describe a 5-D (N,C,D,H,W) tensor view (any strides)."""
    if t.dim() != 5:
        raise RaygunB200Error(f"expected a 5-D tensor, got shape {tuple(t.shape)}")
    s = t.stride()
    return Tensor5(t.shape[0], t.shape[1], t.shape[2], t.shape[3], t.shape[4], s[0], s[1], s[2], s[3], s[4],
                   _dtype_code(t), 0)


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def stream_ptr(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_device(t: torch.Tensor):
    """Fail loudly unless `t` lives on an sm_100 CUDA device."""
    if not t.is_cuda:
        raise RaygunB200Error("raygun_b200 operators run on a CUDA (sm_100a) device only; got a CPU tensor. "
                              "There is no CPU fallback.")
    idx = t.device.index if t.device.index is not None else torch.cuda.current_device()
    if idx not in _checked_devices:
        check(lib().rg_device_check(C.c_int(idx)), "rg_device_check")
        _checked_devices.add(idx)


def count_launch(n: int = 1):
    global launch_count
    launch_count += n


def i3(v):
    return (C.c_int32 * 3)(*[int(x) for x in v])
