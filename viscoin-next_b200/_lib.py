"""ctypes binding of libvcn_b200.so (the C ABI declared in include/vcn_b200.h).

Replaces the reference's plugin loader ``custom_ops.get_plugin`` (stylegan2_ada/torch_utils/
custom_ops.py:46-124) and its ``_init()`` wrappers (bias_act.py:41-51, upfirdn2d.py:26-35).
Unlike the reference there is NO fallback: a missing library, a missing symbol or a non-sm_100
device raises.
"""

import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libvcn_b200.so")

F32, F16, BF16, F64 = 0, 1, 2, 3

_c = ctypes
_vp, _i, _i64, _f, _sz = _c.c_void_p, _c.c_int, _c.c_int64, _c.c_float, _c.c_size_t


class UpfirdnEpilogue(_c.Structure):
    """vcn_upfirdn2d_epilogue"""

    _fields_ = [
        ("scale", _vp), ("noise", _vp), ("noise_stride_n", _i64), ("bias", _vp),
        ("act", _i), ("alpha", _f), ("act_gain", _f), ("clamp", _f),
    ]


class ConvDesc(_c.Structure):
    """vcn_conv_desc"""

    _fields_ = [
        ("kind", _i), ("n", _i), ("h", _i), ("w", _i), ("cin_pad", _i), ("cout", _i), ("kh", _i), ("kw", _i),
        ("out_h", _i), ("out_w", _i), ("y_dtype", _i), ("ys", _i64 * 4), ("w_per_sample", _i), ("tf32", _i),
    ]


CONV_FPROP, CONV_TCONV2, CONV_DGRAD, CONV_SCONV2 = 0, 1, 2, 3

# name -> (restype, argtypes); must list every symbol include/vcn_b200.h declares (tests check this).
SIGNATURES = {
    "vcn_abi_version": (_i, []),
    "vcn_last_error": (_c.c_char_p, []),
    "vcn_launch_count": (_c.c_uint64, []),
    "vcn_conv_profile_counters": (_c.c_int, [_c.POINTER(_c.c_uint64), _c.c_int]),
    "vcn_device_info": (_i, [_i, _c.POINTER(_i), _c.POINTER(_i64), _c.POINTER(_i), _c.POINTER(_i)]),
    "vcn_bias_act_workspace": (_sz, [_i, _i64, _i64, _i64]),
    "vcn_bias_act": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _i, _i, _i, _f, _f, _f, _i64, _i64, _i64, _vp]),
    "vcn_bias_act_static_scratch_bytes": (_sz, []),
    "vcn_bias_act_static": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _f, _f, _i64, _i64, _i64, _vp]),
    "vcn_upfirdn2d": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _c.POINTER(_i64 * 4), _i, _i, _c.POINTER(_i64 * 4),
                           _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _f, _c.POINTER(UpfirdnEpilogue), _vp]),
    "vcn_modulate_pack": (_i, [_vp, _i, _c.POINTER(_i64 * 4), _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "vcn_pack_weights": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _i, _vp, _vp]),
    "vcn_modulate_pack_ex": (_i, [_vp, _i, _c.POINTER(_i64 * 4), _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "vcn_pack_weights_ex": (_i, [_vp, _i, _i, _i, _i, _i, _i, _vp, _i, _i, _vp, _vp]),
    "vcn_modconv_bwd_pack_ex": (_i, [_vp, _vp, _i, _c.POINTER(_i64 * 4), _c.POINTER(_i64 * 4), _c.POINTER(UpfirdnEpilogue),
                                     _vp, _i, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "vcn_demod_coefs": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "vcn_pack_weights_mod": (_i, [_vp, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "vcn_torgb_bwd_fused": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "vcn_set_conv_sm_reserve": (_i, [_i]),
    "vcn_conv_igemm": (_i, [_vp, _vp, _vp, _c.POINTER(ConvDesc), _c.POINTER(UpfirdnEpilogue), _vp]),
    "vcn_modconv_bwd_pack": (_i, [_vp, _vp, _i, _c.POINTER(_i64 * 4), _c.POINTER(_i64 * 4), _c.POINTER(UpfirdnEpilogue),
                                  _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "vcn_modconv_bwd_finish": (_i, [_vp, _c.POINTER(_i64 * 4), _vp, _i, _c.POINTER(_i64 * 4), _vp, _vp, _i, _i, _i, _i, _vp]),
    "vcn_modconv_bwd_finish_cl": (_i, [_vp, _vp, _i, _c.POINTER(_i64 * 4), _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "vcn_lpips_tap_fwd": (_i, [_vp, _vp, _i, _c.POINTER(_i64 * 4), _c.POINTER(_i64 * 4), _vp, _vp, _i, _i, _i, _i, _vp]),
    "vcn_lpips_tap_bwd": (_i, [_vp, _vp, _i, _c.POINTER(_i64 * 4), _c.POINTER(_i64 * 4), _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "vcn_conv_wgrad": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "vcn_demod_grad": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "vcn_maxpool2x2_fwd": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "vcn_image_prep_fwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "vcn_image_prep_bwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "vcn_l1l2_fwd": (_i, [_vp, _vp, _vp, _i64, _vp]),
    "vcn_l1l2_bwd": (_i, [_vp, _vp, _vp, _vp, _i64, _vp]),
    "vcn_conv_first3": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "vcn_maxpool3x3s2_fwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "vcn_maxpool3x3s2_bwd": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "vcn_relu_mask_bwd": (_i, [_vp, _vp, _vp, _vp, _i, _i64, _vp]),
    "vcn_maxpool2x2_relu_bwd": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _vp]),
}

_lock = threading.Lock()
_lib = None
_checked_devices = set()


class VcnError(RuntimeError):
    pass


def load():
    """Load the shared library (once).  Raises VcnError when it is missing - there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise VcnError(
                "libvcn_b200.so not found at %s; build it with `python viscoin-next_b200/build.py` "
                "(or __graft_entry__.build()).  There is no fallback implementation." % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name, None)
            if fn is None:
                raise VcnError("libvcn_b200.so does not export %s (stale build?)" % name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def last_error():
    msg = load().vcn_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(status, what):
    if status != 0:
        raise VcnError("%s failed (status %d): %s" % (what, status, last_error()))


def require_device(index):
    """Hard error unless CUDA device `index` is sm_100 (reference: silent ref fallback, bias_act.py:47-51)."""
    if index in _checked_devices:
        return
    lib = load()
    check(lib.vcn_device_info(int(index), None, None, None, None), "vcn_device_info")
    _checked_devices.add(index)


def dtype_code(torch_dtype):
    import torch
    table = {torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16, torch.float64: F64}
    if torch_dtype not in table:
        raise VcnError("unsupported dtype %s" % torch_dtype)
    return table[torch_dtype]


def ptr(t):
    """Device pointer of a tensor, or None (NULL) for an absent one."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def current_stream(device):
    import torch
    return torch.cuda.current_stream(device).cuda_stream
