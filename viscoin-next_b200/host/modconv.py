"""modconv - host-side mirror of ``modulated_conv2d`` (stylegan2_ada/training/networks.py:27-84) over the
vcn_conv_igemm C ABI (TMA-fed tcgen05/TMEM implicit GEMM, bf16 operands, fp32 accumulation).

Same signature as the reference function plus ``impl=`` (``'cuda'`` = sm_100a kernels, no fallback;
``'ref'`` = the plain-torch formulation of networks.py:27-84 through this package's ref ops).  The CUDA path
computes the reference's non-fused formulation (networks.py:63-72; identical to the fused one):

    xhat = x * s[n,ci]  (vcn_modulate_pack, -> bf16 NHWC)
    y0   = conv(xhat, W)                       up=1: vcn_conv_igemm(FPROP)
                                               up=2: vcn_conv_igemm(TCONV2) -> vcn_upfirdn2d (4x4, pad 1, gain 4)
    y    = y0 * d[n,co] + noise                fused into the GEMM epilogue (up=1) / the FIR epilogue (up=2)

``fused_modconv`` is accepted and ignored (both reference branches are the same function).  ``modconv_layer`` is
the fused SynthesisLayer body (networks.py:286-305): the bias + activation + gain + clamp of the following
``bias_act`` call goes into the same epilogue.
"""

import ctypes
import math

import torch

from .. import _lib
from . import bias_act as _bias_act
from . import upfirdn2d as _upfirdn2d


def _strides4(t):
    return (ctypes.c_int64 * 4)(*[int(s) for s in t.stride()])


def _round_up(v, m):
    return (v + m - 1) // m * m


# ---- thin launch helpers (one C-ABI call each) ---------------------------------------------------------------


def modulate_pack(x, styles):
    """x [N,C,H,W] (any dense-ish strides) * styles [N,C] (fp32 or None) -> bf16 NHWC [N,H,W,C_pad]."""
    lib = _lib.load()
    n, c, h, w = x.shape
    c_pad = _round_up(c, 8)
    out = torch.empty([n, h, w, c_pad], dtype=torch.bfloat16, device=x.device)
    if styles is not None:
        styles = styles.to(torch.float32).contiguous()
        assert styles.shape == (n, c)
    if n == 0:
        return out
    with torch.cuda.device(x.device):
        st = lib.vcn_modulate_pack(_lib.ptr(x), _lib.dtype_code(x.dtype), ctypes.byref(_strides4(x)), _lib.ptr(styles),
                                   _lib.ptr(out), n, c, c_pad, h, w, _lib.current_stream(x.device))
    _lib.check(st, "vcn_modulate_pack")
    return out


def pack_weights(weight, transpose_io=False, want_wsq=False):
    """W [Cout,Cin,kh,kw] -> (wt bf16 [kh*kw, rows, cols_pad], wsq fp32 [Cout,Cin] or None)."""
    lib = _lib.load()
    weight = weight.detach().contiguous()
    cout, cin, kh, kw = weight.shape
    rows, cols = (cin, cout) if transpose_io else (cout, cin)
    cols_pad = _round_up(cols, 8)
    wt = torch.empty([kh * kw, rows, cols_pad], dtype=torch.bfloat16, device=weight.device)
    wsq = torch.empty([cout, cin], dtype=torch.float32, device=weight.device) if want_wsq else None
    with torch.cuda.device(weight.device):
        st = lib.vcn_pack_weights(_lib.ptr(weight), _lib.dtype_code(weight.dtype), cout, cin, kh, kw, int(transpose_io),
                                  _lib.ptr(wt), cols_pad, _lib.ptr(wsq), _lib.current_stream(weight.device))
    _lib.check(st, "vcn_pack_weights")
    return wt, wsq


def demod_coefs(styles, wsq):
    """d[n,co] = rsqrt(sum_ci s^2 wsq[co,ci] + 1e-8) (networks.py:55-58)."""
    lib = _lib.load()
    styles = styles.to(torch.float32).contiguous()
    n, cin = styles.shape
    cout = wsq.shape[0]
    d = torch.empty([n, cout], dtype=torch.float32, device=styles.device)
    if n == 0:
        return d
    with torch.cuda.device(styles.device):
        st = lib.vcn_demod_coefs(_lib.ptr(styles), _lib.ptr(wsq), _lib.ptr(d), n, cin, cout, _lib.current_stream(styles.device))
    _lib.check(st, "vcn_demod_coefs")
    return d


def make_epilogue(scale=None, noise=None, bias=None, act=None, alpha=None, gain=None, clamp=None, keep=None):
    """vcn_upfirdn2d_epilogue from torch tensors.  noise: [N,1,H,W] / [H,W] fp32.  `keep` collects the tensors that
    must outlive the launch call."""
    e = _lib.UpfirdnEpilogue()
    hold = []
    if scale is not None:
        scale = scale.to(torch.float32).contiguous(); hold.append(scale)
        e.scale = scale.data_ptr()
    if noise is not None:
        noise = noise.to(torch.float32).contiguous(); hold.append(noise)
        e.noise = noise.data_ptr()
        e.noise_stride_n = noise.shape[-1] * noise.shape[-2] if noise.ndim == 4 and noise.shape[0] > 1 else 0
    if act is not None:
        spec = _bias_act.activation_funcs[act]
        e.act = spec.cuda_idx
        e.alpha = float(spec.def_alpha if alpha is None else alpha)
        e.act_gain = float(spec.def_gain if gain is None else gain)
        e.clamp = float(-1 if clamp is None else clamp)
        if bias is not None:
            bias = bias.to(torch.float32).contiguous(); hold.append(bias)
            e.bias = bias.data_ptr()
    if keep is not None:
        keep.extend(hold)
    return e, hold


def conv_igemm(xhat, wt, kind, cout, kh, kw, out_dtype=torch.float32, channels_last=False, epilogue=None, out=None):
    """One vcn_conv_igemm launch.  xhat bf16 [N,H,W,Cin_pad]; wt bf16 [kh*kw, cout, Cin_pad].  Returns y as a
    logical [N,cout,OH,OW] tensor (NCHW memory, or channels_last memory when requested)."""
    lib = _lib.load()
    _lib.require_device(xhat.device.index)
    n, h, w, cin_pad = xhat.shape
    assert xhat.dtype == torch.bfloat16 and wt.dtype == torch.bfloat16 and xhat.is_contiguous() and wt.is_contiguous()
    assert wt.shape == (kh * kw, cout, cin_pad), (wt.shape, (kh * kw, cout, cin_pad))
    oh, ow = (2 * h + 1, 2 * w + 1) if kind == _lib.CONV_TCONV2 else (h, w)
    if out is None:
        fmt = torch.channels_last if channels_last else torch.contiguous_format
        out = torch.empty([n, cout, oh, ow], dtype=out_dtype, device=xhat.device, memory_format=fmt)
    assert out.shape == (n, cout, oh, ow)
    if n == 0:
        return out
    d = _lib.ConvDesc(kind=kind, n=n, h=h, w=w, cin_pad=cin_pad, cout=cout, kh=kh, kw=kw, out_h=oh, out_w=ow,
                      y_dtype=_lib.dtype_code(out.dtype), ys=_strides4(out))
    epi = ctypes.byref(epilogue) if epilogue is not None else None
    with torch.cuda.device(xhat.device):
        st = lib.vcn_conv_igemm(_lib.ptr(xhat), _lib.ptr(wt), _lib.ptr(out), ctypes.byref(d), epi,
                                _lib.current_stream(xhat.device))
    _lib.check(st, "vcn_conv_igemm")
    return out
