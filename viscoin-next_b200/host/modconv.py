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
import os

import torch

from .. import _lib
from .misc import profiled_function as _profiled
from . import bias_act as _bias_act
from . import upfirdn2d as _upfirdn2d


# Operand precision of the implicit GEMM for fp32 activations (the reference's layout / dtype):
#   'tf32' (default)  fp32 operands on tcgen05 kind::tf32 - 10-bit mantissa, what the reference's own GPU path computes
#                     (cuDNN TF32 convolutions, torch default; SURVEY 8(c) gate: fwd <= 3e-3), half the bf16 tensor rate;
#   'bf16'            operands rounded to bf16 (8-bit mantissa; gate 1e-2 / 2e-2) at the full rate.
# 16-bit activations (the bf16-resident pipeline the train step and the inference fast path run) always use bf16 operands.
fp32_operands = os.environ.get("VCN_FP32_OPERANDS", "tf32")


def _operand_dtype(x):
    return torch.float32 if (x.dtype == torch.float32 and fp32_operands == "tf32") else torch.bfloat16


def _strides4(t):
    return (ctypes.c_int64 * 4)(*[int(s) for s in t.stride()])


def _round_up(v, m):
    return (v + m - 1) // m * m


# ---- thin launch helpers (one C-ABI call each) ---------------------------------------------------------------


def modulate_pack(x, styles, out_dtype=torch.bfloat16):
    """x [N,C,H,W] (any dense-ish strides) * styles [N,C] (fp32 or None) -> NHWC [N,H,W,C_pad] GEMM operand (bf16, or fp32
    for the kind::tf32 GEMM)."""
    lib = _lib.load()
    n, c, h, w = x.shape
    c_pad = _round_up(c, 8)
    out = torch.empty([n, h, w, c_pad], dtype=out_dtype, device=x.device)
    if styles is not None:
        styles = styles.to(torch.float32).contiguous()
        assert styles.shape == (n, c)
    if n == 0:
        return out
    with torch.cuda.device(x.device):
        st = lib.vcn_modulate_pack_ex(_lib.ptr(x), _lib.dtype_code(x.dtype), ctypes.byref(_strides4(x)), _lib.ptr(styles),
                                      _lib.ptr(out), _lib.dtype_code(out_dtype), n, c, c_pad, h, w, _lib.current_stream(x.device))
    _lib.check(st, "vcn_modulate_pack")
    return out


def pack_weights(weight, transpose_io=False, want_wsq=False, out_dtype=torch.bfloat16):
    """W [Cout,Cin,kh,kw] -> (wt [kh*kw, rows, cols_pad] in the GEMM operand dtype, wsq fp32 [Cout,Cin] or None)."""
    lib = _lib.load()
    weight = weight.detach().contiguous()
    cout, cin, kh, kw = weight.shape
    rows, cols = (cin, cout) if transpose_io else (cout, cin)
    cols_pad = _round_up(cols, 8)
    wt = torch.empty([kh * kw, rows, cols_pad], dtype=out_dtype, device=weight.device)
    wsq = torch.empty([cout, cin], dtype=torch.float32, device=weight.device) if want_wsq else None
    with torch.cuda.device(weight.device):
        st = lib.vcn_pack_weights_ex(_lib.ptr(weight), _lib.dtype_code(weight.dtype), cout, cin, kh, kw, int(transpose_io),
                                     _lib.ptr(wt), _lib.dtype_code(out_dtype), cols_pad, _lib.ptr(wsq), _lib.current_stream(weight.device))
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


def make_epilogue(scale=None, noise=None, bias=None, act=None, alpha=None, gain=None, clamp=None, keep=None, out_shape=None):
    """vcn_upfirdn2d_epilogue from torch tensors.  noise: anything the reference accepts, i.e. broadcastable to
    [N,1,H,W] (networks.py:71,83 `x.add_(noise)`); the kernels index it as a dense [H,W] plane per sample (or one shared
    plane), so with ``out_shape = (N, H, W)`` it is validated and expanded to one of those two forms first.  `keep`
    collects the tensors that must outlive the launch call."""
    e = _lib.UpfirdnEpilogue()
    hold = []
    if scale is not None:
        scale = scale.to(torch.float32).contiguous(); hold.append(scale)
        e.scale = scale.data_ptr()
    if noise is not None:
        if out_shape is not None:
            n_, oh_, ow_ = out_shape
            if noise.ndim == 4 and noise.shape[1] != 1:
                raise ValueError("noise must broadcast to [N,1,H,W]: per-channel noise %s is not supported" % (tuple(noise.shape),))
            if tuple(noise.shape[-2:]) != (oh_, ow_) or (noise.ndim == 4 and noise.shape[0] not in (1, n_)) or noise.ndim not in (2, 4):
                per_sample = noise.ndim == 4 and noise.shape[0] != 1
                noise = noise.expand([n_ if per_sample else 1, 1, oh_, ow_]) if noise.ndim == 4 else noise.expand([oh_, ow_])
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


def conv_igemm(xhat, wt, kind, cout, kh, kw, out_dtype=torch.float32, channels_last=False, epilogue=None, out=None,
               w_per_sample=False):
    """One vcn_conv_igemm launch.  xhat [N,H,W,Cin_pad] and wt [kh*kw, cout, Cin_pad] both bf16 (kind::f16 GEMM) or both
    fp32 (kind::tf32 GEMM).  Returns y as a logical [N,cout,OH,OW] tensor (NCHW memory, or channels_last when requested)."""
    lib = _lib.load()
    _lib.require_device(xhat.device.index)
    n, h, w, cin_pad = xhat.shape
    assert xhat.dtype in (torch.bfloat16, torch.float32) and wt.dtype == xhat.dtype and xhat.is_contiguous() and wt.is_contiguous()
    want = (n, kh * kw, cout, cin_pad) if w_per_sample else (kh * kw, cout, cin_pad)
    assert tuple(wt.shape) == want, (tuple(wt.shape), want)
    oh, ow = (2 * h + 1, 2 * w + 1) if kind == _lib.CONV_TCONV2 else ((h - 1) // 2, (w - 1) // 2) if kind == _lib.CONV_SCONV2 else (h, w)
    if out is None:
        fmt = torch.channels_last if channels_last else torch.contiguous_format
        out = torch.empty([n, cout, oh, ow], dtype=out_dtype, device=xhat.device, memory_format=fmt)
    assert out.shape == (n, cout, oh, ow)
    if n == 0:
        return out
    d = _lib.ConvDesc(kind=kind, n=n, h=h, w=w, cin_pad=cin_pad, cout=cout, kh=kh, kw=kw, out_h=oh, out_w=ow,
                      y_dtype=_lib.dtype_code(out.dtype), ys=_strides4(out), w_per_sample=int(w_per_sample),
                      tf32=int(xhat.dtype == torch.float32))
    epi = ctypes.byref(epilogue) if epilogue is not None else None
    with torch.cuda.device(xhat.device):
        st = lib.vcn_conv_igemm(_lib.ptr(xhat), _lib.ptr(wt), _lib.ptr(out), ctypes.byref(d), epi,
                                _lib.current_stream(xhat.device))
    _lib.check(st, "vcn_conv_igemm")
    return out


def pack_weights_mod(weight, styles, dcoef=None):
    """Per-sample weights of the fused branch (networks.py:54-60): bf16 [N, kh*kw, Cout, Cin_pad] = W * s[n,ci] * d[n,co]."""
    lib = _lib.load()
    weight = weight.detach().contiguous()
    cout, cin, kh, kw = weight.shape
    styles = styles.to(torch.float32).contiguous()
    n = styles.shape[0]
    cin_pad = _round_up(cin, 8)
    wn = torch.empty([n, kh * kw, cout, cin_pad], dtype=torch.bfloat16, device=weight.device)
    if n == 0:
        return wn
    with torch.cuda.device(weight.device):
        st = lib.vcn_pack_weights_mod(_lib.ptr(weight), _lib.dtype_code(weight.dtype), _lib.ptr(styles), _lib.ptr(dcoef),
                                      _lib.ptr(wn), n, cout, cin, cin_pad, kh, kw, _lib.current_stream(weight.device))
    _lib.check(st, "vcn_pack_weights_mod")
    return wn


def bwd_pack(dy, y, epilogue, want_gsum, want_db, out_dtype=torch.bfloat16):
    """vcn_modconv_bwd_pack: (dy0 NHWC [N,H,W,C_pad] in the GEMM operand dtype, gsum [N,C] or None, db [C] or None)."""
    lib = _lib.load()
    n, c, h, w = dy.shape
    c_pad = _round_up(c, 8)
    dy0 = torch.empty([n, h, w, c_pad], dtype=out_dtype, device=dy.device)
    gsum = torch.zeros([n, c], dtype=torch.float32, device=dy.device) if want_gsum else None
    db = torch.zeros([c], dtype=torch.float32, device=dy.device) if want_db else None
    if n == 0:
        return dy0, gsum, db
    if y is not None:
        assert y.shape == dy.shape and y.dtype == dy.dtype
    with torch.cuda.device(dy.device):
        st = lib.vcn_modconv_bwd_pack_ex(_lib.ptr(dy), _lib.ptr(y), _lib.dtype_code(dy.dtype), ctypes.byref(_strides4(dy)),
                                         ctypes.byref(_strides4(y)) if y is not None else None, ctypes.byref(epilogue),
                                         _lib.ptr(dy0), _lib.dtype_code(out_dtype), _lib.ptr(gsum), _lib.ptr(db), n, c, c_pad, h, w,
                                         _lib.current_stream(dy.device))
    _lib.check(st, "vcn_modconv_bwd_pack")
    return dy0, gsum, db


torgb_fused = os.environ.get("VCN_TORGB_FUSED", "1") == "1"


def torgb_bwd_fused(x, dy0, weight, styles, want_dsum, acc=None):
    """vcn_torgb_bwd_fused: (dx bf16 channels_last like x, dsum [N,C] or None) from the packed ToRGB gradient rows dy0 [N,H,W,8];
    acc (bf16 channels_last like x, optional) is added to dx in the same pass."""
    lib = _lib.load()
    n, c, h, w = x.shape
    cout = weight.shape[0]
    wgt = weight.detach().reshape(cout, c).to(torch.float32).contiguous()
    dx = torch.empty_like(x, memory_format=torch.channels_last)
    dsum = torch.zeros([n, c], dtype=torch.float32, device=x.device) if want_dsum else None
    assert dy0.dtype == torch.bfloat16 and tuple(dy0.shape) == (n, h, w, 8) and dy0.is_contiguous()
    with torch.cuda.device(x.device):
        if acc is not None:
            assert acc.shape == x.shape and acc.dtype == x.dtype and acc.is_contiguous(memory_format=torch.channels_last)
        st = lib.vcn_torgb_bwd_fused(_lib.ptr(x), _lib.ptr(dy0), _lib.ptr(wgt), _lib.ptr(styles), _lib.ptr(acc), _lib.ptr(dx), _lib.ptr(dsum),
                                     _lib.dtype_code(x.dtype), n, c, h, w, cout, _lib.current_stream(x.device))
    _lib.check(st, "vcn_torgb_bwd_fused")
    return dx, dsum


def bwd_finish(dxhat, x, styles, want_dsum):
    """vcn_modconv_bwd_finish: dxhat (fp32) *= s in place; returns dsum [N,C] = sum_p dxhat*x (or None)."""
    lib = _lib.load()
    n, c, h, w = dxhat.shape
    assert dxhat.dtype == torch.float32
    dsum = torch.zeros([n, c], dtype=torch.float32, device=dxhat.device) if want_dsum else None
    if n == 0:
        return dsum
    with torch.cuda.device(dxhat.device):
        st = lib.vcn_modconv_bwd_finish(_lib.ptr(dxhat), ctypes.byref(_strides4(dxhat)), _lib.ptr(x) if want_dsum else None,
                                        _lib.dtype_code(x.dtype), ctypes.byref(_strides4(x)), _lib.ptr(styles), _lib.ptr(dsum),
                                        n, c, h, w, _lib.current_stream(dxhat.device))
    _lib.check(st, "vcn_modconv_bwd_finish")
    return dsum


def bwd_finish_cl(dxhat, x, styles, want_dsum):
    """vcn_modconv_bwd_finish_cl on the resident pipeline: dxhat bf16 logical [N,C,H,W] in channels_last memory with C a
    multiple of 8 (the GEMM output), scaled by s in place; returns dsum [N,C] = sum_p dxhat*x (or None)."""
    lib = _lib.load()
    n, c, h, w = dxhat.shape
    assert dxhat.dtype == torch.bfloat16 and dxhat.is_contiguous(memory_format=torch.channels_last) and c % 8 == 0
    dsum = torch.zeros([n, c], dtype=torch.float32, device=dxhat.device) if want_dsum else None
    if n == 0:
        return dsum
    if want_dsum:
        assert x.dtype == torch.bfloat16 and x.stride(1) == 1 and x.shape == dxhat.shape
    with torch.cuda.device(dxhat.device):
        st = lib.vcn_modconv_bwd_finish_cl(_lib.ptr(dxhat), _lib.ptr(x) if want_dsum else None, _lib.dtype_code(dxhat.dtype),
                                           ctypes.byref(_strides4(x)) if want_dsum else None, _lib.ptr(styles), _lib.ptr(dsum),
                                           n, c, c, h, w, _lib.current_stream(dxhat.device))
    _lib.check(st, "vcn_modconv_bwd_finish_cl")
    return dsum


def conv_wgrad(dy0, xhat, cout, cin, kh, kw, transposed=False):
    """vcn_conv_wgrad: dy0 bf16 NHWC [N,OH,OW,Cout_pad], xhat bf16 NHWC [N,H,W,Cin_pad] -> dW fp32 [cout,cin,kh,kw]."""
    lib = _lib.load()
    n, h, w, cin_pad = xhat.shape
    _, oh, ow, cout_pad = dy0.shape
    assert dy0.dtype == torch.bfloat16 and xhat.dtype == torch.bfloat16 and dy0.is_contiguous() and xhat.is_contiguous()
    dw = torch.zeros([cout, cin, kh, kw], dtype=torch.float32, device=xhat.device)
    if n == 0:
        return dw
    with torch.cuda.device(xhat.device):
        st = lib.vcn_conv_wgrad(_lib.ptr(dy0), _lib.ptr(xhat), _lib.ptr(dw), 1 if transposed else 0, n, h, w, cin_pad, cin,
                                cout_pad, cout, kh, kw, oh, ow, _lib.current_stream(xhat.device))
    _lib.check(st, "vcn_conv_wgrad")
    return dw


def demod_grad(ds, styles, gsum, dcoef, wsq):
    lib = _lib.load()
    n, cin = ds.shape
    cout = dcoef.shape[1]
    if n == 0:
        return ds
    with torch.cuda.device(ds.device):
        st = lib.vcn_demod_grad(_lib.ptr(ds), _lib.ptr(styles), _lib.ptr(gsum), _lib.ptr(dcoef), _lib.ptr(wsq), n, cin, cout,
                                _lib.current_stream(ds.device))
    _lib.check(st, "vcn_demod_grad")
    return ds


# ---- packed-weight cache (synthesis weights are frozen in VisCoIN: pack once per weight version) -------------------

_wcache = {}


def _packed(weight, transpose_io, flip, want_wsq, dtype=torch.bfloat16):
    """Packed taps (+ wsq) of `weight` in the GEMM operand dtype, cached per tensor object and version.  The entry holds a
    weak reference to the tensor so that a recycled address or id can never alias a dead tensor's entry."""
    import weakref
    key = (id(weight), transpose_io, flip, dtype)
    hit = _wcache.get(key)
    if hit is not None:
        ref, version, ptr, wt, wsq = hit
        if ref() is weight and version == weight._version and ptr == weight.data_ptr():
            return wt, wsq
    src = weight.detach()
    if flip:
        src = src.flip([2, 3])
    wt, wsq = pack_weights(src, transpose_io=transpose_io, want_wsq=True, out_dtype=dtype)
    if len(_wcache) > 512:
        for k in [k for k, v in _wcache.items() if v[0]() is None]:
            del _wcache[k]
    _wcache[key] = (weakref.ref(weight), weight._version, weight.data_ptr(), wt, wsq)
    return wt, wsq


_FUSABLE_ACTS = (None, "linear", "lrelu")


class _ModConvLayer(torch.autograd.Function):
    """y = bias_act(modulated_conv2d(x, W, s, noise), b)  (act None = plain modulated_conv2d)."""

    @staticmethod
    def forward(ctx, x, weight, styles, noise, bias, resample_filter, cfg):
        up, demodulate, flip_weight, act, alpha, gain, clamp = cfg[:7]
        tap = len(cfg) > 7 and bool(cfg[7])  # also return x (pass-through handle): its gradient is folded into this backward
        _lib.require_device(x.device.index)
        n, cin, h, w = x.shape
        cout, cin_w, kh, kw = weight.shape
        assert cin_w == cin and kh == kw and kh in (1, 3), "modulated_conv2d(impl='cuda'): 1x1 and 3x3 kernels only"
        assert styles is None or tuple(styles.shape) == (n, cin)
        styles32 = styles.to(torch.float32).contiguous() if styles is not None else None
        hold = []
        op = _operand_dtype(x)  # fp32 activations: tf32 operands unless modconv.fp32_operands == 'bf16'
        if up == 1:
            wt, wsq = _packed(weight, False, False, demodulate, op)
        else:
            assert up == 2 and kh == 3, "modulated_conv2d(impl='cuda'): up=2 needs a 3x3 kernel"
            assert resample_filter is not None and tuple(resample_filter.shape) == (4, 4), "up=2 needs a 4x4 resample filter"
            # reference: flip_weight=False -> conv_transpose2d with the weight as is (conv2d_resample.py:127-138)
            wt, wsq = _packed(weight, False, bool(flip_weight), demodulate, op)
        d = demod_coefs(styles32, wsq) if demodulate else None
        xhat = modulate_pack(x, styles32, op)
        epi, hold = make_epilogue(scale=d, noise=noise, bias=bias, act=act, alpha=alpha, gain=gain, clamp=clamp,
                                  out_shape=(n, h * up, w * up))
        # resident pipeline: 16-bit channels-last activations stay 16-bit channels-last (forward and backward); a ToRGB
        # layer (cout < 8) leaves it and writes fp32 NCHW for the image accumulation
        resident = x.dtype == torch.bfloat16 and x.stride(1) == 1 and cin % 8 == 0
        cl_out = resident and cout % 8 == 0
        out_dtype = x.dtype if (cl_out or not resident) else torch.float32
        if up == 1:
            kind = _lib.CONV_FPROP if flip_weight else _lib.CONV_DGRAD
            y = conv_igemm(xhat, wt, kind, cout, kh, kw, out_dtype=out_dtype, channels_last=cl_out, epilogue=epi)
        else:
            z = conv_igemm(xhat, wt, _lib.CONV_TCONV2, cout, 3, 3, out_dtype=torch.bfloat16 if cl_out else torch.float32,
                           channels_last=cl_out)
            f = resample_filter.to(device=x.device, dtype=torch.float32)
            fmt = torch.channels_last if cl_out else torch.contiguous_format
            y = _upfirdn2d.launch(z, f, 1, 1, 1, 1, 1, 1, 1, 1, False, 4.0, epilogue=epi,
                                  out=torch.empty([n, cout, 2 * h, 2 * w], dtype=out_dtype, device=x.device, memory_format=fmt))
        ctx.resident = resident
        ctx.op = op
        ctx.cfg = cfg
        ctx.has = (styles is not None, noise is not None, bias is not None)
        ctx.save_for_backward(x, weight, styles32, noise, bias, resample_filter, d, wsq, y)
        del hold
        if tap:
            return y, x.view_as(x)
        return y

    @staticmethod
    def backward(ctx, dy, dx_pass=None):
        x, weight, styles32, noise, bias, resample_filter, d, wsq, y = ctx.saved_tensors
        up, demodulate, flip_weight, act, alpha, gain, clamp = ctx.cfg[:7]
        if dy is None:  # only the pass-through handle was used downstream
            return dx_pass, None, None, None, None, None, None
        need_x, need_w, need_s, need_noise, need_b = ctx.needs_input_grad[:5]
        n, cin, h, w = x.shape
        cout, _, kh, kw = weight.shape
        resident = ctx.resident
        if dy.dtype != y.dtype:
            dy = dy.to(y.dtype)
        if y.stride(1) == 1 and y.shape[1] > 1:
            dy = dy.contiguous(memory_format=torch.channels_last)
        elif not dy.is_contiguous():
            dy = dy.contiguous()
        epi, hold = make_epilogue(scale=d, noise=noise, bias=bias, act=act, alpha=alpha, gain=gain, clamp=clamp,
                                  out_shape=(n, h * up, w * up))
        want_g = demodulate and (need_s or need_w)
        op = ctx.op
        dy0, gsum, db = bwd_pack(dy, y if (act is not None or want_g) else None, epi, want_g, need_b and bias is not None, op)
        dx = ds = dnoise = dw = None
        dz = None
        if up == 2 and (need_x or need_s or need_w):
            f = resample_filter.to(device=x.device, dtype=torch.float32)
            # FIR^T: same op, flipped filter, padding fw-1-p = 2 (upfirdn2d.py:246-261); channels_last view of dy0
            dz = _upfirdn2d.launch(dy0.permute(0, 3, 1, 2), f, 1, 1, 1, 1, 2, 2, 2, 2, True, 4.0)
            dz = dz.permute(0, 2, 3, 1)
            assert dz.is_contiguous()
        fused_rgb = (torgb_fused and resident and up == 1 and kh == 1 and kw == 1 and cout <= 4 and op == torch.bfloat16 and not demodulate
                     and x.dtype == torch.bfloat16 and cin >= 64 and cin & (cin - 1) == 0 and x.is_contiguous(memory_format=torch.channels_last)
                     and n * h * w * cin >= (1 << 21))
        if fused_rgb and (need_x or need_s):
            # ToRGB: the input-gradient GEMM is a rank-<=4 expansion - one streaming pass (vcn_torgb_bwd_fused)
            acc = None
            if dx_pass is not None and need_x:
                acc = dx_pass.to(x.dtype).contiguous(memory_format=torch.channels_last)
                dx_pass = None
            dx, ds = torgb_bwd_fused(x, dy0, weight, styles32, need_s and styles32 is not None, acc=acc)
            if not need_x:
                dx = None
        elif need_x or need_s:
            if up == 1:
                wt_t, _ = _packed(weight, True, False, False, op)
                kind = _lib.CONV_DGRAD if flip_weight else _lib.CONV_FPROP
                dxhat = conv_igemm(dy0, wt_t, kind, cin, kh, kw, out_dtype=torch.bfloat16 if resident else torch.float32,
                                   channels_last=resident)
            else:
                wt_t, _ = _packed(weight, True, bool(flip_weight), False, op)
                dxhat = conv_igemm(dz, wt_t, _lib.CONV_SCONV2, cin, 3, 3, out_dtype=torch.bfloat16 if resident else torch.float32,
                                   channels_last=resident)
            if resident:
                ds1 = bwd_finish_cl(dxhat, x, styles32, need_s and styles32 is not None)
            else:
                ds1 = bwd_finish(dxhat, x, styles32, need_s and styles32 is not None)
            if need_x:
                dx = dxhat if x.dtype == dxhat.dtype else dxhat.to(x.dtype)
            if need_s and styles32 is not None:
                ds = ds1
                if demodulate:
                    ds = demod_grad(ds, styles32, gsum, d, wsq)
        if need_w:
            # dW = sum_n,p dy0 (x) xhat on the tensor cores (both operands as the other GEMMs use them), plus the path
            # through the demodulation coefficients: d = rsqrt(sum (s W)^2) gives  -W * sum_n gsum d^2 s^2
            # the weight-gradient GEMM (MN-major operands) runs bf16 operands in either mode
            xhat = modulate_pack(x, styles32)
            if op != torch.bfloat16:
                dy0 = dy0.to(torch.bfloat16)
                dz = dz.to(torch.bfloat16) if dz is not None else None
            if up == 1:
                dw = conv_wgrad(dy0, xhat, cout, cin, kh, kw)
                if not flip_weight:
                    dw = dw.flip([2, 3])
            else:
                dw = conv_wgrad(dz, xhat, cout, cin, 3, 3, transposed=True)
                if flip_weight:
                    dw = dw.flip([2, 3])
            if demodulate:
                e = (gsum * d * d).t() @ styles32.square()
                dw = dw - weight.detach().float() * e.reshape(cout, cin, 1, 1)
            dw = dw.to(weight.dtype)
        if need_noise and noise is not None:
            # rare path (noise buffers are constants / fresh randn in the reference): dz summed over channels
            if act is not None:
                dzf = _bias_act._call(dy, None, None, y, None, 1, 1, _bias_act.activation_funcs[act].cuda_idx,
                                      epi.alpha, epi.act_gain, epi.clamp, False)[0]
            else:
                dzf = dy
            dnoise = dzf.sum(dim=1, keepdim=True)  # [N,1,H,W] -> un-broadcast to the shape the caller passed
            dnoise = dnoise.sum_to_size(noise.shape) if noise.ndim == 4 else dnoise.sum(dim=[0, 1]).sum_to_size(noise.shape)
            dnoise = dnoise.to(noise.dtype)
        del hold
        if dx_pass is not None and need_x:  # pass-through gradient on a path without the fused kernel
            dx = dx_pass.to(dx.dtype) + dx if dx is not None else dx_pass
        return dx, dw, ds, dnoise, (db.to(bias.dtype) if db is not None else None), None, None


def _check_cuda(x, what):
    if x.device.type != "cuda":
        raise _lib.VcnError("%s(impl='cuda') needs a CUDA tensor; this build has no CPU fallback (use impl='ref')" % what)


@_profiled
def modconv_layer(x, weight, styles, noise=None, bias=None, up=1, padding=None, resample_filter=None, demodulate=True,
                  flip_weight=True, act="lrelu", alpha=None, gain=None, clamp=None, tap=False):
    """Fused SynthesisLayer / ToRGBLayer body: bias_act(modulated_conv2d(...), bias, act, gain, clamp)
    (networks.py:299-304, 322-323) in one GEMM epilogue.  Activations other than linear / lrelu run bias_act
    as a second kernel."""
    _check_cuda(x, "modconv_layer")
    k = int(weight.shape[-1])
    assert padding is None or padding == k // 2, "modulated_conv2d(impl='cuda'): padding must be k//2"
    assert up in (1, 2)
    fuse = act in _FUSABLE_ACTS and not (act == "lrelu" and alpha == 0)
    cfg = (up, bool(demodulate), bool(flip_weight), act if fuse else None, alpha, gain, clamp)
    if styles is None:
        demodulate = False
        cfg = (up, False, bool(flip_weight), cfg[3], alpha, gain, clamp)
    if tap:
        # `tap`: returns (y, x2) where x2 is a second handle on x - the caller hands x2 to whatever else consumes x, and the sum
        # of the two gradients of x is formed inside this layer's backward (ToRGB: inside vcn_torgb_bwd_fused) instead of by a
        # separate add pass over the full activation
        assert fuse, "tap needs a fusable activation"
        return _ModConvLayer.apply(x, weight, styles, noise, bias, resample_filter, cfg + (True,))
    y = _ModConvLayer.apply(x, weight, styles, noise, bias if fuse else None, resample_filter, cfg)
    if not fuse:
        y = _bias_act.bias_act(y, bias.to(y.dtype) if bias is not None else None, act=act, alpha=alpha, gain=gain, clamp=clamp)
    return y


@_profiled
def modulated_conv2d(x, weight, styles, noise=None, up=1, down=1, padding=0, resample_filter=None, demodulate=True,
                     flip_weight=True, fused_modconv=True, impl="cuda"):
    """Signature of training/networks.py:27-40 (+ impl).  x [N,Cin,H,W], weight [Cout,Cin,k,k], styles [N,Cin],
    noise broadcastable to [N,1,Hout,Wout] or None."""
    assert impl in ["ref", "cuda"]
    if impl == "ref":
        return _modulated_conv2d_ref(x, weight, styles, noise, up, down, padding, resample_filter, demodulate, flip_weight,
                                     fused_modconv)
    _check_cuda(x, "modulated_conv2d")
    if down != 1:
        # not on the synthesis path: the reference's non-fused branch (networks.py:63-72) composed from this package's
        # operators - x*s -> conv2d_resample(down) [FIR + stride-2 tcgen05 GEMM] -> fma(y, dcoefs, noise)
        from . import conv2d_resample as _c2r
        from . import fma as _fma
        n, cin = int(x.shape[0]), int(x.shape[1])
        dcoefs = None
        if demodulate:
            wmod = weight.unsqueeze(0) * styles.reshape(n, 1, cin, 1, 1)
            dcoefs = (wmod.square().sum(dim=[2, 3, 4]) + 1e-8).rsqrt()
        y = _c2r.conv2d_resample(x * styles.to(x.dtype).reshape(n, cin, 1, 1), weight.to(x.dtype), f=resample_filter, up=up,
                                 down=down, padding=padding, flip_weight=flip_weight)
        nz = None
        if noise is not None:
            nz = noise.to(y.dtype)
            nz = nz.reshape(1, 1, *nz.shape) if nz.ndim == 2 else nz
        if demodulate and nz is not None:
            return _fma.fma(y, dcoefs.to(y.dtype).reshape(n, -1, 1, 1), nz)
        if demodulate:
            return y * dcoefs.to(y.dtype).reshape(n, -1, 1, 1)
        return y + nz if nz is not None else y
    return modconv_layer(x, weight, styles, noise=noise, bias=None, up=up, padding=padding, resample_filter=resample_filter,
                         demodulate=demodulate, flip_weight=flip_weight, act=None)


def _modulated_conv2d_ref(x, weight, styles, noise, up, down, padding, resample_filter, demodulate, flip_weight,
                          fused_modconv):
    """Plain torch formulation (non-fused branch of networks.py:63-72; the grouped branch is the same function)."""
    n = x.shape[0]
    cout, cin, kh, kw = weight.shape
    dcoefs = None
    if demodulate:
        w = weight.unsqueeze(0) * styles.reshape(n, 1, -1, 1, 1)
        dcoefs = (w.square().sum(dim=[2, 3, 4]) + 1e-8).rsqrt()
    x = x * styles.to(x.dtype).reshape(n, -1, 1, 1)
    px0, px1, py0, py1 = _upfirdn2d._parse_padding(padding)
    wgt = weight.to(x.dtype)
    if up == 1 and down == 1:
        if not flip_weight:
            wgt = wgt.flip([2, 3])
        x = torch.nn.functional.conv2d(x, wgt, padding=[py0, px0])
    elif up == 2 and down == 1:
        wt = wgt.transpose(0, 1)
        if flip_weight:
            wt = wt.flip([2, 3])
        x = torch.nn.functional.conv_transpose2d(x, wt, stride=2)
        fw, fh = _upfirdn2d._get_filter_size(resample_filter)
        p = [px0 + (fw + 1) // 2 - (kw - 1), px1 + (fw - 2) // 2 - (kw - 2), py0 + (fh + 1) // 2 - (kh - 1), py1 + (fh - 2) // 2 - (kh - 2)]
        x = _upfirdn2d._upfirdn2d_ref(x, resample_filter, padding=p, gain=4)
    else:
        raise NotImplementedError("ref path covers the synthesis configurations (up in {1,2}, down=1)")
    if dcoefs is not None:
        x = x * dcoefs.to(x.dtype).reshape(n, -1, 1, 1)
    if noise is not None:
        x = x + noise.to(x.dtype)
    return x


# ---- resident-activation pipeline (inference): activations stay bf16 NHWC between layers -------------------------------


def layer_nhwc(act, weight, styles, noise=None, bias=None, up=1, resample_filter=None, demodulate=True, act_name="lrelu",
               alpha=None, gain=None, clamp=None, per_sample=None, out_nchw_f32=False):
    """One synthesis layer on a resident activation: act bf16 [N,H,W,Cin_pad] (NHWC, unmodulated) -> bf16 [N,H',W',Cout]
    (or fp32 NCHW [N,Cout,H',W'] with out_nchw_f32, used by ToRGB).  No autograd (inference path).

    per_sample=True  : the reference's fused branch - weights W*s[n]*d[n] per sample (vcn_pack_weights_mod), the GEMM
                       reads the activation as is: no pass over the activation at all (large resolutions).
    per_sample=False : the non-fused branch - xhat = act*s (one channels-last pass), shared weights, d in the epilogue
                       (small resolutions, where N copies of the weights would outweigh the activation)."""
    n, h, w, cin_pad = act.shape
    cout, cin, kh, kw = weight.shape
    assert act.dtype == torch.bfloat16 and act.is_contiguous() and cin_pad == _round_up(cin, 8)
    if per_sample is None:
        per_sample = (kh == 1) or h * w >= 128 * 128
    styles32 = styles.to(torch.float32).contiguous()
    d = None
    if demodulate:
        _, wsq = _packed(weight, False, False, True)
        d = demod_coefs(styles32, wsq)
    if per_sample:
        wt = pack_weights_mod(weight, styles32, d)
        xin, scale = act, None
    else:
        wt, _ = _packed(weight, False, False, demodulate)
        xin, scale = modulate_pack(act.permute(0, 3, 1, 2)[:, :cin], styles32), d
    if up == 1:
        epi, hold = make_epilogue(scale=scale, noise=noise, bias=bias, act=act_name, alpha=alpha, gain=gain, clamp=clamp,
                                  out_shape=(n, h, w))
        y = conv_igemm(xin, wt, _lib.CONV_FPROP, cout, kh, kw, out_dtype=torch.float32 if out_nchw_f32 else torch.bfloat16,
                       channels_last=not out_nchw_f32, epilogue=epi, w_per_sample=per_sample)
        del hold
        return y if out_nchw_f32 else y.permute(0, 2, 3, 1)
    assert up == 2 and kh == 3 and not out_nchw_f32
    z = conv_igemm(xin, wt, _lib.CONV_TCONV2, cout, 3, 3, out_dtype=torch.bfloat16, channels_last=True, w_per_sample=per_sample)
    epi, hold = make_epilogue(scale=scale, noise=noise, bias=bias, act=act_name, alpha=alpha, gain=gain, clamp=clamp,
                              out_shape=(n, 2 * h, 2 * w))
    f = resample_filter.to(device=act.device, dtype=torch.float32)
    y = _upfirdn2d.launch(z, f, 1, 1, 1, 1, 1, 1, 1, 1, False, 4.0, epilogue=epi)
    del hold
    return y.permute(0, 2, 3, 1)
