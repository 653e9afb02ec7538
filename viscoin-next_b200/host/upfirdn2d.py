"""upfirdn2d - host-side mirror of stylegan2_ada/torch_utils/ops/upfirdn2d.py over the vcn_upfirdn2d C ABI.

Public surface identical to the reference module: ``setup_filter`` (upfirdn2d.py:72-116), ``upfirdn2d``
(:120-164), ``filter2d`` (:272-304), ``upsample2d`` (:308-343), ``downsample2d`` (:347-382) and the private
helpers ``_parse_scaling`` / ``_parse_padding`` / ``_get_filter_size`` that conv2d_resample imports
(conv2d_resample.py:16-17).  ``impl='cuda'`` = sm_100a kernels, no fallback; ``impl='ref'`` = plain torch
formulation of upfirdn2d.py:168-208.  Backward is the same op with up/down swapped and the filter flipped
(upfirdn2d.py:246-261), so gradients of any order come from the same kernels.
"""

import ctypes

import numpy as np
import torch

from .. import _lib
from .misc import profiled_function as _profiled


def _parse_scaling(scaling):
    if isinstance(scaling, int):
        scaling = [scaling, scaling]
    assert isinstance(scaling, (list, tuple))
    assert all(isinstance(v, int) for v in scaling)
    sx, sy = scaling
    assert sx >= 1 and sy >= 1
    return sx, sy


def _parse_padding(padding):
    if isinstance(padding, int):
        padding = [padding, padding]
    assert isinstance(padding, (list, tuple))
    assert all(isinstance(v, int) for v in padding)
    if len(padding) == 2:
        px, py = padding
        padding = [px, px, py, py]
    px0, px1, py0, py1 = padding
    return px0, px1, py0, py1


def _get_filter_size(f):
    if f is None:
        return 1, 1
    assert isinstance(f, torch.Tensor) and f.ndim in [1, 2]
    fw, fh = int(f.shape[-1]), int(f.shape[0])
    assert fw >= 1 and fh >= 1
    return fw, fh


def setup_filter(f, device=torch.device("cpu"), normalize=True, flip_filter=False, gain=1, separable=None):
    """FIR filter preparation, semantics of upfirdn2d.py:72-116 (fp32 result; 1-D kept only for >= 8 taps)."""
    if f is None:
        f = 1
    f = torch.as_tensor(f, dtype=torch.float32)
    assert f.ndim in [0, 1, 2]
    assert f.numel() > 0
    if f.ndim == 0:
        f = f[np.newaxis]
    if separable is None:
        separable = f.ndim == 1 and f.numel() >= 8
    if f.ndim == 1 and not separable:
        f = f.ger(f)
    assert f.ndim == (1 if separable else 2)
    if normalize:
        f = f / f.sum()
    if flip_filter:
        f = f.flip(list(range(f.ndim)))
    f = f * (gain ** (f.ndim / 2))
    return f.to(device=device)


def upfirdn2d(x, f, up=1, down=1, padding=0, flip_filter=False, gain=1, impl="cuda"):
    """Pad, upsample, FIR-filter and downsample a batch of images; signature of upfirdn2d.py:120."""
    assert isinstance(x, torch.Tensor)
    assert impl in ["ref", "cuda"]
    if impl == "ref":
        return _upfirdn2d_ref(x, f, up=up, down=down, padding=padding, flip_filter=flip_filter, gain=gain)
    if x.device.type != "cuda":
        raise _lib.VcnError("upfirdn2d(impl='cuda') needs a CUDA tensor; this build has no CPU fallback (use impl='ref')")
    upx, upy = _parse_scaling(up)
    downx, downy = _parse_scaling(down)
    px0, px1, py0, py1 = _parse_padding(padding)
    return _Upfirdn2d.apply(x, f, upx, upy, downx, downy, px0, px1, py0, py1, bool(flip_filter), float(gain))


@_profiled
def _upfirdn2d_ref(x, f, up=1, down=1, padding=0, flip_filter=False, gain=1):
    """Plain torch ops, any device (upfirdn2d.py:168-208)."""
    assert isinstance(x, torch.Tensor) and x.ndim == 4
    if f is None:
        f = torch.ones([1, 1], dtype=torch.float32, device=x.device)
    assert isinstance(f, torch.Tensor) and f.ndim in [1, 2]
    assert f.dtype == torch.float32 and not f.requires_grad
    n, c, h, w = x.shape
    upx, upy = _parse_scaling(up)
    downx, downy = _parse_scaling(down)
    px0, px1, py0, py1 = _parse_padding(padding)
    x = x.reshape([n, c, h, 1, w, 1])
    x = torch.nn.functional.pad(x, [0, upx - 1, 0, 0, 0, upy - 1])
    x = x.reshape([n, c, h * upy, w * upx])
    x = torch.nn.functional.pad(x, [max(px0, 0), max(px1, 0), max(py0, 0), max(py1, 0)])
    x = x[:, :, max(-py0, 0): x.shape[2] - max(-py1, 0), max(-px0, 0): x.shape[3] - max(-px1, 0)]
    f = f * (gain ** (f.ndim / 2))
    f = f.to(x.dtype)
    if not flip_filter:
        f = f.flip(list(range(f.ndim)))
    f = f[np.newaxis, np.newaxis].repeat([c, 1] + [1] * f.ndim)
    if f.ndim == 4:
        x = torch.nn.functional.conv2d(input=x, weight=f, groups=c)
    else:
        x = torch.nn.functional.conv2d(input=x, weight=f.unsqueeze(2), groups=c)
        x = torch.nn.functional.conv2d(input=x, weight=f.unsqueeze(3), groups=c)
    return x[:, :, ::downy, ::downx]


def _strides4(t):
    return (ctypes.c_int64 * 4)(*[int(s) for s in t.stride()])


def launch(x, f2d, upx, upy, downx, downy, px0, px1, py0, py1, flip, gain, epilogue=None, out=None):
    """One vcn_upfirdn2d call.  f2d: fp32 [fh, fw] on x's device.  Returns y in x's memory format."""
    lib = _lib.load()
    _lib.require_device(x.device.index)
    assert x.ndim == 4 and f2d.ndim == 2 and f2d.dtype == torch.float32
    n, c, ih, iw = x.shape
    fh, fw = f2d.shape
    ow = (iw * upx + px0 + px1 - fw + downx) // downx
    oh = (ih * upy + py0 + py1 - fh + downy) // downy
    assert ow >= 1 and oh >= 1, "upfirdn2d: output would be empty"
    channels_last = x.stride()[1] == 1 and c > 1 and x.is_contiguous(memory_format=torch.channels_last)
    if not channels_last and not x.is_contiguous():
        x = x.contiguous()
    if out is None:
        out = torch.empty([n, c, oh, ow], dtype=x.dtype, device=x.device,
                          memory_format=torch.channels_last if channels_last else torch.contiguous_format)
    if out.numel() == 0:
        return out
    f2d = f2d.contiguous()
    epi = ctypes.byref(epilogue) if epilogue is not None else None
    with torch.cuda.device(x.device):
        st = lib.vcn_upfirdn2d(
            _lib.ptr(x), _lib.ptr(f2d), _lib.ptr(out), _lib.dtype_code(x.dtype), n, c, ih, iw, ctypes.byref(_strides4(x)),
            oh, ow, ctypes.byref(_strides4(out)), fh, fw, upx, upy, downx, downy, px0, px1, py0, py1, int(flip), gain,
            epi, _lib.current_stream(x.device))
    _lib.check(st, "vcn_upfirdn2d")
    return out


class _Upfirdn2d(torch.autograd.Function):
    """reference: Upfirdn2dCuda (upfirdn2d.py:228-264)."""

    @staticmethod
    def forward(ctx, x, f, upx, upy, downx, downy, px0, px1, py0, py1, flip, gain):
        assert isinstance(x, torch.Tensor) and x.ndim == 4
        if f is None:
            f = torch.ones([1, 1], dtype=torch.float32, device=x.device)
        assert isinstance(f, torch.Tensor) and f.ndim in [1, 2]
        assert f.dtype == torch.float32
        if f.device != x.device:
            f = f.to(x.device)
        if f.ndim == 2:
            y = launch(x, f, upx, upy, downx, downy, px0, px1, py0, py1, flip, gain)
        else:  # separable: two 1-D passes with sqrt(gain) each (upfirdn2d.py:239-240)
            g = float(np.sqrt(gain))
            y = launch(x, f.unsqueeze(0), upx, 1, downx, 1, px0, px1, 0, 0, flip, g)
            y = launch(y, f.unsqueeze(1), 1, upy, 1, downy, 0, 0, py0, py1, flip, g)
        ctx.save_for_backward(f)
        ctx.x_shape = x.shape
        ctx.cfg = (upx, upy, downx, downy, px0, px1, py0, py1, flip, gain)
        return y

    @staticmethod
    def backward(ctx, dy):
        (f,) = ctx.saved_tensors
        upx, upy, downx, downy, px0, px1, py0, py1, flip, gain = ctx.cfg
        _, _, ih, iw = ctx.x_shape
        _, _, oh, ow = dy.shape
        fw, fh = _get_filter_size(f)
        dx = None
        if ctx.needs_input_grad[0]:
            p = [fw - px0 - 1, iw * upx - ow * downx + px0 - upx + 1, fh - py0 - 1, ih * upy - oh * downy + py0 - upy + 1]
            dx = _Upfirdn2d.apply(dy, f, downx, downy, upx, upy, p[0], p[1], p[2], p[3], not flip, gain)
        assert not ctx.needs_input_grad[1]
        return (dx,) + (None,) * 11


def filter2d(x, f, padding=0, flip_filter=False, gain=1, impl="cuda"):
    """Same-size FIR filtering (upfirdn2d.py:272-304)."""
    px0, px1, py0, py1 = _parse_padding(padding)
    fw, fh = _get_filter_size(f)
    p = [px0 + fw // 2, px1 + (fw - 1) // 2, py0 + fh // 2, py1 + (fh - 1) // 2]
    return upfirdn2d(x, f, padding=p, flip_filter=flip_filter, gain=gain, impl=impl)


def upsample2d(x, f, up=2, padding=0, flip_filter=False, gain=1, impl="cuda"):
    """FIR upsampling; output is a multiple of the input size (upfirdn2d.py:308-343)."""
    upx, upy = _parse_scaling(up)
    px0, px1, py0, py1 = _parse_padding(padding)
    fw, fh = _get_filter_size(f)
    p = [px0 + (fw + upx - 1) // 2, px1 + (fw - upx) // 2, py0 + (fh + upy - 1) // 2, py1 + (fh - upy) // 2]
    return upfirdn2d(x, f, up=up, padding=p, flip_filter=flip_filter, gain=gain * upx * upy, impl=impl)


def downsample2d(x, f, down=2, padding=0, flip_filter=False, gain=1, impl="cuda"):
    """FIR downsampling; output is a fraction of the input size (upfirdn2d.py:347-382)."""
    downx, downy = _parse_scaling(down)
    px0, px1, py0, py1 = _parse_padding(padding)
    fw, fh = _get_filter_size(f)
    p = [px0 + (fw - downx + 1) // 2, px1 + (fw - downx) // 2, py0 + (fh - downy + 1) // 2, py1 + (fh - downy) // 2]
    return upfirdn2d(x, f, down=down, padding=p, flip_filter=flip_filter, gain=gain, impl=impl)
