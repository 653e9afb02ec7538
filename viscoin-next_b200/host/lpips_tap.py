"""lpips_tap - fused LPIPS feature-tap distance over the vcn_lpips_tap_{fwd,bwd} C ABI (csrc/lpips_tap.cu).

    lpips_tap(fa, fb, lin) -> [N]:  mean_p sum_c lin_c (fa/(|fa|+eps) - fb/(|fb|+eps))^2      (lpips v0.1.4, one tap)

fa gets a gradient (the rebuilt-image branch of viscoin/training/losses.py:113-157), fb and lin do not (the original
images' features and the frozen `lin` weights).  ``impl='ref'`` is the plain torch formulation.
"""

import ctypes

import torch

from .. import _lib


def _strides4(t):
    return (ctypes.c_int64 * 4)(*[int(s) for s in t.stride()])


def lpips_tap_ref(fa, fb, lin, eps=1e-10):
    def unit(f):
        f = f.float()
        return f / ((f.square().sum(dim=1, keepdim=True) + 1e-30).sqrt() + eps)
    d = (unit(fa) - unit(fb)).square()
    return (d * lin.reshape(1, -1, 1, 1)).sum(dim=1).mean(dim=[1, 2])


def supported(fa):
    c = fa.shape[1]
    v = c // 8
    return fa.is_cuda and c % 8 == 0 and v <= 64 and (v & (v - 1) == 0 or v % 32 == 0)


class _LpipsTap(torch.autograd.Function):
    @staticmethod
    def forward(ctx, fa, fb, lin):
        lib = _lib.load()
        _lib.require_device(fa.device.index)
        n, c, h, w = fa.shape
        assert fb.shape == fa.shape and fb.dtype == fa.dtype
        fa = fa.contiguous(memory_format=torch.channels_last)
        fb = fb.contiguous(memory_format=torch.channels_last)
        lin = lin.detach().reshape(-1).to(torch.float32).contiguous()
        assert lin.numel() == c
        val = torch.zeros([n], dtype=torch.float32, device=fa.device)
        with torch.cuda.device(fa.device):
            st = lib.vcn_lpips_tap_fwd(_lib.ptr(fa), _lib.ptr(fb), _lib.dtype_code(fa.dtype), ctypes.byref(_strides4(fa)),
                                       ctypes.byref(_strides4(fb)), _lib.ptr(lin), _lib.ptr(val), n, c, h, w,
                                       _lib.current_stream(fa.device))
        _lib.check(st, "vcn_lpips_tap_fwd")
        ctx.save_for_backward(fa, fb, lin)
        return val

    @staticmethod
    def backward(ctx, gval):
        fa, fb, lin = ctx.saved_tensors
        if not ctx.needs_input_grad[0]:
            return None, None, None
        lib = _lib.load()
        n, c, h, w = fa.shape
        gval = gval.to(torch.float32).contiguous()
        da = torch.empty_like(fa, memory_format=torch.channels_last)
        with torch.cuda.device(fa.device):
            st = lib.vcn_lpips_tap_bwd(_lib.ptr(fa), _lib.ptr(fb), _lib.dtype_code(fa.dtype), ctypes.byref(_strides4(fa)),
                                       ctypes.byref(_strides4(fb)), _lib.ptr(lin), _lib.ptr(gval), _lib.ptr(da), n, c, h, w,
                                       _lib.current_stream(fa.device))
        _lib.check(st, "vcn_lpips_tap_bwd")
        return da, None, None


def lpips_tap(fa, fb, lin, impl="cuda"):
    assert impl in ["ref", "cuda"]
    if impl == "ref":
        return lpips_tap_ref(fa, fb, lin)
    if fa.device.type != "cuda":
        raise _lib.VcnError("lpips_tap(impl='cuda') needs a CUDA tensor; this build has no CPU fallback (use impl='ref')")
    return _LpipsTap.apply(fa, fb.detach(), lin)
