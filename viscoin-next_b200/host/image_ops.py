"""image_ops - fused elementwise passes over the [N,3,H,W] fp32 images of the VisCoIN iteration (csrc/image_ops.cu):

  image_prep(x, shift, scale)   bf16 channels_last ((x - shift) / scale): the input conversion of the frozen classifier f and
                                LPIPS's ScalingLayer (lpips v0.1.4; viscoin/training/losses.py:111-123) in one pass each way
  l1_l2(a, b)                   (mean |a-b|, mean (a-b)^2): the first two terms of reconstruction_loss (losses.py:125-157)
"""

import torch

from .. import _lib


def supported(x):
    return x.is_cuda and x.dtype == torch.float32 and x.ndim == 4 and x.shape[1] == 3 and x.is_contiguous()


class _ImagePrep(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, shift, inv_scale):
        n, _, h, w = x.shape
        y = torch.empty([n, 3, h, w], dtype=torch.bfloat16, device=x.device, memory_format=torch.channels_last)
        with torch.cuda.device(x.device):
            st = _lib.load().vcn_image_prep_fwd(_lib.ptr(x), _lib.ptr(y), _lib.ptr(shift), _lib.ptr(inv_scale), n, h, w,
                                                _lib.current_stream(x.device))
        _lib.check(st, "vcn_image_prep_fwd")
        ctx.inv_scale = inv_scale
        return y

    @staticmethod
    def backward(ctx, gy):
        n, _, h, w = gy.shape
        gy = gy.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        gx = torch.empty([n, 3, h, w], dtype=torch.float32, device=gy.device)
        with torch.cuda.device(gy.device):
            st = _lib.load().vcn_image_prep_bwd(_lib.ptr(gy), _lib.ptr(gx), _lib.ptr(ctx.inv_scale), n, h, w, _lib.current_stream(gy.device))
        _lib.check(st, "vcn_image_prep_bwd")
        return gx, None, None


def image_prep(x, shift=None, inv_scale=None):
    """x fp32 NCHW [N,3,H,W] -> bf16 channels_last (x - shift[c]) * inv_scale[c]; shift / inv_scale fp32 [3] or None."""
    if torch.is_grad_enabled() and x.requires_grad:
        return _ImagePrep.apply(x, shift, inv_scale)
    return _ImagePrep.forward(_Ctx(), x, shift, inv_scale)


class _Ctx:
    pass


class _L1L2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        out = torch.zeros(2, dtype=torch.float32, device=a.device)
        with torch.cuda.device(a.device):
            st = _lib.load().vcn_l1l2_fwd(_lib.ptr(a), _lib.ptr(b), _lib.ptr(out), a.numel(), _lib.current_stream(a.device))
        _lib.check(st, "vcn_l1l2_fwd")
        ctx.save_for_backward(a, b)
        return out / a.numel()

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        g2 = (g.to(torch.float32) / a.numel()).contiguous()
        ga = torch.empty_like(a)
        with torch.cuda.device(a.device):
            st = _lib.load().vcn_l1l2_bwd(_lib.ptr(a), _lib.ptr(b), _lib.ptr(g2), _lib.ptr(ga), a.numel(), _lib.current_stream(a.device))
        _lib.check(st, "vcn_l1l2_bwd")
        return ga, None


def l1_l2_supported(a, b):
    return (a.is_cuda and a.dtype == torch.float32 and b.dtype == torch.float32 and a.shape == b.shape and a.is_contiguous()
            and b.is_contiguous() and a.numel() % 4 == 0 and not b.requires_grad)


def l1_l2(a, b):
    """[mean |a-b|, mean (a-b)^2] as one fp32 tensor of two elements; differentiable w.r.t. a."""
    return _L1L2.apply(a, b.detach())
