"""fma - host-side mirror of stylegan2_ada/torch_utils/ops/fma.py:15-45: ``a * b + c`` with broadcasting.

The reference only uses it inside the non-fused modulated_conv2d (networks.py:69) as
``fma(x[N,C,H,W], dcoefs[N,C,1,1], noise[N,1,H,W])``; that shape class runs as ONE launch of the FIR kernel's fused
epilogue (identity 1x1 filter, ``scale`` = b, ``noise`` = c).  Other broadcast patterns use torch.addcmul (the
reference's own forward, fma.py:22).  Gradients come from autograd of the composing ops.
"""

import torch

from . import upfirdn2d as _upfirdn2d
from .. import _lib


class _FmaModconv(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, c):
        n, ch, h, w = a.shape
        e = _lib.UpfirdnEpilogue()
        b32 = b.reshape(n, ch).to(torch.float32).contiguous()
        c32 = c.reshape(-1, h, w).to(torch.float32).contiguous()
        e.scale, e.noise = b32.data_ptr(), c32.data_ptr()
        e.noise_stride_n = h * w if c32.shape[0] > 1 else 0
        one = torch.ones([1, 1], dtype=torch.float32, device=a.device)
        y = _upfirdn2d.launch(a, one, 1, 1, 1, 1, 0, 0, 0, 0, False, 1.0, epilogue=e)
        ctx.save_for_backward(a, b)
        ctx.c_shape = c.shape
        return y

    @staticmethod
    def backward(ctx, dy):
        a, b = ctx.saved_tensors
        da = dy * b if ctx.needs_input_grad[0] else None
        db = (dy * a).sum(dim=[2, 3], keepdim=True) if ctx.needs_input_grad[1] else None
        dc = None
        if ctx.needs_input_grad[2]:
            dc = dy.sum(dim=1, keepdim=True)
            if ctx.c_shape[0] == 1:
                dc = dc.sum(dim=0, keepdim=True)
        return da, db, dc


def fma(a, b, c, impl="cuda"):
    assert impl in ["ref", "cuda"]
    if impl == "cuda" and a.device.type == "cuda" and a.ndim == 4 and tuple(b.shape) == (a.shape[0], a.shape[1], 1, 1) \
            and c.ndim == 4 and c.shape[1] == 1 and tuple(c.shape[2:]) == tuple(a.shape[2:]) and c.shape[0] in (1, a.shape[0]):
        return _FmaModconv.apply(a, b, c)
    if impl == "cuda" and a.device.type != "cuda":
        raise _lib.VcnError("fma(impl='cuda') needs CUDA tensors; this build has no CPU fallback (use impl='ref')")
    return torch.addcmul(c, a, b)
