"""bias_act - host-side mirror of stylegan2_ada/torch_utils/ops/bias_act.py over the vcn_bias_act C ABI.

Same public surface as the reference module: ``activation_funcs`` (bias_act.py:23-33) and
``bias_act(x, b, dim, act, alpha, gain, clamp, impl)`` (bias_act.py:55-89).  ``impl='cuda'`` runs the
sm_100a kernels (no fallback: CPU tensors or a missing library raise); ``impl='ref'`` is the
plain-torch formulation of bias_act.py:93-123, kept because the switch is part of the operator API.
First and second order gradients are supported like the reference (bias_act.py:144-205); the bias
gradient is produced by the kernel call itself instead of a separate ``dx.sum``.
"""

import math

import torch

from .. import _lib
from .misc import profiled_function as _profiled


class _Spec(dict):
    """Attribute-style dict (stands in for dnnlib.EasyDict used by the reference table)."""

    __getattr__ = dict.__getitem__


_SQRT2 = math.sqrt(2.0)

activation_funcs = {
    "linear": _Spec(func=lambda x, **_: x, def_alpha=0, def_gain=1, cuda_idx=1, ref="", has_2nd_grad=False),
    "relu": _Spec(func=lambda x, **_: torch.nn.functional.relu(x), def_alpha=0, def_gain=_SQRT2, cuda_idx=2, ref="y", has_2nd_grad=False),
    "lrelu": _Spec(func=lambda x, alpha, **_: torch.nn.functional.leaky_relu(x, alpha), def_alpha=0.2, def_gain=_SQRT2, cuda_idx=3, ref="y", has_2nd_grad=False),
    "tanh": _Spec(func=lambda x, **_: torch.tanh(x), def_alpha=0, def_gain=1, cuda_idx=4, ref="y", has_2nd_grad=True),
    "sigmoid": _Spec(func=lambda x, **_: torch.sigmoid(x), def_alpha=0, def_gain=1, cuda_idx=5, ref="y", has_2nd_grad=True),
    "elu": _Spec(func=lambda x, **_: torch.nn.functional.elu(x), def_alpha=0, def_gain=1, cuda_idx=6, ref="y", has_2nd_grad=True),
    "selu": _Spec(func=lambda x, **_: torch.nn.functional.selu(x), def_alpha=0, def_gain=1, cuda_idx=7, ref="y", has_2nd_grad=True),
    "softplus": _Spec(func=lambda x, **_: torch.nn.functional.softplus(x), def_alpha=0, def_gain=1, cuda_idx=8, ref="y", has_2nd_grad=True),
    "swish": _Spec(func=lambda x, **_: torch.sigmoid(x) * x, def_alpha=0, def_gain=_SQRT2, cuda_idx=9, ref="x", has_2nd_grad=True),
}


def _resolve(act, alpha, gain, clamp):
    assert clamp is None or clamp >= 0
    spec = activation_funcs[act]
    alpha = float(spec.def_alpha if alpha is None else alpha)
    gain = float(spec.def_gain if gain is None else gain)
    clamp = float(-1 if clamp is None else clamp)
    return spec, alpha, gain, clamp


def bias_act(x, b=None, dim=1, act="linear", alpha=None, gain=None, clamp=None, impl="cuda"):
    """Fused ``clamp(act(x + b) * gain)``; signature and defaults of bias_act.py:55."""
    assert isinstance(x, torch.Tensor)
    assert impl in ["ref", "cuda"]
    if impl == "ref":
        return _bias_act_ref(x=x, b=b, dim=dim, act=act, alpha=alpha, gain=gain, clamp=clamp)
    if x.device.type != "cuda":
        raise _lib.VcnError("bias_act(impl='cuda') needs a CUDA tensor; this build has no CPU fallback (use impl='ref')")
    spec, alpha, gain, clamp = _resolve(act, alpha, gain, clamp)
    if b is not None:
        assert isinstance(b, torch.Tensor) and b.ndim == 1
        assert 0 <= dim < x.ndim
        assert b.shape[0] == x.shape[dim]
        assert b.dtype == x.dtype, "x and b must have the same dtype"
    return _BiasAct.apply(x, b, dim, spec.cuda_idx, alpha, gain, clamp, spec.ref, spec.has_2nd_grad)


@_profiled
def _bias_act_ref(x, b=None, dim=1, act="linear", alpha=None, gain=None, clamp=None):
    """Plain torch ops, any device (bias_act.py:93-123)."""
    spec, alpha, gain, clamp = _resolve(act, alpha, gain, clamp)
    if b is not None:
        assert isinstance(b, torch.Tensor) and b.ndim == 1
        assert 0 <= dim < x.ndim
        assert b.shape[0] == x.shape[dim]
        x = x + b.reshape([-1 if i == dim else 1 for i in range(x.ndim)])
    x = spec.func(x, alpha=alpha)
    if gain != 1:
        x = x * gain
    if clamp >= 0:
        x = x.clamp(-clamp, clamp)
    return x


def _memory_format(t):
    return torch.channels_last if t.ndim == 4 and t.stride()[1] == 1 and t.shape[1] > 1 else torch.contiguous_format


def _dense(t, fmt):
    if t.ndim == 4:
        return t.contiguous(memory_format=fmt)
    return t.contiguous()


def _call(x, b, xref, yref, dy, grad, dim, act_idx, alpha, gain, clamp, want_db):
    """One vcn_bias_act launch.  x decides layout; returns (y, db or None)."""
    lib = _lib.load()
    _lib.require_device(x.device.index)
    y = torch.empty_like(x)  # preserve_format keeps the dense layout of x
    assert y.stride() == x.stride()
    size_b = x.shape[dim] if x.ndim else 1
    step_b = x.stride(dim) if x.ndim else 1
    if x.numel() and step_b == 0:  # size-1 dims may report odd strides
        step_b = 1
    db = ws = None
    ws_bytes = 0
    if want_db:
        db = torch.empty([size_b], dtype=x.dtype, device=x.device)
        ws_bytes = lib.vcn_bias_act_workspace(_lib.dtype_code(x.dtype), x.numel(), size_b, max(step_b, 1))
        if ws_bytes:
            ws = torch.empty([ws_bytes], dtype=torch.uint8, device=x.device)
        if x.numel() == 0:
            db.zero_()
            return y, db
    with torch.cuda.device(x.device):
        st = lib.vcn_bias_act(
            _lib.ptr(x), _lib.ptr(b), _lib.ptr(xref), _lib.ptr(yref), _lib.ptr(dy), _lib.ptr(y), _lib.ptr(db),
            _lib.ptr(ws), ws_bytes, _lib.dtype_code(x.dtype), grad, act_idx, alpha, gain, clamp,
            x.numel(), size_b, max(step_b, 1), _lib.current_stream(x.device))
    _lib.check(st, "vcn_bias_act")
    return y, db


def _same_layout(t, like):
    if t is None:
        return None
    if t.stride() != like.stride():
        t = t.contiguous(memory_format=_memory_format(like)) if like.ndim == 4 else t.contiguous()
    return t


class _BiasAct(torch.autograd.Function):
    """Forward op (reference: BiasActCuda, bias_act.py:144-175)."""

    @staticmethod
    def forward(ctx, x, b, dim, act_idx, alpha, gain, clamp, ref, has_2nd):
        fmt = _memory_format(x)
        x = _dense(x, fmt)
        b = b.contiguous() if b is not None else None
        trivial = act_idx == 1 and gain == 1 and clamp < 0 and b is None
        y = x if trivial else _call(x, b, None, None, None, 0, dim, act_idx, alpha, gain, clamp, False)[0]
        keep_x = "x" in ref or has_2nd
        # The reference's CUDA wiring saves y only for ref='y' activations (bias_act.py:154-157), which leaves the
        # clamp mask of `linear` without a yref; the impl='ref' path (the parity target) masks, so y is kept
        # whenever a clamp is active.
        keep_y = "y" in ref or (clamp >= 0 and "x" not in ref)
        ctx.save_for_backward(x if keep_x else None, b if keep_x else None, y if keep_y else None)
        ctx.cfg = (dim, act_idx, alpha, gain, clamp, ref, has_2nd, trivial or (act_idx == 1 and gain == 1 and clamp < 0))
        ctx.has_b = b is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, b, y = ctx.saved_tensors
        dim, act_idx, alpha, gain, clamp, ref, has_2nd, passthrough = ctx.cfg
        need_dx, need_db = ctx.needs_input_grad[0], ctx.has_b and ctx.needs_input_grad[1]
        dx = db = None
        if need_dx or need_db:
            if passthrough:
                dx = dy
                if need_db:
                    other = [i for i in range(dy.ndim) if i != dim]
                    db = dy.sum(other) if other else dy
            else:
                dx, db = _BiasActGrad.apply(dy, x, b, y, dim, act_idx, alpha, gain, clamp, has_2nd, need_db)
                if not need_db:
                    db = None
        return dx, db, None, None, None, None, None, None, None


class _BiasActGrad(torch.autograd.Function):
    """First-order backward op, itself differentiable (reference: BiasActCudaGrad, bias_act.py:178-205)."""

    @staticmethod
    def forward(ctx, dy, x, b, y, dim, act_idx, alpha, gain, clamp, has_2nd, want_db):
        like = y if y is not None else x
        fmt = _memory_format(like) if like is not None else _memory_format(dy)
        dy = _dense(dy, fmt)
        if like is not None and dy.stride() != like.stride():
            dy = _same_layout(dy, like)
        dx, db = _call(dy, b, x, y, None, 1, dim, act_idx, alpha, gain, clamp, want_db)
        ctx.save_for_backward(dy if has_2nd else None, x, b, y)
        ctx.cfg = (dim, act_idx, alpha, gain, clamp, has_2nd)
        if db is None:
            db = dy.new_empty([0])
            ctx.mark_non_differentiable(db)
        return dx, db

    @staticmethod
    def backward(ctx, d_dx, d_db):
        dy, x, b, y = ctx.saved_tensors
        dim, act_idx, alpha, gain, clamp, has_2nd = ctx.cfg
        d_dy = d_x = d_b = None
        # d_db (gradient flowing into the bias-gradient output) is a broadcast add onto d_dx.
        if d_db is not None and d_db.numel() and d_dx is not None:
            shape = [-1 if i == dim else 1 for i in range(d_dx.ndim)]
            d_dx = d_dx + d_db.reshape(shape)
        elif d_db is not None and d_db.numel() and d_dx is None:
            like = y if y is not None else x
            shape = [-1 if i == dim else 1 for i in range(like.ndim)]
            d_dx = d_db.reshape(shape).expand_as(like)
        if d_dx is None:
            return (None,) * 11
        like = y if y is not None else x
        d_dx = _same_layout(d_dx, like) if like is not None else _dense(d_dx, _memory_format(d_dx))
        if ctx.needs_input_grad[0]:
            d_dy, _ = _BiasActGrad.apply(d_dx, x, b, y, dim, act_idx, alpha, gain, clamp, has_2nd, False)
        if has_2nd and (ctx.needs_input_grad[1] or ctx.needs_input_grad[2]):
            want_db = b is not None and ctx.needs_input_grad[2]
            d_x, d_b = _call(d_dx, b, x, y, dy, 2, dim, act_idx, alpha, gain, clamp, want_db)
        return d_dy, d_x, d_b, None, None, None, None, None, None, None, None
