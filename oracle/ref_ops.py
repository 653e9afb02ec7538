"""Oracle: torch/CPU restatement of the reference's ``impl='ref'`` operator path.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Every function cites the
reference lines it follows; paths are relative to ``/root/reference/stylegan2_ada``.
All functions are plain differentiable torch code, so first/second order
gradients of the oracle come from autograd exactly as they do for the
reference's own ref path.
"""

import math

import torch
import torch.nn.functional as F

# ---------------------------------------------------------------------------
# bias_act  (torch_utils/ops/bias_act.py:23-33 table, :93-123 ref function)
# ---------------------------------------------------------------------------

_SQRT2 = math.sqrt(2.0)

# name -> (callable(x, alpha), default alpha, default gain, kernel id, saved ref, has 2nd grad)
ACTIVATIONS = {
    "linear": (lambda x, a: x, 0.0, 1.0, 1, "", False),
    "relu": (lambda x, a: F.relu(x), 0.0, _SQRT2, 2, "y", False),
    "lrelu": (lambda x, a: F.leaky_relu(x, a), 0.2, _SQRT2, 3, "y", False),
    "tanh": (lambda x, a: torch.tanh(x), 0.0, 1.0, 4, "y", True),
    "sigmoid": (lambda x, a: torch.sigmoid(x), 0.0, 1.0, 5, "y", True),
    "elu": (lambda x, a: F.elu(x), 0.0, 1.0, 6, "y", True),
    "selu": (lambda x, a: F.selu(x), 0.0, 1.0, 7, "y", True),
    "softplus": (lambda x, a: F.softplus(x), 0.0, 1.0, 8, "y", True),
    "swish": (lambda x, a: torch.sigmoid(x) * x, 0.0, _SQRT2, 9, "x", True),
}


def act_defaults(act):
    """(def_alpha, def_gain) of an activation, bias_act.py:23-33."""
    _, a, g, _, _, _ = ACTIVATIONS[act]
    return a, g


def bias_act(x, b=None, dim=1, act="linear", alpha=None, gain=None, clamp=None):
    """bias_act.py:93-123: (x + b) -> act -> *gain -> clamp(-c, c)."""
    fn, def_alpha, def_gain, _, _, _ = ACTIVATIONS[act]
    alpha = float(def_alpha if alpha is None else alpha)
    gain = float(def_gain if gain is None else gain)
    clamp = float(-1 if clamp is None else clamp)
    if b is not None:
        if b.ndim != 1 or not (0 <= dim < x.ndim) or b.shape[0] != x.shape[dim]:
            raise AssertionError("bias shape does not match x along dim")
        view = [1] * x.ndim
        view[dim] = -1
        x = x + b.reshape(view)
    x = fn(x, alpha)
    if gain != 1:
        x = x * gain
    if clamp >= 0:
        x = x.clamp(-clamp, clamp)
    return x


# ---------------------------------------------------------------------------
# upfirdn2d family  (torch_utils/ops/upfirdn2d.py)
# ---------------------------------------------------------------------------


def _pair(v):
    """upfirdn2d.py:37-44 (_parse_scaling)."""
    if isinstance(v, int):
        return v, v
    sx, sy = v
    assert isinstance(sx, int) and isinstance(sy, int) and sx >= 1 and sy >= 1
    return sx, sy


def parse_padding(padding):
    """upfirdn2d.py:46-55 -> (padx0, padx1, pady0, pady1)."""
    if isinstance(padding, int):
        padding = [padding, padding]
    padding = list(padding)
    assert all(isinstance(p, int) for p in padding)
    if len(padding) == 2:
        px, py = padding
        padding = [px, px, py, py]
    px0, px1, py0, py1 = padding
    return px0, px1, py0, py1


def filter_size(f):
    """upfirdn2d.py:57-68 -> (fw, fh); None filter is 1x1."""
    if f is None:
        return 1, 1
    assert f.ndim in (1, 2)
    return int(f.shape[-1]), int(f.shape[0])


def setup_filter(f, normalize=True, flip_filter=False, gain=1, separable=None):
    """upfirdn2d.py:72-116.  1-D taps with < 8 entries become a 2-D outer product."""
    if f is None:
        f = 1
    f = torch.as_tensor(f, dtype=torch.float32).clone()
    assert f.ndim in (0, 1, 2) and f.numel() > 0
    if f.ndim == 0:
        f = f[None]
    if separable is None:
        separable = f.ndim == 1 and f.numel() >= 8
    if f.ndim == 1 and not separable:
        f = torch.outer(f, f)
    assert f.ndim == (1 if separable else 2)
    if normalize:
        f = f / f.sum()
    if flip_filter:
        f = f.flip(list(range(f.ndim)))
    return f * (gain ** (f.ndim / 2))


def upfirdn2d(x, f, up=1, down=1, padding=0, flip_filter=False, gain=1):
    """upfirdn2d.py:168-208: zero-insert, pad/crop, FIR (true convolution unless
    flip_filter), decimate."""
    assert x.ndim == 4
    if f is None:
        f = torch.ones([1, 1], dtype=torch.float32)
    assert f.dtype == torch.float32 and f.ndim in (1, 2)
    n, c, h, w = x.shape
    upx, upy = _pair(up)
    dnx, dny = _pair(down)
    px0, px1, py0, py1 = parse_padding(padding)

    # zero insertion (:184-186)
    z = x.new_zeros(n, c, h * upy, w * upx)
    z[:, :, ::upy, ::upx] = x
    # positive pads, then negative pads as crops (:189-190)
    z = F.pad(z, [max(px0, 0), max(px1, 0), max(py0, 0), max(py1, 0)])
    z = z[:, :, max(-py0, 0): z.shape[2] - max(-py1, 0), max(-px0, 0): z.shape[3] - max(-px1, 0)]

    # filter set-up (:193-196): gain split over the passes; conv2d correlates, so a
    # true convolution needs the flipped taps.
    k = (f * (gain ** (f.ndim / 2))).to(x.dtype)
    if not flip_filter:
        k = k.flip(list(range(k.ndim)))
    if k.ndim == 2:
        z = F.conv2d(z, k[None, None].expand(c, 1, -1, -1), groups=c)
    else:  # separable: rows of taps along x, then along y (:202-204)
        z = F.conv2d(z, k[None, None, None, :].expand(c, 1, 1, -1), groups=c)
        z = F.conv2d(z, k[None, None, :, None].expand(c, 1, -1, 1), groups=c)
    return z[:, :, ::dny, ::dnx]


def filter2d(x, f, padding=0, flip_filter=False, gain=1):
    """upfirdn2d.py:272-304: same-size filtering."""
    px0, px1, py0, py1 = parse_padding(padding)
    fw, fh = filter_size(f)
    p = [px0 + fw // 2, px1 + (fw - 1) // 2, py0 + fh // 2, py1 + (fh - 1) // 2]
    return upfirdn2d(x, f, padding=p, flip_filter=flip_filter, gain=gain)


def upsample2d(x, f, up=2, padding=0, flip_filter=False, gain=1):
    """upfirdn2d.py:308-343: output is `up` x the input; gain scaled by up^2."""
    upx, upy = _pair(up)
    px0, px1, py0, py1 = parse_padding(padding)
    fw, fh = filter_size(f)
    p = [px0 + (fw + upx - 1) // 2, px1 + (fw - upx) // 2,
         py0 + (fh + upy - 1) // 2, py1 + (fh - upy) // 2]
    return upfirdn2d(x, f, up=up, padding=p, flip_filter=flip_filter, gain=gain * upx * upy)


def downsample2d(x, f, down=2, padding=0, flip_filter=False, gain=1):
    """upfirdn2d.py:347-382: output is 1/`down` of the input."""
    dnx, dny = _pair(down)
    px0, px1, py0, py1 = parse_padding(padding)
    fw, fh = filter_size(f)
    p = [px0 + (fw - dnx + 1) // 2, px1 + (fw - dnx) // 2,
         py0 + (fh - dny + 1) // 2, py1 + (fh - dny) // 2]
    return upfirdn2d(x, f, down=down, padding=p, flip_filter=flip_filter, gain=gain)


# ---------------------------------------------------------------------------
# conv2d_resample  (torch_utils/ops/conv2d_resample.py)
# ---------------------------------------------------------------------------


def _conv(x, w, stride=1, padding=0, groups=1, transpose=False, flip_weight=True):
    """conv2d_resample.py:29-54 minus the cuDNN channels_last work-around (a pure
    performance path that computes the same thing).  ATen conv correlates, so
    flip_weight=False flips the taps first (:35-36)."""
    if not flip_weight:
        w = w.flip([2, 3])
    if transpose:
        return F.conv_transpose2d(x, w, stride=stride, padding=padding, groups=groups)
    return F.conv2d(x, w, stride=stride, padding=padding, groups=groups)


def conv2d_resample(x, w, f=None, up=1, down=1, padding=0, groups=1, flip_weight=True, flip_filter=False):
    """conv2d_resample.py:59-154, same branch order."""
    assert x.ndim == 4 and w.ndim == 4 and w.dtype == x.dtype
    assert f is None or (f.dtype == torch.float32 and f.ndim in (1, 2))
    cout, cin_g, kh, kw = (int(s) for s in w.shape)
    fw, fh = filter_size(f)
    px0, px1, py0, py1 = parse_padding(padding)

    if up > 1:  # :89-93
        px0 += (fw + up - 1) // 2
        px1 += (fw - up) // 2
        py0 += (fh + up - 1) // 2
        py1 += (fh - up) // 2
    if down > 1:  # :94-98
        px0 += (fw - down + 1) // 2
        px1 += (fw - down) // 2
        py0 += (fh - down + 1) // 2
        py1 += (fh - down) // 2

    if kw == 1 and kh == 1 and down > 1 and up == 1:  # :101-104
        x = upfirdn2d(x, f, down=down, padding=[px0, px1, py0, py1], flip_filter=flip_filter)
        return _conv(x, w, groups=groups, flip_weight=flip_weight)

    if kw == 1 and kh == 1 and up > 1 and down == 1:  # :107-110
        x = _conv(x, w, groups=groups, flip_weight=flip_weight)
        return upfirdn2d(x, f, up=up, padding=[px0, px1, py0, py1], gain=up ** 2, flip_filter=flip_filter)

    if down > 1 and up == 1:  # :113-116
        x = upfirdn2d(x, f, padding=[px0, px1, py0, py1], flip_filter=flip_filter)
        return _conv(x, w, stride=down, groups=groups, flip_weight=flip_weight)

    if up > 1:  # :119-136 transpose conv + FIR
        if groups == 1:
            wt = w.transpose(0, 1)
        else:
            wt = w.reshape(groups, cout // groups, cin_g, kh, kw).transpose(1, 2)
            wt = wt.reshape(groups * cin_g, cout // groups, kh, kw)
        px0 -= kw - 1
        px1 -= kw - up
        py0 -= kh - 1
        py1 -= kh - up
        pxt = max(min(-px0, -px1), 0)
        pyt = max(min(-py0, -py1), 0)
        x = _conv(x, wt, stride=up, padding=[pyt, pxt], groups=groups, transpose=True, flip_weight=(not flip_weight))
        x = upfirdn2d(x, f, padding=[px0 + pxt, px1 + pxt, py0 + pyt, py1 + pyt], gain=up ** 2, flip_filter=flip_filter)
        if down > 1:
            x = upfirdn2d(x, f, down=down, flip_filter=flip_filter)
        return x

    if up == 1 and down == 1 and px0 == px1 and py0 == py1 and px0 >= 0 and py0 >= 0:  # :139-141
        return _conv(x, w, padding=[py0, px0], groups=groups, flip_weight=flip_weight)

    # generic fallback (:144-148)
    x = upfirdn2d(x, (f if up > 1 else None), up=up, padding=[px0, px1, py0, py1], gain=up ** 2, flip_filter=flip_filter)
    x = _conv(x, w, groups=groups, flip_weight=flip_weight)
    if down > 1:
        x = upfirdn2d(x, f, down=down, flip_filter=flip_filter)
    return x


# ---------------------------------------------------------------------------
# fma  (torch_utils/ops/fma.py:15-27)
# ---------------------------------------------------------------------------


def fma(a, b, c):
    """a * b + c (the reference's custom backward is the autograd of this)."""
    return torch.addcmul(c, a, b)


# ---------------------------------------------------------------------------
# modulated_conv2d  (training/networks.py:27-84)
# ---------------------------------------------------------------------------


def modulated_conv2d(x, weight, styles, noise=None, up=1, down=1, padding=0, resample_filter=None,
                     demodulate=True, flip_weight=True, fused_modconv=True):
    n = int(x.shape[0])
    cout, cin, kh, kw = weight.shape
    assert x.shape[1] == cin and tuple(styles.shape) == (n, cin)

    if x.dtype == torch.float16 and demodulate:  # :47-49 fp16 pre-normalisation
        weight = weight * (1 / math.sqrt(cin * kh * kw) / weight.norm(float("inf"), dim=[1, 2, 3], keepdim=True))
        styles = styles / styles.norm(float("inf"), dim=1, keepdim=True)

    wmod = dcoefs = None
    if demodulate or fused_modconv:  # :54-56
        wmod = weight.unsqueeze(0) * styles.reshape(n, 1, cin, 1, 1)
    if demodulate:  # :57-58
        dcoefs = (wmod.square().sum(dim=[2, 3, 4]) + 1e-8).rsqrt()
    if demodulate and fused_modconv:  # :59-60
        wmod = wmod * dcoefs.reshape(n, cout, 1, 1, 1)

    if not fused_modconv:  # :63-72 scale activations before / after a shared-weight conv
        x = x * styles.to(x.dtype).reshape(n, cin, 1, 1)
        x = conv2d_resample(x, weight.to(x.dtype), f=resample_filter, up=up, down=down, padding=padding,
                            flip_weight=flip_weight)
        if demodulate and noise is not None:
            x = fma(x, dcoefs.to(x.dtype).reshape(n, cout, 1, 1), noise.to(x.dtype))
        elif demodulate:
            x = x * dcoefs.to(x.dtype).reshape(n, cout, 1, 1)
        elif noise is not None:
            x = x + noise.to(x.dtype)
        return x

    # :75-84 one grouped conv with groups = batch
    x = x.reshape(1, n * cin, *x.shape[2:])
    wg = wmod.reshape(n * cout, cin, kh, kw)
    x = conv2d_resample(x, wg.to(x.dtype), f=resample_filter, up=up, down=down, padding=padding, groups=n,
                        flip_weight=flip_weight)
    x = x.reshape(n, cout, *x.shape[2:])
    if noise is not None:
        x = x + noise
    return x
