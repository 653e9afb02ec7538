"""Oracle: numpy restatement of the arithmetic inside the reference's two CUDA plugins.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  The torch ref path
(``ref_ops.py``) pins forward values and autograd gradients; this file pins the
*kernel-level* contract the C ABI mirrors - ``grad`` = 0/1/2 of bias_act with the
``xref`` / ``yref`` / ``dy`` operands, and the gather index math of upfirdn2d -
for small inputs.  Paths relative to ``/root/reference/stylegan2_ada/torch_utils/ops``.
"""

import numpy as np

ACT_ID = {"linear": 1, "relu": 2, "lrelu": 3, "tanh": 4, "sigmoid": 5, "elu": 6, "selu": 7, "softplus": 8,
          "swish": 9}

_SELU_SCALE = 1.0507009873554804934193349852946
_SELU_ALPHA = 1.6732632423543772848170429916717


def bias_act_kernel(x, b, xref, yref, dy, grad, dim, act, alpha, gain, clamp):
    """bias_act.cu:23-147 in float64 numpy.  `x` is the primary operand (input for grad 0,
    incoming gradient for grad >= 1).  None = absent operand.  Returns y."""
    A = ACT_ID[act] if isinstance(act, str) else int(act)
    x = np.asarray(x, dtype=np.float64).copy()
    shape = [1] * x.ndim
    if b is not None:
        shape[dim] = -1
        bb = np.asarray(b, dtype=np.float64).reshape(shape)
    else:
        bb = 0.0
    xr = np.asarray(xref, dtype=np.float64).copy() if xref is not None else np.zeros_like(x)
    yr = np.asarray(yref, dtype=np.float64).copy() if yref is not None else np.zeros_like(x)
    d = np.asarray(dy, dtype=np.float64) if dy is not None else 1.0
    yy = yr / gain if gain != 0 else np.zeros_like(yr)

    if grad == 0:  # cu:52
        x = x + bb
    else:
        xr = xr + bb

    with np.errstate(over="ignore", invalid="ignore", divide="ignore"):
        y = np.zeros_like(x)
        if A == 1:  # cu:55-59
            if grad in (0, 1):
                y = x
        elif A == 2:  # cu:62-66
            if grad == 0:
                y = np.where(x > 0, x, 0.0)
            if grad == 1:
                y = np.where(yy > 0, x, 0.0)
        elif A == 3:  # cu:69-73
            if grad == 0:
                y = np.where(x > 0, x, x * alpha)
            if grad == 1:
                y = np.where(yy > 0, x, x * alpha)
        elif A == 4:  # cu:76-81
            if grad == 0:
                c = np.exp(x)
                dd = 1.0 / c
                y = np.where(x < -80, -1.0, np.where(x > 80, 1.0, (c - dd) / (c + dd)))
            if grad == 1:
                y = x * (1 - yy * yy)
            if grad == 2:
                y = x * (1 - yy * yy) * (-2 * yy)
        elif A == 5:  # cu:84-89
            if grad == 0:
                y = np.where(x < -80, 0.0, 1.0 / (np.exp(-x) + 1))
            if grad == 1:
                y = x * yy * (1 - yy)
            if grad == 2:
                y = x * yy * (1 - yy) * (1 - 2 * yy)
        elif A == 6:  # cu:92-97
            if grad == 0:
                y = np.where(x >= 0, x, np.exp(x) - 1)
            if grad == 1:
                y = np.where(yy >= 0, x, x * (yy + 1))
            if grad == 2:
                y = np.where(yy >= 0, 0.0, x * (yy + 1))
        elif A == 7:  # cu:100-105
            sa = _SELU_SCALE * _SELU_ALPHA
            if grad == 0:
                y = np.where(x >= 0, _SELU_SCALE * x, sa * (np.exp(x) - 1))
            if grad == 1:
                y = np.where(yy >= 0, x * _SELU_SCALE, x * (yy + sa))
            if grad == 2:
                y = np.where(yy >= 0, 0.0, x * (yy + sa))
        elif A == 8:  # cu:108-113
            if grad == 0:
                y = np.where(x > 80, x, np.log(np.exp(x) + 1))
            if grad == 1:
                y = x * (1 - np.exp(-yy))
            if grad == 2:
                c = np.exp(-yy)
                y = x * c * (1 - c)
        elif A == 9:  # cu:116-130
            if grad == 0:
                y = np.where(x < -80, 0.0, x / (np.exp(-x) + 1))
            else:
                c = np.exp(xr)
                dd = c + 1
                if grad == 1:
                    y = np.where(xr > 40, x, x * c * (xr + dd) / (dd * dd))
                else:
                    y = np.where(xr > 40, 0.0, x * c * (xr * (2 - dd) + 2 * dd) / (dd * dd * dd))
                yr = np.where(xr < -80, 0.0, xr / (np.exp(-xr) + 1) * gain)
        else:
            raise ValueError(f"unknown activation id {A}")

        y = y * (gain * d)  # cu:133
        if clamp >= 0:  # cu:136-142, strict inequalities
            if grad == 0:
                y = np.where((y > -clamp) & (y < clamp), y, np.where(y >= 0, clamp, -clamp))
            else:
                y = np.where((yr > -clamp) & (yr < clamp), y, 0.0)
    return y


def _floor_div(a, b):
    return a // b  # python ints already floor; upfirdn2d.cu:20-24


def upfirdn2d_out_size(in_size, up, pad0, pad1, ftaps, down):
    """upfirdn2d.cpp:32-33 (C integer division of a non-negative numerator)."""
    return (in_size * up + pad0 + pad1 - ftaps + down) // down


def upfirdn2d_kernel(x, f, upx, upy, downx, downy, padx0, padx1, pady0, pady1, flip, gain):
    """Gather form of the reference kernel, upfirdn2d.cu:29-92 (float64, python loops:
    small inputs only).  x: [N,C,H,W]; f: [fh,fw] float32 taps."""
    x = np.asarray(x, dtype=np.float64)
    f = np.asarray(f, dtype=np.float64)
    n, c, ih, iw = x.shape
    fh, fw = f.shape
    ow = upfirdn2d_out_size(iw, upx, padx0, padx1, fw, downx)
    oh = upfirdn2d_out_size(ih, upy, pady0, pady1, fh, downy)
    assert ow >= 1 and oh >= 1
    y = np.zeros((n, c, oh, ow), dtype=np.float64)
    for oy in range(oh):
        mid_y = oy * downy + upy - 1 - pady0
        in_y = min(max(_floor_div(mid_y, upy), 0), ih)
        h = min(max(_floor_div(mid_y + fh, upy), 0), ih) - in_y
        fy0 = mid_y + fh - (in_y + 1) * upy
        if flip:
            fy0 = fh - 1 - fy0
        step_y = upy if flip else -upy
        for ox in range(ow):
            mid_x = ox * downx + upx - 1 - padx0
            in_x = min(max(_floor_div(mid_x, upx), 0), iw)
            w = min(max(_floor_div(mid_x + fw, upx), 0), iw) - in_x
            fx0 = mid_x + fw - (in_x + 1) * upx
            if flip:
                fx0 = fw - 1 - fx0
            step_x = upx if flip else -upx
            acc = np.zeros((n, c), dtype=np.float64)
            for j in range(h):
                for i in range(w):
                    acc += x[:, :, in_y + j, in_x + i] * f[fy0 + j * step_y, fx0 + i * step_x]
            y[:, :, oy, ox] = acc * gain
    return y
