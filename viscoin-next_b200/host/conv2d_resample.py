"""conv2d_resample - mirror of ``stylegan2_ada/torch_utils/ops/conv2d_resample.py`` (:59-154) on the sm_100a kernels.

Same signature and the same branch order as the reference:

    conv2d_resample(x, w, f=None, up=1, down=1, padding=0, groups=1, flip_weight=True, flip_filter=False, impl='cuda')

``impl='cuda'`` composes the operators of this package (every convolution is the tcgen05 implicit GEMM ``vcn_conv_igemm``
with bf16 operands and fp32 accumulation, every FIR is ``vcn_upfirdn2d``), forward and backward (dx, dw):

    branch of the reference (conv2d_resample.py)               here
    ---------------------------------------------------------  ------------------------------------------------------------
    1x1, down>1          (:101-104)  FIR-decimate -> 1x1 conv   upfirdn2d(down) -> FPROP
    1x1, up>1            (:107-110)  1x1 conv -> FIR-upsample   FPROP -> upfirdn2d(up, gain up^2)
    kxk, down>1, up=1    (:113-116)  FIR -> strided conv        upfirdn2d -> SCONV2 (3x3, stride 2; dx = TCONV2, dw = wgrad)
    kxk, up>1            (:119-142)  transposed conv -> FIR     TCONV2 -> upfirdn2d (3x3, up=2: the SynthesisLayer path)
    no resampling        (:145-147)  conv                       FPROP / DGRAD (flipped taps)

Covered: 1x1 and 3x3 kernels, factors 1 and 2, symmetric padding k//2, groups == 1 or the per-sample grouping of the
fused-modconv branch (``groups = N`` with x = [1, N*Cin, H, W], networks.py:78-81 -> the GEMM's per-sample-weight mode) -
everything ``training/networks.py`` asks for (SynthesisBlock in train and eval mode, the 'resnet' skips,
DiscriminatorBlock's down=2 convolutions).  Anything else (other groupings, asymmetric padding, other factors / kernel
sizes) raises ``VcnError``: the reference's generic fallback
(:150-154) has no sm_100a counterpart here and there is no silent library path.  ``impl='ref'`` is the reference's
formulation in plain torch (any device) - the checker, not a fallback."""

import torch

from .. import _lib
from .misc import profiled_function as _profiled
from . import modconv as _modconv
from . import upfirdn2d as _upfirdn2d


def _get_weight_shape(w):
    return [int(s) for s in w.shape]


class _StridedConv3x3(torch.autograd.Function):
    """y[n,co,i,j] = sum_{ci,ky,kx} z[n,ci,2i+ky,2j+kx] * W[co,ci,ky,kx]  (3x3, stride 2, no padding: (2h+1, 2w+1) -> (h, w)),
    the strided convolution of conv2d_resample.py:113-116.  Forward = SCONV2, dz = TCONV2 with the transposed pack,
    dW = the transposed-conv weight-gradient GEMM with the operand roles exchanged."""

    @staticmethod
    def forward(ctx, z, weight):
        n, cin, hz, wz = z.shape
        cout = weight.shape[0]
        zhat = _modconv.modulate_pack(z, None)
        wt, _ = _modconv.pack_weights(weight)
        y = _modconv.conv_igemm(zhat, wt, _lib.CONV_SCONV2, cout, 3, 3, out_dtype=torch.float32)
        ctx.save_for_backward(zhat if weight.requires_grad else None, weight)
        ctx.zshape = tuple(z.shape)
        ctx.zdtype = z.dtype
        return y.to(z.dtype)

    @staticmethod
    def backward(ctx, dy):
        zhat, weight = ctx.saved_tensors
        n, cin, hz, wz = ctx.zshape
        cout = weight.shape[0]
        dyh = _modconv.modulate_pack(dy, None)  # bf16 NHWC [n, h, w, cout_pad]
        dz = dw = None
        if ctx.needs_input_grad[0]:
            wt_t, _ = _modconv.pack_weights(weight, transpose_io=True)
            dz = _modconv.conv_igemm(dyh, wt_t, _lib.CONV_TCONV2, cin, 3, 3, out_dtype=torch.float32)
            assert tuple(dz.shape) == ctx.zshape
            dz = dz.to(ctx.zdtype)
        if ctx.needs_input_grad[1]:
            # kind-1 wgrad: out[a][b][k] = sum big[n,2i+ky,2j+kx,a] * small[n,i,j,b]; here big = z (ci), small = dy (co)
            dw = _modconv.conv_wgrad(zhat, dyh, cin, cout, 3, 3, transposed=True).transpose(0, 1).contiguous().to(weight.dtype)
        return dz, dw


class _GroupedConv(torch.autograd.Function):
    """conv2d_resample(..., groups=G) as the reference's fused-modconv branch calls it (networks.py:78-81):
    x = [1, G*Cin, H, W], w = [G*Cout, Cin, k, k] - G independent convolutions, one per sample.  That is exactly the
    per-sample-weight mode of the implicit GEMM (``w_per_sample``: the weight tile's 4th TMA coordinate is the sample):
    up=1 -> FPROP / DGRAD taps, up=2 -> TCONV2 + the 4x4 FIR.  Backward: dx = the per-sample transposed GEMM, dw = one
    weight-gradient GEMM per group.  Unpickled generators keep their own ``modulated_conv2d`` (persistence.py:179-227) and
    reach the sm_100a kernels through this entry in eval mode."""

    @staticmethod
    def _pack(w, groups, flip, transpose_io):
        gco, cin, kh, kw = w.shape
        cout = gco // groups
        src = w.detach()
        if flip:
            src = src.flip([2, 3])
        if transpose_io:  # [G,Cout,Cin,k,k] -> [G*Cin, Cout, k, k]: rows = in-channels of the forward
            src = src.reshape(groups, cout, cin, kh, kw).transpose(1, 2).reshape(groups * cin, cout, kh, kw)
        rows = src.shape[0] // groups
        wt, _ = _modconv.pack_weights(src.contiguous())  # [taps, G*rows, cols_pad]
        return wt.reshape(kh * kw, groups, rows, wt.shape[-1]).permute(1, 0, 2, 3).contiguous()  # [G, taps, rows, cols_pad]

    @staticmethod
    def forward(ctx, x, w, f, groups, up, flip_weight):
        _, gc, h, wd = x.shape
        cin, cout, k = gc // groups, w.shape[0] // groups, w.shape[-1]
        xs = x.reshape(groups, cin, h, wd)
        xhat = _modconv.modulate_pack(xs, None)
        if up == 1:
            wt = _GroupedConv._pack(w, groups, False, False)
            kind = _lib.CONV_FPROP if flip_weight else _lib.CONV_DGRAD
            y = _modconv.conv_igemm(xhat, wt, kind, cout, k, k, out_dtype=torch.float32, w_per_sample=True)
        else:
            wt = _GroupedConv._pack(w, groups, bool(flip_weight), False)
            z = _modconv.conv_igemm(xhat, wt, _lib.CONV_TCONV2, cout, 3, 3, out_dtype=torch.float32, w_per_sample=True)
            y = _upfirdn2d.launch(z, f.to(device=x.device, dtype=torch.float32), 1, 1, 1, 1, 1, 1, 1, 1, False, 4.0)
        ctx.save_for_backward(xhat if w.requires_grad else None, w, f)
        ctx.cfg = (groups, up, bool(flip_weight), cin, cout, k, x.dtype)
        return y.reshape(1, groups * cout, y.shape[2], y.shape[3]).to(x.dtype)

    @staticmethod
    def backward(ctx, dy):
        xhat, w, f = ctx.saved_tensors
        groups, up, flip_weight, cin, cout, k, xdtype = ctx.cfg
        dy = dy.reshape(groups, cout, dy.shape[2], dy.shape[3])
        dy0 = _modconv.modulate_pack(dy, None)
        if up == 2:  # FIR^T: the same FIR with the flipped filter, padding fw-1-p = 2 (upfirdn2d.py:246-261)
            dy0 = _upfirdn2d.launch(dy0.permute(0, 3, 1, 2), f.to(device=dy.device, dtype=torch.float32), 1, 1, 1, 1, 2, 2, 2, 2, True, 4.0)
            dy0 = dy0.permute(0, 2, 3, 1)
            assert dy0.is_contiguous()
        dx = dw = None
        if ctx.needs_input_grad[0]:
            if up == 1:
                wt_t = _GroupedConv._pack(w, groups, False, True)
                kind = _lib.CONV_DGRAD if flip_weight else _lib.CONV_FPROP
                dx = _modconv.conv_igemm(dy0, wt_t, kind, cin, k, k, out_dtype=torch.float32, w_per_sample=True)
            else:
                wt_t = _GroupedConv._pack(w, groups, flip_weight, True)
                dx = _modconv.conv_igemm(dy0, wt_t, _lib.CONV_SCONV2, cin, 3, 3, out_dtype=torch.float32, w_per_sample=True)
            dx = dx.reshape(1, groups * cin, dx.shape[2], dx.shape[3]).to(xdtype)
        if ctx.needs_input_grad[1]:
            parts = []
            for g in range(groups):
                d = _modconv.conv_wgrad(dy0[g:g + 1], xhat[g:g + 1], cout, cin, k, k, transposed=(up == 2))
                if (up == 1 and not flip_weight) or (up == 2 and flip_weight):
                    d = d.flip([2, 3])
                parts.append(d)
            dw = torch.cat(parts).to(w.dtype)
        return dx, dw, None, None, None, None


def _conv(x, w, flip_weight=True):
    """Stride-1 convolution, padding k//2 (1x1 or 3x3), no modulation, through the fused layer op (dx, dw in its backward)."""
    return _modconv.modconv_layer(x, w, None, demodulate=False, flip_weight=flip_weight, act=None)


def _unsupported(what):
    raise _lib.VcnError("conv2d_resample(impl='cuda'): %s is not covered by the sm_100a path (use impl='ref' to check a "
                        "configuration; there is no library fallback)" % what)


@_profiled
def conv2d_resample(x, w, f=None, up=1, down=1, padding=0, groups=1, flip_weight=True, flip_filter=False, impl="cuda"):
    """conv2d_resample.py:59-154.  x [N,Cin,H,W]; w [Cout,Cin/groups,kh,kw]; f = setup_filter() result or None."""
    assert impl in ["ref", "cuda"]
    assert isinstance(x, torch.Tensor) and x.ndim == 4
    assert isinstance(w, torch.Tensor) and w.ndim == 4 and w.dtype == x.dtype
    assert f is None or (isinstance(f, torch.Tensor) and f.ndim in [1, 2] and f.dtype == torch.float32)
    assert isinstance(up, int) and up >= 1
    assert isinstance(down, int) and down >= 1
    assert isinstance(groups, int) and groups >= 1
    if impl == "ref":
        return _conv2d_resample_ref(x, w, f, up, down, padding, groups, flip_weight, flip_filter)
    if x.device.type != "cuda":
        raise _lib.VcnError("conv2d_resample(impl='cuda') needs a CUDA tensor; this build has no CPU fallback (use impl='ref')")
    cout, cin_g, kh, kw = _get_weight_shape(w)
    fw, fh = _upfirdn2d._get_filter_size(f)
    px0, px1, py0, py1 = _upfirdn2d._parse_padding(padding)
    if groups != 1:
        # the fused-modconv call pattern (networks.py:78-81): one group per sample, no resampling except up=2 with the 4x4 filter
        ok = x.shape[0] == 1 and x.shape[1] % groups == 0 and cout % groups == 0 and x.shape[1] // groups == cin_g \
            and kh == kw and kh in (1, 3) and down == 1 and px0 == px1 == py0 == py1 == kh // 2 \
            and (up == 1 or (up == 2 and kh == 3 and f is not None and tuple(f.shape) == (4, 4) and not flip_filter))
        if not ok:
            _unsupported("groups = %d outside the per-sample pattern x=[1,G*Cin,H,W], w=[G*Cout,Cin,k,k], k in {1,3}, up in {1,2}" % groups)
        return _GroupedConv.apply(x, w, f, groups, up, flip_weight)
    if kh != kw or kh not in (1, 3):
        _unsupported("a %dx%d kernel" % (kh, kw))
    if up not in (1, 2) or down not in (1, 2) or (up == 2 and down == 2):
        _unsupported("up = %d, down = %d" % (up, down))
    if not (px0 == px1 == py0 == py1 == kh // 2):
        _unsupported("padding %s with a %dx%d kernel (only k//2 on every side)" % ([px0, px1, py0, py1], kh, kw))

    # padding adjusted for the resampling filter (:89-98)
    if up > 1:
        px0 += (fw + up - 1) // 2; px1 += (fw - up) // 2; py0 += (fh + up - 1) // 2; py1 += (fh - up) // 2
    if down > 1:
        px0 += (fw - down + 1) // 2; px1 += (fw - down) // 2; py0 += (fh - down + 1) // 2; py1 += (fh - down) // 2

    if kh == 1 and down > 1:  # :101-104
        x = _upfirdn2d.upfirdn2d(x, f, down=down, padding=[px0, px1, py0, py1], flip_filter=flip_filter)
        return _conv(x, w, flip_weight)
    if kh == 1 and up > 1:  # :107-110
        x = _conv(x, w, flip_weight)
        return _upfirdn2d.upfirdn2d(x, f, up=up, padding=[px0, px1, py0, py1], gain=up ** 2, flip_filter=flip_filter)
    if down > 1:  # :113-116
        z = _upfirdn2d.upfirdn2d(x, f, padding=[px0, px1, py0, py1], flip_filter=flip_filter)
        if z.shape[2] % 2 != 1 or z.shape[3] % 2 != 1 or z.shape[2] < 3 or z.shape[3] < 3:
            _unsupported("a stride-2 3x3 convolution over a %dx%d filtered input (needs odd extents >= 3)" % (z.shape[2], z.shape[3]))
        return _StridedConv3x3.apply(z, w if flip_weight else w.flip([2, 3]))
    if up > 1:  # :119-142 transposed conv + FIR (the fused op carries both, and their backward)
        if f is None or tuple(f.shape) != (4, 4) or flip_filter:
            _unsupported("up = 2 with a filter other than an unflipped 4x4")
        return _modconv.modconv_layer(x, w, None, up=2, resample_filter=f, demodulate=False, flip_weight=flip_weight, act=None)
    return _conv(x, w, flip_weight)  # :145-147


def _conv2d_resample_ref(x, w, f, up, down, padding, groups, flip_weight, flip_filter):
    """The reference's branches in plain torch (conv2d_resample.py:29-54 wrapper + :59-154), any device."""
    F = torch.nn.functional
    U = lambda t, **kw: _upfirdn2d.upfirdn2d(t, f, impl="ref", **kw)  # noqa: E731
    cout, cin_g, kh, kw = _get_weight_shape(w)
    fw, fh = _upfirdn2d._get_filter_size(f)
    px0, px1, py0, py1 = _upfirdn2d._parse_padding(padding)

    def conv(t, wt, stride=1, pad=(0, 0), transpose=False, flip=True):
        if not flip:  # conv2d_resample.py:35-36: the wrapper flips when flip_weight is False
            wt = wt.flip([2, 3])
        op = F.conv_transpose2d if transpose else F.conv2d
        return op(t, wt.to(t.dtype), stride=stride, padding=list(pad), groups=groups)

    if up > 1:
        px0 += (fw + up - 1) // 2; px1 += (fw - up) // 2; py0 += (fh + up - 1) // 2; py1 += (fh - up) // 2
    if down > 1:
        px0 += (fw - down + 1) // 2; px1 += (fw - down) // 2; py0 += (fh - down + 1) // 2; py1 += (fh - down) // 2
    if kw == 1 and kh == 1 and (down > 1 and up == 1):
        return conv(U(x, down=down, padding=[px0, px1, py0, py1], flip_filter=flip_filter), w, flip=flip_weight)
    if kw == 1 and kh == 1 and (up > 1 and down == 1):
        return U(conv(x, w, flip=flip_weight), up=up, padding=[px0, px1, py0, py1], gain=up ** 2, flip_filter=flip_filter)
    if down > 1 and up == 1:
        return conv(U(x, padding=[px0, px1, py0, py1], flip_filter=flip_filter), w, stride=down, flip=flip_weight)
    if up > 1:
        if groups == 1:
            wt = w.transpose(0, 1)
        else:
            wt = w.reshape(groups, cout // groups, cin_g, kh, kw).transpose(1, 2).reshape(groups * cin_g, cout // groups, kh, kw)
        px0 -= kw - 1; px1 -= kw - up; py0 -= kh - 1; py1 -= kh - up
        pxt, pyt = max(min(-px0, -px1), 0), max(min(-py0, -py1), 0)
        x = conv(x, wt, stride=up, pad=(pyt, pxt), transpose=True, flip=(not flip_weight))
        x = U(x, padding=[px0 + pxt, px1 + pxt, py0 + pyt, py1 + pyt], gain=up ** 2, flip_filter=flip_filter)
        if down > 1:
            x = U(x, down=down, flip_filter=flip_filter)
        return x
    if up == 1 and down == 1 and px0 == px1 and py0 == py1 and px0 >= 0 and py0 >= 0:
        return conv(x, w, pad=(py0, px0), flip=flip_weight)
    x = U(x, up=up, padding=[px0, px1, py0, py1], gain=up ** 2, flip_filter=flip_filter) if f is not None or up > 1 or (px0, px1, py0, py1) != (0, 0, 0, 0) else x
    x = conv(x, w, flip=flip_weight)
    if down > 1:
        x = U(x, down=down, flip_filter=flip_filter)
    return x
