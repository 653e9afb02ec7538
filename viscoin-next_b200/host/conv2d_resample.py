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

Covered: groups == 1, 1x1 and 3x3 kernels, factors 1 and 2, symmetric padding k//2 - everything ``training/networks.py``
asks for (SynthesisBlock, the 'resnet' skips, DiscriminatorBlock's down=2 convolutions).  Anything else (grouped
convolutions, asymmetric padding, other factors / kernel sizes) raises ``VcnError``: the reference's generic fallback
(:150-154) has no sm_100a counterpart here and there is no silent library path.  ``impl='ref'`` is the reference's
formulation in plain torch (any device) - the checker, not a fallback."""

import torch

from .. import _lib
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


def _conv(x, w, flip_weight=True):
    """Stride-1 convolution, padding k//2 (1x1 or 3x3), no modulation, through the fused layer op (dx, dw in its backward)."""
    return _modconv.modconv_layer(x, w, None, demodulate=False, flip_weight=flip_weight, act=None)


def _unsupported(what):
    raise _lib.VcnError("conv2d_resample(impl='cuda'): %s is not covered by the sm_100a path (use impl='ref' to check a "
                        "configuration; there is no library fallback)" % what)


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
        _unsupported("groups = %d" % groups)
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
