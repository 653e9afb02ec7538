"""frozen_nets - the two FROZEN networks of the VisCoIN iteration on a fused 16-bit channels-last path:

  FusedVGGFeatures   the five LPIPS taps of VGG-16 (third-party ``lpips`` v0.1.4 on torchvision's vgg16.features;
                     viscoin/training/losses.py:111-123)
  FusedResNet        the HF ResNet-50 of ``Classifier`` after ``fold_batchnorm`` (viscoin/models/classifiers.py:83-92)

Both are frozen and in eval mode in the reference's loop (viscoin/training/viscoin.py:114-116): only input gradients are
ever needed.  Under autocast the step spent ~7 of its ~22 ms in ATen elementwise kernels around their convolutions (one
bias-add, one ReLU and, in ResNet, one residual-add pass per convolution, weight casts every call, max-pool forward at
~1 TB/s and a four-pass max-pool backward).  Here:

  * weights are cast to bf16 channels_last ONCE (cached per parameter version), activations stay bf16 NHWC;
  * convolution + bias + ReLU (+ residual add) is ONE library call - cuDNN's fused ``cudnn_convolution_relu`` /
    ``cudnn_convolution_add_relu`` (the convolutions of f / VGG-16 are library code in the reference too)
    - or, for VGG's 3x3 convolutions with ``impl='vcn'``, this package's tcgen05 implicit GEMM with the same epilogue;
  * the passes between them are this package's kernels (csrc/frozen_ops.cu): 2x2 max-pool forward, and ONE backward pass
    per layer that does max-pool backward + tap-gradient add + ReLU mask (``vcn_maxpool2x2_relu_bwd`` / ``vcn_relu_mask_bwd``);
  * the input gradient of each convolution is ``aten::convolution_backward`` with only the input mask set (cuDNN dgrad), or
    the implicit GEMM's DGRAD with ``impl='vcn'``.

Numerics = the bf16 autocast path they replace (bf16 operands, fp32 accumulation, bf16 storage); parity tests compare with
the plain fp32 modules (tests/test_gpu_frozen.py).
"""

import ctypes
import os
import weakref

import torch
import torch.nn.functional as F

from .. import _lib
from . import image_ops as _image_ops


def _cl(t):
    return t.contiguous(memory_format=torch.channels_last)


class _WeightCache:
    """bf16 channels_last copies of frozen parameters, re-made when a parameter's version or storage changes."""

    def __init__(self):
        self._c = {}

    def get(self, p, dtype=torch.bfloat16):
        key = id(p)
        hit = self._c.get(key)
        # the entry holds a weak reference to ITS parameter: a recycled id (+ recycled storage address, same version) of a dead
        # parameter must not serve another network's weights
        if hit is not None and hit[2]() is p and hit[0] == (p._version, p.data_ptr(), dtype):
            return hit[1]
        if p.is_cuda and torch.cuda.is_current_stream_capturing():
            raise _lib.VcnError("frozen_nets: weight cache miss inside a CUDA graph capture (run one eager step first)")
        with torch.no_grad():
            q = p.detach().to(dtype)
            if q.ndim == 4:
                q = _cl(q)
        if len(self._c) > 1024:
            for k in [k for k, v in self._c.items() if v[2]() is None]:
                del self._c[k]
        self._c[key] = ((p._version, p.data_ptr(), dtype), q, weakref.ref(p))
        return q


_weights = _WeightCache()

# first (3-channel) convolutions of VGG-16 and of the ResNet stem: "direct" = csrc/first_conv.cu (default), "cudnn" = the
# library kernel, "vcn" = the tcgen05 implicit GEMM on an 8-channel NHWC copy of the image (VGG only).  Measured in the step
# (profiles/r2l_ab.txt): vcn 16.13 ms vs cudnn 15.55 ms - a 64-channel k-block carrying 8 real channels wastes 7/8 of the
# operand staging.
first_layer_impl = os.environ.get("VCN_VGG_FIRST", "direct")
_first_packs = {}
fused_image_ops = os.environ.get("VCN_FUSED_IMAGE_OPS", "1") == "1"
resnet_fork = os.environ.get("VCN_RESNET_FORK", "1") == "1"


def conv_first3(x, w, b, ks, stride, pad, relu=True):
    """y = [relu](conv(x, w) + b) for a 3-channel bf16 channels_last image and a [64,3,ks,ks] weight (vcn_conv_first3)."""
    n, _, h, wd = x.shape
    key = (id(w), w._version, w.data_ptr())
    hit = _first_packs.get(id(w))
    if hit is None or hit[0] != key:
        if torch.cuda.is_current_stream_capturing():
            raise _lib.VcnError("frozen_nets: weight pack miss inside a CUDA graph capture (run one eager step first)")
        k = ks * ks * 3
        kp = (k + 15) // 16 * 16
        wp = torch.zeros([kp, 64], dtype=torch.bfloat16, device=w.device)
        wp[:k] = w.detach().permute(2, 3, 1, 0).reshape(k, 64).to(torch.bfloat16)
        hit = (key, wp, b.detach().float().contiguous() if b is not None else None, w)
        _first_packs[id(w)] = hit
    _, wp, bias, _ = hit
    oh, ow = (h + 2 * pad - ks) // stride + 1, (wd + 2 * pad - ks) // stride + 1
    y = torch.empty([n, 64, oh, ow], dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    with torch.cuda.device(x.device):
        st = _lib.load().vcn_conv_first3(_lib.ptr(x), _lib.ptr(wp), _lib.ptr(bias), _lib.ptr(y), _lib.dtype_code(x.dtype), ks, stride,
                                         pad, int(relu), n, h, wd, _lib.current_stream(x.device))
    _lib.check(st, "vcn_conv_first3")
    return y
# MaxPool2d(3, 2, 1) of the ResNet stem: "vcn" = csrc/frozen_ops.cu, "aten" = torch's NHWC kernels
stem_pool_impl = os.environ.get("VCN_STEM_POOL", "vcn")


# ---- this package's kernels ---------------------------------------------------------------------------------------------


def maxpool2x2(x):
    """x: 16-bit channels_last [N,C,H,W] -> [N,C,H/2,W/2] (channels_last).  No autograd (used inside the fused functions)."""
    n, c, h, w = x.shape
    assert x.is_contiguous(memory_format=torch.channels_last) and x.dtype in (torch.bfloat16, torch.float16)
    y = torch.empty([n, c, h // 2, w // 2], dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    with torch.cuda.device(x.device):
        st = _lib.load().vcn_maxpool2x2_fwd(_lib.ptr(x), _lib.ptr(y), _lib.dtype_code(x.dtype), n, h, w, c, _lib.current_stream(x.device))
    _lib.check(st, "vcn_maxpool2x2_fwd")
    return y


def relu_mask_bwd(y, g, g2=None):
    """dz = (g + g2) * (y > 0), all 16-bit with y's (dense) layout."""
    assert g.shape == y.shape and g.dtype == y.dtype and g.stride() == y.stride()
    if g2 is not None:
        assert g2.shape == y.shape and g2.dtype == y.dtype and g2.stride() == y.stride()
    dz = torch.empty_like(y)
    with torch.cuda.device(y.device):
        st = _lib.load().vcn_relu_mask_bwd(_lib.ptr(y), _lib.ptr(g), _lib.ptr(g2), _lib.ptr(dz), _lib.dtype_code(y.dtype), y.numel(),
                                           _lib.current_stream(y.device))
    _lib.check(st, "vcn_relu_mask_bwd")
    return dz


def maxpool2x2_relu_bwd(y, gpool, gtap=None):
    """dz = (gtap + maxpool_backward(gpool; y)) * (y > 0);  y channels_last [N,C,H,W], gpool [N,C,H/2,W/2]."""
    n, c, h, w = y.shape
    assert y.is_contiguous(memory_format=torch.channels_last) and gpool.is_contiguous(memory_format=torch.channels_last)
    assert gpool.shape == (n, c, h // 2, w // 2) and gpool.dtype == y.dtype
    if gtap is not None:
        assert gtap.shape == y.shape and gtap.dtype == y.dtype and gtap.is_contiguous(memory_format=torch.channels_last)
    dz = torch.empty_like(y)
    with torch.cuda.device(y.device):
        st = _lib.load().vcn_maxpool2x2_relu_bwd(_lib.ptr(y), _lib.ptr(gpool), _lib.ptr(gtap), _lib.ptr(dz), _lib.dtype_code(y.dtype),
                                                 n, h, w, c, _lib.current_stream(y.device))
    _lib.check(st, "vcn_maxpool2x2_relu_bwd")
    return dz


class _MaxPool3x3s2(torch.autograd.Function):
    """torch.nn.MaxPool2d(3, 2, 1) on 16-bit channels_last activations (the ResNet stem's pooler) on this package's kernels."""

    @staticmethod
    def forward(ctx, x):
        n, c, h, w = x.shape
        oh, ow = (h - 1) // 2 + 1, (w - 1) // 2 + 1
        y = torch.empty([n, c, oh, ow], dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        need = ctx is not None and x.requires_grad
        idx = torch.empty([n, oh, ow, c], dtype=torch.uint8, device=x.device) if need else None
        with torch.cuda.device(x.device):
            st = _lib.load().vcn_maxpool3x3s2_fwd(_lib.ptr(x), _lib.ptr(y), _lib.ptr(idx), _lib.dtype_code(x.dtype), n, h, w, c,
                                                  _lib.current_stream(x.device))
        _lib.check(st, "vcn_maxpool3x3s2_fwd")
        if need:
            ctx.save_for_backward(idx)
            ctx.shape = (n, c, h, w)
        return y

    @staticmethod
    def backward(ctx, gy):
        (idx,) = ctx.saved_tensors
        n, c, h, w = ctx.shape
        gy = _cl(gy)
        gx = torch.empty([n, c, h, w], dtype=gy.dtype, device=gy.device, memory_format=torch.channels_last)
        with torch.cuda.device(gy.device):
            st = _lib.load().vcn_maxpool3x3s2_bwd(_lib.ptr(gy), _lib.ptr(idx), _lib.ptr(gx), _lib.dtype_code(gy.dtype), n, h, w, c,
                                                  _lib.current_stream(gy.device))
        _lib.check(st, "vcn_maxpool3x3s2_bwd")
        return gx


def maxpool3x3s2(x):
    """MaxPool2d(3, 2, 1) of a 16-bit channels_last tensor; differentiable."""
    assert x.is_contiguous(memory_format=torch.channels_last) and x.dtype in (torch.bfloat16, torch.float16) and x.shape[1] % 8 == 0
    if torch.is_grad_enabled() and x.requires_grad:
        return _MaxPool3x3s2.apply(x)
    return _MaxPool3x3s2.forward(None, x)


# ---- convolution + bias + ReLU (+ residual), input gradient only ----------------------------------------------------------


def _conv_fwd(x, w, b, stride, padding, relu=True, residual=None, impl="cudnn"):
    """y = [relu](conv(x, w) + b [+ residual]) in ONE launch.  x, w, residual bf16 channels_last; b bf16."""
    if impl == "vcn" and relu and residual is None and w.shape[-1] == 3 and stride == (1, 1) and padding == (1, 1) and x.shape[1] % 8 == 0:
        from . import modconv as M
        xh = x.permute(0, 2, 3, 1)
        wt, _ = M._packed(w, False, False, False)
        out = torch.empty([x.shape[0], w.shape[0], x.shape[2], x.shape[3]], dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        # relu = lrelu with alpha 0: the compile-time lrelu epilogue of the GEMM (the generic activation switch is slower)
        epi, hold = M.make_epilogue(bias=b, act="lrelu", alpha=0.0, gain=1.0)
        M.conv_igemm(xh, wt, _lib.CONV_FPROP, w.shape[0], 3, 3, out=out, epilogue=epi)
        del hold
        return out
    if (first_layer_impl == "direct" and residual is None and x.shape[1] == 3 and w.shape[0] == 64 and x.dtype == torch.bfloat16
            and ((w.shape[-1] == 3 and stride == (1, 1) and padding == (1, 1)) or (w.shape[-1] == 7 and stride == (2, 2) and padding == (3, 3)))
            and x.is_contiguous(memory_format=torch.channels_last)):
        return conv_first3(x, w, b, int(w.shape[-1]), int(stride[0]), int(padding[0]), relu)
    if (first_layer_impl == "vcn" and relu and residual is None and x.shape[1] < 8 and w.shape[-1] == 3 and stride == (1, 1)
            and padding == (1, 1) and w.shape[0] % 64 == 0 and x.dtype == torch.bfloat16):
        # 3-channel image -> 64 channels (VGG-16 conv1_1): cuDNN pads the input to 8 channels and runs a legacy sm80 kernel
        # at 160-180 us for [16,3,256,256]; the implicit GEMM on an 8-channel NHWC copy of the image is write-bound
        from . import modconv as M
        xh = M.modulate_pack(x, None)  # bf16 NHWC [N,H,W,8], pad channels zero
        wt, _ = M._packed(w, False, False, False)
        out = torch.empty([x.shape[0], w.shape[0], x.shape[2], x.shape[3]], dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        epi, hold = M.make_epilogue(bias=b, act="lrelu", alpha=0.0, gain=1.0)
        M.conv_igemm(xh, wt, _lib.CONV_FPROP, w.shape[0], 3, 3, out=out, epilogue=epi)
        del hold
        return out
    if relu and residual is None:
        return torch.cudnn_convolution_relu(x, w, b, stride, padding, (1, 1), 1)
    if relu:
        return torch.cudnn_convolution_add_relu(x, w, residual, 1.0, b, stride, padding, (1, 1), 1)
    y = F.conv2d(x, w, b, stride, padding)
    return y if residual is None else y + residual


def _conv_dgrad(dz, x_shape, x_like, w, stride, padding, impl="cudnn"):
    """Input gradient of conv(x, w) given d(output) = dz (frozen weights: no dw, no db)."""
    if impl == "vcn" and w.shape[-1] == 3 and stride == (1, 1) and padding == (1, 1) and w.shape[0] % 8 == 0 and w.shape[1] % 8 == 0:
        from . import modconv as M
        wt_t, _ = M._packed(w, True, False, False)
        out = torch.empty(x_shape, dtype=dz.dtype, device=dz.device, memory_format=torch.channels_last)
        M.conv_igemm(dz.permute(0, 2, 3, 1), wt_t, _lib.CONV_DGRAD, w.shape[1], 3, 3, out=out)
        return out
    return torch.ops.aten.convolution_backward(dz, x_like, w, None, list(stride), list(padding), [1, 1], False, [0, 0], 1,
                                               [True, False, False])[0]


class _ConvReLU(torch.autograd.Function):
    """One frozen convolution + bias + [residual] + [ReLU]; backward = ReLU mask (this package's kernel) + library dgrad."""

    @staticmethod
    def forward(ctx, x, w, b, residual, stride, padding, relu, fork=False):
        y = _conv_fwd(x, w, b, stride, padding, relu=relu, residual=residual)
        ctx.cfg = (stride, padding, relu, residual is not None, fork)
        ctx.save_for_backward(x, w, y if relu else None)
        if fork:
            # two handles on the same activation (the next bottleneck reads it as convolution input AND as residual): their
            # gradients come back separately and are summed inside the ReLU-mask pass instead of by a torch add kernel
            return y, y.view_as(y)
        return y

    @staticmethod
    def backward(ctx, dy, dy2=None):
        x, w, y = ctx.saved_tensors
        stride, padding, relu, has_res, fork = ctx.cfg
        prep = lambda t: None if t is None else (_cl(t) if t.dtype == x.dtype else _cl(t.to(x.dtype)))
        dy, dy2 = prep(dy), prep(dy2)
        if dy is None:
            dy, dy2 = dy2, None
        if relu:
            dz = relu_mask_bwd(y, dy, dy2)
        else:
            dz = dy if dy2 is None else dy + dy2
        dx = _conv_dgrad(dz, x.shape, x, w, stride, padding) if ctx.needs_input_grad[0] else None
        dres = dz if (has_res and ctx.needs_input_grad[3]) else None
        return dx, None, None, dres, None, None, None, None


_bias_sums = {}


def _summed_bias(p, q):
    """bf16(p + q) of two frozen bias vectors, cached per parameter versions."""
    key = (id(p), id(q))
    ver = (p._version, q._version, p.data_ptr(), q.data_ptr())
    hit = _bias_sums.get(key)
    if hit is None or hit[0] != ver:
        if p.is_cuda and torch.cuda.is_current_stream_capturing():
            raise _lib.VcnError("frozen_nets: bias cache miss inside a CUDA graph capture (run one eager step first)")
        with torch.no_grad():
            hit = (ver, (p.detach().float() + q.detach().float()).to(torch.bfloat16), p, q)
        _bias_sums[key] = hit
    return hit[1]


def conv_relu(x, conv, relu=True, residual=None, no_bias=False, add_bias_of=None, fork=False):
    """`conv`: a frozen torch.nn.Conv2d with bias (BatchNorm already folded).  Differentiable w.r.t. x / residual only.
    no_bias: leave the bias out (it is added by the consumer); add_bias_of: another frozen conv whose bias is added too;
    fork: return the output twice (see _ConvReLU.forward)."""
    w = _weights.get(conv.weight)
    b = None if no_bias else (_weights.get(conv.bias) if add_bias_of is None else _summed_bias(conv.bias, add_bias_of.bias))
    stride, padding = tuple(conv.stride), tuple(conv.padding)
    if torch.is_grad_enabled() and (x.requires_grad or (residual is not None and residual.requires_grad)):
        return _ConvReLU.apply(x, w, b, residual, stride, padding, relu, fork)
    y = _conv_fwd(x, w, b, stride, padding, relu=relu, residual=residual)
    return (y, y) if fork else y


# ---- ResNet-50 (HF ResNetModel after fold_batchnorm) ------------------------------------------------------------------------


class FusedResNet:
    """Runs ``Classifier.forward`` (classifiers.py:83-92) over the folded HF modules: logits (fp32) and the tuple of hidden
    states (bf16 channels_last), as ``ResNetModel(..., output_hidden_states=True)`` returns them."""

    def __init__(self, classifier):
        self.m = classifier
        r = classifier.resnet
        emb = r.embedder.embedder
        if not (isinstance(emb.normalization, torch.nn.Identity) and emb.convolution.bias is not None):
            raise _lib.VcnError("FusedResNet needs a BatchNorm-folded classifier (viscoin_models.fold_batchnorm)")

    def _bottleneck(self, layer, x, x_res=None, fork=False):
        """x: the block input as the convolutions read it; x_res: the same activation as the identity shortcut reads it (the
        second handle of the previous block's forked output); fork: return this block's output twice."""
        x_res = x if x_res is None else x_res
        convs = list(layer.layer)
        if len(convs) != 3:
            raise _lib.VcnError("FusedResNet: only bottleneck layers are implemented")
        sc = layer.shortcut
        proj = not isinstance(sc, torch.nn.Identity)
        # projection shortcut: its output only ever meets the third convolution's fused (+ residual, + bias, ReLU) epilogue,
        # so its bias rides along there (b3 + b_sc, summed once in fp32) instead of costing a separate bias-add pass per call
        res = conv_relu(x_res, sc.convolution, relu=False, no_bias=True) if proj else x_res
        h = conv_relu(x, convs[0].convolution)
        h = conv_relu(h, convs[1].convolution)
        return conv_relu(h, convs[2].convolution, relu=True, residual=res,  # += residual, then the block's ReLU
                         add_bias_of=sc.convolution if proj else None, fork=fork)

    def __call__(self, images):
        r = self.m.resnet
        if fused_image_ops and _image_ops.supported(images):
            x = _image_ops.image_prep(images)  # fp32 NCHW -> bf16 channels_last in one pass (and one pass back)
        else:
            x = _cl(images.to(torch.bfloat16))
        x = conv_relu(x, r.embedder.embedder.convolution)
        pool = r.embedder.pooler
        if (stem_pool_impl == "vcn" and isinstance(pool, torch.nn.MaxPool2d) and pool.kernel_size in (3, (3, 3)) and pool.stride in (2, (2, 2))
                and pool.padding in (1, (1, 1)) and not pool.ceil_mode and x.shape[1] % 8 == 0):
            x = maxpool3x3s2(x)
        else:
            x = pool(x)
        hidden = []
        # every bottleneck output but the last is read twice by the next block (convolution input + shortcut): forked outputs
        # let the two gradients meet inside the ReLU-mask kernel (frozen_ops.cu, HAS2) instead of a torch add pass each
        layers = [(si, layer) for si, stage in enumerate(r.encoder.stages) for layer in stage.layers]
        x_res, last_stage = None, -1
        for k, (si, layer) in enumerate(layers):
            if si != last_stage:
                hidden.append(x)
                last_stage = si
            fork = resnet_fork and k + 1 < len(layers) and x.requires_grad
            out = self._bottleneck(layer, x, x_res, fork=fork)
            x, x_res = out if fork else (out, None)
        hidden.append(x)
        pooled = x.float().mean(dim=[2, 3])
        lin = self.m.linear
        logits = F.linear(pooled, lin.weight.float(), lin.bias.float())
        return logits, tuple(hidden)


# ---- VGG-16 feature taps ---------------------------------------------------------------------------------------------------


class _VggTaps(torch.autograd.Function):
    """img (bf16 channels_last, already shifted / scaled) -> the 5 taps relu1_2, relu2_2, relu3_3, relu4_3, relu5_3.
    Saves the 13 post-ReLU activations; backward walks the stack with ONE elementwise pass per layer."""

    @staticmethod
    def forward(ctx, x, plan, impl, *wb):
        acts, taps, inputs = [], [], []
        k = 0
        for op in plan:  # 'c' = conv + bias + relu, 'p' = 2x2 max pool, 't' = tap
            if op == "c":
                inputs.append(x)
                x = _conv_fwd(x, wb[2 * k], wb[2 * k + 1], (1, 1), (1, 1), impl=impl if x.shape[1] % 8 == 0 else "cudnn")
                acts.append(x)
                k += 1
            elif op == "p":
                x = maxpool2x2(x)
            else:
                taps.append(x)
        ctx.plan, ctx.impl, ctx.n_layers = plan, impl, k
        ctx.save_for_backward(*inputs, *acts, *wb)
        return tuple(taps)

    @staticmethod
    def backward(ctx, *dtaps):
        L = ctx.n_layers
        saved = ctx.saved_tensors
        inputs, acts, wb = saved[:L], saved[L:2 * L], saved[2 * L:]
        plan = ctx.plan
        g = None          # gradient w.r.t. the tensor flowing out of the current position (None = zero)
        pooled = False    # g is at pooled resolution and must go through the pool backward of the layer below
        k, t = L, len(dtaps)
        dtap = None
        for op in reversed(plan):
            if op == "t":
                t -= 1
                dtap = dtaps[t]
                if dtap is not None:
                    dtap = _cl(dtap.to(acts[0].dtype))
            elif op == "p":
                pooled = True
            else:
                k -= 1
                y = acts[k]
                if pooled:
                    if g is None:
                        dz = relu_mask_bwd(y, dtap) if dtap is not None else None
                    else:
                        dz = maxpool2x2_relu_bwd(y, g, dtap)
                elif g is None:
                    dz = relu_mask_bwd(y, dtap) if dtap is not None else None
                else:
                    dz = relu_mask_bwd(y, g, dtap)
                pooled, dtap = False, None
                if dz is None:
                    g = None
                    continue
                w = wb[2 * k]
                if k == 0 and not ctx.needs_input_grad[0]:
                    g = None
                    break
                g = _conv_dgrad(dz, inputs[k].shape, inputs[k], w, (1, 1), (1, 1), impl=ctx.impl)
        return (g, None, None) + (None,) * len(wb)


class FusedVGGFeatures:
    """The LPIPS feature extractor over ``LPIPS.slices`` (torchvision vgg16.features split at the five taps)."""

    def __init__(self, lpips, impl="cudnn"):
        self.lpips, self.impl = lpips, impl
        plan, convs = [], []
        for s in lpips.slices:
            for layer in s:
                if isinstance(layer, torch.nn.Conv2d):
                    assert tuple(layer.kernel_size) == (3, 3) and tuple(layer.stride) == (1, 1) and tuple(layer.padding) == (1, 1)
                    plan.append("c")
                    convs.append(layer)
                elif isinstance(layer, torch.nn.MaxPool2d):
                    plan.append("p")
                elif not isinstance(layer, (torch.nn.ReLU, torch.nn.Identity)):
                    raise _lib.VcnError("FusedVGGFeatures: unexpected layer %r" % (layer,))
            plan.append("t")
        self.plan, self.convs = tuple(plan), convs

    def _shift_scale(self):
        lp = self.lpips
        key = (lp.shift._version, lp.scale._version, lp.shift.data_ptr())
        if getattr(self, "_ss", None) is None or self._ss[0] != key:
            self._ss = (key, lp.shift.detach().reshape(3).float().contiguous(), (1.0 / lp.scale.detach().reshape(3).float()).contiguous())
        return self._ss[1], self._ss[2]

    def __call__(self, x):
        """x: fp32 / bf16 [N,3,H,W] image batch -> list of 5 bf16 channels_last feature maps."""
        lp = self.lpips
        if fused_image_ops and _image_ops.supported(x):
            x = _image_ops.image_prep(x, self._shift_scale()[0], self._shift_scale()[1])
        else:
            x = _cl(((x.float() - lp.shift) / lp.scale).to(torch.bfloat16))
        wb = []
        for c in self.convs:
            wb += [_weights.get(c.weight), _weights.get(c.bias)]
        if torch.is_grad_enabled() and x.requires_grad:
            return list(_VggTaps.apply(x, self.plan, self.impl, *wb))
        with torch.no_grad():
            return list(_VggTaps.forward(_NoCtx(), x, self.plan, self.impl, *wb))


class _NoCtx:
    """Stand-in ctx for running _VggTaps.forward without autograd."""

    def save_for_backward(self, *a):
        pass
