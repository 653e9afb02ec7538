"""The frozen networks of the VisCoIN step on the fused bf16 channels_last path (viscoin_next_b200.frozen_nets; SURVEY
8(f) N2): this package's pool / mask kernels bit-exactly against ATen, and FusedResNet / FusedVGGFeatures against the plain
fp32 modules they stand in for (viscoin/models/classifiers.py:83-92; lpips VGG-16, viscoin/training/losses.py:111-123).

Tolerances: the kernels of csrc/frozen_ops.cu select / mask / add bf16 values with one rounding -> bit-exact vs ATen's bf16
ops.  The fused networks compute what bf16 autocast computes (bf16 operands and storage, fp32 accumulation) and are compared
with the fp32 modules: activations rel-Linf <= 3e-2, logits <= 3e-2 (of max |logit|), input gradients rel-L2 <= 1.5e-1."""

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _fz():
    import viscoin_next_b200.host.frozen_nets as fz
    return fz


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _rel2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _relu_like(shape, gen, ties=True):
    """bf16 channels_last tensor like a ReLU output: ~half zeros, plus exact positive ties inside 2x2 windows."""
    y = torch.relu(torch.randn(shape, device="cuda", generator=gen))
    if ties:
        y[:, :, 0::2, 1::2] = torch.where(torch.rand(y[:, :, 0::2, 1::2].shape, device="cuda", generator=gen) < 0.2, y[:, :, 0::2, 0::2], y[:, :, 0::2, 1::2])
        y[:, :, 1::2, 0::2] = torch.where(torch.rand(y[:, :, 1::2, 0::2].shape, device="cuda", generator=gen) < 0.1, y[:, :, 0::2, 1::2], y[:, :, 1::2, 0::2])
    return y.bfloat16().contiguous(memory_format=torch.channels_last)


@pytest.mark.parametrize("shape", [(2, 64, 32, 48), (3, 8, 2, 2), (1, 512, 16, 16), (2, 24, 6, 10)])
def test_maxpool_forward_and_fused_backward_bit_exact(shape):
    fz = _fz()
    gen = torch.Generator(device="cuda").manual_seed(sum(shape))
    y = _relu_like(shape, gen)
    assert torch.equal(fz.maxpool2x2(y), F.max_pool2d(y, 2, 2))
    n, c, h, w = shape
    gpool = torch.randn(n, c, h // 2, w // 2, device="cuda", generator=gen).bfloat16().contiguous(memory_format=torch.channels_last)
    gtap = torch.randn(shape, device="cuda", generator=gen).bfloat16().contiguous(memory_format=torch.channels_last)
    for tap in (None, gtap):
        # ATen: pre-activation z with relu(z) = y; d/dz [ <maxpool(relu(z)), gpool> + <relu(z), gtap> ]
        z = y.clone().requires_grad_(True)
        r = torch.relu(z)
        loss = (F.max_pool2d(r, 2, 2).float() * gpool.float()).sum()
        if tap is not None:
            loss = loss + (r.float() * tap.float()).sum()
        (want,) = torch.autograd.grad(loss, [z])
        got = fz.maxpool2x2_relu_bwd(y, gpool, tap)
        assert torch.equal(got, want), (shape, tap is not None, float((got.float() - want.float()).abs().max()))
    g2 = torch.randn(shape, device="cuda", generator=gen).bfloat16().contiguous(memory_format=torch.channels_last)
    assert torch.equal(fz.relu_mask_bwd(y, gtap), torch.where(y > 0, gtap, torch.zeros_like(gtap)))
    assert torch.equal(fz.relu_mask_bwd(y, gtap, g2), torch.where(y > 0, gtap + g2, torch.zeros_like(gtap)))


@pytest.mark.parametrize("shape", [(2, 64, 32, 48), (3, 8, 7, 5), (1, 16, 1, 9), (2, 64, 128, 128)])
def test_maxpool3x3s2_forward_and_backward_bit_exact(shape):
    """MaxPool2d(3, 2, 1) of the ResNet stem (odd sizes, borders, exact ties): values and input gradient == ATen in bf16."""
    fz = _fz()
    gen = torch.Generator(device="cuda").manual_seed(5)
    x = _relu_like(shape, gen, ties=shape[2] % 2 == 0 and shape[3] % 2 == 0)
    y_ref = F.max_pool2d(x, 3, 2, 1)
    assert torch.equal(fz.maxpool3x3s2(x), y_ref)  # no-grad path (no index tensor)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ya, yb = fz.maxpool3x3s2(xa), F.max_pool2d(xb, 3, 2, 1)
    assert torch.equal(ya, yb) and ya.is_contiguous(memory_format=torch.channels_last)
    gy = torch.randn(yb.shape, device="cuda", generator=gen).bfloat16().contiguous(memory_format=torch.channels_last)
    ya.backward(gy)
    yb.backward(gy)
    assert torch.equal(xa.grad, xb.grad)


@pytest.mark.parametrize("ks,stride,pad,shape", [(3, 1, 1, (2, 3, 40, 56)), (3, 1, 1, (1, 3, 19, 21)), (7, 2, 3, (2, 3, 64, 96)),
                                                 (7, 2, 3, (1, 3, 37, 45)), (3, 1, 1, (4, 3, 256, 256)), (7, 2, 3, (4, 3, 256, 256))])
@pytest.mark.parametrize("relu", [True, False])
def test_conv_first3_matches_fp32_convolution(ks, stride, pad, shape, relu):
    """vcn_conv_first3 (VGG conv1_1 / ResNet stem on mma.sync) vs F.conv2d in fp32 on the same bf16-rounded operands: the only
    differences are the fp32 accumulation order and the final bf16 rounding (2^-9 relative): rel-Linf <= 6e-3."""
    fz = _fz()
    gen = torch.Generator(device="cuda").manual_seed(ks * 100 + shape[2])
    x = torch.randn(shape, device="cuda", generator=gen).bfloat16().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(64, 3, ks, ks, device="cuda", generator=gen) / (ks * 1.7)).bfloat16().contiguous(memory_format=torch.channels_last)
    b = torch.randn(64, device="cuda", generator=gen).bfloat16()
    y = fz.conv_first3(x, w, b, ks, stride, pad, relu)
    ref = F.conv2d(x.float(), w.float(), b.float(), stride, pad)
    if relu:
        ref = torch.relu(ref)
    assert y.shape == ref.shape and y.dtype == torch.bfloat16 and y.is_contiguous(memory_format=torch.channels_last)
    assert _rel(y, ref) <= 6e-3


def test_frozen_ops_reject_bad_arguments():
    import viscoin_next_b200 as V
    fz = _fz()
    x = torch.zeros(1, 8, 3, 4, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    with pytest.raises(V._lib.VcnError):
        fz.maxpool2x2(x)  # odd height
    with pytest.raises(AssertionError):
        fz.maxpool2x2(torch.zeros(1, 8, 4, 4, device="cuda"))  # fp32


def _classifier(conditioned):
    import viscoin_next_b200 as V
    VM = V.viscoin_models
    torch.manual_seed(0)
    cls = VM.Classifier(200).cuda().eval().requires_grad_(False)
    if conditioned:
        VM.condition_random_init(cls)
    g = torch.Generator().manual_seed(1)
    VM.calibrate_batchnorm(cls, torch.randn(8, 3, 256, 256, generator=g).cuda())
    return cls, VM


def test_fused_resnet_blocks_teacher_forced():
    """Every bottleneck of the (HF-default, unconditioned) random-init ResNet-50 on the SAME input as the fp32 module:
    output rel-Linf <= 2e-2 (measured 1.0e-2) and input gradient rel-L2 <= 2e-1 per block (measured 1.2e-1: a bf16 activation
    within rounding of zero flips its ReLU mask, and a fraction p of flipped masks costs sqrt(p) in L2 whatever computes it).
    This isolates the arithmetic of the fused path from the error amplification of the network itself (next test)."""
    fz = _fz()
    cls, VM = _classifier(False)
    VM.fold_batchnorm(cls)
    fused = fz.FusedResNet(cls)
    x = torch.randn(2, 3, 256, 256, device="cuda", generator=torch.Generator(device="cuda").manual_seed(2))
    h = cls.resnet.embedder(x)
    worst = [0.0, 0.0]
    for stage in cls.resnet.encoder.stages:
        for layer in stage.layers:
            hr = h.detach().clone().requires_grad_(True)
            out_ref = layer(hr)
            dy = torch.randn_like(out_ref)
            (g_ref,) = torch.autograd.grad(out_ref, [hr], dy)
            hb = h.detach().bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
            out = fused._bottleneck(layer, hb)
            (g,) = torch.autograd.grad(out, [hb], dy.bfloat16().contiguous(memory_format=torch.channels_last))
            worst = [max(worst[0], _rel(out, out_ref)), max(worst[1], _rel2(g, g_ref))]
            h = out_ref.detach()
    print("FusedResNet per block (same input): out %.2e, dx rel-L2 %.2e" % tuple(worst))
    assert worst[0] <= 2e-2 and worst[1] <= 2e-1


@pytest.mark.parametrize("conditioned", [True, False])
def test_fused_resnet_matches_plain_module(conditioned):
    """Whole network: logits, the five hidden states and the input gradient against the fp32 module, next to the SAME
    comparison for the path the fused one replaces (the plain module under bf16 autocast).  With the harness' conditioned
    initialisation (viscoin_models.condition_random_init) the gates are tight; with HF's default initialisation the
    network amplifies any rounding ~1.3x per block, so the gate is relative: no worse than 1.5x the autocast path."""
    fz = _fz()
    cls, VM = _classifier(conditioned)
    x = torch.randn(4, 3, 256, 256, device="cuda", generator=torch.Generator(device="cuda").manual_seed(2))
    xr = x.clone().requires_grad_(True)
    logits_ref, hidden_ref = cls(xr)  # fp32 modules, BatchNorm in eval mode
    dl = torch.randn_like(logits_ref)
    dh = [torch.randn_like(h) * 0.1 for h in hidden_ref[-3:]]
    (gx_ref,) = torch.autograd.grad([logits_ref] + list(hidden_ref[-3:]), [xr], [dl] + dh)
    xa = x.clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        logits_ac, hidden_ac = cls(xa)
    (gx_ac,) = torch.autograd.grad([logits_ac.float()] + list(hidden_ac[-3:]), [xa], [dl] + [d.to(h.dtype) for d, h in zip(dh, hidden_ac[-3:])])
    ac = (float((logits_ac.float() - logits_ref).abs().max() / logits_ref.abs().max()), max(_rel(a, b) for a, b in zip(hidden_ac, hidden_ref)), _rel2(gx_ac, gx_ref))
    VM.fold_batchnorm(cls)
    fused = fz.FusedResNet(cls)
    xf = x.clone().requires_grad_(True)
    logits, hidden = fused(xf)
    assert len(hidden) == len(hidden_ref) == 5 and logits.dtype == torch.float32
    e_logits = float((logits - logits_ref).abs().max() / logits_ref.abs().max())
    e_hidden = [_rel(a, b) for a, b in zip(hidden, hidden_ref)]
    (gx,) = torch.autograd.grad([logits] + list(hidden[-3:]), [xf], [dl] + [d.to(h.dtype) for d, h in zip(dh, hidden[-3:])])
    e_g = _rel2(gx, gx_ref)
    print("FusedResNet[conditioned=%s]: logits %.2e hidden %s dx rel-L2 %.2e | autocast path: logits %.2e hidden %.2e dx %.2e" % (
        conditioned, e_logits, ["%.1e" % e for e in e_hidden], e_g, ac[0], ac[1], ac[2]))
    if conditioned:
        assert e_logits <= 3e-2 and max(e_hidden) <= 8e-2  # measured 1.2e-2 / 5.5e-2 (autocast path: 1.3e-2 / 5.8e-2)
        assert e_logits <= 1.5 * ac[0] + 1e-2 and max(e_hidden) <= 1.5 * ac[1] + 1e-2 and e_g <= 1.5 * ac[2] + 2e-2
    else:
        # HF-default initialisation: chaotic (two bf16 implementations - or two cuDNN algorithm choices - end up O(1) apart at
        # the last stage); informational beyond the first two hidden states, which the amplification has not reached yet
        assert e_hidden[0] <= 1e-2 and e_hidden[1] <= 5e-2 and all(bool(torch.isfinite(h).all()) for h in hidden)
    with torch.no_grad():  # the no-grad path (f on the constant real + fake images) gives the same numbers
        l2, h2 = fused(x)
    assert torch.equal(l2, logits) and all(torch.equal(a, b) for a, b in zip(h2, hidden))


@pytest.mark.parametrize("impl", ["cudnn", "vcn"])
def test_fused_vgg_taps_and_lpips_gradient(impl):
    import viscoin_next_b200 as V
    fz = _fz()
    torch.manual_seed(3)
    lp = V.viscoin_models.LPIPS().cuda()
    g = torch.Generator(device="cuda").manual_seed(4)
    a = (torch.rand(2, 3, 128, 128, device="cuda", generator=g) * 2 - 1)
    b = (torch.rand(2, 3, 128, 128, device="cuda", generator=g) * 2 - 1)
    ar = a.clone().requires_grad_(True)
    lp.fused = False
    want = lp(ar, b).sum()  # plain torch formulation, fp32
    (ga_ref,) = torch.autograd.grad(want, [ar])
    taps_ref = lp.features(a)
    lp.fused = True
    lp.feature_fn = fz.FusedVGGFeatures(lp, impl=impl)
    taps = lp.features(a)
    e_taps = [_rel(x, y) for x, y in zip(taps, taps_ref)]
    assert all(t.dtype == torch.bfloat16 and t.is_contiguous(memory_format=torch.channels_last) for t in taps)
    af = a.clone().requires_grad_(True)
    got = lp(af, b).sum()
    (ga,) = torch.autograd.grad(got, [af])
    e_v, e_g = abs(float(got) - float(want)) / abs(float(want)), _rel2(ga, ga_ref)
    print("FusedVGG[%s]: taps %s value %.2e d(img) rel-L2 %.2e" % (impl, ["%.1e" % e for e in e_taps], e_v, e_g))
    assert max(e_taps) <= 3e-2 and e_v <= 2e-2 and e_g <= 1.5e-1


def test_image_prep_and_l1l2_match_torch():
    """csrc/image_ops.cu: (x - shift) / scale -> bf16 channels_last and its adjoint; mean |a-b|, mean (a-b)^2 and their gradient."""
    from viscoin_next_b200.host import image_ops as io
    gen = torch.Generator(device="cuda").manual_seed(3)
    x = torch.randn(3, 3, 37, 52, device="cuda", generator=gen).requires_grad_(True)
    shift = torch.tensor([-0.030, -0.088, -0.188], device="cuda")
    scale = torch.tensor([0.458, 0.448, 0.450], device="cuda")
    y = io.image_prep(x, shift, 1.0 / scale)
    ref = ((x.detach() - shift.view(1, 3, 1, 1)) / scale.view(1, 3, 1, 1)).bfloat16()
    assert y.is_contiguous(memory_format=torch.channels_last) and _rel(y, ref) <= 8e-3  # 1 bf16 ulp: (x-s)*(1/c) vs (x-s)/c
    gy = torch.randn(3, 3, 37, 52, device="cuda", generator=gen).bfloat16()
    y.backward(gy)
    assert _rel(x.grad, gy.float() / scale.view(1, 3, 1, 1)) <= 1e-6
    assert torch.equal(io.image_prep(x.detach()), x.detach().bfloat16())
    a = torch.randn(4, 3, 64, 64, device="cuda", generator=gen).requires_grad_(True)
    b = torch.randn(4, 3, 64, 64, device="cuda", generator=gen)
    a2 = a.detach().clone().requires_grad_(True)
    v = io.l1_l2(a, b)
    d = a2 - b
    vr = torch.stack([d.abs().mean(), d.square().mean()])
    assert _rel(v, vr) <= 1e-5
    (v[0] * 0.7 + v[1] * 1.3).backward()
    (vr[0] * 0.7 + vr[1] * 1.3).backward()
    assert _rel(a.grad, a2.grad) <= 1e-5
