"""GPU parity: modulated_conv2d / fused SynthesisLayer body on the tcgen05 implicit GEMM vs the oracle
(oracle/ref_ops.py = the reference's impl='ref' path) and the golden vectors recorded from the reference.

Stated tolerances (rel-L-inf = max|a-b| / max|b|):
  * operands exactly representable in bf16 (isolates the GEMM machinery, fp32 accumulation): <= 2e-5
  * fp32 inputs, bf16 tensor-core operands (x*s and W are rounded to bf16, 2^-9 relative each): forward <= 1e-2,
    gradients (dx, dstyles, dbias) <= 2e-2.  SURVEY 8(c) gate for bf16 is 2e-2.
"""

import math

import pytest
import torch

from oracle import ref_ops as R

pytestmark = pytest.mark.gpu

TOL_FWD, TOL_GRAD, TOL_EXACT = 1e-2, 2e-2, 2e-5


def _M():
    import viscoin_next_b200 as V
    return V.modconv


def _rel(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()


def _keys(g):
    return sorted({k.rsplit(".", 1)[0] for k in g.keys if k.startswith("mod.")})


TOL_TF32_FWD, TOL_TF32_GRAD = 3e-3, 5e-3  # fp32 I/O on TF32 tensor cores


@pytest.fixture(params=["tf32", "bf16"])
def operand_mode(request):
    """fp32 activations through the API: 'tf32' = fp32 operands on tcgen05 kind::tf32 (the default, the reference's own GPU
    arithmetic), 'bf16' = operands rounded to bf16 (round 1's only mode).  Yields (mode, forward gate, gradient gate)."""
    M = _M()
    old = M.fp32_operands
    M.fp32_operands = request.param
    yield (request.param,) + ((TOL_TF32_FWD, TOL_TF32_GRAD) if request.param == "tf32" else (TOL_FWD, TOL_GRAD))
    M.fp32_operands = old


def test_golden_modconv(golden, operand_mode):
    """The 36 reference-recorded cases (up 1/2, k 3/1, demod on/off, noise none/const/random, fused flag both), in both
    operand precisions: TF32 operands fwd <= 3e-3 / grads <= 5e-3, bf16 operands 1e-2 / 2e-2."""
    M = _M()
    mode, TOL_FWD, TOL_GRAD = operand_mode
    g = golden("conv_modconv")
    f4 = g.t("f4", device="cuda")
    worst = {}
    for key in _keys(g):
        body = key[len("mod."):]
        up, k, demod, fused = int(body[2]), int(body[4]), bool(int(body[6])), bool(int(body[-1]))
        x, w, s = (g.t(f"{key}.{n}", device="cuda").requires_grad_(n != "w") for n in ("x", "w", "s"))
        noise = g.t(key + ".noise", device="cuda")
        y = M.modulated_conv2d(x, w, s, noise=noise, up=up, padding=k // 2, resample_filter=f4, demodulate=demod,
                               flip_weight=(up == 1), fused_modconv=fused)
        dx, ds = torch.autograd.grad(y, [x, s], g.t(key + ".dy", device="cuda"))
        e = (_rel(y, g.t(key + ".y")), _rel(dx, g.t(key + ".dx")), _rel(ds, g.t(key + ".ds")))
        worst[key] = e
        assert e[0] <= TOL_FWD and e[1] <= TOL_GRAD and e[2] <= TOL_GRAD, (key, e)
    print("[%s operands] worst fwd/dx/ds:" % mode, [max(v[i] for v in worst.values()) for i in range(3)])


CASES = [
    # n, cin, cout, res, k, up, demod, noise, act
    (2, 64, 64, 16, 3, 1, True, "rand", None),
    (3, 128, 64, 8, 3, 2, True, "const", None),
    (2, 64, 3, 32, 1, 1, False, None, None),          # ToRGB shape
    (2, 96, 40, 12, 3, 1, True, "rand", "lrelu"),     # odd sizes, fused layer
    (1, 72, 136, 9, 3, 2, True, "rand", "lrelu"),
    (4, 512, 512, 4, 3, 1, True, "const", "lrelu"),   # b4.conv1
    (2, 256, 128, 32, 3, 2, True, "rand", "lrelu"),
    (2, 128, 3, 16, 1, 1, False, None, "linear"),
    (2, 64, 64, 8, 3, 1, True, None, "sigmoid"),      # not fusable: second kernel
    (1, 192, 160, 8, 3, 1, True, "rand", "lrelu"),    # few M tiles: several narrow out-channel tiles per pixel tile
]


@pytest.mark.parametrize("case", CASES, ids=[str(c) for c in CASES])
def test_layer_vs_oracle(case):
    M = _M()
    n, cin, cout, res, k, up, demod, noise_kind, act = case
    g = torch.Generator().manual_seed(sum(v for v in case[:6]))
    x = torch.randn(n, cin, res, res, generator=g).requires_grad_(True)
    w = torch.randn(cout, cin, k, k, generator=g)
    s = (torch.randn(n, cin, generator=g) * 0.5 + 1).requires_grad_(True)
    b = (torch.randn(cout, generator=g) * 0.3).requires_grad_(True)
    ro = res * up
    noise = None
    if noise_kind == "rand":
        noise = torch.randn(n, 1, ro, ro, generator=g) * 0.4
    elif noise_kind == "const":
        noise = torch.randn(ro, ro, generator=g) * 0.4
    f4 = R.setup_filter([1, 3, 3, 1])
    clamp = 2.5 if act == "lrelu" else None
    sm = s * (1 / math.sqrt(cin)) if not demod else s  # ToRGB weight_gain keeps magnitudes O(1)
    y_ref = R.modulated_conv2d(x, w, sm, noise=noise, up=up, padding=k // 2, resample_filter=f4, demodulate=demod,
                               flip_weight=(up == 1), fused_modconv=False)
    if act is not None:
        y_ref = R.bias_act(y_ref, b, act=act, clamp=clamp)
    dy = torch.randn(y_ref.shape, generator=g)
    want = [x, s] + ([b] if act is not None else [])
    grads_ref = torch.autograd.grad(y_ref, want, dy)

    xc = x.detach().cuda().requires_grad_(True)
    sc = s.detach().cuda().requires_grad_(True)
    bc = b.detach().cuda().requires_grad_(True)
    smc = sc * (1 / math.sqrt(cin)) if not demod else sc
    nc = noise.cuda() if noise is not None else None
    if act is None:
        y = M.modulated_conv2d(xc, w.cuda(), smc, noise=nc, up=up, padding=k // 2, resample_filter=f4.cuda(),
                               demodulate=demod, flip_weight=(up == 1))
    else:
        y = M.modconv_layer(xc, w.cuda(), smc, noise=nc, bias=bc, up=up, padding=k // 2, resample_filter=f4.cuda(),
                            demodulate=demod, flip_weight=(up == 1), act=act, clamp=clamp)
    grads = torch.autograd.grad(y, [xc, sc] + ([bc] if act is not None else []), dy.cuda())
    assert _rel(y, y_ref) <= TOL_FWD
    if act == "lrelu":
        # lrelu / clamp are not smooth: pre-activations inside the forward rounding band take the other branch, and a
        # flipped branch changes that element's gradient by a factor 5 (or to 0 at the clamp).  The activation pattern
        # is pinned through the forward tolerance above; the gradients are compared AT THE PATTERN OF THE CUDA FORWARD,
        # i.e. against the oracle with the piecewise-linear activation frozen to the branches the kernel took.
        yc = y.detach().cpu()
        slope = torch.where(yc > 0, 1.0, 0.2) * math.sqrt(2) * (yc.abs() < clamp)
        z_ref = R.modulated_conv2d(x, w, sm, noise=noise, up=up, padding=k // 2, resample_filter=f4, demodulate=demod,
                                   flip_weight=(up == 1), fused_modconv=False) + b.reshape(1, -1, 1, 1)
        flipped = ((z_ref.detach() > 0) != (yc > 0)).float().mean().item()
        assert flipped < 0.01, flipped
        grads_ref = torch.autograd.grad(z_ref * slope, want, dy)
    for gname, a, r in zip(("dx", "ds", "db"), grads, grads_ref):
        assert _rel(a, r) <= TOL_GRAD, (gname, _rel(a, r))


@pytest.mark.parametrize("up", [1, 2])
def test_exact_on_bf16_representable_operands(up):
    """x, W bf16-representable and styles powers of two: x*s and W are exact in bf16, so only fp32 accumulation
    order separates the tensor-core path from the fp32 oracle."""
    M = _M()
    g = torch.Generator().manual_seed(up)
    n, cin, cout, res = 2, 128, 96, 16
    x = torch.randn(n, cin, res, res, generator=g).bfloat16().float().requires_grad_(True)
    w = (torch.randn(cout, cin, 3, 3, generator=g) / 8).bfloat16().float()
    s = (2.0 ** torch.randint(-2, 3, (n, cin), generator=g)).requires_grad_(True)
    f4 = R.setup_filter([1, 3, 3, 1])
    y_ref = R.modulated_conv2d(x, w, s, up=up, padding=1, resample_filter=f4, demodulate=False, flip_weight=(up == 1),
                               fused_modconv=False)
    xc = x.detach().cuda().requires_grad_(True)
    y = M.modulated_conv2d(xc, w.cuda(), s.detach().cuda(), up=up, padding=1, resample_filter=f4.cuda(),
                           demodulate=False, flip_weight=(up == 1))
    assert _rel(y, y_ref) <= TOL_EXACT
    # dgrad with a bf16-representable upstream gradient
    dy = torch.randn(y_ref.shape, generator=g).bfloat16().float()
    (dx_ref,) = torch.autograd.grad(y_ref, [x], dy)
    (dx,) = torch.autograd.grad(y, [xc], dy.cuda())
    assert _rel(dx, dx_ref) <= (TOL_EXACT if up == 1 else 5e-3)  # up=2: the FIR^T output is re-rounded to bf16


def test_full_size_linearity():
    """Config sizes (N=8, 128->128 at 256^2): additivity in x (exact in the accumulator up to bf16 rounding of the
    operand) and a strided spot check against the oracle on a crop that does not touch the border."""
    M = _M()
    g = torch.Generator(device="cuda").manual_seed(0)
    n, c, r = 8, 128, 256
    x = torch.randn(n, c, r, r, device="cuda", generator=g)
    w = torch.randn(c, c, 3, 3, device="cuda", generator=g)
    s = torch.randn(n, c, device="cuda", generator=g) * 0.3 + 1
    y = M.modulated_conv2d(x, w, s, padding=1)
    assert y.shape == (n, c, r, r) and torch.isfinite(y).all()
    # crop: output rows 100..107, cols 60..75 of sample 3 need input rows 99..108, cols 59..76
    xs = x[3:4, :, 99:109, 59:77].cpu()
    y_ref = R.modulated_conv2d(xs, w.cpu(), s[3:4].cpu(), padding=1, fused_modconv=False)[:, :, 1:-1, 1:-1]
    assert _rel(y[3:4, :, 100:108, 60:76], y_ref) <= TOL_FWD
    y2 = M.modulated_conv2d(2 * x, w, s, padding=1)
    assert _rel(y2, 2 * y) <= 1e-6  # scaling by 2 commutes with bf16 rounding


def test_errors():
    M = _M()
    from viscoin_next_b200 import _lib
    x = torch.zeros(1, 8, 4, 4)
    with pytest.raises(_lib.VcnError):
        M.modulated_conv2d(x, torch.zeros(8, 8, 3, 3), torch.ones(1, 8), padding=1)  # CPU tensor: no fallback
    with pytest.raises(_lib.VcnError):  # down = 3 has no sm_100a counterpart (down = 2 is covered, see the down2 tests)
        M.modulated_conv2d(x.cuda(), torch.zeros(8, 8, 3, 3).cuda(), torch.ones(1, 8).cuda(), down=3, padding=1,
                           resample_filter=torch.ones(4, 4).cuda())
    w = torch.zeros(8, 8, 3, 3, device="cuda", requires_grad=True)
    y = M.modulated_conv2d(x.cuda(), w, torch.ones(1, 8, device="cuda"), padding=1)
    y.sum().backward()  # weight gradient (vcn_conv_wgrad): x = 0 gives exactly 0
    assert w.grad is not None and float(w.grad.abs().max()) == 0.0


RESIDENT_CASES = [
    # n, cin, cout, res, k, up, demod, noise, act        (cin % 8 == 0: the resident pipeline's requirement)
    (2, 64, 64, 16, 3, 1, True, "rand", "lrelu"),
    (3, 128, 64, 8, 3, 2, True, "const", "lrelu"),
    (2, 64, 3, 32, 1, 1, False, None, "linear"),       # ToRGB: leaves the pipeline, fp32 NCHW out
    (1, 72, 136, 9, 3, 2, True, "rand", "lrelu"),
    (2, 256, 128, 32, 3, 2, True, "rand", "lrelu"),
    (4, 512, 512, 4, 3, 1, True, "const", "lrelu"),
    (2, 96, 40, 12, 3, 1, True, "rand", None),
    # 65..128 out channels with >= 256 pixels per sample: the swapped-operand GEMM (M = channels, N = 256 pixels)
    (2, 128, 128, 32, 3, 1, True, "rand", "lrelu"),    # forward and input-gradient GEMM both swapped
    (1, 96, 72, 24, 3, 1, True, "const", "lrelu"),     # partial channel tile, ragged pixel tiles
    (2, 128, 96, 20, 3, 2, True, "rand", "lrelu"),     # transposed conv (4 phases) / stride-2 gather backward
    (2, 80, 128, 17, 1, 1, False, None, "linear"),     # 1x1
]
TOL_RES_FWD, TOL_RES_GRAD = 1.5e-2, 3e-2  # + one bf16 rounding (2^-9) of the stored output / input gradient


@pytest.mark.parametrize("case", RESIDENT_CASES, ids=[str(c) for c in RESIDENT_CASES])
def test_resident_bf16_channels_last_layer_vs_oracle(case):
    """Training-mode resident pipeline: bf16 channels_last activations in, bf16 channels_last out (fp32 NCHW for ToRGB),
    gradients dx (bf16 channels_last), dstyles, dbias - against the fp32 oracle fed the same bf16-rounded input."""
    M = _M()
    n, cin, cout, res, k, up, demod, noise_kind, act = case
    g = torch.Generator().manual_seed(7 + sum(v for v in case[:6]))
    x = torch.randn(n, cin, res, res, generator=g).bfloat16().float().requires_grad_(True)
    w = torch.randn(cout, cin, k, k, generator=g)
    s = (torch.randn(n, cin, generator=g) * 0.5 + 1).requires_grad_(True)
    b = (torch.randn(cout, generator=g) * 0.3).requires_grad_(True)
    ro = res * up
    noise = None
    if noise_kind == "rand":
        noise = torch.randn(n, 1, ro, ro, generator=g) * 0.4
    elif noise_kind == "const":
        noise = torch.randn(ro, ro, generator=g) * 0.4
    f4 = R.setup_filter([1, 3, 3, 1])
    clamp = 2.5 if act == "lrelu" else None
    sm = s * (1 / math.sqrt(cin)) if not demod else s
    z_ref = R.modulated_conv2d(x, w, sm, noise=noise, up=up, padding=k // 2, resample_filter=f4, demodulate=demod,
                               flip_weight=(up == 1), fused_modconv=False)
    y_ref = R.bias_act(z_ref, b, act=act, clamp=clamp) if act is not None else z_ref
    dy = torch.randn(y_ref.shape, generator=g).bfloat16().float()
    want = [x, s] + ([b] if act is not None else [])

    xc = x.detach().cuda().bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    sc = s.detach().cuda().requires_grad_(True)
    bc = b.detach().cuda().requires_grad_(True)
    smc = sc * (1 / math.sqrt(cin)) if not demod else sc
    nc = noise.cuda() if noise is not None else None
    y = M.modconv_layer(xc, w.cuda(), smc, noise=nc, bias=bc if act is not None else None, up=up, padding=k // 2,
                        resample_filter=f4.cuda(), demodulate=demod, flip_weight=(up == 1), act=act, clamp=clamp)
    if cout % 8 == 0:
        assert y.dtype == torch.bfloat16 and y.is_contiguous(memory_format=torch.channels_last)
        dyc = dy.cuda().bfloat16().contiguous(memory_format=torch.channels_last)
    else:
        assert y.dtype == torch.float32 and y.is_contiguous()
        dyc = dy.cuda()
    grads = torch.autograd.grad(y, [xc, sc] + ([bc] if act is not None else []), dyc)
    assert grads[0].dtype == torch.bfloat16 and grads[0].shape == xc.shape
    assert _rel(y, y_ref) <= TOL_RES_FWD, _rel(y, y_ref)
    if act == "lrelu":  # gradients at the activation pattern of the CUDA forward (see test_layer_vs_oracle)
        yc = y.detach().float().cpu()
        slope = torch.where(yc > 0, 1.0, 0.2) * math.sqrt(2) * (yc.abs() < clamp)
        grads_ref = torch.autograd.grad((z_ref + b.reshape(1, -1, 1, 1)) * slope, want, dy)
    else:
        grads_ref = torch.autograd.grad(y_ref, want, dy)
    for gname, a, r in zip(("dx", "ds", "db"), grads, grads_ref):
        assert _rel(a, r) <= TOL_RES_GRAD, (gname, _rel(a, r))


WGRAD_CASES = [
    # n, cin, cout, res, k, up, demod, resident
    (2, 64, 64, 16, 3, 1, True, False),
    (3, 128, 64, 8, 3, 2, True, False),
    (2, 64, 3, 32, 1, 1, False, False),        # ToRGB
    (2, 96, 40, 12, 3, 1, True, True),
    (1, 72, 136, 9, 3, 2, True, True),
    (4, 512, 512, 4, 3, 1, True, True),        # b4.conv1: 2 co tiles, 2 ci tiles
    (2, 256, 128, 32, 3, 2, False, True),
    (2, 320, 200, 20, 3, 1, True, False),      # ragged channel tiles, pixel boxes overhanging the image
]


@pytest.mark.parametrize("case", WGRAD_CASES, ids=[str(c) for c in WGRAD_CASES])
def test_weight_gradient_vs_oracle(case):
    """dL/dW of modulated_conv2d (config 3 of SURVEY 8(d): Generator training) on the MN-major tcgen05 wgrad GEMM, including
    the path through the demodulation coefficients, against autograd of the fp32 oracle.  bf16 operands: <= 2e-2."""
    M = _M()
    n, cin, cout, res, k, up, demod, resident = case
    g = torch.Generator().manual_seed(31 + sum(int(v) for v in case))
    x = torch.randn(n, cin, res, res, generator=g).bfloat16().float()
    w = torch.randn(cout, cin, k, k, generator=g).requires_grad_(True)
    s = torch.randn(n, cin, generator=g) * 0.5 + 1
    ro = res * up
    noise = torch.randn(n, 1, ro, ro, generator=g) * 0.4
    f4 = R.setup_filter([1, 3, 3, 1])
    sm = s * (1 / math.sqrt(cin)) if not demod else s
    y_ref = R.modulated_conv2d(x, w, sm, noise=noise, up=up, padding=k // 2, resample_filter=f4, demodulate=demod,
                               flip_weight=(up == 1), fused_modconv=False)
    dy = torch.randn(y_ref.shape, generator=g).bfloat16().float()
    (dw_ref,) = torch.autograd.grad(y_ref, [w], dy)

    xc = x.cuda()
    if resident:
        xc = xc.bfloat16().contiguous(memory_format=torch.channels_last)
    wc = w.detach().cuda().requires_grad_(True)
    y = M.modulated_conv2d(xc, wc, sm.cuda(), noise=noise.cuda(), up=up, padding=k // 2, resample_filter=f4.cuda(),
                           demodulate=demod, flip_weight=(up == 1))
    dyc = dy.cuda().to(y.dtype)
    if y.stride(1) == 1:
        dyc = dyc.contiguous(memory_format=torch.channels_last)
    (dw,) = torch.autograd.grad(y, [wc], dyc)
    assert dw.shape == w.shape and dw.dtype == torch.float32
    e = _rel(dw, dw_ref)
    assert e <= TOL_GRAD, e


def test_conv2d_resample_golden(golden):
    """conv2d_resample(impl='cuda') - every branch networks.py uses (plain 3x3 / 1x1, up=2 3x3 = transposed conv + FIR, up=2 1x1,
    down=2 3x3 = FIR + stride-2 conv, down=2 1x1) - against the outputs and gradients recorded from the reference
    (conv2d_resample.py:59-154, CONV_CASES).  bf16 tensor-core operands: y <= 1e-2, dx / dw <= 2e-2 (rel-L-inf).  The grouped case
    (groups = N in the fused-modconv pattern) runs on the per-sample-weight GEMM; asymmetric padding is not on the sm_100a
    path: loud VcnError, no library fallback."""
    import viscoin_next_b200 as V
    from oracle.make_golden import CONV_CASES
    g = golden("conv_modconv")
    f4 = g.t("f4", device="cuda")
    worst = {}
    for name, (xs, ws, kw, use_f) in CONV_CASES.items():
        x = g.t(f"conv.{name}.x", device="cuda").requires_grad_(True)
        w = g.t(f"conv.{name}.w", device="cuda").requires_grad_(True)
        if name == "asym_pad":
            with pytest.raises(V._lib.VcnError):
                V.conv2d_resample.conv2d_resample(x, w, f=(f4 if use_f else None), **kw)
            continue
        y = V.conv2d_resample.conv2d_resample(x, w, f=(f4 if use_f else None), **kw)
        dx, dw = torch.autograd.grad(y, [x, w], g.t(f"conv.{name}.dy", device="cuda"))
        e = (_rel(y, g.t(f"conv.{name}.y")), _rel(dx, g.t(f"conv.{name}.dx")), _rel(dw, g.t(f"conv.{name}.dw")))
        worst[name] = e
        assert e[0] <= TOL_FWD and e[1] <= TOL_GRAD and e[2] <= TOL_GRAD, (name, e)
    print("conv2d_resample worst y/dx/dw:", {k: tuple("%.1e" % v for v in e) for k, e in worst.items()})


@pytest.mark.parametrize("n,cin,cout,res", [(2, 64, 128, 32), (1, 24, 40, 18)])
def test_conv2d_resample_down2_vs_oracle(n, cin, cout, res):
    """DiscriminatorBlock-shaped down=2 3x3 convolution (networks.py:571 path: FIR, then the stride-2 tcgen05 GEMM; dx through the
    transposed-conv kind, dw through the MN-major weight-gradient GEMM with exchanged operand roles) vs the fp32 oracle."""
    import viscoin_next_b200 as V
    from oracle import ref_ops as R
    g = torch.Generator().manual_seed(n * 100 + cin)
    x = torch.randn(n, cin, res, res, generator=g).requires_grad_(True)
    w = (torch.randn(cout, cin, 3, 3, generator=g) / (3 * cin ** 0.5)).requires_grad_(True)
    f = R.setup_filter([1, 3, 3, 1])
    for flip_weight in (True, False):
        y_ref = R.conv2d_resample(x, w, f=f, down=2, padding=1, flip_weight=flip_weight)
        dy = torch.randn(y_ref.shape, generator=g)
        dx_ref, dw_ref = torch.autograd.grad(y_ref, [x, w], dy)
        xc, wc = x.detach().cuda().requires_grad_(True), w.detach().cuda().requires_grad_(True)
        y = V.conv2d_resample.conv2d_resample(xc, wc, f=f.cuda(), down=2, padding=1, flip_weight=flip_weight)
        dx, dw = torch.autograd.grad(y, [xc, wc], dy.cuda())
        e = (_rel(y, y_ref), _rel(dx, dx_ref), _rel(dw, dw_ref))
        assert e[0] <= TOL_FWD and e[1] <= TOL_GRAD and e[2] <= TOL_GRAD, (flip_weight, e)


@pytest.mark.parametrize("demod,use_noise", [(True, True), (True, False), (False, True)])
def test_modulated_conv2d_down2_vs_oracle(demod, use_noise):
    """modulated_conv2d(down=2) (networks.py:27-84 with down>1; not used by the synthesis network, kept for API parity):
    x*s -> FIR + stride-2 GEMM -> fma(dcoefs, noise), forward and dx / dstyles against the fp32 oracle."""
    import viscoin_next_b200 as V
    from oracle import ref_ops as R
    g = torch.Generator().manual_seed(5)
    n, cin, cout, res = 2, 32, 48, 16
    x = torch.randn(n, cin, res, res, generator=g).requires_grad_(True)
    w = torch.randn(cout, cin, 3, 3, generator=g)
    s = (torch.randn(n, cin, generator=g) * 0.5 + 1).requires_grad_(True)
    noise = torch.randn(n, 1, res // 2, res // 2, generator=g) * 0.2 if use_noise else None
    f = R.setup_filter([1, 3, 3, 1])
    y_ref = R.modulated_conv2d(x, w, s, noise=noise, down=2, padding=1, resample_filter=f, demodulate=demod, fused_modconv=False)
    dy = torch.randn(y_ref.shape, generator=g)
    dx_ref, ds_ref = torch.autograd.grad(y_ref, [x, s], dy)
    xc, sc = x.detach().cuda().requires_grad_(True), s.detach().cuda().requires_grad_(True)
    y = V.modconv.modulated_conv2d(xc, w.cuda(), sc, noise=noise.cuda() if use_noise else None, down=2, padding=1,
                                   resample_filter=f.cuda(), demodulate=demod)
    dx, ds = torch.autograd.grad(y, [xc, sc], dy.cuda())
    e = (_rel(y, y_ref), _rel(dx, dx_ref), _rel(ds, ds_ref))
    assert e[0] <= TOL_FWD and e[1] <= TOL_GRAD and e[2] <= TOL_GRAD, e


@pytest.mark.parametrize("shape", [(3, 128, 37, 21), (2, 512, 4, 4), (2, 64, 16, 16), (2, 24, 9, 9), (1, 2048, 3, 5)])
@pytest.mark.parametrize("with_styles", [True, False])
def test_modulate_pack_channels_last_exact(shape, with_styles):
    """vcn_modulate_pack on a resident bf16 channels_last activation (rows kernel for power-of-two channel counts, generic
    kernel otherwise): xhat[n,h,w,c] = bf16(x * s[n,c]) - one fp32 multiply and one rounding, bit-exact against torch."""
    M = _M()
    g = torch.Generator().manual_seed(sum(shape))
    n, c, h, w = shape
    x = torch.randn(shape, generator=g).bfloat16().cuda().contiguous(memory_format=torch.channels_last)
    s = (torch.randn(n, c, generator=g) + 1).cuda() if with_styles else None
    got = M.modulate_pack(x, s)
    want = x.float() * (s[:, :, None, None] if with_styles else 1.0)
    want = want.bfloat16().permute(0, 2, 3, 1).contiguous()
    c_pad = (c + 7) // 8 * 8
    assert got.shape == (n, h, w, c_pad)
    assert torch.equal(got[..., :c], want)
    if c_pad > c:
        assert float(got[..., c:].abs().max()) == 0.0


@pytest.mark.parametrize("g,cin,cout,res,k,up,flip_weight", [(3, 64, 48, 16, 3, 1, True), (2, 128, 3, 32, 1, 1, True),
                                                            (2, 64, 64, 8, 3, 2, False), (4, 24, 40, 9, 3, 1, False)])
def test_conv2d_resample_grouped_per_sample(g, cin, cout, res, k, up, flip_weight):
    """conv2d_resample(groups=N) in the call pattern of the reference's FUSED modconv branch (networks.py:78-81:
    x = [1, N*Cin, H, W], w = [N*Cout, Cin, k, k]) -> the per-sample-weight mode of the implicit GEMM; y, dx and dw vs the
    reference formulation in plain torch (grouped F.conv2d / F.conv_transpose2d + FIR)."""
    import viscoin_next_b200 as V
    C = V.conv2d_resample
    gen = torch.Generator().manual_seed(g * 100 + cin + up)
    x = torch.randn(1, g * cin, res, res, generator=gen)
    w = torch.randn(g * cout, cin, k, k, generator=gen) / math.sqrt(cin * k * k)
    f = R.setup_filter([1, 3, 3, 1])
    xr, wr = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    y_ref = C.conv2d_resample(xr, wr, f=f, up=up, padding=k // 2, groups=g, flip_weight=flip_weight, impl="ref")
    dy = torch.randn(y_ref.shape, generator=gen)
    dx_ref, dw_ref = torch.autograd.grad(y_ref, [xr, wr], dy)
    xc, wc = x.cuda().requires_grad_(True), w.cuda().requires_grad_(True)
    y = C.conv2d_resample(xc, wc, f=f.cuda(), up=up, padding=k // 2, groups=g, flip_weight=flip_weight)
    dx, dw = torch.autograd.grad(y, [xc, wc], dy.cuda())
    e = (_rel(y, y_ref), _rel(dx, dx_ref), _rel(dw, dw_ref))
    print("grouped", (g, cin, cout, res, k, up), "y/dx/dw", e)
    assert e[0] <= TOL_FWD and e[1] <= TOL_GRAD and e[2] <= TOL_GRAD, e


def test_torgb_fused_backward_matches_gemm_path():
    """vcn_torgb_bwd_fused (one streaming pass) against the GEMM + finish path it replaces, on the resident bf16 ToRGB layer of
    the 128^2 block: dx and dstyles agree to bf16 rounding (the fused pass keeps W in fp32 and does not round dxhat to bf16,
    so it is the more accurate of the two: both are gated against an fp32 torch formulation)."""
    import viscoin_next_b200 as V
    M = V.modconv
    g = torch.Generator(device="cuda").manual_seed(21)
    n, c, r = 4, 256, 64
    x = torch.randn(n, c, r, r, device="cuda", generator=g).bfloat16().contiguous(memory_format=torch.channels_last)
    w = torch.randn(3, c, 1, 1, device="cuda", generator=g) / c ** 0.5
    b = torch.randn(3, device="cuda", generator=g)
    s0 = torch.randn(n, c, device="cuda", generator=g) + 1
    gy = torch.randn(n, 3, r, r, device="cuda", generator=g)
    outs = {}
    for fused in (True, False):
        M.torgb_fused = fused
        try:
            xi, si = x.clone().requires_grad_(True), s0.clone().requires_grad_(True)
            y = M.modconv_layer(xi, w, si, bias=b, demodulate=False, act="linear", clamp=256.0)
            y.backward(gy.to(y.dtype))
            outs[fused] = (y.detach().float(), xi.grad.float(), si.grad.float())
        finally:
            M.torgb_fused = True
    xr, sr = x.float().requires_grad_(True), s0.clone().requires_grad_(True)
    yr = torch.clamp(torch.nn.functional.conv2d(xr * sr.view(n, c, 1, 1), w, b), -256, 256)
    yr.backward(gy)
    rel = lambda a, ref: float((a - ref).norm() / ref.norm())
    assert torch.equal(outs[True][0], outs[False][0])
    for k, (name, ref) in enumerate((("y", yr.detach()), ("dx", xr.grad), ("dstyles", sr.grad))):
        e_f, e_g = rel(outs[True][k], ref), rel(outs[False][k], ref)
        print("%-8s fused %.2e  gemm %.2e" % (name, e_f, e_g))
        assert e_f <= 1e-2 and e_g <= 1.5e-2, (name, e_f, e_g)


def test_conv_sm_reserve_does_not_change_results():
    """vcn_set_conv_sm_reserve only shrinks the persistent grid: the tiles are the same."""
    import viscoin_next_b200 as V
    from viscoin_next_b200 import _lib
    M = V.modconv
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(4, 256, 32, 32, device="cuda", generator=g).bfloat16().contiguous(memory_format=torch.channels_last)
    w = torch.randn(256, 256, 3, 3, device="cuda", generator=g)
    s = torch.randn(4, 256, device="cuda", generator=g) + 1
    lib = _lib.load()
    y0 = M.modconv_layer(x, w, s)
    try:
        _lib.check(lib.vcn_set_conv_sm_reserve(20), "vcn_set_conv_sm_reserve")
        y1 = M.modconv_layer(x, w, s)
    finally:
        _lib.check(lib.vcn_set_conv_sm_reserve(0), "vcn_set_conv_sm_reserve")
    assert torch.equal(y0, y1)
    assert lib.vcn_set_conv_sm_reserve(1000) != 0


def test_torgb_tap_folds_the_second_gradient():
    """modconv_layer(..., tap=True) returns (rgb, x2): the gradient arriving through x2 is added to the ToRGB input gradient
    inside the fused backward (vcn_torgb_bwd_fused with `acc`).  Against the same graph without the tap (autograd's own add):
    equal to one bf16 rounding of the sum."""
    import viscoin_next_b200 as V
    M = V.modconv
    g = torch.Generator(device="cuda").manual_seed(33)
    n, c, r = 4, 128, 64
    x0 = torch.randn(n, c, r, r, device="cuda", generator=g).bfloat16().contiguous(memory_format=torch.channels_last)
    w = torch.randn(3, c, 1, 1, device="cuda", generator=g) / c ** 0.5
    b = torch.randn(3, device="cuda", generator=g)
    s0 = torch.randn(n, c, device="cuda", generator=g) + 1
    gy = torch.randn(n, 3, r, r, device="cuda", generator=g)
    gx2 = torch.randn(n, c, r, r, device="cuda", generator=g).bfloat16().contiguous(memory_format=torch.channels_last)
    res = {}
    for tap in (True, False):
        x, s = x0.clone().requires_grad_(True), s0.clone().requires_grad_(True)
        if tap:
            y, x2 = M.modconv_layer(x, w, s, bias=b, demodulate=False, act="linear", clamp=256.0, tap=True)
            assert x2.data_ptr() == x.data_ptr()
        else:
            y, x2 = M.modconv_layer(x, w, s, bias=b, demodulate=False, act="linear", clamp=256.0), x
        torch.autograd.backward([y, x2 * 1.0], [gy.to(y.dtype), gx2])
        res[tap] = (y.detach().float(), x.grad.float(), s.grad.float())
    assert torch.equal(res[True][0], res[False][0])
    # the style gradient is a sum of per-CTA partials added with atomics: equal up to the fp32 summation order
    assert float((res[True][2] - res[False][2]).abs().max() / res[False][2].abs().max()) <= 1e-5
    e = float((res[True][1] - res[False][1]).abs().max() / res[False][1].abs().max())
    print("tap vs autograd add: dx rel-Linf %.2e" % e)
    assert e <= 8e-3  # the reference path rounds the ToRGB gradient to bf16 before the add, the fused pass adds in fp32
