"""GPU parity: vcn_upfirdn2d (through the host mirror / C ABI) vs the oracle and the golden vectors recorded
from the reference's impl='ref' path.  Tolerance: fp32 exact-FMA kernel, rel-L-inf <= 1e-5;
bf16/fp16: <= 2 ulp of the storage type against the fp32 oracle on the rounded inputs.
"""

import pytest
import torch

from oracle import ref_ops as R
from oracle.make_golden import UPFIRDN_CASES

pytestmark = pytest.mark.gpu


def _ops():
    import viscoin_next_b200 as V
    return V.upfirdn2d


def _rel(a, b):
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.numel() == 0:
        return 0.0
    return ((a - b).abs().max() / b.abs().max().clamp_min(1.0)).item()


@pytest.mark.parametrize("name", list(UPFIRDN_CASES))
def test_golden(golden, name):
    U = _ops()
    g = golden("upfirdn2d")
    _, _, kw = UPFIRDN_CASES[name]
    x = g.t(name + ".x", device="cuda").requires_grad_(True)
    f = g.t(name + ".f", device="cuda")
    y = U.upfirdn2d(x, f, impl="cuda", **kw)
    assert _rel(y, g.t(name + ".y")) <= 1e-5
    (dx,) = torch.autograd.grad(y, [x], g.t(name + ".dy", device="cuda"))
    assert _rel(dx, g.t(name + ".dx")) <= 1e-5


def test_golden_wrappers(golden):
    U = _ops()
    g = golden("upfirdn2d")
    x, f4 = g.t("wrap.x", device="cuda"), g.t("wrap.f", device="cuda")
    assert _rel(U.filter2d(x, f4), g.t("wrap.filter2d")) <= 1e-5
    assert _rel(U.upsample2d(x, f4), g.t("wrap.upsample2d")) <= 1e-5
    assert _rel(U.downsample2d(x, f4), g.t("wrap.downsample2d")) <= 1e-5
    assert _rel(U.upsample2d(x, f4, up=[2, 1], padding=[1, 0, 0, 2], gain=0.5), g.t("wrap.upsample2d_p")) <= 1e-5
    assert _rel(U.setup_filter([1, 3, 3, 1]), g.t("setup.a")) <= 1e-7
    assert _rel(U.setup_filter(list(range(1, 10)), gain=3), g.t("setup.c")) <= 1e-7


# hot-path patterns of SURVEY 8(a6) at odd and even sizes (A: after transpose conv, B: RGB skip, primes: backward)
PATTERNS = {
    "A": dict(up=1, down=1, padding=[1, 1, 1, 1], gain=4),
    "A_bwd": dict(up=1, down=1, padding=[2, 2, 2, 2], flip_filter=True, gain=4),
    "B": dict(up=2, down=1, padding=[2, 1, 2, 1], gain=4),
    "B_bwd": dict(up=1, down=2, padding=[1, 1, 1, 1], flip_filter=True, gain=4),
    "down2": dict(up=1, down=2, padding=[1, 1, 1, 1]),
    "up2_p0": dict(up=2, down=1, padding=[1, 2, 1, 2]),
    "up2_mixed": dict(up=2, down=1, padding=[3, 0, 2, 1]),
}


@pytest.mark.parametrize("pat", list(PATTERNS))
@pytest.mark.parametrize("shape", [(2, 3, 9, 9), (1, 5, 17, 17), (2, 4, 33, 33), (1, 2, 65, 65), (2, 3, 16, 16),
                                   (1, 3, 64, 64), (1, 2, 129, 129), (1, 1, 6, 70), (3, 2, 40, 12), (1, 3, 128, 128)])
def test_patterns_vs_oracle(pat, shape):
    U = _ops()
    kw = PATTERNS[pat]
    g = torch.Generator().manual_seed(sum(shape) * 7 + len(pat))
    x = torch.randn(shape, generator=g).requires_grad_(True)
    f = R.setup_filter([1, 3, 3, 1])
    y_ref = R.upfirdn2d(x, f, **kw)
    dy = torch.randn(y_ref.shape, generator=g)
    (dx_ref,) = torch.autograd.grad(y_ref, [x], dy)
    xc = x.detach().cuda().requires_grad_(True)
    y = U.upfirdn2d(xc, f.cuda(), impl="cuda", **kw)
    (dx,) = torch.autograd.grad(y, [xc], dy.cuda())
    assert _rel(y, y_ref) <= 1e-5
    assert _rel(dx, dx_ref) <= 1e-5


@pytest.mark.parametrize("fshape", [(1, 1), (2, 2), (3, 3), (4, 4), (5, 5), (8, 8), (3, 5), (1, 4)])
@pytest.mark.parametrize("up,down,pad", [(1, 1, 0), (2, 1, [3, 2, 3, 2]), (1, 2, [2, 2, 1, 3]), (3, 2, [4, 3, 5, 2]),
                                         ([2, 1], [1, 3], [4, 4, 4, 4]), (1, 1, [-1, 5, 6, -2])])
@pytest.mark.parametrize("flip", [False, True])
def test_generic_vs_oracle(fshape, up, down, pad, flip):
    U = _ops()
    g = torch.Generator().manual_seed(fshape[0] * 13 + fshape[1])
    x = torch.randn(2, 3, 13, 11, generator=g)
    f = torch.randn(fshape, generator=g)
    y_ref = R.upfirdn2d(x, f, up=up, down=down, padding=pad, flip_filter=flip, gain=1.3)
    y = U.upfirdn2d(x.cuda(), f.cuda(), up=up, down=down, padding=pad, flip_filter=flip, gain=1.3, impl="cuda")
    assert _rel(y, y_ref) <= 1e-5


def test_separable_and_channels_last():
    U = _ops()
    g = torch.Generator().manual_seed(5)
    x = torch.randn(2, 6, 20, 18, generator=g)
    f12 = R.setup_filter([0.1, -0.3, 0.7, 1.1, 2.0, 3.1, 2.9, 1.9, 1.2, 0.6, -0.2, 0.05])
    assert f12.ndim == 1
    y_ref = R.upfirdn2d(x, f12, up=2, padding=[6, 5, 6, 5], gain=4)
    y = U.upfirdn2d(x.cuda(), f12.cuda(), up=2, padding=[6, 5, 6, 5], gain=4, impl="cuda")
    assert _rel(y, y_ref) <= 1e-5
    f4 = R.setup_filter([1, 3, 3, 1])
    xc = x.cuda().contiguous(memory_format=torch.channels_last)
    y = U.upfirdn2d(xc, f4.cuda(), up=2, padding=[2, 1, 2, 1], gain=4, impl="cuda")
    assert y.is_contiguous(memory_format=torch.channels_last)
    assert _rel(y, R.upfirdn2d(x, f4, up=2, padding=[2, 1, 2, 1], gain=4)) <= 1e-5


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float64])
def test_other_dtypes(dtype):
    U = _ops()
    g = torch.Generator().manual_seed(9)
    x = torch.randn(2, 4, 32, 32, generator=g).to(dtype)
    f4 = R.setup_filter([1, 3, 3, 1])
    for kw in (PATTERNS["A"], PATTERNS["B"], PATTERNS["B_bwd"]):
        y_ref = R.upfirdn2d(x.double(), f4, **kw) if dtype == torch.float64 else R.upfirdn2d(x.float(), f4, **kw)
        y = U.upfirdn2d(x.cuda(), f4.cuda(), impl="cuda", **kw)
        assert y.dtype == dtype
        tol = 1e-6 if dtype == torch.float64 else 2 * torch.finfo(dtype).eps
        assert _rel(y, y_ref) <= tol


def test_errors_and_empty():
    U = _ops()
    from viscoin_next_b200 import _lib
    f4 = R.setup_filter([1, 3, 3, 1])
    with pytest.raises(_lib.VcnError):
        U.upfirdn2d(torch.zeros(1, 1, 4, 4), f4, impl="cuda")
    e = torch.zeros(0, 3, 8, 8, device="cuda")
    assert U.upfirdn2d(e, f4.cuda(), up=2, padding=[2, 1, 2, 1], impl="cuda").shape == (0, 3, 16, 16)
    with pytest.raises(AssertionError):
        U.upfirdn2d(torch.zeros(1, 1, 2, 2, device="cuda"), f4.cuda(), impl="cuda")  # empty output


def test_full_size_roundtrip_properties():
    """Config-2 sizes: pattern A on [8,128,257,257] and B on [8,3,128,128]: adjoint identity <U x, y> = <x, U^T y>
    (size-independent; U^T is the swapped-factor call, upfirdn2d.py:246-261) and a strided pointwise check."""
    U = _ops()
    g = torch.Generator(device="cuda").manual_seed(1)
    f4 = R.setup_filter([1, 3, 3, 1]).cuda()
    for shape, kw in (((8, 128, 257, 257), PATTERNS["A"]), ((8, 3, 128, 128), PATTERNS["B"]),
                      ((4, 64, 256, 256), PATTERNS["B_bwd"])):
        x = torch.randn(shape, device="cuda", generator=g).requires_grad_(True)
        y = U.upfirdn2d(x, f4, impl="cuda", **kw)
        w = torch.randn(y.shape, device="cuda", generator=g)
        (xt,) = torch.autograd.grad(y, [x], w)
        lhs = (y.detach().double() * w.double()).sum()
        rhs = (x.detach().double() * xt.double()).sum()
        assert abs(lhs - rhs).item() <= 1e-6 * max(abs(lhs).item(), 1.0) + 1e-3
        # one plane against the oracle
        xs = x.detach()[1:2, 1:2].cpu()
        assert _rel(y.detach()[1:2, 1:2], R.upfirdn2d(xs, f4.cpu(), **kw)) <= 1e-5


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(2, 64, 17, 17), (1, 72, 33, 40), (2, 8, 9, 9), (1, 128, 65, 65), (3, 16, 5, 70)])
@pytest.mark.parametrize("pat", ["A", "A_bwd"])
def test_channels_last_16bit_fast_path(dtype, shape, pat):
    """The TMA-staged channels-last kernel (16-bit storage, up = down = 1): the FIR after the transposed convolution
    and its adjoint.  Oracle in fp32 on the rounded inputs; <= 2 ulp of the storage type."""
    U = _ops()
    kw = PATTERNS[pat]
    g = torch.Generator().manual_seed(sum(shape))
    x = torch.randn(shape, generator=g).to(dtype)
    f4 = R.setup_filter([1, 3, 3, 1])
    y_ref = R.upfirdn2d(x.float(), f4, **kw)
    xc = x.cuda().contiguous(memory_format=torch.channels_last)
    y = U.upfirdn2d(xc, f4.cuda(), impl="cuda", **kw)
    assert y.dtype == dtype and y.is_contiguous(memory_format=torch.channels_last)
    assert _rel(y, y_ref) <= 2 * torch.finfo(dtype).eps


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("act,alpha,gain,clamp", [(None, None, None, None), ("lrelu", None, None, None), ("lrelu", 0.2, 1.7, 0.9),
                                                  ("linear", None, 0.5, 0.4), ("lrelu", 1.5, 1.0, None), ("tanh", None, None, None),
                                                  ("lrelu", 0.2, -1.0, 0.5)])
@pytest.mark.parametrize("sep", [True, False])
def test_channels_last_fused_epilogue(dtype, act, alpha, gain, clamp, sep):
    """Channels-last 16-bit FIR with the fused epilogue y = bias_act(fir(x) * scale[n,c] + noise[n,:,y,x], bias) - the second
    half of an up=2 SynthesisLayer (networks.py:299-304 after conv2d_resample.py:125-142) - for the packed
    (lrelu / linear, gain > 0, 0 <= alpha <= 1) and generic epilogues, with a separable filter (the in-kernel rank-1
    factorisation + sliding-window path) and a non-separable one (direct path).  fp32 oracle on the rounded inputs;
    <= 2 ulp of the storage type relative to the output range."""
    from viscoin_next_b200 import modconv as M
    U = _ops()
    g = torch.Generator().manual_seed(11)
    n, c, h, w = 2, 72, 37, 45
    x = torch.randn(n, c, h, w, generator=g).to(dtype)
    f = R.setup_filter([1, 3, 3, 1])
    if not sep:
        f = f + 0.05 * torch.randn(4, 4, generator=g)
    scale = torch.rand(n, c, generator=g) + 0.5
    noise = torch.randn(n, 1, h - 1, w - 1, generator=g) * 0.3
    bias = torch.randn(c, generator=g) * 0.3
    z = R.upfirdn2d(x.float(), f, padding=[1, 1, 1, 1], gain=4) * scale[:, :, None, None] + noise
    y_ref = R.bias_act(z, bias, act=act, alpha=alpha, gain=gain, clamp=clamp) if act is not None else z
    epi, hold = M.make_epilogue(scale=scale.cuda(), noise=noise.cuda(), bias=bias.cuda() if act else None, act=act, alpha=alpha,
                                gain=gain, clamp=clamp)
    xc = x.cuda().contiguous(memory_format=torch.channels_last)
    y = U.launch(xc, f.cuda(), 1, 1, 1, 1, 1, 1, 1, 1, False, 4.0, epilogue=epi)
    assert y.dtype == dtype and y.is_contiguous(memory_format=torch.channels_last)
    assert _rel(y, y_ref) <= 2 * torch.finfo(dtype).eps


SYM6 = [0.015404109327027373, 0.0034907120842174702, -0.11799011114819057, -0.048311742585633, 0.4910559419267466,
        0.787641141030194, 0.3379294217276218, -0.07263752278646252, -0.021060292512300564, 0.04472490177066578,
        0.0017677118642428036, -0.007800708325034148]  # the 12-tap wavelet AugmentPipe resamples with (augment.py:27)


@pytest.mark.parametrize("taps", [5, 8, 12, 13, 24, 31])
@pytest.mark.parametrize("up,down", [(1, 1), (2, 1), (1, 2)])
@pytest.mark.parametrize("flip", [False, True])
def test_separable_long_filter_1d_kernels(taps, up, down, flip):
    """1-D filters of 5..32 taps (the reference's separable specialisations, upfirdn2d.cu:219-246,264-338)
    run two `upfirdn1d` passes (x then y).  Exact-FMA fp32 kernels: rel-Linf <= 1e-5 vs the oracle; odd sizes, asymmetric
    and negative padding, both factor directions, forward and the swapped-factor backward."""
    U = _ops()
    g = torch.Generator().manual_seed(taps * 7 + up * 3 + down)
    f = torch.tensor(SYM6) if taps == 12 else torch.randn(taps, generator=g)
    x = torch.randn(2, 3, 37, 141, generator=g, requires_grad=True)
    pad = [taps // 2 + 1, taps // 2 - 2, taps // 2 - 1, taps // 2] if taps != 13 else [-1, 9, 7, -2]
    y_ref = R.upfirdn2d(x, f, up=up, down=down, padding=pad, flip_filter=flip, gain=float(up * up))
    dy = torch.randn(y_ref.shape, generator=g)
    (dx_ref,) = torch.autograd.grad(y_ref, [x], dy)
    xc = x.detach().cuda().requires_grad_(True)
    import viscoin_next_b200 as V
    before = V._lib.load().vcn_launch_count()
    y = U.upfirdn2d(xc, f.cuda(), up=up, down=down, padding=pad, flip_filter=flip, gain=float(up * up), impl="cuda")
    assert V._lib.load().vcn_launch_count() - before == 2  # one launch per 1-D pass
    (dx,) = torch.autograd.grad(y, [xc], dy.cuda())
    assert _rel(y, y_ref) <= 1e-5 and _rel(dx, dx_ref) <= 1e-5


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_separable_long_filter_16bit_and_full_size(dtype):
    """AugmentPipe-sized call (images at twice the 256^2 resolution, 12-tap sym6, up then down by 2): the 16-bit kernels
    against the fp32 oracle at the storage tolerance, and the up-then-down round trip at full size against the same
    composition through the generic gather kernel's arithmetic (linearity: f(2x) = 2 f(x), a size-independent property)."""
    U = _ops()
    g = torch.Generator().manual_seed(3)
    f = torch.tensor(SYM6)
    x = torch.randn(2, 3, 64, 80, generator=g)
    y_ref = R.upfirdn2d(x, f, up=2, padding=[6, 5, 6, 5], gain=4)
    y = U.upfirdn2d(x.cuda().to(dtype), f.cuda(), up=2, padding=[6, 5, 6, 5], gain=4, impl="cuda")
    assert y.dtype == dtype and _rel(y.float(), y_ref) <= 2e-2
    big = torch.randn(8, 3, 512, 512, device="cuda", generator=torch.Generator(device="cuda").manual_seed(4))
    a = U.upfirdn2d(U.upfirdn2d(big, f.cuda(), up=2, padding=[6, 5, 6, 5], gain=4), f.cuda(), down=2, padding=[5, 5, 5, 5])
    b = U.upfirdn2d(U.upfirdn2d(2 * big, f.cuda(), up=2, padding=[6, 5, 6, 5], gain=4), f.cuda(), down=2, padding=[5, 5, 5, 5])
    assert a.shape == big.shape and torch.allclose(2 * a, b, rtol=1e-5, atol=1e-6)
