"""GPU parity: fused LPIPS feature tap (vcn_lpips_tap_fwd / _bwd) vs the plain torch formulation of lpips v0.1.4's
normalize_tensor / squared difference / lin / spatial mean, and vs the CPU oracle (oracle/ref_step.lpips_vgg) for the
whole LPIPS module.  fp32: value rel <= 1e-5, gradient rel-Linf <= 1e-4.  bf16 features: value <= 1e-4 (same rounded
inputs on both sides), gradient <= 1e-2 (one bf16 rounding of the stored gradient)."""

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("c,h", [(64, 32), (128, 16), (256, 9), (512, 4), (512, 7)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_tap_matches_torch(c, h, dtype):
    import viscoin_next_b200 as V
    T = V.lpips_tap
    g = torch.Generator().manual_seed(c + h)
    n = 3
    fa = torch.relu(torch.randn(n, c, h, h + 1, generator=g)).to(dtype).cuda().contiguous(memory_format=torch.channels_last)
    fb = torch.relu(torch.randn(n, c, h, h + 1, generator=g)).to(dtype).cuda().contiguous(memory_format=torch.channels_last)
    fa[0, :, 0, 0] = 0  # a zero feature vector: value finite, gradient finite (0)
    lin = (torch.rand(1, c, 1, 1, generator=g) / c).cuda()
    gv = torch.randn(n, generator=g).cuda()
    fa1 = fa.clone().requires_grad_(True)
    val = T.lpips_tap(fa1, fb, lin)
    (da,) = torch.autograd.grad(val, fa1, gv)
    fa2 = fa.clone().float().requires_grad_(True)
    ref = T.lpips_tap(fa2, fb.float(), lin, impl="ref")
    (dr,) = torch.autograd.grad(ref, fa2, gv)
    assert val.shape == (n,) and da.shape == fa.shape and da.dtype == dtype
    assert bool(torch.isfinite(da.float()).all())
    tv, tg = (1e-5, 1e-4) if dtype == torch.float32 else (1e-4, 1e-2)
    assert _rel(val, ref) <= tv, _rel(val, ref)
    assert _rel(da, dr) <= tg, _rel(da, dr)


def test_lpips_module_fused_matches_oracle():
    import viscoin_next_b200 as V
    from oracle import ref_step as RS
    torch.manual_seed(5)
    m = V.viscoin_models.LPIPS()
    convs = [(l.weight.detach().clone(), l.bias.detach().clone()) for s in m.slices for l in s if isinstance(l, torch.nn.Conv2d)]
    lins = [p.detach().clone() for p in m.lins]
    a = torch.randn(2, 3, 64, 64).requires_grad_(True)
    b = torch.randn(2, 3, 64, 64)
    want = RS.lpips_vgg(convs, lins, a, b)
    (dw,) = torch.autograd.grad(want.sum(), a)
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        m = m.cuda()
        ac = a.detach().cuda().requires_grad_(True)
        got = m(ac, b.cuda())
        (dg,) = torch.autograd.grad(got.sum(), ac)
    finally:
        torch.backends.cudnn.allow_tf32 = old
    assert got.shape == (2, 1, 1, 1)
    assert _rel(got, want) <= 1e-4, _rel(got, want)
    assert _rel(dg, dw) <= 1e-3, _rel(dg, dw)
