"""CPU: the oracle (oracle/ref_ops.py, ref_net.py, kernel_math.py) replayed against the golden vectors
that oracle/make_golden.py recorded from the REAL reference (`impl='ref'`, /root/reference).  This is
what pins the oracle; the GPU parity tests then compare the CUDA path with the oracle and with the same
fixtures."""

import numpy as np
import pytest
import torch

from oracle import kernel_math as KM
from oracle import ref_net as RN
from oracle import ref_ops as R
from oracle.make_golden import CONV_CASES, UPFIRDN_CASES

ACTS = list(R.ACTIVATIONS)


def _close(a, b, tol=1e-5):
    a = torch.as_tensor(a, dtype=torch.float64)
    b = torch.as_tensor(b, dtype=torch.float64)
    assert a.shape == b.shape
    err = (a - b).abs().max().item() if a.numel() else 0.0
    scale = max(b.abs().max().item(), 1.0) if b.numel() else 1.0
    assert err <= tol * scale, f"max-abs {err:.3e} > {tol:.0e} * {scale:.3g}"


def _bias_cases(g):
    return sorted({k.split(".")[0] for k in g.keys})


def test_bias_act_forward_and_grads(golden):
    g = golden("bias_act")
    for key in _bias_cases(g):
        act = key.rsplit("_", 1)[0]
        dim, clamp, gain, alpha = g.np(key + ".meta")
        kw = dict(dim=int(dim), act=act, alpha=None if alpha < 0 else float(alpha),
                  gain=None if gain < 0 else float(gain), clamp=None if clamp < 0 else float(clamp))
        x = g.t(key + ".x").requires_grad_(True)
        b = g.t(key + ".b").requires_grad_(True)
        dy = g.t(key + ".dy").requires_grad_(True)
        y = R.bias_act(x, b, **kw)
        _close(y, g.np(key + ".y"))
        dx, db = torch.autograd.grad(y, [x, b], dy, create_graph=True)
        _close(dx, g.np(key + ".dx"))
        _close(db, g.np(key + ".db"))
        d_dy, d_x = torch.autograd.grad(dx, [dy, x], g.t(key + ".ddx"), allow_unused=True)
        _close(d_dy, g.np(key + ".d_dy"))
        if key + ".d_x" in g:
            _close(d_x, g.np(key + ".d_x"))


def test_bias_act_kernel_table_matches_ref_path(golden):
    """kernel_math.bias_act_kernel (the grad=0/1/2 contract of the C ABI) vs reference ref-path outputs."""
    g = golden("bias_act")
    for key in _bias_cases(g):
        act = key.rsplit("_", 1)[0]
        dim, clamp, gain, alpha = g.np(key + ".meta")
        da, dg = R.act_defaults(act)
        alpha = da if alpha < 0 else float(alpha)
        gain = dg if gain < 0 else float(gain)
        x, b, dy, y = (g.np(key + s) for s in (".x", ".b", ".dy", ".y"))
        y0 = KM.bias_act_kernel(x, b, None, None, None, 0, int(dim), act, alpha, gain, clamp)
        np.testing.assert_allclose(y0, y, rtol=2e-5, atol=2e-6)
        # exclude elements sitting exactly on a kink / clamp edge (ref and kernel conventions differ there)
        dx = KM.bias_act_kernel(dy, b, x, y, None, 1, int(dim), act, alpha, gain, clamp)
        np.testing.assert_allclose(dx, g.np(key + ".dx"), rtol=5e-5, atol=5e-6)
        ddx = g.np(key + ".ddx")
        d_dy = KM.bias_act_kernel(ddx, b, x, y, None, 1, int(dim), act, alpha, gain, clamp)
        np.testing.assert_allclose(d_dy, g.np(key + ".d_dy"), rtol=5e-5, atol=5e-6)
        if R.ACTIVATIONS[act][5] and key + ".d_x" in g:
            d_x = KM.bias_act_kernel(ddx, b, x, y, dy, 2, int(dim), act, alpha, gain, clamp)
            np.testing.assert_allclose(d_x, g.np(key + ".d_x"), rtol=2e-4, atol=2e-5)


@pytest.mark.parametrize("name", list(UPFIRDN_CASES))
def test_upfirdn2d(golden, name):
    g = golden("upfirdn2d")
    _, _, kw = UPFIRDN_CASES[name]
    x = g.t(name + ".x").requires_grad_(True)
    f = g.t(name + ".f")
    y = R.upfirdn2d(x, f, **kw)
    _close(y, g.np(name + ".y"))
    (dx,) = torch.autograd.grad(y, [x], g.t(name + ".dy"))
    _close(dx, g.np(name + ".dx"))
    # gather-form kernel math (2-D filters only)
    if f is None or f.ndim == 2:
        up = kw.get("up", 1)
        down = kw.get("down", 1)
        upx, upy = (up, up) if isinstance(up, int) else up
        dnx, dny = (down, down) if isinstance(down, int) else down
        px0, px1, py0, py1 = R.parse_padding(kw.get("padding", 0))
        ff = np.ones((1, 1), np.float32) if f is None else f.numpy()
        yk = KM.upfirdn2d_kernel(x.detach().numpy(), ff, upx, upy, dnx, dny, px0, px1, py0, py1,
                                 bool(kw.get("flip_filter", False)), float(kw.get("gain", 1)))
        np.testing.assert_allclose(yk, g.np(name + ".y"), rtol=1e-5, atol=1e-6)


def test_upfirdn2d_backward_is_swapped_forward(golden):
    """SURVEY App. B / upfirdn2d.py:246-261: dx = upfirdn2d(dy, up<->down, flipped, padding p)."""
    g = golden("upfirdn2d")
    for name in ("A", "B", "B_bwd", "down2", "generic_u3d2", "crop"):
        _, _, kw = UPFIRDN_CASES[name]
        x, f, dy = g.t(name + ".x"), g.t(name + ".f"), g.t(name + ".dy")
        up, down = kw.get("up", 1), kw.get("down", 1)
        upx, upy = (up, up) if isinstance(up, int) else up
        dnx, dny = (down, down) if isinstance(down, int) else down
        px0, px1, py0, py1 = R.parse_padding(kw.get("padding", 0))
        fw, fh = R.filter_size(f)
        _, _, ih, iw = x.shape
        _, _, oh, ow = dy.shape
        p = [fw - px0 - 1, iw * upx - ow * dnx + px0 - upx + 1, fh - py0 - 1, ih * upy - oh * dny + py0 - upy + 1]
        dx = R.upfirdn2d(dy, f, up=[dnx, dny], down=[upx, upy], padding=p,
                         flip_filter=not kw.get("flip_filter", False), gain=kw.get("gain", 1))
        _close(dx, g.np(name + ".dx"))


def test_upfirdn2d_wrappers_and_setup_filter(golden):
    g = golden("upfirdn2d")
    x, f4 = g.t("wrap.x"), g.t("wrap.f")
    _close(R.filter2d(x, f4), g.np("wrap.filter2d"))
    _close(R.upsample2d(x, f4), g.np("wrap.upsample2d"))
    _close(R.downsample2d(x, f4), g.np("wrap.downsample2d"))
    _close(R.upsample2d(x, f4, up=[2, 1], padding=[1, 0, 0, 2], gain=0.5), g.np("wrap.upsample2d_p"))
    _close(R.setup_filter([1, 3, 3, 1]), g.np("setup.a"))
    _close(R.setup_filter([1, 2, 3], normalize=False, flip_filter=True, gain=2), g.np("setup.b"))
    _close(R.setup_filter(list(range(1, 10)), gain=3), g.np("setup.c"))
    _close(R.setup_filter(None), g.np("setup.d"))


@pytest.mark.parametrize("name", list(CONV_CASES))
def test_conv2d_resample(golden, name):
    g = golden("conv_modconv")
    _, _, kw, use_f = CONV_CASES[name]
    x = g.t(f"conv.{name}.x").requires_grad_(True)
    w = g.t(f"conv.{name}.w").requires_grad_(True)
    y = R.conv2d_resample(x, w, f=(g.t("f4") if use_f else None), **kw)
    _close(y, g.np(f"conv.{name}.y"))
    dx, dw = torch.autograd.grad(y, [x, w], g.t(f"conv.{name}.dy"))
    _close(dx, g.np(f"conv.{name}.dx"))
    _close(dw, g.np(f"conv.{name}.dw"))


def test_fma(golden):
    g = golden("conv_modconv")
    a, b, c = (g.t("fma." + k).requires_grad_(True) for k in "abc")
    y = R.fma(a, b, c)
    _close(y, g.np("fma.y"))
    da, db, dc = torch.autograd.grad(y, [a, b, c], g.t("fma.dy"))
    _close(da, g.np("fma.da")); _close(db, g.np("fma.db")); _close(dc, g.np("fma.dc"))


def modconv_keys(g):
    return sorted({k.rsplit(".", 1)[0] for k in g.keys if k.startswith("mod.")})


def parse_modconv_key(key):
    body = key[len("mod."):]  # up{u}k{k}d{0|1}{none|const|rand}f{0|1}
    up, k, demod = int(body[2]), int(body[4]), bool(int(body[6]))
    fused = bool(int(body[-1]))
    return up, k, demod, fused


def test_modulated_conv2d(golden):
    g = golden("conv_modconv")
    keys = modconv_keys(g)
    assert len(keys) == 36
    for key in keys:
        up, k, demod, fused = parse_modconv_key(key)
        x, w, s = (g.t(f"{key}.{n}").requires_grad_(True) for n in ("x", "w", "s"))
        y = R.modulated_conv2d(x, w, s, noise=g.t(key + ".noise"), up=up, padding=k // 2, resample_filter=g.t("f4"),
                               demodulate=demod, flip_weight=(up == 1), fused_modconv=fused)
        _close(y, g.np(key + ".y"), 2e-5)
        dx, dw, ds = torch.autograd.grad(y, [x, w, s], g.t(key + ".dy"))
        _close(dx, g.np(key + ".dx"), 2e-5); _close(dw, g.np(key + ".dw"), 2e-5); _close(ds, g.np(key + ".ds"), 2e-5)


def test_generator_networks(golden):
    g = golden("networks")
    sd = g.sub("G.sd.")
    z = g.t("G.z")
    kw = dict(channel_base=512, channel_max=16)
    for mode in ("eval", "train"):
        ws = g.t(f"G.{mode}.ws").requires_grad_(True)
        img = RN.synthesis_network(ws, sd, 32, prefix="synthesis.", training=(mode == "train"), noise_mode="const")
        _close(img, g.np(f"G.{mode}.img"), 2e-5)
        (dws,) = torch.autograd.grad(img, [ws], g.t(f"G.{mode}.dimg"))
        _close(dws, g.np(f"G.{mode}.dws"), 5e-5)
    _close(RN.generator(z, sd, 32, mapping_layers=2, noise_mode="const"), g.np("G.full.img"), 2e-5)
    _close(RN.generator(z, sd, 32, mapping_layers=2, truncation_psi=0.7, noise_mode="none"), g.np("G.trunc.img"), 2e-5)
    noises = {"synthesis." + k: v for k, v in g.sub("G.random.noise.").items()}
    ws = RN.mapping_network(z, sd, "mapping.", 2, RN.synthesis_num_ws(32))
    img = RN.synthesis_network(ws, sd, 32, prefix="synthesis.", noise_mode="random", noises=noises)
    _close(img, g.np("G.random.img"), 2e-5)


@pytest.mark.parametrize("arch", ["orig", "resnet"])
def test_other_architectures(golden, arch):
    g = golden("networks")
    sd = g.sub(f"S_{arch}.sd.")
    img = RN.synthesis_network(g.t(f"S_{arch}.ws"), sd, 16, noise_mode="const", architecture=arch, conv_clamp=0.8)
    _close(img, g.np(f"S_{arch}.img"), 2e-5)


def test_generator_adapted(golden):
    g = golden("networks")
    sd = g.sub("GA.sd.")
    z1 = g.t("GA.z1").requires_grad_(True)
    z2 = g.t("GA.z2").requires_grad_(True)
    img, ws = RN.generator_adapted(z1, z2, sd, img_resolution=32, z_dim=8, training=True, return_latents=True,
                                   noise_mode="const")
    _close(ws, g.np("GA.ws"), 2e-5)
    _close(img, g.np("GA.img"), 2e-5)
    dz1, dz2 = torch.autograd.grad(img, [z1, z2], g.t("GA.dimg"))
    _close(dz1, g.np("GA.dz1"), 5e-5); _close(dz2, g.np("GA.dz2"), 5e-5)
