"""CPU: host-side logic that needs no GPU - the impl='ref' formulations of the mirrors against the oracle, the
drop-in installer's attribute wiring on the real reference modules (only where /root/reference exists), and the
state_dict compatibility of the network mirrors with the reference-recorded fixtures."""

import os
import sys

import pytest
import torch

from oracle import ref_ops as R

REF = "/root/reference"


def _close(a, b, tol=1e-5):
    assert a.shape == b.shape
    assert ((a - b).abs().max() / b.abs().max().clamp_min(1)).item() <= tol


def test_ref_mirrors_match_oracle():
    import viscoin_next_b200 as V
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 6, 9, 9, generator=g)
    b = torch.randn(6, generator=g)
    f = R.setup_filter([1, 3, 3, 1])
    for act in R.ACTIVATIONS:
        _close(V.bias_act.bias_act(x, b, act=act, clamp=1.5, impl="ref"), R.bias_act(x, b, act=act, clamp=1.5))
    _close(V.upfirdn2d.upfirdn2d(x, f, up=2, padding=[2, 1, 2, 1], gain=4, impl="ref"), R.upfirdn2d(x, f, up=2, padding=[2, 1, 2, 1], gain=4))
    _close(V.upfirdn2d.upsample2d(x, f, impl="ref"), R.upsample2d(x, f))
    _close(V.upfirdn2d.downsample2d(x[:, :, :8, :8], f, impl="ref"), R.downsample2d(x[:, :, :8, :8], f))
    w = torch.randn(5, 6, 3, 3, generator=g)
    s = torch.randn(2, 6, generator=g) + 1
    for up in (1, 2):
        noise = torch.randn(2, 1, 9 * up, 9 * up, generator=g)
        y = V.modconv.modulated_conv2d(x, w, s, noise=noise, up=up, padding=1, resample_filter=f, flip_weight=(up == 1), impl="ref")
        _close(y, R.modulated_conv2d(x, w, s, noise=noise, up=up, padding=1, resample_filter=f, flip_weight=(up == 1)), 2e-5)
    a, bb, c = torch.randn(2, 3, 4, 4), torch.randn(2, 3, 1, 1), torch.randn(2, 1, 4, 4)
    _close(V.fma.fma(a, bb, c, impl="ref"), R.fma(a, bb, c))


def test_network_mirrors_load_reference_state_dicts(golden):
    from viscoin_next_b200 import networks as N
    g = golden("networks")
    G = N.Generator(z_dim=32, c_dim=0, w_dim=32, img_resolution=32, img_channels=3, mapping_kwargs=dict(num_layers=2),
                    synthesis_kwargs=dict(channel_base=512, channel_max=16))
    G.load_state_dict(g.sub("G.sd."), strict=True)
    GA = N.GeneratorAdapted(z_dim=8, w_dim=32, img_resolution=32, synthesis_kwargs=dict(channel_base=512, channel_max=16),
                            mapping_kwargs=dict(coarse_layer=2, mid_layer=10, num_layers=1))
    GA.load_state_dict(g.sub("GA.sd."), strict=True)
    for arch in ("orig", "resnet"):
        S = N.SynthesisNetwork(w_dim=32, img_resolution=16, img_channels=3, channel_base=256, channel_max=16, architecture=arch, conv_clamp=0.8)
        S.load_state_dict(g.sub(f"S_{arch}.sd."), strict=True)
    # default VisCoIN shapes: 14 ws, 28.5 M parameters in the generator
    S = N.SynthesisNetwork(w_dim=512, img_resolution=256, img_channels=3)
    assert S.num_ws == 14 and abs(sum(p.numel() for p in S.parameters()) / 1e6 - 27.93) < 0.2


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree only exists in the build container")
def test_patch_install_on_reference_modules():
    sys.path.insert(0, os.path.join(REF, "stylegan2_ada"))
    try:
        from torch_utils.ops import bias_act as rb, upfirdn2d as ru, fma as rf, conv2d_resample as rc  # noqa
        from training import networks as rn
        import viscoin_next_b200 as V
        orig = rb.bias_act
        done = V.patch.install()
        try:
            assert "torch_utils.ops.bias_act.bias_act" in done and "training.networks.modulated_conv2d" in done
            assert rb.bias_act is V.bias_act.bias_act and ru.upsample2d is V.upfirdn2d.upsample2d
            assert rn.modulated_conv2d is V.modconv.modulated_conv2d and rf.fma is V.fma.fma
            assert rc.conv2d_resample is V.conv2d_resample.conv2d_resample
            assert "lrelu" in rb.activation_funcs  # table kept
            # a reference layer now routes through the replacement: on CPU tensors that is a loud error, not a fallback
            layer = rn.SynthesisLayer(8, 8, w_dim=16, resolution=8)
            with pytest.raises(V._lib.VcnError):
                layer(torch.randn(1, 8, 8, 8), torch.randn(1, 16), noise_mode="const")
        finally:
            V.patch.uninstall()
        assert rb.bias_act is orig
    finally:
        sys.path.remove(os.path.join(REF, "stylegan2_ada"))


def test_conv2d_resample_ref_matches_reference_golden():
    """The mirror's impl='ref' formulation of conv2d_resample (every branch: plain, 1x1/3x3 up and down, grouped, asymmetric
    padding through the generic path) reproduces the outputs and gradients recorded from the reference
    (tests/golden/conv_modconv.npz, oracle/make_golden.py CONV_CASES); impl='cuda' on CPU tensors is a loud error."""
    import numpy as np
    import viscoin_next_b200 as V
    from oracle.make_golden import CONV_CASES
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "conv_modconv.npz"))
    f4 = torch.from_numpy(g["f4"])
    for name, (xs, ws, kw, use_f) in CONV_CASES.items():
        x = torch.from_numpy(g[f"conv.{name}.x"]).requires_grad_(True)
        w = torch.from_numpy(g[f"conv.{name}.w"]).requires_grad_(True)
        y = V.conv2d_resample.conv2d_resample(x, w, f=(f4 if use_f else None), impl="ref", **kw)
        dx, dw = torch.autograd.grad(y, [x, w], torch.from_numpy(g[f"conv.{name}.dy"]))
        for got, key in ((y, "y"), (dx, "dx"), (dw, "dw")):
            want = torch.from_numpy(g[f"conv.{name}.{key}"])
            assert got.shape == want.shape and (got.detach() - want).abs().max() <= 1e-5 * want.abs().max().clamp_min(1), (name, key)
    with pytest.raises(V._lib.VcnError):
        V.conv2d_resample.conv2d_resample(torch.randn(1, 4, 8, 8), torch.randn(4, 4, 3, 3), padding=1)
