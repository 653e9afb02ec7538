"""Generate the golden input/output vectors that pin the oracle and the CUDA path.

TEST INFRASTRUCTURE ONLY.  Runs ONLY in the build container, where the real
reference is mounted read-only at /root/reference: it imports the reference's own
modules (``torch_utils.ops.*``, ``training.networks``, ``viscoin.models.gan``),
executes their ``impl='ref'`` CPU path on seeded inputs and stores inputs,
outputs and gradients as small ``.npz`` files under ``tests/golden/``.  The GPU
box has no /root/reference; tests there read only the committed fixtures.

    python -m oracle.make_golden            # rewrites tests/golden/*.npz
"""

import os
import sys
import warnings

import numpy as np
import torch

REF_ROOT = os.environ.get("VCN_REFERENCE_ROOT", "/root/reference")
OUT_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _import_reference():
    if not os.path.isdir(REF_ROOT):
        raise SystemExit(f"{REF_ROOT} not found: golden vectors can only be regenerated in the build container")
    sys.path.insert(0, os.path.join(REF_ROOT, "stylegan2_ada"))
    sys.path.insert(0, REF_ROOT)
    warnings.filterwarnings("ignore")
    from torch_utils.ops import bias_act, conv2d_resample, fma, upfirdn2d  # noqa
    from training import networks  # noqa
    return bias_act, upfirdn2d, conv2d_resample, fma, networks


def _np(t):
    return None if t is None else t.detach().cpu().numpy()


def _save(name, **arrays):
    os.makedirs(OUT_DIR, exist_ok=True)
    path = os.path.join(OUT_DIR, name + ".npz")
    np.savez_compressed(path, **{k: v for k, v in arrays.items() if v is not None})
    print(f"  wrote {os.path.relpath(path)}  ({os.path.getsize(path) / 1024:.1f} KiB, {len(arrays)} arrays)")


# ---------------------------------------------------------------------------


def golden_bias_act(bias_act):
    g = torch.Generator().manual_seed(1234)
    out = {}
    cases = []
    for act in bias_act.activation_funcs:
        for ci, (clamp, gain, alpha) in enumerate([(None, None, None), (0.6, 0.75, 0.3)]):
            cases.append((act, ci, clamp, gain, alpha, (2, 5, 3, 4), 1))
    cases.append(("lrelu", 2, None, None, None, (3, 7), 1))  # 2-D, mapping-FC shape
    cases.append(("lrelu", 3, 1.5, None, None, (4, 6, 2), 2))  # bias on the last dim
    cases.append(("linear", 4, None, None, None, (2, 3, 4, 4), 1))
    for act, ci, clamp, gain, alpha, shape, dim in cases:
        key = f"{act}_{ci}"
        x = (torch.randn(shape, generator=g) * 1.5).requires_grad_(True)
        b = torch.randn([shape[dim]], generator=g).requires_grad_(True)
        dy = torch.randn(shape, generator=g).requires_grad_(True)
        y = bias_act.bias_act(x, b, dim=dim, act=act, alpha=alpha, gain=gain, clamp=clamp, impl="ref")
        dx, db = torch.autograd.grad(y, [x, b], dy, create_graph=True)
        out.update({f"{key}.x": _np(x), f"{key}.b": _np(b), f"{key}.dy": _np(dy), f"{key}.y": _np(y),
                    f"{key}.dx": _np(dx), f"{key}.db": _np(db)})
        out[f"{key}.meta"] = np.array([dim, -1 if clamp is None else clamp, -1 if gain is None else gain,
                                       -1 if alpha is None else alpha], dtype=np.float64)
        # second order: d/d{dy,x} of <dx, ddx>
        ddx = torch.randn(shape, generator=g)
        d_dy, d_x = torch.autograd.grad(dx, [dy, x], ddx, allow_unused=True)
        out.update({f"{key}.ddx": _np(ddx), f"{key}.d_dy": _np(d_dy)})
        if d_x is not None:
            out[f"{key}.d_x"] = _np(d_x)
    _save("bias_act", **out)


UPFIRDN_CASES = {
    # name: (shape, filter spec, kwargs)  -- A/A'/B/B' are the hot-path patterns of SURVEY 8(a6)
    "A": ((2, 3, 9, 9), [1, 3, 3, 1], dict(up=1, down=1, padding=[1, 1, 1, 1], gain=4)),
    "A_bwd": ((2, 3, 8, 8), [1, 3, 3, 1], dict(up=1, down=1, padding=[2, 2, 2, 2], flip_filter=True, gain=4)),
    "B": ((2, 3, 8, 8), [1, 3, 3, 1], dict(up=2, down=1, padding=[2, 1, 2, 1], gain=4)),
    "B_bwd": ((2, 3, 16, 16), [1, 3, 3, 1], dict(up=1, down=2, padding=[1, 1, 1, 1], flip_filter=True, gain=4)),
    "down2": ((1, 4, 12, 10), [1, 3, 3, 1], dict(up=1, down=2, padding=[1, 1, 1, 1])),
    "generic_u3d2": ((1, 2, 7, 6), [[1, 2, 0.5], [0.25, -1, 3], [2, 1, 1], [0.5, 0.5, -2]],
                     dict(up=[3, 2], down=[2, 3], padding=[2, -1, 3, 1], gain=1.7)),
    "crop": ((1, 2, 10, 11), [1, 2, 1], dict(padding=[-1, -2, -1, 0], flip_filter=True)),
    "sep12": ((1, 2, 20, 18), [0.1, -0.3, 0.7, 1.1, 2.0, 3.1, 2.9, 1.9, 1.2, 0.6, -0.2, 0.05],
              dict(up=2, down=1, padding=[6, 5, 6, 5], gain=4)),
    "identity_f": ((1, 2, 5, 5), None, dict(up=2, padding=[0, 1, 0, 1])),
}


def golden_upfirdn2d(upfirdn2d):
    g = torch.Generator().manual_seed(4321)
    out = {}
    for name, (shape, fspec, kw) in UPFIRDN_CASES.items():
        x = torch.randn(shape, generator=g).requires_grad_(True)
        f = None
        if fspec is not None:
            f = upfirdn2d.setup_filter(fspec, normalize=(name != "generic_u3d2"))
            out[f"{name}.f"] = _np(f)
        y = upfirdn2d.upfirdn2d(x, f, impl="ref", **kw)
        dy = torch.randn(y.shape, generator=g)
        (dx,) = torch.autograd.grad(y, [x], dy)
        out.update({f"{name}.x": _np(x), f"{name}.y": _np(y), f"{name}.dy": _np(dy), f"{name}.dx": _np(dx)})
    # wrappers (padding arithmetic of filter2d / upsample2d / downsample2d)
    f4 = upfirdn2d.setup_filter([1, 3, 3, 1])
    x = torch.randn(2, 3, 8, 8, generator=g)
    out["wrap.x"], out["wrap.f"] = _np(x), _np(f4)
    out["wrap.filter2d"] = _np(upfirdn2d.filter2d(x, f4, impl="ref"))
    out["wrap.upsample2d"] = _np(upfirdn2d.upsample2d(x, f4, impl="ref"))
    out["wrap.downsample2d"] = _np(upfirdn2d.downsample2d(x, f4, impl="ref"))
    out["wrap.upsample2d_p"] = _np(upfirdn2d.upsample2d(x, f4, up=[2, 1], padding=[1, 0, 0, 2], gain=0.5, impl="ref"))
    # setup_filter variants
    out["setup.a"] = _np(upfirdn2d.setup_filter([1, 3, 3, 1]))
    out["setup.b"] = _np(upfirdn2d.setup_filter([1, 2, 3], normalize=False, flip_filter=True, gain=2))
    out["setup.c"] = _np(upfirdn2d.setup_filter(list(range(1, 10)), gain=3))  # separable (>= 8 taps)
    out["setup.d"] = _np(upfirdn2d.setup_filter(None))
    _save("upfirdn2d", **out)


CONV_CASES = {
    # name: (x shape, w shape, kwargs, use filter)
    "plain3": ((2, 4, 8, 8), (6, 4, 3, 3), dict(padding=1), False),
    "plain1": ((2, 4, 8, 8), (3, 4, 1, 1), dict(), False),
    "up2_3": ((2, 4, 8, 8), (6, 4, 3, 3), dict(up=2, padding=1, flip_weight=False), True),
    "up2_1": ((2, 4, 8, 8), (6, 4, 1, 1), dict(up=2), True),
    "down2_3": ((2, 4, 8, 8), (6, 4, 3, 3), dict(down=2, padding=1), True),
    "down2_1": ((2, 4, 8, 8), (6, 4, 1, 1), dict(down=2), True),
    "grouped_up2": ((1, 8, 6, 6), (12, 4, 3, 3), dict(up=2, padding=1, groups=2, flip_weight=False), True),
    "asym_pad": ((1, 3, 7, 7), (2, 3, 3, 3), dict(padding=[2, 0, 1, 1]), False),
}


def golden_conv_modconv(upfirdn2d, conv2d_resample, fma, networks):
    g = torch.Generator().manual_seed(777)
    out = {}
    f4 = upfirdn2d.setup_filter([1, 3, 3, 1])
    out["f4"] = _np(f4)
    for name, (xs, ws, kw, use_f) in CONV_CASES.items():
        x = torch.randn(xs, generator=g).requires_grad_(True)
        w = torch.randn(ws, generator=g).requires_grad_(True)
        y = conv2d_resample.conv2d_resample(x, w, f=(f4 if use_f else None), **kw)
        dy = torch.randn(y.shape, generator=g)
        dx, dw = torch.autograd.grad(y, [x, w], dy)
        out.update({f"conv.{name}.x": _np(x), f"conv.{name}.w": _np(w), f"conv.{name}.y": _np(y),
                    f"conv.{name}.dy": _np(dy), f"conv.{name}.dx": _np(dx), f"conv.{name}.dw": _np(dw)})

    a = torch.randn(2, 5, 4, 4, generator=g).requires_grad_(True)
    b = torch.randn(2, 5, 1, 1, generator=g).requires_grad_(True)
    c = torch.randn(2, 1, 4, 4, generator=g).requires_grad_(True)
    y = fma.fma(a, b, c)
    dy = torch.randn(y.shape, generator=g)
    da, db, dc = torch.autograd.grad(y, [a, b, c], dy)
    out.update({"fma.a": _np(a), "fma.b": _np(b), "fma.c": _np(c), "fma.y": _np(y), "fma.dy": _np(dy),
                "fma.da": _np(da), "fma.db": _np(db), "fma.dc": _np(dc)})

    n, cin, cout, r = 2, 8, 6, 8
    for up in (1, 2):
        for k in (3, 1):
            if up == 2 and k == 1:
                continue
            for demod in (True, False):
                for noise_kind in ("none", "const", "rand"):
                    for fused in (True, False):
                        key = f"mod.up{up}k{k}d{int(demod)}{noise_kind}f{int(fused)}"
                        x = torch.randn(n, cin, r, r, generator=g).requires_grad_(True)
                        w = torch.randn(cout, cin, k, k, generator=g).requires_grad_(True)
                        s = (torch.randn(n, cin, generator=g) + 1).requires_grad_(True)
                        ro = r * up
                        noise = None
                        if noise_kind == "const":
                            noise = torch.randn(ro, ro, generator=g) * 0.3
                        elif noise_kind == "rand":
                            noise = torch.randn(n, 1, ro, ro, generator=g) * 0.3
                        y = networks.modulated_conv2d(x=x, weight=w, styles=s, noise=noise, up=up, padding=k // 2,
                                                      resample_filter=f4, demodulate=demod, flip_weight=(up == 1),
                                                      fused_modconv=fused)
                        dy = torch.randn(y.shape, generator=g)
                        dx, dw, ds = torch.autograd.grad(y, [x, w, s], dy)
                        out.update({f"{key}.x": _np(x), f"{key}.w": _np(w), f"{key}.s": _np(s),
                                    f"{key}.noise": _np(noise), f"{key}.y": _np(y), f"{key}.dy": _np(dy),
                                    f"{key}.dx": _np(dx), f"{key}.dw": _np(dw), f"{key}.ds": _np(ds)})
    _save("conv_modconv", **out)


def _randomize(module, gen):
    """Make every zero-initialised parameter non-trivial (noise_strength, biases) so the
    whole path is exercised."""
    with torch.no_grad():
        for name, p in module.named_parameters():
            if name.endswith("noise_strength"):
                p.copy_(torch.rand([], generator=gen) * 0.2 + 0.05)
            elif name.endswith("bias") and "affine" not in name:
                p.copy_(torch.randn(p.shape, generator=gen) * 0.1)
            elif name.endswith("fixed_w_avg"):
                p.copy_(torch.randn(p.shape, generator=gen) * 0.1)


def golden_networks(networks):
    torch.manual_seed(99)
    g = torch.Generator().manual_seed(99)
    out = {}
    # --- Generator, 32x32, narrow channels -------------------------------------------
    G = networks.Generator(z_dim=32, c_dim=0, w_dim=32, img_resolution=32, img_channels=3,
                           mapping_kwargs=dict(num_layers=2),
                           synthesis_kwargs=dict(channel_base=512, channel_max=16))
    _randomize(G, g)
    for k, v in G.state_dict().items():
        out["G.sd." + k] = _np(v)
    z = torch.randn(3, 32, generator=g)
    out["G.z"] = _np(z)
    for mode in ("eval", "train"):
        G.synthesis.train(mode == "train")
        G.mapping.eval()
        ws = G.mapping(z, None).detach().requires_grad_(True)
        img = G.synthesis(ws, noise_mode="const")
        dimg = torch.randn(img.shape, generator=g)
        (dws,) = torch.autograd.grad(img, [ws], dimg)
        out.update({f"G.{mode}.ws": _np(ws), f"G.{mode}.img": _np(img), f"G.{mode}.dimg": _np(dimg),
                    f"G.{mode}.dws": _np(dws)})
    G.eval()
    out["G.full.img"] = _np(G(z, None, noise_mode="const"))
    out["G.trunc.img"] = _np(G(z, None, truncation_psi=0.7, noise_mode="none"))
    # explicit-noise run: replay of noise_mode='random' with a recorded RNG stream
    torch.manual_seed(5)
    out["G.random.img"] = _np(G.synthesis(G.mapping(z, None), noise_mode="random"))
    torch.manual_seed(5)
    for res in G.synthesis.block_resolutions:
        for lname in (["conv1"] if res == 4 else ["conv0", "conv1"]):
            out[f"G.random.noise.b{res}.{lname}."] = _np(torch.randn([3, 1, res, res]))

    # --- other architectures / conv_clamp ---------------------------------------------
    for arch in ("orig", "resnet"):
        S = networks.SynthesisNetwork(w_dim=32, img_resolution=16, img_channels=3, channel_base=256, channel_max=16,
                                      architecture=arch, conv_clamp=0.8)
        _randomize(S, g)
        for k, v in S.state_dict().items():
            out[f"S_{arch}.sd." + k] = _np(v)
        ws = torch.randn(2, S.num_ws, 32, generator=g)
        S.eval()
        out[f"S_{arch}.ws"] = _np(ws)
        out[f"S_{arch}.img"] = _np(S(ws, noise_mode="const"))

    # --- VisCoIN adapted generator ---------------------------------------------------
    from viscoin.models.gan import GeneratorAdapted
    GA = GeneratorAdapted(z_dim=8, w_dim=32, img_resolution=32,
                          synthesis_kwargs=dict(channel_base=512, channel_max=16))
    _randomize(GA, g)
    for k, v in GA.state_dict().items():
        out["GA.sd." + k] = _np(v)
    z1 = torch.randn(2, 8, 3, 3, generator=g).requires_grad_(True)
    z2 = torch.randn(2, 72, generator=g).requires_grad_(True)
    GA.train()  # the VisCoIN loop runs it in train mode (non-fused modconv)
    img, ws = GA(z1, z2, return_latents=True, noise_mode="const")
    dimg = torch.randn(img.shape, generator=g)
    dz1, dz2 = torch.autograd.grad(img, [z1, z2], dimg)
    out.update({"GA.z1": _np(z1), "GA.z2": _np(z2), "GA.img": _np(img), "GA.ws": _np(ws), "GA.dimg": _np(dimg),
                "GA.dz1": _np(dz1), "GA.dz2": _np(dz2)})
    _save("networks", **out)


def golden_step():
    """VisCoIN step components: Psi, Theta and the loss terms of viscoin/training/losses.py, from the REAL reference
    modules (``lpips`` is not installable offline: the module is stubbed and the LPIPS term patched to zero, so the
    perceptual term stays unpinned - see oracle/ref_step.py)."""
    import types
    stub = types.ModuleType("lpips")
    stub.LPIPS = object
    sys.modules.setdefault("lpips", stub)
    from viscoin.models.concept_extractors import ConceptExtractor
    from viscoin.models.explainers import Explainer
    from viscoin.training import losses as L
    L.lpips_loss = lambda a, b: torch.zeros(())
    g = torch.Generator().manual_seed(4321)
    torch.manual_seed(4321)
    psi = ConceptExtractor(n_concepts=8, latent_channels=16, input_channels1=16, input_channels2=32, input_channels3=64)
    theta = Explainer(n_concepts=8, n_classes=10).eval()
    h = [torch.randn(2, 16, 28, 28, generator=g).requires_grad_(True), torch.randn(2, 32, 14, 14, generator=g).requires_grad_(True),
         torch.randn(2, 64, 7, 7, generator=g).requires_grad_(True)]
    phi, phi2 = psi(tuple(h))
    expl = theta(phi)
    dphi, dphi2 = torch.randn(phi.shape, generator=g), torch.randn(phi2.shape, generator=g)
    dh = torch.autograd.grad([phi, phi2], h, [dphi, dphi2])
    out = {"psi." + k: _np(v) for k, v in psi.state_dict().items()}
    out.update({"theta." + k: _np(v) for k, v in theta.state_dict().items()})
    out.update({"h0": _np(h[0]), "h1": _np(h[1]), "h2": _np(h[2]), "phi": _np(phi), "phi2": _np(phi2), "expl": _np(expl),
                "dphi": _np(dphi), "dphi2": _np(dphi2), "dh0": _np(dh[0]), "dh1": _np(dh[1]), "dh2": _np(dh[2])})
    a, b = torch.randn(4, 10, generator=g), torch.randn(4, 10, generator=g)
    out.update({"cls_a": _np(a), "cls_b": _np(b), "loss.cce": _np(L.cross_cross_entropy_loss(a, b)),
                "loss.of": _np(L.output_fidelity_loss(a, b)),
                "loss.cr": _np(L.concept_regularization_loss(phi)), "loss.ortho": _np(L.concept_orthogonality_loss(psi))})
    x0, x1 = torch.randn(4, 3, 16, 16, generator=g), torch.randn(4, 3, 16, 16, generator=g)
    out.update({"img_a": _np(x0), "img_b": _np(x1),
                "loss.rec_nolpips": _np(L.reconstruction_loss(x0, x1, a, b, 0.1, 3.0))})
    ws = torch.randn(4, 14, 32, generator=g)
    w_avg = torch.randn(32, generator=g)
    model = types.SimpleNamespace(mapping=types.SimpleNamespace(fixed_w_avg=w_avg))
    out.update({"ws": _np(ws), "w_avg": _np(w_avg), "loss.gan": _np(L.gan_regularization_loss(ws, model))})
    _save("step", **out)


def main():
    bias_act, upfirdn2d, conv2d_resample, fma, networks = _import_reference()
    torch.set_num_threads(max(1, len(os.sched_getaffinity(0))))
    print(f"reference at {REF_ROOT}, torch {torch.__version__}")
    if len(sys.argv) > 1 and sys.argv[1] == "step":
        golden_step()
        return
    golden_bias_act(bias_act)
    golden_upfirdn2d(upfirdn2d)
    golden_conv_modconv(upfirdn2d, conv2d_resample, fma, networks)
    golden_networks(networks)
    golden_step()


if __name__ == "__main__":
    main()
