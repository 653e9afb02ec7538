"""The drop-in, exercised on a B200 through the reference's OWN modules: the unmodified
``stylegan2_ada.training.networks`` / ``viscoin.models.gan`` classes from the staged copy in the git-ignored oracle/_ref
(``python -m oracle.stage_ref`` in the build container; absent -> skipped) run once as they are (reference plugins if they
load, else the reference's own torch fallback) and once after ``viscoin_next_b200.patch.install()``.

Tolerance: the patched path computes bf16 tensor-core operands (fp32 accumulation, fp32 storage): image rel-Linf <= 2e-2,
latent / input gradients rel-L2 <= 1e-1 (the chained-network gates of tests/test_gpu_networks.py)."""

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref():
    from oracle import stage_ref
    if not stage_ref.available():
        pytest.skip("oracle/_ref not staged")
    try:
        return stage_ref.load_plugins(), True
    except Exception as e:
        print("reference plugins did not load (%s): unpatched run = the reference's torch fallback" % e)
        return stage_ref.import_reference(), False


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _rel2(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.parametrize("mode", ["train", "eval"])
def test_patch_install_on_reference_synthesis_network(mode):
    """Reference SynthesisNetwork at 64x64 (train mode = non-fused branch networks.py:63-72; eval mode = the fused grouped
    branch :74-84 -> conv2d_resample(groups=N) -> per-sample-weight GEMM), image and d(img)/d(ws), patched vs unpatched."""
    import viscoin_next_b200 as V
    ref, plugins = _ref()
    torch.manual_seed(3)
    S = ref.networks.SynthesisNetwork(w_dim=64, img_resolution=64, img_channels=3, channel_base=2048, channel_max=64)
    with torch.no_grad():
        for name, p in S.named_parameters():
            if name.endswith("noise_strength"):
                p.fill_(0.1)
    S = S.cuda().requires_grad_(False).train(mode == "train")
    ws = torch.randn(3, S.num_ws, 64, device="cuda").requires_grad_(True)
    dimg = torch.randn(3, 3, 64, 64, device="cuda")
    img0 = S(ws, noise_mode="const")
    (g0,) = torch.autograd.grad(img0, [ws], dimg)
    lib = V._lib.load()
    done = V.patch.install()
    try:
        assert any(d.endswith("bias_act.bias_act") for d in done) and any(d.endswith("modulated_conv2d") for d in done), done
        before = lib.vcn_launch_count()
        img1 = S(ws, noise_mode="const")
        (g1,) = torch.autograd.grad(img1, [ws], dimg)
        launched = lib.vcn_launch_count() - before
    finally:
        V.patch.uninstall()
    e, eg = _rel(img1, img0), _rel2(g1, g0)
    print("reference SynthesisNetwork %s mode, plugins=%s: patched vs unpatched img %.2e dws %.2e, libvcn launches %d" % (mode, plugins, e, eg, launched))
    assert launched > 20, "the patched reference modules did not reach the sm_100a kernels"
    assert e <= 2e-2 and eg <= 1e-1


def test_unpickled_style_module_reaches_kernels_through_op_patches():
    """A generator that carries its OWN copy of ``modulated_conv2d`` (what persistence.py:179-227 gives every unpickled
    network: the op modules are late-bound and patched, the function is not): emulated by restoring the reference function
    on the module after install().  Eval mode takes the fused branch -> conv2d_resample(groups=N)."""
    import viscoin_next_b200 as V
    ref, plugins = _ref()
    torch.manual_seed(4)
    S = ref.networks.SynthesisNetwork(w_dim=32, img_resolution=32, img_channels=3, channel_base=1024, channel_max=64)
    S = S.cuda().requires_grad_(False).eval()
    ws = torch.randn(2, S.num_ws, 32, device="cuda")
    img0 = S(ws, noise_mode="const")
    original = ref.networks.modulated_conv2d
    lib = V._lib.load()
    V.patch.install()
    try:
        ref.networks.modulated_conv2d = original  # the module keeps its own function, only the ops are replaced
        before = lib.vcn_launch_count()
        for m in ("eval", "train"):
            S.train(m == "train")
            img1 = S(ws, noise_mode="const")
            assert _rel(img1, img0) <= 2e-2, (m, _rel(img1, img0))
        assert lib.vcn_launch_count() - before > 20
    finally:
        V.patch.uninstall()


def test_reference_plugins_vs_vcn_ops():
    """Per op, on the GPU: the reference's CUDA plugins (when they load) against this package's kernels, fp32 - both are
    exact-FMA kernels: rel-Linf <= 1e-5."""
    import viscoin_next_b200 as V
    ref, plugins = _ref()
    if not plugins:
        pytest.skip("reference plugins did not load")
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(4, 64, 65, 65, device="cuda", generator=g)
    b = torch.randn(64, device="cuda", generator=g)
    f = ref.upfirdn2d.setup_filter([1, 3, 3, 1], device=x.device)
    for kw in (dict(padding=[1, 1, 1, 1], gain=4), dict(up=2, padding=[2, 1, 2, 1], gain=4), dict(down=2, padding=[1, 1, 1, 1], flip_filter=True)):
        a, r = V.upfirdn2d.upfirdn2d(x, f, **kw), ref.upfirdn2d.upfirdn2d(x, f, impl="cuda", **kw)
        assert _rel(a, r) <= 1e-5, kw
    for act in ("lrelu", "linear", "relu", "swish", "tanh"):
        a, r = V.bias_act.bias_act(x, b, act=act, clamp=1.5), ref.bias_act.bias_act(x, b, act=act, clamp=1.5, impl="cuda")
        assert _rel(a, r) <= 1e-5, act
