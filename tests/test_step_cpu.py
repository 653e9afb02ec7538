"""VisCoIN step components on CPU: the oracle restatement (oracle/ref_step.py) and the product's host-side modules
(plain torch, no kernels involved) against vectors produced by the REAL reference modules (tests/golden/step.npz,
written by oracle/make_golden.py::golden_step)."""

import types

import torch

from oracle import ref_step as RS

TOL = dict(rtol=1e-5, atol=1e-6)


def test_oracle_psi_theta_match_reference(golden):
    g = golden("step")
    psi, theta = g.sub("psi."), g.sub("theta.")
    h = [g.t("h%d" % i).requires_grad_(True) for i in range(3)]
    phi, phi2 = RS.concept_extractor(psi, h)
    assert torch.allclose(phi, g.t("phi"), **TOL) and torch.allclose(phi2, g.t("phi2"), **TOL)
    assert torch.allclose(RS.explainer(theta, phi), g.t("expl"), **TOL)
    dh = torch.autograd.grad([phi, phi2], h, [g.t("dphi"), g.t("dphi2")])
    for i in range(3):
        assert torch.allclose(dh[i], g.t("dh%d" % i), rtol=1e-4, atol=1e-6)


def test_oracle_losses_match_reference(golden):
    g = golden("step")
    a, b = g.t("cls_a"), g.t("cls_b")
    assert torch.allclose(RS.cross_cross_entropy_loss(a, b), g.t("loss.cce"), **TOL)
    assert torch.allclose(RS.cross_cross_entropy_loss(b, a), g.t("loss.of"), **TOL)  # output_fidelity(orig, expl) swaps
    assert torch.allclose(RS.concept_regularization_loss(g.t("phi")), g.t("loss.cr"), **TOL)
    assert torch.allclose(RS.concept_orthogonality_loss(g.t("psi.conv5.weight")), g.t("loss.ortho"), **TOL)
    rec = RS.reconstruction_loss(g.t("img_a"), g.t("img_b"), a, b, lambda x, y: torch.zeros(()), 0.1, 3.0)
    assert torch.allclose(rec, g.t("loss.rec_nolpips"), **TOL)
    ws, w_avg = g.t("ws"), g.t("w_avg")
    gan = torch.nn.functional.mse_loss(ws, w_avg.repeat([ws.shape[0], ws.shape[1], 1]))
    assert torch.allclose(gan, g.t("loss.gan"), **TOL)


def test_product_modules_match_reference(golden):
    from viscoin_next_b200 import viscoin_models as VM
    from viscoin_next_b200 import viscoin_step as VS
    g = golden("step")
    psi = VM.ConceptExtractor(n_concepts=8, latent_channels=16, input_channels1=16, input_channels2=32, input_channels3=64)
    psi.load_state_dict(g.sub("psi."), strict=True)
    theta = VM.Explainer(n_concepts=8, n_classes=10).eval()
    theta.load_state_dict(g.sub("theta."), strict=True)
    h = [g.t("h%d" % i).requires_grad_(True) for i in range(3)]
    phi, phi2 = psi(tuple(h))
    assert torch.allclose(phi, g.t("phi"), **TOL) and torch.allclose(phi2, g.t("phi2"), **TOL)
    assert torch.allclose(theta(phi), g.t("expl"), **TOL)
    dh = torch.autograd.grad([phi, phi2], h, [g.t("dphi"), g.t("dphi2")])
    for i in range(3):
        assert torch.allclose(dh[i], g.t("dh%d" % i), rtol=1e-4, atol=1e-6)
    a, b = g.t("cls_a"), g.t("cls_b")
    assert torch.allclose(VS.cross_cross_entropy_loss(a, b), g.t("loss.cce"), **TOL)
    assert torch.allclose(VS.output_fidelity_loss(a, b), g.t("loss.of"), **TOL)
    assert torch.allclose(VS.concept_regularization_loss(g.t("phi")), g.t("loss.cr"), **TOL)
    assert torch.allclose(VS.concept_orthogonality_loss(psi), g.t("loss.ortho"), **TOL)
    rec = VS.reconstruction_loss(g.t("img_a"), g.t("img_b"), a, b, lambda x, y: torch.zeros(()), 0.1, 3.0)
    assert torch.allclose(rec, g.t("loss.rec_nolpips"), **TOL)
    model = types.SimpleNamespace(mapping=types.SimpleNamespace(fixed_w_avg=g.t("w_avg")))
    assert torch.allclose(VS.gan_regularization_loss(g.t("ws"), model), g.t("loss.gan"), **TOL)


def test_lpips_product_matches_oracle():
    """LPIPS (restated from the published algorithm on both sides; the real package is absent): product module vs oracle."""
    from viscoin_next_b200 import viscoin_models as VM
    torch.manual_seed(3)
    m = VM.LPIPS()
    convs = []
    for s in m.slices:
        for layer in s:
            if isinstance(layer, torch.nn.Conv2d):
                convs.append((layer.weight.detach(), layer.bias.detach()))
    assert len(convs) == 13
    lins = [p.detach() for p in m.lins]
    a, b = torch.randn(2, 3, 32, 32), torch.randn(2, 3, 32, 32)
    got = m(a, b)
    want = RS.lpips_vgg(convs, lins, a, b)
    assert got.shape == want.shape == (2, 1, 1, 1)
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-7)
