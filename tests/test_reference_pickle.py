"""Reference on-disk formats -> this package's modules (viscoin_next_b200.persistence).

CPU part: the committed fixtures are byte streams produced by the REAL reference (oracle/make_golden_pickle.py, build
container) with only the embedded ``module_src`` text replaced by a marker; they load without executing anything, into
this package's classes, with exactly the recorded parameters (pinned by tests/golden/networks.npz).  When oracle/_ref is
staged, unstripped pickles of the reference's own objects (incl. the VisCoIN whole-object bundle) are made on the fly and
round-tripped too.  GPU part: the loaded generators produce the reference's images."""

import io
import os
import pickle

import numpy as np
import pytest
import torch

from conftest import GOLDEN_DIR


def _P():
    import viscoin_next_b200 as V
    return V


def test_training_snapshot_pickle_loads_into_package_modules(golden):
    V = _P()
    snap = V.load_reference_pickle(os.path.join(GOLDEN_DIR, "ref_snapshot_stripped.pkl"))
    assert set(snap) == {"G", "D", "G_ema", "augment_pipe", "training_set_kwargs"}
    assert isinstance(snap["G_ema"], V.networks.Generator) and isinstance(snap["G"], V.networks.Generator)
    assert isinstance(snap["D"], V.persistence.OpaqueObject) and snap["D"].class_name == "Discriminator"  # not on the path
    g = golden("networks")
    want = g.sub("G.sd.")
    for name, G in (("G", snap["G"]), ("G_ema", snap["G_ema"])):
        sd = G.state_dict()
        assert set(sd) == set(want)
        assert all(torch.equal(sd[k], want[k]) for k in want), name
    assert not snap["G_ema"].training and not any(p.requires_grad for p in snap["G_ema"].parameters())
    assert snap["G"].training and all(p.requires_grad for p in snap["G"].parameters())
    assert snap["G_ema"].synthesis.b32.conv0.up == 2 and snap["G_ema"].mapping.num_layers == 2


def test_torch_saved_gan_pickle_and_fp16_kwargs(golden):
    V = _P()
    G = V.load_generator(os.path.join(GOLDEN_DIR, "ref_gan_stripped.pt"))
    want = golden("networks").sub("G.sd.")
    assert all(torch.equal(v, want[k]) for k, v in G.state_dict().items())
    G16 = V.load_generator(os.path.join(GOLDEN_DIR, "ref_gan_fp16_stripped.pt"))
    z = np.load(os.path.join(GOLDEN_DIR, "ref_pickles.npz"))
    assert all(np.array_equal(v.numpy(), z["fp16.sd." + k]) for k, v in G16.state_dict().items())
    # num_fp16_res = 2 at 32x32: blocks 16 and 32 are 16-bit (networks.py:442,448), conv_clamp reaches every layer
    assert [getattr(G16.synthesis, "b%d" % r).use_fp16 for r in (4, 8, 16, 32)] == [False, False, True, True]
    assert G16.synthesis.b32.conv1.conv_clamp == 256 and G16.synthesis.b32.torgb.conv_clamp == 256


def test_restricted_unpickler_refuses_foreign_globals():
    V = _P()

    class Evil:
        def __reduce__(self):
            return (os.system, ("true",))

    with pytest.raises(pickle.UnpicklingError):
        V.load_reference_pickle(io.BytesIO(pickle.dumps({"G_ema": Evil()})))


def test_unstripped_reference_objects_roundtrip():
    """Pickles of the reference's OWN objects made here (needs the staged reference): a persistent-class Generator with its
    embedded source, and VisCoIN's whole-object bundle (viscoin/models/utils.py:211-233)."""
    from oracle import stage_ref
    if not stage_ref.available():
        pytest.skip("oracle/_ref not staged")
    V = _P()
    ref = stage_ref.import_reference()
    torch.manual_seed(1)
    G = ref.networks.Generator(z_dim=16, c_dim=0, w_dim=16, img_resolution=16, img_channels=3, mapping_kwargs=dict(num_layers=2),
                               synthesis_kwargs=dict(channel_base=256, channel_max=16, architecture="orig"))
    buf = io.BytesIO()
    torch.save(G, buf)
    assert b"def modulated_conv2d" in buf.getvalue()  # the source really is in there - and is not executed:
    mine = V.load_generator(io.BytesIO(buf.getvalue()))
    assert type(mine).__module__.startswith("viscoin_next_b200") and mine.synthesis.b8.architecture == "orig"
    assert all(torch.equal(v, G.state_dict()[k]) for k, v in mine.state_dict().items())
    # the VisCoIN bundle: plain classes pickled by qualified name + the shared persistent synthesis network
    GA = ref.gan.GeneratorAdapted(z_dim=4, w_dim=16, img_resolution=16, synthesis_kwargs=dict(channel_base=256, channel_max=16))
    psi = ref.concept_extractors.ConceptExtractor(n_concepts=4, latent_channels=8, input_channels1=8, input_channels2=8, input_channels3=8)
    theta = ref.explainers.Explainer(n_concepts=4, n_classes=5)
    bundle = {"gan": GA, "concept_extractor": psi, "explainer": theta}
    buf = io.BytesIO()
    torch.save(bundle, buf)
    out = V.load_reference_pickle(io.BytesIO(buf.getvalue()))
    assert isinstance(out["gan"], V.networks.GeneratorAdapted) and isinstance(out["gan"].mapping, V.networks.MappingNetworkAdapted)
    assert isinstance(out["concept_extractor"], V.viscoin_models.ConceptExtractor) and isinstance(out["explainer"], V.viscoin_models.Explainer)
    for a, b in ((out["gan"], GA), (out["concept_extractor"], psi), (out["explainer"], theta)):
        sa, sb = a.state_dict(), b.state_dict()
        assert set(sa) == set(sb) and all(torch.equal(sa[k], sb[k]) for k in sa)
    # .pth checkpoint (viscoin/models/utils.py:177-208)
    ck = io.BytesIO()
    cls = torch.nn.Linear(2, 2)
    torch.save({"classifier": cls.state_dict(), "concept_extractor": psi.state_dict(), "explainer": theta.state_dict(), "gan": GA.state_dict()}, ck)
    ck.seek(0)
    fresh = V.networks.GeneratorAdapted(z_dim=4, w_dim=16, img_resolution=16, synthesis_kwargs=dict(channel_base=256, channel_max=16))
    V.load_viscoin_checkpoint(ck, torch.nn.Linear(2, 2), V.viscoin_models.ConceptExtractor(4, 8, 8, 8, 8), V.viscoin_models.Explainer(4, 5), fresh)
    assert all(torch.equal(v, GA.state_dict()[k]) for k, v in fresh.state_dict().items())


@pytest.mark.gpu
def test_loaded_generator_reproduces_reference_images(golden):
    V = _P()
    g = golden("networks")
    G = V.load_generator(os.path.join(GOLDEN_DIR, "ref_gan_stripped.pt"), map_location="cuda")
    img = G(g.t("G.z", device="cuda"), None, noise_mode="const")
    ref = g.t("G.full.img")
    e = float((img.cpu() - ref).abs().max() / ref.abs().max())
    assert e <= 2e-2, e
    # fp16-configured network: 16-bit blocks run the bf16-resident pipeline, conv_clamp in the epilogues; the target is the
    # reference's force_fp32 image (fp32 arithmetic, same clamps)
    z = np.load(os.path.join(GOLDEN_DIR, "ref_pickles.npz"))
    G16 = V.load_generator(os.path.join(GOLDEN_DIR, "ref_gan_fp16_stripped.pt"), map_location="cuda")
    img16 = G16(torch.from_numpy(z["fp16_z"]).cuda(), None, noise_mode="const")
    ref16 = torch.from_numpy(z["fp16_img"])
    e16 = float((img16.float().cpu() - ref16).abs().max() / ref16.abs().max())
    img32 = G16(torch.from_numpy(z["fp16_z"]).cuda(), None, noise_mode="const", force_fp32=True)
    e32 = float((img32.float().cpu() - ref16).abs().max() / ref16.abs().max())
    print("loaded generator: fp32 %.2e, fp16-config (bf16 blocks) %.2e, force_fp32 %.2e" % (e, e16, e32))
    assert e16 <= 4e-2 and e32 <= 2e-2
