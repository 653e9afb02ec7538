"""Generate reference-format pickle fixtures for viscoin_next_b200.persistence.

TEST INFRASTRUCTURE ONLY; runs ONLY in the build container (imports the real reference from /root/reference).

The reference's network pickles embed the full source text of ``training/networks.py`` (persistence.py:118-127).  The
fixtures committed here must not carry reference sources, so every ``module_src`` field is replaced by a one-line marker
before the bytes are written - the byte stream is otherwise exactly what ``pickle.dump`` / ``torch.save`` of the
reference objects produce (``_reconstruct_persistent_obj(meta)`` reduce calls, ``_init_args/_init_kwargs``, tensors).
``tests/test_reference_pickle.py`` additionally round-trips UNSTRIPPED pickles made on the fly when ``oracle/_ref`` is
staged.

    python -m oracle.make_golden_pickle     # writes tests/golden/ref_snapshot_stripped.pkl, ref_gan_stripped.pt,
                                            #        ref_gan_fp16_stripped.pt, ref_pickles.npz
"""

import copy
import io
import os
import pickle
import sys

import numpy as np
import torch

from .make_golden import OUT_DIR, REF_ROOT, _import_reference, _randomize

MARKER = "# module source removed by oracle/make_golden_pickle.py (the loader under test never executes it)\n"


def _strip(data_bytes, persistence, zip_format):
    """Re-serialise a reference pickle with every persistent object's `module_src` replaced by MARKER."""

    class Stripped:
        def __init__(self, meta):
            self.meta = meta

        def __reduce__(self):
            return (persistence._reconstruct_persistent_obj, (self.meta,))

    def rebuild(meta):
        meta = dict(meta)
        meta["module_src"] = MARKER
        return Stripped(meta)

    class U(pickle.Unpickler):
        def find_class(self, module, name):
            if name == "_reconstruct_persistent_obj":
                return rebuild
            return super().find_class(module, name)

    class PM:
        __name__ = "strip"
        Unpickler = U
        Pickler = pickle.Pickler
        load = staticmethod(lambda f, **kw: U(f, **kw).load())

    buf = io.BytesIO()
    if zip_format:
        obj = torch.load(io.BytesIO(data_bytes), pickle_module=PM, weights_only=False)
        torch.save(obj, buf)
    else:
        obj = U(io.BytesIO(data_bytes)).load()
        pickle.dump(obj, buf)
    out = buf.getvalue()
    assert b"def modulated_conv2d" not in out and MARKER.encode() in out
    return out


def main():
    bias_act, upfirdn2d, conv2d_resample, fma, networks = _import_reference()
    from torch_utils import persistence
    golden = np.load(os.path.join(OUT_DIR, "networks.npz"))
    # the SAME generator as golden_networks() (make_golden.py:203-210): its parameters are pinned in networks.npz
    torch.manual_seed(99)
    g = torch.Generator().manual_seed(99)
    G = networks.Generator(z_dim=32, c_dim=0, w_dim=32, img_resolution=32, img_channels=3, mapping_kwargs=dict(num_layers=2),
                           synthesis_kwargs=dict(channel_base=512, channel_max=16))
    _randomize(G, g)
    for k, v in G.state_dict().items():
        assert np.array_equal(golden["G.sd." + k], v.numpy()), k
    G_ema = copy.deepcopy(G).eval().requires_grad_(False)
    D = networks.Discriminator(c_dim=0, img_resolution=32, img_channels=3, channel_base=512, channel_max=16)
    snap = io.BytesIO()
    pickle.dump(dict(G=G, D=D, G_ema=G_ema, augment_pipe=None, training_set_kwargs=dict(path="none")), snap)  # training_loop.py:361-371
    with open(os.path.join(OUT_DIR, "ref_snapshot_stripped.pkl"), "wb") as fh:
        fh.write(_strip(snap.getvalue(), persistence, zip_format=False))
    gan = io.BytesIO()
    torch.save(G_ema, gan)  # how VisCoIN's gan-cub.pkl is stored (viscoin/cli/train.py:204 reads it with torch.load)
    with open(os.path.join(OUT_DIR, "ref_gan_stripped.pt"), "wb") as fh:
        fh.write(_strip(gan.getvalue(), persistence, zip_format=True))

    # a `--cfg=auto`-style network: num_fp16_res + conv_clamp=256 (stylegan2_ada/train.py:181-182)
    torch.manual_seed(7)
    g = torch.Generator().manual_seed(7)
    G16 = networks.Generator(z_dim=32, c_dim=0, w_dim=32, img_resolution=32, img_channels=3, mapping_kwargs=dict(num_layers=2),
                             synthesis_kwargs=dict(channel_base=512, channel_max=16, num_fp16_res=2, conv_clamp=256))
    _randomize(G16, g)
    G16.eval().requires_grad_(False)
    z = torch.randn(3, 32, generator=g)
    img32 = G16(z, None, noise_mode="const", force_fp32=True)  # fp32 arithmetic with the clamp: the parity target
    buf = io.BytesIO()
    torch.save(G16, buf)
    with open(os.path.join(OUT_DIR, "ref_gan_fp16_stripped.pt"), "wb") as fh:
        fh.write(_strip(buf.getvalue(), persistence, zip_format=True))
    np.savez_compressed(os.path.join(OUT_DIR, "ref_pickles.npz"), fp16_z=z.numpy(), fp16_img=img32.numpy(),
                        **{"fp16.sd." + k: v.numpy() for k, v in G16.state_dict().items()})
    for f in ("ref_snapshot_stripped.pkl", "ref_gan_stripped.pt", "ref_gan_fp16_stripped.pt", "ref_pickles.npz"):
        print("  wrote tests/golden/%s (%.1f KiB)" % (f, os.path.getsize(os.path.join(OUT_DIR, f)) / 1024))


if __name__ == "__main__":
    main()
