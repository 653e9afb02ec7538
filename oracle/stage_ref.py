"""Stage the UNMODIFIED reference sources the secondary GPU baseline needs into the git-ignored ``oracle/_ref/``.

TEST / MEASUREMENT INFRASTRUCTURE ONLY.  ``/root/reference`` exists only in the build container; ``oracle/_ref/`` is
listed in ``.gitignore`` (the sources never enter this repository's history) but NOT in ``.gpurunignore``, so the copy
travels to the GPU box with the snapshot, exactly like the built ``libvcn_b200.so``.  What is staged (SURVEY 8(d)
"secondary GPU baseline": the reference's own ``impl='cuda'`` plugins + cuDNN convolutions on the same B200):

    stylegan2_ada/torch_utils/**          ops/{bias_act,upfirdn2d}.{py,cpp,cu,h}, conv2d_resample, conv2d_gradfix, fma,
                                          custom_ops (the JIT plugin loader), misc, persistence
    stylegan2_ada/dnnlib/**               EasyDict (bias_act.activation_funcs) + util
    stylegan2_ada/training/{__init__,networks}.py
    viscoin/models/{gan,concept_extractors,explainers}.py

Consumers: ``bench.py --impl reference-gpu``, ``tests/test_gpu_reference_dropin.py`` (patch.install() onto the real
modules on a B200), ``tests/test_reference_pickle.py`` (source-carrying pickles).  Nothing in the product imports it.

    python -m oracle.stage_ref            # (re)creates oracle/_ref/
"""

import os
import shutil
import sys

REF_ROOT = os.environ.get("VCN_REFERENCE_ROOT", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")

TREES = ["stylegan2_ada/torch_utils", "stylegan2_ada/dnnlib"]
FILES = ["stylegan2_ada/training/__init__.py", "stylegan2_ada/training/networks.py",
         "viscoin/models/gan.py", "viscoin/models/concept_extractors.py", "viscoin/models/explainers.py"]


def stage(verbose=True):
    """Copy the listed files.  Returns DEST, or None when the reference is not mounted (GPU box: uses what travelled)."""
    if not os.path.isdir(REF_ROOT):
        return DEST if os.path.isdir(DEST) else None
    for tree in TREES:
        dst = os.path.join(DEST, tree)
        if os.path.isdir(dst):
            shutil.rmtree(dst)
        shutil.copytree(os.path.join(REF_ROOT, tree), dst, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    for f in FILES:
        dst = os.path.join(DEST, f)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(os.path.join(REF_ROOT, f), dst)
    with open(os.path.join(DEST, "STAGED_FROM"), "w") as fh:
        fh.write("unmodified copies from %s made by oracle/stage_ref.py; git-ignored, never committed\n" % REF_ROOT)
    if verbose:
        n = sum(len(fs) for _, _, fs in os.walk(DEST))
        print("staged %d reference files into %s" % (n, os.path.relpath(DEST)))
    return DEST


def available():
    return os.path.isfile(os.path.join(DEST, "stylegan2_ada", "training", "networks.py"))


def import_reference():
    """Put the staged tree on sys.path the way viscoin/models/gan.py:18 does and import the reference modules.
    Returns a namespace with the op modules (top-level ``torch_utils.ops.*``: the ones the networks see, SURVEY section 1
    quirk), ``networks`` (= ``stylegan2_ada.training.networks``) and the VisCoIN model modules."""
    if not available():
        raise RuntimeError("oracle/_ref is not staged (run `python -m oracle.stage_ref` in the build container)")
    for p in (os.path.join(DEST, "stylegan2_ada"), DEST):
        if p not in sys.path:
            sys.path.insert(0, p)
    import types
    import warnings
    warnings.filterwarnings("ignore", category=UserWarning)
    from torch_utils.ops import bias_act, conv2d_resample, fma, upfirdn2d
    from viscoin.models import concept_extractors, explainers, gan
    import stylegan2_ada.training.networks as networks
    return types.SimpleNamespace(bias_act=bias_act, upfirdn2d=upfirdn2d, conv2d_resample=conv2d_resample, fma=fma,
                                 networks=networks, gan=gan, concept_extractors=concept_extractors, explainers=explainers)


def load_plugins(verbose=False):
    """JIT-build / load the reference's two CUDA plugins for sm_100a (custom_ops.py:46-124) and FAIL when they do not load
    (the reference would warn and fall back to its ref path, bias_act.py:47-51 - a silent fallback would void the baseline).
    The build directory is oracle/_ref/torch_extensions so that a build made once on a box is found again."""
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    dest = DEST
    canon = "/root/repo/oracle/_ref"  # the path the plugins were pre-built under (ninja re-builds when paths change);
    if os.path.isdir(canon) and os.path.realpath(canon) == os.path.realpath(DEST):  # on the GPU box /root/repo is a symlink
        dest = canon
    os.environ.setdefault("TORCH_EXTENSIONS_DIR", os.path.join(dest, "torch_extensions"))
    ref = import_reference()
    import torch.utils.cpp_extension as _ext
    from torch_utils import custom_ops
    custom_ops.verbosity = "full" if verbose else "none"
    # torch >= 2.x: cpp_extension.load() no longer registers the built module in sys.modules, so the reference's
    # `importlib.import_module(module_name)` (custom_ops.py:111) raises and the UNMODIFIED reference silently runs its
    # impl='ref' path on the GPU.  Making the build directories importable is an environment fix, not a source edit.
    for name in ("bias_act_plugin", "upfirdn2d_plugin"):
        d = _ext._get_build_directory(name, verbose=False)
        if d not in sys.path:
            sys.path.append(d)
    ok_b = ref.bias_act._init()
    ok_u = ref.upfirdn2d._init()
    if not (ok_b and ok_u and ref.bias_act._plugin is not None and ref.upfirdn2d._plugin is not None):
        raise RuntimeError("reference CUDA plugins failed to build/load (bias_act=%s upfirdn2d=%s)" % (ok_b, ok_u))
    return ref


if __name__ == "__main__":
    if stage() is None:
        raise SystemExit("%s not found and nothing staged" % REF_ROOT)
