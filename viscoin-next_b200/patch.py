"""patch - install the sm_100a operators into the reference's modules (the drop-in of SURVEY 8(b)).

The reference's networks resolve ``bias_act.bias_act`` / ``upfirdn2d.upsample2d`` / ``modulated_conv2d`` … as module
attributes at call time (networks.py:11-16 imports the op MODULES), so replacing attributes on the module objects is
enough; no reference file is edited.  Only modules that are already imported are patched (import the reference
first).  ``uninstall()`` restores the originals."""

import sys

from .host import bias_act as _bias_act
from .host import conv2d_resample as _conv2d_resample
from .host import fma as _fma
from .host import modconv as _modconv
from .host import upfirdn2d as _upfirdn2d

_saved = []

_TARGETS = {
    "bias_act": (_bias_act, ["bias_act"]),
    "upfirdn2d": (_upfirdn2d, ["upfirdn2d", "filter2d", "upsample2d", "downsample2d", "setup_filter"]),
    "fma": (_fma, ["fma"]),
    "conv2d_resample": (_conv2d_resample, ["conv2d_resample"]),
}


def _set(mod, name, value):
    _saved.append((mod, name, getattr(mod, name)))
    setattr(mod, name, value)


def install():
    """Returns the list of 'module.attr' names that were replaced."""
    done = []
    for prefix in ("torch_utils.ops.", "stylegan2_ada.torch_utils.ops."):
        for short, (src, names) in _TARGETS.items():
            mod = sys.modules.get(prefix + short)
            if mod is None:
                continue
            for n in names:
                if hasattr(mod, n) and getattr(mod, n) is not getattr(src, n):
                    _set(mod, n, getattr(src, n))
                    done.append(prefix + short + "." + n)
    for name in ("stylegan2_ada.training.networks", "training.networks"):
        mod = sys.modules.get(name)
        if mod is not None and hasattr(mod, "modulated_conv2d") and mod.modulated_conv2d is not _modconv.modulated_conv2d:
            _set(mod, "modulated_conv2d", _modconv.modulated_conv2d)
            done.append(name + ".modulated_conv2d")
    return done


def uninstall():
    while _saved:
        mod, name, value = _saved.pop()
        setattr(mod, name, value)
