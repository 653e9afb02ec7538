"""viscoin-next_b200: B200-native (sm_100a) StyleGAN2-ADA synthesis hot path for VisCoIN-Next.

Only what the path needs lives here: ``csrc/`` (hand-written CUDA kernels behind the C ABI of
``include/vcn_b200.h``), ``_lib`` (ctypes binding, no fallback) and ``host/`` (the Python mirror of the
reference operator API: bias_act, upfirdn2d, conv2d_resample, fma, modulated_conv2d and the drop-in
installer).  Import as ``viscoin_next_b200`` (the hyphenated directory is not an identifier).
"""

from . import _lib  # noqa: F401
from .host import bias_act, conv2d_resample, dp, fma, lpips_tap, misc, modconv, networks, persistence, upfirdn2d, viscoin_models, viscoin_step  # noqa: F401
from .host.persistence import load_generator, load_reference_pickle, load_viscoin_checkpoint  # noqa: F401
from . import patch  # noqa: F401

__all__ = ["_lib", "bias_act", "conv2d_resample", "dp", "fma", "lpips_tap", "modconv", "load_generator", "load_reference_pickle", "load_viscoin_checkpoint", "misc", "networks", "patch", "persistence", "upfirdn2d", "viscoin_models", "viscoin_step"]
