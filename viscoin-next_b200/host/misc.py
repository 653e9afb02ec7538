"""misc - the two helpers of ``stylegan2_ada/torch_utils/misc.py`` the hot path uses: ``profiled_function``
(misc.py:98-103: wraps a function in ``torch.autograd.profiler.record_function`` so that torch-profiler / NVTX traces
of the patched reference show the same range names: ``modulated_conv2d``, ``conv2d_resample``, ``_bias_act_ref``,
``_upfirdn2d_ref``, ``normalize_2nd_moment`` - networks.py:20,26, conv2d_resample.py:58, bias_act.py:92, upfirdn2d.py:167)
and ``assert_shape`` (misc.py:80-93)."""

import functools

import torch


def profiled_function(fn):
    @functools.wraps(fn)
    def decorator(*args, **kwargs):
        with torch.autograd.profiler.record_function(fn.__name__):
            return fn(*args, **kwargs)
    return decorator


def assert_shape(tensor, ref_shape):
    if tensor.ndim != len(ref_shape):
        raise AssertionError(f"Wrong number of dimensions: got {tensor.ndim}, expected {len(ref_shape)}")
    for idx, (size, ref_size) in enumerate(zip(tensor.shape, ref_shape)):
        if ref_size is None:
            continue
        if size != ref_size:
            raise AssertionError(f"Wrong size for dimension {idx}: got {size}, expected {ref_size}")
