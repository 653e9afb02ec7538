"""CPU: the C-ABI library builds, loads and exports every symbol include/vcn_b200.h declares (no compute calls)."""

import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as G
    G._load_build_module().build()
    import viscoin_next_b200 as V
    return V._lib


def _declared():
    src = open(os.path.join(ROOT, "include", "vcn_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vcn_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(lib):
    names = _declared()
    assert "vcn_bias_act" in names and "vcn_upfirdn2d" in names
    cdll = ctypes.CDLL(lib.LIB_PATH)
    for n in names:
        assert hasattr(cdll, n), "libvcn_b200.so does not export %s" % n
    # the ctypes table binds exactly the declared set
    assert sorted(lib.SIGNATURES) == names


def test_abi_version_and_errors(lib):
    l = lib.load()
    assert l.vcn_abi_version() >= 1
    # argument validation happens before any CUDA call: usable without a GPU
    st = l.vcn_bias_act(None, None, None, None, None, None, None, None, 0, 0, 0, 3, 0.2, 1.0, -1.0, 16, 1, 1, None)
    assert st != 0 and b"NULL" in l.vcn_last_error()
    st = l.vcn_bias_act(None, None, None, None, None, None, None, None, 0, 0, 7, 3, 0.2, 1.0, -1.0, 16, 1, 1, None)
    assert st != 0


def test_no_cpu_fallback(lib):
    import torch
    import viscoin_next_b200 as V
    with pytest.raises(lib.VcnError):
        V.bias_act.bias_act(torch.zeros(2, 3), torch.zeros(3), impl="cuda")
    with pytest.raises(lib.VcnError):
        V.upfirdn2d.upfirdn2d(torch.zeros(1, 1, 8, 8), None, impl="cuda")
    # impl='ref' stays available on CPU (operator API of the reference)
    assert V.bias_act.bias_act(torch.ones(2, 3), torch.ones(3), impl="ref").sum().item() == 12.0


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "viscoin-next_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith(".py"):
                txt = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
