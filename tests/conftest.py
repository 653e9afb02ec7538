"""pytest configuration: `gpu` marker, repo root on sys.path, golden-fixture loader."""

import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_cuda = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_cuda = False
    if has_cuda:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


class Golden:
    """Dictionary-like view of one tests/golden/*.npz file with prefix helpers."""

    def __init__(self, name):
        self._z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
        self.keys = list(self._z.keys())

    def __contains__(self, k):
        return k in self._z

    def np(self, k):
        return self._z[k]

    def t(self, k, **kw):
        import torch
        return torch.from_numpy(self._z[k]).clone().to(**kw) if k in self._z else None

    def sub(self, prefix):
        """{suffix: tensor} for every key that starts with prefix."""
        import torch
        return {k[len(prefix):]: torch.from_numpy(self._z[k]).clone() for k in self.keys if k.startswith(prefix)}


@pytest.fixture(scope="session")
def golden():
    cache = {}

    def load(name):
        if name not in cache:
            cache[name] = Golden(name)
        return cache[name]

    return load
