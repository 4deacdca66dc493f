import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import uuo_mocap_b200  # noqa: E402,F401  (import shim for the hyphenated package directory)

GOLD = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


_MODELS = {}


@pytest.fixture(scope="session")
def get_model():
    from uuo_mocap_b200 import synthetic

    def _get(kind):
        if kind not in _MODELS:
            _MODELS[kind] = synthetic.make_model(kind)
        return _MODELS[kind]
    return _get


@pytest.fixture(scope="session")
def golden():
    def _load(name):
        return np.load(os.path.join(GOLD, name), allow_pickle=False)
    return _load


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(1e-30, np.abs(b).max()))


def frame_rel_err(a, b):
    """Per-frame-normalised error of [F, ...] arrays: max over frames f of max|a_f - b_f| / max|b_f|.  Stricter than
    ``rel_err`` (one normaliser for the whole tensor): a frame whose values are small cannot hide behind a large one."""
    a = np.asarray(a, dtype=np.float64).reshape(np.shape(a)[0], -1)
    b = np.asarray(b, dtype=np.float64).reshape(np.shape(b)[0], -1)
    num = np.abs(a - b).max(axis=1)
    den = np.maximum(np.abs(b).max(axis=1), 1e-30)
    return float((num / den).max())
