import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HERE = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, HERE):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built_libraries():
    """Both shared libraries are built in-tree before any test (and travel to the GPU box prebuilt)."""
    from oracle import orclib
    orclib.build()
    from dsd2dxd_b200 import _capi
    if not os.path.exists(_capi.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()


@pytest.fixture(scope="session")
def fixtures():
    """name -> (payload uint8 array, meta dict) of the committed fixture excerpts."""
    data = np.load(os.path.join(HERE, "golden", "fixtures.npz"))
    meta = json.load(open(os.path.join(HERE, "golden", "fixtures_meta.json")))
    return {k: (data[k], meta[k]) for k in meta}


@pytest.fixture(scope="session")
def ref_files(tmp_path_factory):
    """The reference's DSD fixtures, whole (tests/golden/ref_fixtures.tar.xz, made by make_ref_fixtures.sh):
    relative path -> absolute path of the unpacked file."""
    import tarfile
    root = tmp_path_factory.mktemp("ref_fixtures")
    with tarfile.open(os.path.join(HERE, "golden", "ref_fixtures.tar.xz"), "r:xz") as tf:
        tf.extractall(root, filter="data")
    out = {}
    for d, _, files in os.walk(root):
        for f in files:
            full = os.path.join(d, f)
            out[os.path.relpath(full, root)] = full
    return out


@pytest.fixture(scope="session")
def golden_digests():
    return json.load(open(os.path.join(HERE, "golden", "golden_digests.json")))


def have_gpu():
    try:
        import dsd2dxd_b200 as d2p
        return d2p.device_count() > 0
    except Exception:
        return False
