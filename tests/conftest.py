import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu under gpurun)")


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle_lib():
    """The CPU oracle (test infrastructure only)."""
    odir = os.path.join(ROOT, "oracle")
    so = os.path.join(odir, "liboracle_vsann.so")
    src = os.path.join(odir, "oracle_vsann.cpp")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.run(["make", "-C", odir], check=True, capture_output=True)
    import vector_store_text_b200 as vst
    return vst.load_library(so, prefix="ovs_")


@pytest.fixture(scope="session")
def vst():
    import vector_store_text_b200 as m
    return m


def make_data(n, d, seed, clustered=False, centres=64):
    rng = np.random.default_rng(seed)
    if not clustered:
        return rng.standard_normal((n, d), dtype=np.float32)
    c = rng.standard_normal((centres, d), dtype=np.float32)
    a = rng.integers(0, centres, n)
    return (c[a] + 0.3 * rng.standard_normal((n, d), dtype=np.float32)).astype(np.float32)


def recall_at_k(found_keys, truth_keys):
    hits = 0
    for f, t in zip(found_keys, truth_keys):
        hits += len(set(f.tolist()) & set(t.tolist()))
    return hits / truth_keys.size
