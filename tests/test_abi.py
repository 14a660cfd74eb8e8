"""CPU test: the C-ABI library loads and exports every symbol include/vsann.h declares
(no compute calls without a GPU), and fails loudly — no CPU fallback — when no device exists."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, has_gpu


def header_symbols():
    src = open(os.path.join(ROOT, "include", "vsann.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vs_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(vst):
    import importlib.util
    spec = importlib.util.spec_from_file_location("vsbuild", os.path.join(ROOT, "vector-store-text_b200", "build.py"))
    b = importlib.util.module_from_spec(spec); spec.loader.exec_module(b)
    path = b.build()
    lib = C.CDLL(path)
    syms = header_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in vsann.h but not exported"
    assert sorted(vst.ABI_SYMBOLS) == syms
    lib.vs_version.restype = C.c_char_p
    assert b"sm_100a" in lib.vs_version()


def test_oracle_exports_same_abi(oracle_lib):
    for s in header_symbols():
        if s in ("vs_add_batch_device", "vs_search_batch_device", "vs_search_batch_fanout_device", "vs_merge_topk_device", "vs_merge_topk_peers_device", "vs_merge_topk_flags_device", "vs_exact_search_noalloc_device", "vs_export_level"):
            continue   # device-pointer entry points have no CPU meaning
        if s in ("vs_index_save", "vs_index_load"):
            continue   # persistence is not part of the reference's algorithm (it has none): nothing to restate
        assert hasattr(oracle_lib, "o" + s), f"oracle misses o{s}"


@pytest.mark.skipif(has_gpu(), reason="checks the no-device failure mode")
def test_no_cpu_fallback(vst):
    with pytest.raises(vst.VsError) as e:
        vst.VsIndex(8)
    assert e.value.status == vst.VS_ERR_CUDA
    assert "no CPU fallback" in e.value.message
