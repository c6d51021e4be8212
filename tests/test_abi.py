"""The C-ABI library loads and exports every symbol include/sspp_gpu.h declares (no compute calls)."""
import ctypes
import os
import re

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(REPO, "include", "sspp_gpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sspp_gpu_\w+)\s*\(", text)))


def test_header_symbols_exported(built):
    from sspp_b200.build import LIB_GPU
    lib = ctypes.CDLL(LIB_GPU)
    names = declared_symbols()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_engine_refuses_without_library(built, monkeypatch):
    """no CPU fallback: a missing CUDA library is an ImportError, not a silent slow path"""
    import sspp_b200.engine as engine
    monkeypatch.setattr(engine, "_lib", None)
    monkeypatch.setattr(engine, "LIB_GPU", "/nonexistent/libsspp_gpu.so")
    try:
        engine.load_library()
    except ImportError as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("load_library succeeded without the library")


def test_product_does_not_touch_oracle():
    """nothing under sspp_b200/ or include/ may reference the oracle"""
    bad = []
    for root in ("sspp_b200", "include"):
        for d, _, files in os.walk(os.path.join(REPO, root)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                    t = open(os.path.join(d, f), errors="ignore").read()
                    if re.search(r"oracle_lib|liboracle|libsspp_ref|octomap_restated|oracle/", t):
                        bad.append(os.path.join(d, f))
    assert not bad, bad
