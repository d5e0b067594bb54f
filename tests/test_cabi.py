"""CPU: the C-ABI library loads and exports every symbol include/pings_b200.h declares (no compute calls)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "pings_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pg_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from pings_b200 import _lib
    L = _lib.lib()
    names = declared_symbols()
    assert len(names) >= 15
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing
    assert L.pg_abi_version() >= 2


def test_workspace_size_queries_are_pure_host_calls():
    from pings_b200 import _lib
    L = _lib.lib()
    g = L.pg_raster_geom_bytes(0, 1000)
    i = L.pg_raster_image_bytes(0, 640, 480)
    b = L.pg_raster_binning_bytes(0, 640, 480, 12345)
    assert g > 1000 * 48 and i > 640 * 480 * 8 and b > 12345 * 6


def test_missing_library_fails_loudly(monkeypatch):
    from pings_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libpings_b200.so")
    try:
        _lib.lib()
    except _lib.PingsB200LibraryMissing as e:
        assert "no CPU fallback" in str(e)
    else:
        raise AssertionError("expected PingsB200LibraryMissing")


def test_cpu_tensors_are_rejected():
    import pytest
    import torch
    from pings_b200.neural_points import NeuralPoints, default_config
    npm = NeuralPoints(default_config(device="cpu", buffer_size=1000))
    with pytest.raises(RuntimeError):
        npm.radius_neighborhood_search(torch.zeros(4, 3))
