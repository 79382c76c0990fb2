"""CPU: the C-ABI library loads and exports every symbol include/trase_b200.h
declares; host-only entry points behave (no compute calls without a GPU)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "trase_b200.h")).read()
    return sorted(set(re.findall(r"^TRASE_API [^\n(]*?\b(trase_b200_\w+)\(", txt, flags=re.M)))


def test_header_and_binding_list_agree():
    from trase_b200 import _lib
    assert header_symbols() == sorted(_lib.EXPORTS)


def test_library_exports_every_declared_symbol(native_lib):
    for sym in header_symbols():
        assert hasattr(native_lib, sym), sym


def test_version_and_error_string(native_lib):
    assert native_lib.trase_b200_version() >= 100
    assert native_lib.trase_b200_last_error() is not None


def test_workspace_layouts_are_aligned_and_ordered(native_lib):
    from trase_b200 import _lib
    for P in (0, 1, 255, 256, 257, 20_000, 300_000):
        off = _lib.geom_offsets(P)
        vals = list(off.values())
        assert vals == sorted(vals) and all(v % 256 == 0 for v in vals)
        assert native_lib.trase_b200_geom_bytes(P) >= vals[-1] + 32
    off = _lib.img_offsets(1352, 1014)
    assert off["n_contrib"] - off["final_T"] >= 1352 * 1014 * 4
    assert native_lib.trase_b200_img_bytes(1352, 1014) >= off["ranges"] + 85 * 64 * 8
    boff = _lib.binning_offsets(1000)
    assert boff["point_list_keys"] >= 4000 and boff["point_list_keys"] % 256 == 0
    assert native_lib.trase_b200_binning_bytes(1000) >= boff["point_list_keys"] + 8000
    assert native_lib.trase_b200_scratch_bytes(10, 1000) >= 12000 + native_lib.trase_b200_sort_scratch_bytes(1000)


def _ref_higher_msb(n):
    # the procedure of rasterizer_impl.cu:38-53, restated
    msb, step = 16, 16
    while step > 1:
        step //= 2
        msb = msb + step if (n >> msb) else msb - step
    return msb + 1 if (n >> msb) else msb


@pytest.mark.parametrize("n", [0, 1, 2, 3, 63, 64, 65, 2040, 5440, 8160, 65535, 65536, (1 << 31) - 1, 1 << 31])
def test_higher_msb_matches_reference_procedure(native_lib, n):
    assert native_lib.trase_b200_higher_msb(n) == _ref_higher_msb(n)


def test_argument_validation_without_gpu(native_lib):
    from trase_b200 import _lib
    a = _lib.FwdArgs()
    a.P, a.W, a.H = 10, 0, 16
    assert native_lib.trase_b200_rasterize_forward(C.byref(a)) != 0
    assert b"bad sizes" in native_lib.trase_b200_last_error()
    a.W = 16
    assert native_lib.trase_b200_rasterize_forward(C.byref(a)) != 0
    assert b"required" in native_lib.trase_b200_last_error()
    b = _lib.BwdArgs()
    b.P, b.W, b.H = 10, 16, 16
    assert native_lib.trase_b200_rasterize_backward(C.byref(b)) != 0
    assert native_lib.trase_b200_sort_pairs(None, None, None, None, None, 10, 40, None, 0, None) != 0
    assert native_lib.trase_b200_mark_visible(5, None, None, None, None, None) != 0


def test_missing_library_fails_loudly(monkeypatch):
    from trase_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libtrase_b200.so")
    with pytest.raises(RuntimeError, match="no CPU or PyTorch fallback"):
        _lib.load()


def test_sort_scratch_is_monotone_in_capacity(native_lib):
    """A workspace sized for `capacity` must serve every smaller sort (the depth sort of the P Gaussians borrows the
    one sized for the instance capacity), across the switch from 1024- to 4096-key tiles at 2^20 keys."""
    f = native_lib.trase_b200_sort_scratch_bytes
    caps = [1, 1000, 262_144, 300_000, 1 << 20, (1 << 20) + 1, 1_054_096, 2_000_000, 4_000_000, 1 << 22, 10_000_000]
    sizes = [f(c) for c in caps]
    assert sizes == sorted(sizes)
    assert native_lib.trase_b200_scratch_bytes(1_000_000, 1_054_096) >= f(1_000_000)
    assert native_lib.trase_b200_scratch_bytes(2_000_000, 1_500_000) >= native_lib.trase_b200_scratch_bytes(0, 2_000_000)
