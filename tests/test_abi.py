"""The C-ABI library loads and exports every symbol that include/ob200.h declares (no compute calls)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "ob200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ob200_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_surface():
    names = declared_functions()
    for must in ("ob200_ctx_create", "ob200_space_set", "ob200_sample_transform", "ob200_gp_fit", "ob200_score",
                 "ob200_score_host", "ob200_acq_argmax", "ob200_gp_export", "ob200_gp_import"):
        assert must in names


def test_library_builds_and_exports_every_declared_symbol():
    from metaopt_b200 import build
    lib_path = build.build(force=False)
    lib = ctypes.CDLL(lib_path)
    for name in declared_functions():
        assert hasattr(lib, name), "libob200.so does not export %s" % name


def test_ctypes_binding_covers_the_header():
    from metaopt_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_functions()
    lib = _lib.load()
    assert lib.ob200_version() >= 1


def test_struct_layouts_match_the_header():
    from metaopt_b200 import _lib
    assert ctypes.sizeof(_lib.Slot) == 72 and ctypes.sizeof(_lib.Best) == 32
    assert _lib.Slot.p0.offset == 24 and _lib.Slot.nhi.offset == 64


def test_missing_library_is_a_loud_error(monkeypatch, tmp_path):
    from metaopt_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        _lib.load()


def test_no_gpu_means_no_engine():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from metaopt_b200.engine import get_engine
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        get_engine(0)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "metaopt_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
