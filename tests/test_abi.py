"""CPU-only checks of the drop-in boundary: the shared library loads without a GPU driver,
exports every symbol include/omnia_recall.h declares, and refuses to open a store with a clear
message instead of falling back to anything."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def _declared():
    text = (ROOT / "include" / "omnia_recall.h").read_text()
    return sorted(set(re.findall(r"OMR_API\s+[\w\s\*]+?\b(omr_\w+)\s*\(", text)))


def test_library_builds_and_exports_every_declared_symbol():
    from omnia_b200 import _lib, build
    build.build()
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"
    assert sorted(_lib.SIGNATURES) == names, "ctypes table and header disagree"


def test_header_cites_reference_for_every_entry_point():
    text = (ROOT / "include" / "omnia_recall.h").read_text()
    assert text.count("internal/memory") >= 12


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from omnia_b200 import GPUMemoryStore, RecallError
    with pytest.raises(RecallError, match="no CPU fallback"):
        GPUMemoryStore(16, 100)


def test_product_never_imports_oracle():
    for p in (ROOT / "omnia_b200").rglob("*"):
        if p.suffix in {".py", ".cu", ".cuh", ".cpp", ".h"}:
            assert "oracle" not in p.read_text().lower().replace("oracle's", ""), f"{p} mentions the oracle"


def test_bad_config_is_rejected_before_touching_cuda():
    from omnia_b200 import _lib as L
    lib = L.load()
    h = ctypes.c_void_p()
    cfg = L.Config(5000, 0, 10, 0, 1, 0, 0, 0, 0)
    assert lib.omr_open(ctypes.byref(cfg), ctypes.byref(h)) == L.E_DIM     # embedding_schema.go:64-68
    assert b"2000" in lib.omr_last_error()
    cfg = L.Config(16, 0, 10, 0, 1, 0, 48, 0, 0)
    assert lib.omr_open(ctypes.byref(cfg), ctypes.byref(h)) == L.E_INVALID
