"""CPU: the C-ABI shared library builds, loads, and exports every symbol include/avdiff_b200.h declares.
No compute call is made here (there is no GPU in the build container)."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def header_symbols():
    text = (ROOT / "include" / "avdiff_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(avd_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_bound_symbols(lib):
    from multimodal_diffusion_b200 import _lib
    assert header_symbols() == sorted(_lib.EXPORTED_SYMBOLS)


def test_library_exports_every_declared_symbol(lib):
    raw = C.CDLL(str(ROOT / "multimodal_diffusion_b200" / "libavdiff_b200.so"))
    for name in header_symbols():
        assert hasattr(raw, name), f"{name} is declared in include/avdiff_b200.h but not exported"


def test_struct_layouts_match(lib):
    from multimodal_diffusion_b200 import _lib
    for i, st in enumerate(_lib.ABI_STRUCTS):
        assert lib.avd_struct_size(i) == C.sizeof(st), st.__name__
    assert lib.avd_struct_size(99) == -1
    assert lib.avd_version() >= 100
    assert [lib.avd_nseg(p) for p in (1, 2, 3, 6, 4)] == [1, 1, 2, 3, -1]
    assert lib.avd_prof_num_kinds() >= 8 and lib.avd_prof_kind_name(2) == b"gemm_tcgen05"


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    """The product path has no fallback: without the .so every operator raises AvdLibraryError."""
    from multimodal_diffusion_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "libavdiff_b200.so")
    with pytest.raises(_lib.AvdLibraryError):
        _lib.load()


def test_product_never_imports_the_oracle():
    pkg = ROOT / "multimodal_diffusion_b200"
    for py in pkg.rglob("*.py"):
        src = py.read_text()
        assert "oracle" not in re.sub(r'""".*?"""', "", src, flags=re.S).replace("# oracle", ""), py
    for cu in (pkg / "csrc").glob("*"):
        assert "oracle" not in cu.read_text(), cu
