"""CPU: libbrag.so loads, exports every symbol include/brag.h declares, the ctypes table covers
the header, and the engine refuses to come up without a B200 (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "brag.h")


def header_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(brag_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    from rag_application_b200 import _native
    if not os.path.exists(_native.LIB_PATH):
        from rag_application_b200.build import build_native
        build_native()
    return _native.load()


def test_header_declares_the_expected_surface():
    syms = header_symbols()
    for need in ("brag_create", "brag_destroy", "brag_last_error", "brag_dense_set", "brag_sparse_set_csr",
                 "brag_mask_set", "brag_dense_topk", "brag_dense_rescore", "brag_sparse_topk", "brag_rrf",
                 "brag_merge_topk", "brag_hybrid_topk"):
        assert need in syms


def test_library_exports_every_header_symbol(lib):
    for s in header_symbols():
        assert hasattr(lib, s), f"libbrag.so lacks {s}"


def test_ctypes_table_matches_header():
    from rag_application_b200 import _native
    assert sorted(_native.PROTOTYPES) == header_symbols()


def test_prototype_arity_matches_header():
    """Argument counts in the ctypes table equal the header's parameter counts."""
    from rag_application_b200 import _native
    src = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for name, (_, args) in _native.PROTOTYPES.items():
        m = re.search(r"\b%s\s*\(([^;]*?)\)\s*;" % name, src, flags=re.S)
        assert m, name
        params = m.group(1).strip()
        n = 0 if params in ("", "void") else params.count(",") + 1
        assert n == len(args), f"{name}: header has {n} parameters, ctypes table {len(args)}"


def test_version_and_no_cpu_fallback(lib):
    assert lib.brag_version() >= 1
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu tests")
    h = C.c_void_p()
    rc = lib.brag_create(0, C.byref(h))
    assert rc != 0 and not h.value
    msg = lib.brag_last_error().decode()
    assert "no CPU fallback" in msg or "CUDA" in msg
    from rag_application_b200 import Engine
    from rag_application_b200._native import BragError
    with pytest.raises(BragError):
        Engine(0)


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under rag_application_b200/ may reference it."""
    pkg = os.path.join(ROOT, "rag_application_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "libbm25_cpu" not in txt, f


def test_every_tuning_knob_is_documented():
    """Every environment knob the library or the host code reads (getenv / os.environ) is listed in DESIGN.md."""
    import re
    pkg = os.path.join(ROOT, "rag_application_b200")
    knobs = set()
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                knobs |= set(re.findall(r'getenv\("(BRAG_[A-Z0-9_]+)"\)', txt))
                knobs |= set(re.findall(r'environ(?:\.get)?[\[(]\s*"(BRAG_[A-Z0-9_]+)"', txt))
    assert knobs, "no knobs found: the patterns above no longer match the code"
    design = open(os.path.join(ROOT, "DESIGN.md")).read()
    missing = sorted(k for k in knobs if k not in design)
    assert not missing, f"undocumented knobs: {missing}"
