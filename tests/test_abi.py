"""The C-ABI library loads and exports every symbol include/clsm_b200.h declares (CPU; no
compute calls).  Also checks that the Python binding table covers the header exactly."""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def header_symbols():
    src = open(os.path.join(ROOT, "include", "clsm_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(clsm_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from dnn_page_vectors_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = header_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), "missing export: " + n


def test_binding_table_matches_header():
    from dnn_page_vectors_b200 import _lib
    assert sorted(_lib.SIGNATURES) == header_symbols()


def test_version_and_error_string_without_gpu():
    from dnn_page_vectors_b200 import _lib
    assert _lib.lib.clsm_version() >= 100
    # argument validation happens before any CUDA call, so this is safe on a CPU-only box
    rc = _lib.lib.clsm_emb_gather(None, 4, None, 0, 0, 0, 7, None, 0, 1.0, 0, None)
    assert rc == _lib.EINVAL
    assert b"dtype" in _lib.lib.clsm_last_error()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "dnn_page_vectors_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(dirpath, f), encoding="utf-8").read()
                assert re.search(r"^\s*(from|import)\s+oracle", text, flags=re.M) is None, f
