"""CPU: the C-ABI shared library loads and exports every symbol the headers declare
(no compute call is made without a GPU)."""
import ctypes
import glob
import os
import re

import pytest

from cugraph_forked_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_functions(path):
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    text = re.sub(r"#if 0.*?#endif", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b((?:b200|cugraph)_[a-z0-9_]+)\s*\(", text)))


def test_library_is_built():
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() first"


def test_raw_abi_symbols_exported():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared_functions(os.path.join(ROOT, "include", "b200_cugraph.h"))
    assert len(names) >= 20
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    # and the ctypes table of the Python mirror covers exactly the header
    assert sorted(_lib.SIGNATURES) == names


def test_cugraph_c_symbols_exported():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    headers = glob.glob(os.path.join(ROOT, "include", "cugraph_c", "*.h"))
    if not headers:
        pytest.skip("cugraph_c layer not present")
    missing = []
    for hdr in headers:
        missing += [n for n in _declared_functions(hdr) if not hasattr(lib, n)]
    assert not missing, missing


def test_build_info_mentions_sm_100a():
    assert b"sm_100a" in _lib.lib().b200_build_info()


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libcugraph_b200.so")
    with pytest.raises(ImportError, match="no fallback"):
        _lib.lib()
