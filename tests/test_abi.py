"""The C-ABI library builds, loads, and exports exactly the symbols include/probtalk_b200.h declares
(no compute calls: runs without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "probtalk_b200.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(pt_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    from probtalk_b200 import build
    path = build.build()
    return ctypes.CDLL(path)


def test_header_symbols_exported(lib):
    syms = declared_symbols()
    assert len(syms) >= 12
    for s in syms:
        assert hasattr(lib, s), "symbol %s declared in the header is not exported" % s


def test_ctypes_table_matches_header():
    from probtalk_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    # argument counts agree with the header prototypes
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    for name, (_, args) in _lib.SIGNATURES.items():
        m = re.search(r"\b%s\s*\(([^;]*?)\)\s*;" % name, text, flags=re.S)
        assert m, name
        params = [p for p in m.group(1).split(",") if p.strip() and p.strip() != "void"]
        assert len(params) == len(args), "%s: header has %d params, ctypes table %d" % (name, len(params), len(args))


def test_version_and_error_text(lib):
    lib.pt_version.restype = ctypes.c_int
    lib.pt_last_error.restype = ctypes.c_char_p
    assert lib.pt_version() >= 100
    assert isinstance(lib.pt_last_error(), bytes)


def test_gemm_desc_layout_matches_header():
    """ctypes mirror of pt_gemm_desc: field order and count follow the header struct."""
    from probtalk_b200._lib import GemmDesc
    text = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    body = re.search(r"typedef struct pt_gemm_desc \{(.*?)\} pt_gemm_desc;", text, flags=re.S).group(1)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):
            names.append(re.findall(r"[A-Za-z_][A-Za-z0-9_]*", part)[-1])
    assert names == [f[0] for f in GemmDesc._fields_]


def test_argument_validation_without_gpu(lib):
    """Entry points reject bad arguments before touching the device."""
    lib.pt_pq_search.restype = ctypes.c_int
    lib.pt_last_error.restype = ctypes.c_char_p
    rc = lib.pt_pq_search(None, ctypes.c_longlong(0), ctypes.c_longlong(0), ctypes.c_longlong(0), 1, 1, 4, 64, 96, None,
                          None, None, None, None, None)
    assert rc != 0 and b"D must be 64 or 128" in lib.pt_last_error()
    rc = lib.pt_pq_search(None, ctypes.c_longlong(0), ctypes.c_longlong(0), ctypes.c_longlong(0), 1, 1, 4, 100, 64, None,
                          None, None, None, None, None)
    assert rc != 0 and b"multiple of 16" in lib.pt_last_error()
    rc = lib.pt_pq_search(None, ctypes.c_longlong(0), ctypes.c_longlong(0), ctypes.c_longlong(0), 1, 1, 1, 2048, 64, None,
                          None, None, None, None, None)
    assert rc != 0 and b"null pointer" in lib.pt_last_error()          # G = 1, K = 2048, D = 64 is a valid shape
    lib.pt_attention.restype = ctypes.c_int
    rc = lib.pt_attention(ctypes.c_void_p(16), ctypes.c_longlong(8), ctypes.c_void_p(16), ctypes.c_longlong(8),
                          ctypes.c_void_p(16), ctypes.c_longlong(8), ctypes.c_void_p(16), ctypes.c_longlong(8), 1, 8, 32,
                          4, 4, ctypes.c_float(1.0), None)
    assert rc != 0 and b"head_dim" in lib.pt_last_error()


def test_no_cpu_fallback():
    import torch
    from probtalk_b200 import ops
    from probtalk_b200._lib import PtError
    with pytest.raises(PtError):
        ops.pq_search(torch.zeros(1, 512, 2), torch.zeros(4, 128, 128))
