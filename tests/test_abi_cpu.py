"""
The drop-in boundary without a GPU: libpyccb200.so loads, exports exactly the entry points include/pyccb200.h
declares (and pycc_b200/_lib.py binds), and the product path fails loudly - status PCCB_ERR_NODEV with a message
from the C side, EngineError from the Python side - instead of falling back to anything on the CPU.
No compute entry point is called here.
"""
import ctypes
import os
import re
import shutil
import subprocess

import pytest
import torch

from pycc_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pyccb200.h")
PCCB_ERR_NODEV = -3          # include/pyccb200.h, the status enum


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    return re.findall(r"^\s*(?:int|int64_t|const\s+char\s*\*)\s+(pccb_[a-z0-9_]+)\s*\(", src, flags=re.M)


def test_header_and_binding_table_agree():
    names = _declared()
    assert len(names) == len(set(names)) and len(names) >= 30
    assert set(names) == set(_lib.EXPORTS)
    assert "extern \"C\"" in open(HEADER).read()          # a C ABI for C++ callers as well


def test_library_exports_every_declared_symbol():
    assert os.path.exists(_lib.LIB_PATH), "build the library first: __graft_entry__.build()"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared():
        getattr(lib, name)                                  # AttributeError = the symbol is missing
    lib.pccb_abi_version.restype = ctypes.c_int
    assert lib.pccb_abi_version() == _lib.ABI_VERSION
    # and nothing beyond the header is exported under the library's prefix
    if shutil.which("nm"):
        out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True, check=True).stdout
        exported = {line.split()[-1] for line in out.splitlines() if " T pccb_" in line}
        assert exported == set(_declared())


def test_signatures_carry_no_torch_types():
    """Plain pointers and sizes only: the header includes nothing but the C standard headers."""
    src = open(HEADER).read()
    includes = re.findall(r"#include\s+[<\"]([^>\"]+)[>\"]", src)
    assert includes and all(h in ("stddef.h", "stdint.h") for h in includes), includes
    code = re.sub(r"//[^\n]*", "", re.sub(r"/\*.*?\*/", "", src, flags=re.S))          # declarations without the comments
    assert "torch" not in code.lower() and "at::" not in code and "std::" not in code


@pytest.mark.skipif(torch.cuda.is_available(), reason="needs a machine WITHOUT a CUDA device")
def test_no_device_is_an_error_not_a_fallback():
    lib = _lib.load()
    assert lib.pccb_device_count() == 0
    ctx = ctypes.c_void_p()
    rc = lib.pccb_create(0, ctypes.byref(ctx))
    assert rc == PCCB_ERR_NODEV and not ctx.value
    lib.pccb_last_error.restype = ctypes.c_char_p
    assert b"no CPU fallback" in lib.pccb_last_error(None)
    from pycc_b200.backend import CudaBackend
    with pytest.raises(_lib.EngineError, match="no CPU fallback"):
        CudaBackend()
    # the public constructor goes through the same back end
    import numpy as np
    import pycc_b200
    from pycc_b200.synthetic import SyntheticSystem
    s = SyntheticSystem(2, 3, seed=1)
    with pytest.raises(_lib.EngineError, match="no CPU fallback"):
        pycc_b200.ccwfn.from_arrays(s.F, s.blocks, 2)
    assert np.isfinite(s.F).all()


def test_product_never_imports_the_oracle():
    """oracle/ is the checker (tests, smoke(), bench.py's cpu_baseline / --impl reference legs): nothing under
    pycc_b200/ may import, load or execute it, and no module there falls back to NumPy einsum for a contraction."""
    pkg = os.path.join(ROOT, "pycc_b200")
    bad = []
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if not f.endswith(".py"):
                continue
            text = open(os.path.join(dirpath, f)).read()
            for n, line in enumerate(text.splitlines(), 1):
                code = line.split("#", 1)[0]
                if re.search(r"^\s*(from|import)\s+oracle\b|import_module\(\s*[\"']oracle|[\"']oracle[./]", code):
                    bad.append("%s:%d: %s" % (f, n, line.strip()))
                if re.search(r"\b(np|numpy|torch)\.einsum\(", code) and f not in ("synthetic.py",):
                    bad.append("%s:%d: %s" % (f, n, line.strip()))
    assert not bad, "\n".join(bad)
