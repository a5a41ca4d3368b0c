"""The C-ABI library loads without a GPU and exports every symbol include/b200recon.h declares."""
import ctypes
import os
import re

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "b200recon.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2r_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    from b200recon import _lib
    names = _declared()
    assert len(names) >= 20
    lib = ctypes.CDLL(_lib.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_binding_covers_header():
    from b200recon import _lib
    assert sorted(_lib.SIGNATURES) == _declared()
    L = _lib.load()
    assert L.b2r_version() >= 100


def test_no_gpu_fails_loudly():
    """Without a CUDA device the product path must raise, never fall back to the CPU."""
    import torch
    if torch.cuda.is_available():
        return
    import pytest
    from b200recon import _lib, runtime
    with pytest.raises(_lib.B2RError):
        runtime.ctx(0)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "3d-reconstruction-planning_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "oracle" not in txt.lower() or f in ("synthetic.py", "raycast.cu"), (dp, f)
