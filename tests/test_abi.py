"""The C-ABI library loads on a machine without a GPU and exports every function include/*.h declares."""
import ctypes
import glob
import os
import re

from tests.conftest import ROOT


def _declared():
    names = []
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names += re.findall(r"\b((?:bltcu|blt)_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_header_symbols_exported():
    import __graft_entry__ as g

    g.build()
    libs = [ctypes.CDLL(p) for p in glob.glob(os.path.join(ROOT, "blt_b200", "lib", "*.so"))]
    assert libs
    names = _declared()
    assert "bltcu_closest" in names
    for nm in names:
        assert any(hasattr(l, nm) for l in libs), f"{nm} declared in include/ but not exported"


def test_no_gpu_fails_loudly():
    """Without a CUDA device the product returns BLT_STATUS_ERROR (no CPU fallback)."""
    import numpy as np
    import torch

    import blt_b200

    if torch.cuda.is_available():
        return
    r = blt_b200.closest(np.array([[1, 0], [-1, 2]]), np.array([[2, 3], [2, 3]]))
    assert r["rc"] == blt_b200.BLT_STATUS_ERROR
