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


def test_ctypes_mirrors_have_the_header_layout(tmp_path):
    """The Python mirrors of bltcu_opts and bltcu_stats must have the sizes and field offsets the C header gives them
    (a drifted mirror would make every option silently mean something else)."""
    import subprocess

    from blt_b200 import search

    fields_o = [f[0] for f in search.Opts._fields_]
    fields_s = [f[0] for f in search.Stats._fields_]
    src = tmp_path / "sz.c"
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "bltcu.h"', 'int main(void) {',
             '  printf("%zu %zu\\n", sizeof(bltcu_opts), sizeof(bltcu_stats));']
    lines += [f'  printf("%zu\\n", offsetof(bltcu_opts, {f}));' for f in fields_o]
    lines += [f'  printf("%zu\\n", offsetof(bltcu_stats, {f}));' for f in fields_s]
    lines += ['  return 0;', '}']
    src.write_text("\n".join(lines))
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)])
    out = subprocess.check_output([str(exe)], text=True).split()
    assert int(out[0]) == ctypes.sizeof(search.Opts) and int(out[1]) == ctypes.sizeof(search.Stats)
    offs = list(map(int, out[2:]))
    assert offs[:len(fields_o)] == [getattr(search.Opts, f).offset for f in fields_o]
    assert offs[len(fields_o):] == [getattr(search.Stats, f).offset for f in fields_s]


def test_reference_c_client_compiles_against_our_headers():
    """The reference's test_blt.c builds unchanged against include/blt.h (ENVIRONMENT_64 and all twelve prototypes);
    it is run on the GPU box by tests/test_gpu_blt_api.py."""
    from blt_b200 import build

    if not os.path.exists(build.REF_CLIENT_SRC):
        return
    exe = build.build_ref_client()
    assert exe and os.path.exists(exe)
    import subprocess

    syms = subprocess.check_output(["nm", "-D", "--undefined-only", exe], text=True)
    for nm in ("blt_init", "blt_var", "blt_assume", "blt_check", "blt_model", "blt_free"):
        assert nm in syms
