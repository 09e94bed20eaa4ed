"""Builds blt_b200/lib/libbltcu.so with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = [os.path.join(HERE, "csrc", "search.cu"), os.path.join(HERE, "csrc", "peaks.cu"),
       os.path.join(HERE, "csrc", "lll_batch.cu"), os.path.join(HERE, "csrc", "host", "lll.cpp"),
       os.path.join(HERE, "csrc", "host", "sound.cpp")]
HOST_SRC = [os.path.join(HERE, "csrc", "host", f) for f in ("context.cpp", "blt_api.cpp")]
HOST_HDR = [os.path.join(HERE, "csrc", "host", "context.h"), os.path.join(HERE, "..", "include", "blt.h"),
            os.path.join(HERE, "..", "include", "codes.h")]
LIBBLT = os.path.join(HERE, "lib", "libblt.so")
RUN_BLT = os.path.join(HERE, "lib", "run_blt")
OUT = os.path.join(HERE, "lib", "libbltcu.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def _obj(src: str) -> str:
    return os.path.join(HERE, "lib", "obj", os.path.basename(src) + ".o")


def _compile(src: str, verbose: bool) -> None:
    cmd = ["nvcc"] + [f for f in NVCC_FLAGS if f != "-shared"] + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", _obj(src), src]
    subprocess.check_call(cmd)


def build(force: bool = False, verbose: bool = False) -> str:
    """Every source is compiled to its own object (in parallel, only when stale), then linked into libbltcu.so."""
    from concurrent.futures import ThreadPoolExecutor

    os.makedirs(os.path.join(HERE, "lib", "obj"), exist_ok=True)
    hdrs = [os.path.join(HERE, "..", "include", "bltcu.h"), os.path.join(HERE, "csrc", "lll_fixed.h"), os.path.join(HERE, "csrc", "lll_fp.h")]
    newest_hdr = max(os.path.getmtime(h) for h in hdrs)
    stale = [src for src in SRC if force or not os.path.exists(_obj(src))
             or os.path.getmtime(_obj(src)) < max(os.path.getmtime(src), newest_hdr)]
    if stale:
        with ThreadPoolExecutor(max_workers=len(stale)) as ex:
            list(ex.map(lambda src: _compile(src, verbose), stale))
    if stale or not os.path.exists(OUT) or any(os.path.getmtime(_obj(src)) > os.path.getmtime(OUT) for src in SRC):
        subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", OUT] + [_obj(src) for src in SRC])
        build_host(force=True)
    else:
        build_host()
    build_ref_client()
    return OUT


def build_host(force: bool = False) -> str:
    """libblt.so (the blt.h API: Context, Expr, parser) and the run_blt CLI, linked against libbltcu.so."""
    deps = HOST_SRC + HOST_HDR + [os.path.join(HERE, "csrc", "host", "run_blt.cpp"), OUT]
    if not force and os.path.exists(LIBBLT) and os.path.exists(RUN_BLT) and all(
            os.path.getmtime(LIBBLT) >= os.path.getmtime(d) for d in deps):
        return LIBBLT
    libdir = os.path.dirname(OUT)
    cxx = ["g++", "-O2", "-std=c++17", "-fPIC", "-Wall"]
    link = ["-L" + libdir, "-lbltcu", "-Wl,-rpath,$ORIGIN"]
    subprocess.check_call(cxx + ["-shared", "-o", LIBBLT] + HOST_SRC + link)
    subprocess.check_call(cxx + ["-o", RUN_BLT, os.path.join(HERE, "csrc", "host", "run_blt.cpp")] +
                          ["-L" + libdir, "-lblt", "-lbltcu", "-Wl,-rpath,$ORIGIN"])
    return LIBBLT


REF_CLIENT_SRC = "/root/reference/libblt/test_blt.c"
REF_CLIENT = os.path.join(HERE, "..", "oracle", "_ref", "test_blt")


def build_ref_client() -> str | None:
    """The reference's own C client test (libblt/test_blt.c) compiled UNCHANGED, from where it lies, against this
    repository's include/blt.h + include/codes.h and linked with libblt.so.  The source is fed through stdin so that
    its `#include "blt.h"` resolves to include/ (a file's own directory would win otherwise); only test_blt.h comes
    from the reference tree.  Built where /root/reference exists (the GPU box runs the prebuilt binary)."""
    if not os.path.exists(REF_CLIENT_SRC):
        return REF_CLIENT if os.path.exists(REF_CLIENT) else None
    os.makedirs(os.path.dirname(REF_CLIENT), exist_ok=True)
    if os.path.exists(REF_CLIENT) and os.path.getmtime(REF_CLIENT) >= max(os.path.getmtime(LIBBLT), os.path.getmtime(REF_CLIENT_SRC)):
        return REF_CLIENT
    inc = os.path.join(HERE, "..", "include")
    with open(REF_CLIENT_SRC, "rb") as f:
        subprocess.run(["gcc", "-x", "c", "-O1", "-I", inc, "-I", os.path.dirname(REF_CLIENT_SRC), "-o", REF_CLIENT, "-",
                        "-L" + os.path.dirname(OUT), "-lblt", "-lbltcu", "-Wl,-rpath,$ORIGIN/../../blt_b200/lib"],
                       stdin=f, check=True)
    return REF_CLIENT


if __name__ == "__main__":
    print(build(force=True, verbose=True))
