"""Developer probe (not a test): idle warps / queued records over time of one search (BLTCU_TRACE lines on stderr).

    python tests/trace_probe.py [fixture.npz]
"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import blt_b200

fx = sys.argv[1] if len(sys.argv) > 1 else "bench_dense60_w140.npz"
z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", fx))
lat, con = z["lat"], z["con"]
blt_b200.closest(lat, con, timeout_ms=120000)
os.environ["BLTCU_TRACE"] = "1"
os.environ["BLTCU_VERBOSE"] = "1"
os.environ["BLTCU_FEED_DIV"] = "64"
for world in (8, 1):
    word = (ctypes.c_int * 16)()
    print(f"--- rank 0 of {world}", file=sys.stderr, flush=True)
    r = blt_b200.closest(lat, con, rank=0, world=world, stop_flag=ctypes.addressof(word), no_early_exit=1, timeout_ms=120000)
    print(f"    kernel {r['kernel_ms']:.1f} ms nodes {r['nodes']} busy {r['busy_frac']:.3f}", file=sys.stderr, flush=True)
