"""Developer probe (not a test): how early does ONE rank of eight touch the on-demand subtrees?  One device.

    python tests/feed_policy_probe.py [fixture.npz]

Rank 0 of a world of 8 runs alone with a shared block, so everything it takes from the on-demand part before its own
share has run dry is the fault of the hunger test (BLTCU_VERBOSE lines on stderr say how many and when).
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
blt_b200.closest(lat, con, rank=0, world=8, timeout_ms=120000)
os.environ["BLTCU_VERBOSE"] = "1"
for polls, per_idle, hold, fdiv in [(1, 0, 150, 8), (1, 32, 150, 8), (1, 0, 150, 3), (1, 32, 150, 3), (1, 16, 150, 3), (1, 8, 150, 3),
                                    (1, 32, 50, 3), (1, 32, 150, 2), (1, 16, 150, 2)]:
    os.environ["BLTCU_FEED_POLLS"] = str(polls)
    os.environ["BLTCU_FEED_PER_IDLE"] = str(per_idle)
    os.environ["BLTCU_FEED_HOLD_US"] = str(hold)
    os.environ["BLTCU_FEED_DIV"] = str(fdiv)
    word = (ctypes.c_int * 16)()
    print(f"--- polls {polls} one subtree per {per_idle} idle warps, hold-off {hold} us, on-demand tail 1/{fdiv}", file=sys.stderr, flush=True)
    r = blt_b200.closest(lat, con, rank=0, world=8, stop_flag=ctypes.addressof(word), no_early_exit=1, timeout_ms=120000)
    print(f"    kernel {r['kernel_ms']:.1f} ms nodes {r['nodes']} busy {r['busy_frac']:.3f} -> {r['nodes'] / r['kernel_ms'] / 1e3:.1f} Mnodes/s", file=sys.stderr, flush=True)
