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
K = ("BLTCU_FEED_BACKLOG", "BLTCU_FEED_PCT", "BLTCU_FEED_DIV", "BLTCU_FEED_PER_IDLE", "BLTCU_FEED_HOLD_US", "BLTCU_FEED_POLLS")
for env in [{}, {"BLTCU_FEED_BACKLOG": "64"}, {"BLTCU_FEED_BACKLOG": "256"},
            {"BLTCU_FEED_PCT": "50"}, {"BLTCU_FEED_PCT": "50", "BLTCU_FEED_BACKLOG": "32"}, {"BLTCU_FEED_PCT": "50", "BLTCU_FEED_BACKLOG": "64"},
            {"BLTCU_FEED_PCT": "50", "BLTCU_FEED_BACKLOG": "128"}, {"BLTCU_FEED_PCT": "50", "BLTCU_FEED_BACKLOG": "256"},
            {"BLTCU_FEED_PCT": "50", "BLTCU_FEED_BACKLOG": "512"},
            {"BLTCU_FEED_PCT": "70", "BLTCU_FEED_BACKLOG": "128"}, {"BLTCU_FEED_PCT": "70", "BLTCU_FEED_BACKLOG": "512"}]:
    for k in K:
        os.environ.pop(k, None)
    os.environ.update(env)
    word = (ctypes.c_int * 16)()
    print(f"--- {env}", file=sys.stderr, flush=True)
    r = blt_b200.closest(lat, con, rank=0, world=8, stop_flag=ctypes.addressof(word), no_early_exit=1, timeout_ms=120000)
    print(f"    kernel {r['kernel_ms']:.1f} ms nodes {r['nodes']} busy {r['busy_frac']:.3f} -> {r['nodes'] / r['kernel_ms'] / 1e3:.1f} Mnodes/s", file=sys.stderr, flush=True)
