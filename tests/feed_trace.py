"""Developer probe (not a test): per-rank launch and feed timing of one multi-GPU search (BLTCU_VERBOSE lines on stderr).

    python tests/feed_trace.py [fixture.npz] [gpus]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import blt_b200

fx = sys.argv[1] if len(sys.argv) > 1 else "bench_dense60_w140.npz"
gpus = int(sys.argv[2]) if len(sys.argv) > 2 else 2
z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", fx))
lat, con = z["lat"], z["con"]
blt_b200.closest(lat, con, gpus=gpus, timeout_ms=120000)
blt_b200.closest(lat, con, gpus=1, timeout_ms=120000)
os.environ["BLTCU_VERBOSE"] = "1"
for g in (1, gpus):
    r = blt_b200.closest(lat, con, gpus=g, timeout_ms=120000)
    print(f"gpus {r['gpus']} kernel {r['kernel_ms']:.1f} ms total {r['total_ms']:.1f} ms busy {r['busy_frac']:.3f} nodes {r['nodes']}", file=sys.stderr, flush=True)
