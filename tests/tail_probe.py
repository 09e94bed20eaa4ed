"""Developer probe (not a test): the end of a search under the donation knobs (BLTCU_GIVE_MAX, BLTCU_CHAIN_FIRST).

    python tests/tail_probe.py [fixture.npz]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import blt_b200

fx = sys.argv[1] if len(sys.argv) > 1 else "bench_dense60_w140.npz"
z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", fx))
lat, con = z["lat"], z["con"]
blt_b200.closest(lat, con, timeout_ms=120000)
for give, chain in [(1, 0), (1, 1), (2, 1), (4, 1), (8, 1), (16, 1), (8, 0)]:
    os.environ["BLTCU_GIVE_MAX"] = str(give)
    os.environ["BLTCU_CHAIN_FIRST"] = str(chain)
    out = []
    for world in (1, 8):
        r = blt_b200.closest(lat, con, rank=0, world=world, no_early_exit=1, timeout_ms=120000)
        out.append(f"world {world}: {r['kernel_ms']:.1f} ms busy {r['busy_frac']:.3f} donated {r['donated']} nodes {r['nodes']}")
    print(f"give {give} chain_first {chain}: " + " | ".join(out), flush=True)
