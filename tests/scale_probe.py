"""Developer probe (not a test): predicted multi-GPU scaling from one GPU -- runs every rank's share of a static split
one after the other and reports the slowest rank.

    python tests/scale_probe.py [fixture.npz] [split_depth ...]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import blt_b200

fx = sys.argv[1] if len(sys.argv) > 1 else "bench_dense60_w130.npz"
z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", fx))
lat, con = z["lat"], z["con"]
depths = [int(x) for x in sys.argv[2:]] or [6]
r1 = blt_b200.closest(lat, con)
r1 = blt_b200.closest(lat, con)
print(f"1 GPU: {r1['kernel_ms']:.1f} ms, {r1['nodes'] / r1['kernel_ms'] / 1e3:.1f} Mnodes/s")
for depth in depths:
    for world in (2, 4, 8):
        ms, nodes = [], 0
        for rank in range(world):
            r = blt_b200.closest(lat, con, rank=rank, world=world, split_depth=depth, timeout_ms=300000)
            assert r["rc"] == 0
            ms.append(r["kernel_ms"])
            nodes += r["nodes"]
        print(f"depth {depth} world {world}: slowest rank {max(ms):.1f} ms fastest {min(ms):.1f} ms nodes {nodes} "
              f"ok {nodes == r1['nodes']} -> {nodes / max(ms) / 1e3:.1f} Mnodes/s, speed-up {r1['kernel_ms'] / max(ms):.2f}", flush=True)
