"""Developer probe (not a test): warp-count scaling of the search kernel on a bench fixture.

    python tests/perf_probe.py [fixture.npz] [wpc ...]
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import blt_b200

fx = sys.argv[1] if len(sys.argv) > 1 else "bench_dense60.npz"
z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", fx))
lat, con = z["lat"], z["con"]
wpcs = [int(x) for x in sys.argv[2:]] or [1, 2, 4, 8, 10, 12, 14, 16]
blt_b200.closest(lat, con)  # warm-up, allocations
for wpc in wpcs:
    r = blt_b200.closest(lat, con, timeout_ms=120000, warps_per_cta=wpc)
    busy_us = r["kernel_ms"] * 1e3 * r["warps"] * r["busy_frac"] / r["nodes"]
    print(f"wpc {wpc:2d} rc {r['rc']} nodes {r['nodes']} warps {r['warps']} ms {r['kernel_ms']:.2f} "
          f"Mnodes/s {r['nodes'] / r['kernel_ms'] / 1e3:.2f} busy {r['busy_frac']:.2f} busy-us/node/warp {busy_us:.1f} "
          f"donated {r['donated']} counts_ok {r['nodes'] == int(z['nodes']) and r['pivots'] == int(z['pivots']) and (r['sol'] == z['sol']).all()}", flush=True)
