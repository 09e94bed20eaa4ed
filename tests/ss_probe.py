"""Developer probe (not a test): the dimension-60 subset-sum fixtures under a time limit.

    python tests/ss_probe.py [timeout_ms] [seeds...]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import blt_b200

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
tmo = int(sys.argv[1]) if len(sys.argv) > 1 else 30000
seeds = [int(x) for x in sys.argv[2:]] or [0, 3]
z = np.load(f"{G}/subset_sum60.npz")
for s in seeds:
    lat, con = z[f"lat{s}"], z[f"con{s}"]
    t = time.time()
    g = blt_b200.closest(lat, con, timeout_ms=tmo)
    hist = g["hist"]
    deep = [int(i) for i in np.nonzero(hist)[0][:5]]
    print(f"seed {s}: rc {g['rc']} found {g['found']} nodes {g['nodes']} kernel_ms {g['kernel_ms']:.1f} wall {time.time() - t:.1f}s "
          f"busy {g['busy_frac']:.2f} deepest levels reached {deep}", flush=True)
