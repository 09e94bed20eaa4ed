"""Developer probe (not a test): one search over all GPUs of the box in one process, with and without work sharing.

    python tests/share_probe.py [fixture.npz] [gpus]
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import blt_b200

fx = sys.argv[1] if len(sys.argv) > 1 else "bench_dense60_w140.npz"
gpus = int(sys.argv[2]) if len(sys.argv) > 2 else 2
z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", fx))
lat, con = z["lat"], z["con"]
blt_b200.closest(lat, con, gpus=gpus, timeout_ms=60000)
for name, env, verbose in [("no sharing", {}, 0), ("sharing", {"BLTCU_SHARE": "1"}, 0), ("sharing", {"BLTCU_SHARE": "1"}, 1), ("no sharing", {}, 0), ("sharing", {"BLTCU_SHARE": "1"}, 0)]:
    os.environ.pop("BLTCU_SHARE", None)
    os.environ.pop("BLTCU_VERBOSE", None)
    os.environ.update(env)
    if verbose:
        os.environ["BLTCU_VERBOSE"] = "1"
    best = None
    for _ in range(1 if verbose else 3):
        t0 = time.perf_counter()
        r = blt_b200.closest(lat, con, gpus=gpus, timeout_ms=60000)
        ms = (time.perf_counter() - t0) * 1e3
        best = ms if best is None else min(best, ms)
    ok = r["rc"] == 0 and r["nodes"] == int(z["nodes"]) and r["pivots"] == int(z["pivots"]) and (r["sol"] == z["sol"]).all()
    print(f"{name:12s} gpus {r['gpus']} wall {best:.1f} ms kernel {r['kernel_ms']:.1f} ms -> {r['nodes'] / best / 1e3:.1f} Mnodes/s "
          f"busy {r['busy_frac']:.3f} rc {r['rc']} nodes {r['nodes']} counts_ok {ok}", flush=True)
