"""Developer probe (not a test): the on-demand part of the multi-GPU deal, one process driving every GPU.

    python tests/feed_probe.py [fixture.npz] [gpus]

Times closest(gpus=N) under several BLTCU_FEED_* settings (read per call by the library).
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
blt_b200.closest(lat, con, gpus=gpus, timeout_ms=120000)  # warm-up, allocations on every device
K = ("BLTCU_FEED_STRIDED", "BLTCU_FEED_DIV", "BLTCU_FEED_PER_IDLE", "BLTCU_FEED_HOLD_US", "BLTCU_FEED_RAMP_GUARD",
     "BLTCU_FEED_BACKLOG", "BLTCU_FEED_PCT", "BLTCU_FEED_EAGER")
for name, env in [("shipped", {}),
                  ("backlog 128, 12 %", {"BLTCU_FEED_BACKLOG": "128"}),
                  ("backlog 128, 35 %", {"BLTCU_FEED_BACKLOG": "128", "BLTCU_FEED_PCT": "35"}),
                  ("backlog 128, 50 %", {"BLTCU_FEED_BACKLOG": "128", "BLTCU_FEED_PCT": "50"}),
                  ("backlog 512, 50 %", {"BLTCU_FEED_BACKLOG": "512", "BLTCU_FEED_PCT": "50"}),
                  ("backlog 32, 50 %", {"BLTCU_FEED_BACKLOG": "32", "BLTCU_FEED_PCT": "50"}),
                  ("backlog 128, 70 %", {"BLTCU_FEED_BACKLOG": "128", "BLTCU_FEED_PCT": "70"}),
                  ("backlog 128, 85 %", {"BLTCU_FEED_BACKLOG": "128", "BLTCU_FEED_PCT": "85"}),
                  ("eager 128, 50 %", {"BLTCU_FEED_BACKLOG": "128", "BLTCU_FEED_PCT": "50", "BLTCU_FEED_EAGER": "1"})]:
    for k in K:
        os.environ.pop(k, None)
    os.environ.update(env)
    best = None
    for _ in range(4):
        t0 = time.perf_counter()
        r = blt_b200.closest(lat, con, gpus=gpus, timeout_ms=120000)
        ms = (time.perf_counter() - t0) * 1e3
        best = ms if best is None else min(best, ms)
    ok = r["nodes"] == int(z["nodes"]) and r["pivots"] == int(z["pivots"]) and (r["sol"] == z["sol"]).all()
    print(f"{name:18s} gpus {r['gpus']} wall {best:.1f} ms kernel {r['kernel_ms']:.1f} ms -> {r['nodes'] / best / 1e3:.1f} Mnodes/s "
          f"busy {r['busy_frac']:.3f} counts_ok {ok}", flush=True)
