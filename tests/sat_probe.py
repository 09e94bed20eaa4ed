"""Developer probe (not a test): GPU time-to-decision on the committed SAT fixtures (no CPU work here).

    python tests/sat_probe.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import blt_b200

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
z = np.load(f"{G}/jpeg_reduced.npz")
for name in ("test_95", "test_50", "test_50_8"):
    lat, con = z[f"lat_{name}"], z[f"con_{name}"]
    blt_b200.closest(lat, con, timeout_ms=60000)
    line = f"{name:10s}"
    for eager in (0, 1, 4):
        got = blt_b200.closest(lat, con, timeout_ms=60000, eager=eager)
        line += f" | eager {eager}: rc {got['rc']} found {got['found']} nodes {got['nodes']} {got['kernel_ms']:.2f} ms"
    print(line, flush=True)
