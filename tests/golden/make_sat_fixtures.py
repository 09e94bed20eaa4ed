"""Generates the SAT-side search inputs used by the time-to-decision tests and probes:

  jpeg_reduced.npz   the three 64 x 64 JPEG fixtures of the reference (test_95, test_50, test_50_8 .lat/.con,
                     libblt/test_regression.cc:84-106) after the oracle's exact LLL (oracle/lll.py), as closest() sees them
  subset_sum60.npz   SURVEY.md section 8d config 3: 60 weights of 45 bits, hidden 0/1 solution with 30 ones, seeds 0..3,
                     after convert/stretch (oracle/context.py) and the exact LLL.  The 60 x 61 reduction takes the Python
                     oracle minutes per instance, so this file is produced with the product's C++ LLL (bltcu_lll), which
                     tests/test_host_layer.py pins bit for bit against oracle/lll.py.

    python tests/golden/make_sat_fixtures.py
"""
import os
import sys
from fractions import Fraction

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import bltio, context  # noqa: E402
from tests.instances import subset_sum  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")
out = {}
for name in ("test_95", "test_50", "test_50_8"):
    lat0 = bltio.read_boost(f"{G}/{name}.lat")
    con0 = bltio.read_boost(f"{G}/{name}.con")
    r = context.solve_raw(lat0, con0)
    assert r["rc"] == 10
    out[f"lat_{name}"] = np.array(r["lat"], dtype=np.int64)
    out[f"con_{name}"] = np.array(con0, dtype=np.int64)
    out[f"nodes_{name}"] = r["nodes"]
    out[f"sol_{name}"] = r["sol"]
np.savez_compressed(f"{G}/jpeg_reduced.npz", **out)

import blt_b200  # noqa: E402  (only its host-side exact LLL)

out = {}
for seed in range(4):
    A, C, xs = subset_sum(60, 45, seed)
    L0 = [[Fraction(int(A[i][j])) for i in range(A.shape[0])] for j in range(A.shape[1])]
    C0 = [[Fraction(int(b[0])), Fraction(int(b[1]))] for b in C]
    L, Ci = context.convert_to_integral(L0, C0)
    assert context.stretch_to_cube(L, Ci) == 0
    rank, Lr, U = blt_b200.lll(np.array(L, dtype=np.int64))
    out[f"lat{seed}"] = np.array(Lr[len(Lr) - rank:], dtype=np.int64)
    out[f"con{seed}"] = np.array(Ci, dtype=np.int64)
    out[f"xs{seed}"] = xs
np.savez_compressed(f"{G}/subset_sum60.npz", **out)

# The same lattice with 26-bit weights (density 2.3: many subsets hit the target, so a search that is worth the name
# finds one): seeds whose instance the oracle decides SAT without an LP failure; counts of the oracle stored.
import oracle  # noqa: E402

out = {}
for seed in range(12):
    A, C, xs = subset_sum(60, 26, seed)
    L0 = [[Fraction(int(A[i][j])) for i in range(A.shape[0])] for j in range(A.shape[1])]
    C0 = [[Fraction(int(b[0])), Fraction(int(b[1]))] for b in C]
    L, Ci = context.convert_to_integral(L0, C0)
    assert context.stretch_to_cube(L, Ci) == 0
    rank, Lr, U = blt_b200.lll(np.array(L, dtype=np.int64))
    lat, con = np.array(Lr[len(Lr) - rank:], dtype=np.int64), np.array(Ci, dtype=np.int64)
    r = oracle.closest(lat, con, max_nodes=2000000)
    if r["rc"] == 0 and oracle.check_soln(lat, con, r["sol"]) and r["nodes"] > 1000:
        out[f"lat{seed}"], out[f"con{seed}"], out[f"nodes{seed}"] = lat, con, r["nodes"]
        print("b26 seed", seed, "SAT after", r["nodes"], "nodes")
out["seeds"] = np.array(sorted(int(k[3:]) for k in out if k.startswith("lat")))
np.savez_compressed(f"{G}/subset_sum60_b26.npz", **out)
