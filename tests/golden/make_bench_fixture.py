"""Generates tests/golden/bench_dense60.npz (width 110, the full-size parity fixture) and bench_dense60_w<width>.npz
(width 120: the bench.py workload; larger widths take the CPU oracle an hour and more): the bench.py workload (SURVEY.md section 8d, config 4 at the metric's
shape n = m = 60).  The LLL-reduced basis comes from the oracle's exact LLL (oracle/lll.py); the expected counts
come from the oracle's closest() and are stored so that the GPU run can be checked at full size.

    python tests/golden/make_bench_fixture.py [width] [seed]
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from tests.instances import dense_bilp, reduced_search_input  # noqa: E402

n = 60
w = int(sys.argv[1]) if len(sys.argv) > 1 else 110
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
A, C = dense_bilp(n, w, seed)
lat, con, U = reduced_search_input(A, C)
t = time.time()
r = oracle.closest(lat, con)
print(f"oracle: rc={r['rc']} niter={r['niter']} nprune={r['nprune']} nodes={r['nodes']} pivots={r['pivots']} "
      f"{time.time() - t:.1f}s")
assert r["rc"] == 0 and not oracle.check_soln(lat, con, r["sol"]), "bench instance must be UNSAT"
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "bench_dense60.npz" if w == 110 else f"bench_dense60_w{w}.npz"), lat=lat, con=con, A=A, C=C, w=w,
                    seed=seed, niter=r["niter"], nprune=r["nprune"], nodes=r["nodes"], pivots=r["pivots"],
                    hist=r["hist"], sol=r["sol"])
