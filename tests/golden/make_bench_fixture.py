"""Generates tests/golden/bench_dense60.npz (width 110, the full-size parity fixture) and bench_dense60_w<width>.npz
(width 120: the bench.py workload; larger widths take the CPU oracle an hour and more): the bench.py workload (SURVEY.md section 8d, config 4 at the metric's
shape n = m = 60).  The LLL-reduced basis comes from the oracle's exact LLL (oracle/lll.py); the expected counts
come from the oracle's closest() and are stored so that the GPU run can be checked at full size.

    python tests/golden/make_bench_fixture.py [width] [seed] [n]

n = 80 (BASELINE.json config 4): bench_dense80_w75.npz (93,413 nodes, full parity) and bench_dense80_w105.npz (~1e7
nodes, twelve minutes of oracle time).  The pure-Python LLL of the oracle needs several minutes at 80 x 80, so for
n > 60 the basis is reduced with the product's host LLL (bltcu_lll, pinned bit for bit against oracle/lll.py by
tests/test_host_layer.py, no GPU involved) unless BLT_FIXTURE_ORACLE_LLL=1 asks for the oracle's.
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from tests.instances import dense_bilp, reduced_search_input  # noqa: E402

w = int(sys.argv[1]) if len(sys.argv) > 1 else 110
seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
n = int(sys.argv[3]) if len(sys.argv) > 3 else 60
A, C = dense_bilp(n, w, seed)
if n > 60 and not os.environ.get("BLT_FIXTURE_ORACLE_LLL"):
    import blt_b200  # host-side exact LLL only (common width: stretch_to_cube is the identity, context.cc:455-501)

    rank, Lr, U = blt_b200.lll(A.T.copy())
    lat, con = Lr[n - rank:], C
else:
    lat, con, U = reduced_search_input(A, C)
t = time.time()
r = oracle.closest(lat, con)
print(f"oracle: rc={r['rc']} niter={r['niter']} nprune={r['nprune']} nodes={r['nodes']} pivots={r['pivots']} "
      f"{time.time() - t:.1f}s")
assert r["rc"] == 0 and not oracle.check_soln(lat, con, r["sol"]), "bench instance must be UNSAT"
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "bench_dense60.npz" if (w == 110 and n == 60) else f"bench_dense{n}_w{w}.npz"), lat=lat, con=con, A=A, C=C, w=w,
                    seed=seed, niter=r["niter"], nprune=r["nprune"], nodes=r["nodes"], pivots=r["pivots"],
                    hist=r["hist"], sol=r["sol"])
