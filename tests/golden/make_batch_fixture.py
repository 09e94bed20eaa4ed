"""Generates tests/golden/batch20.npz: SURVEY.md section 8d config 5 at reduced count -- `count` instances,
A in Z^{20x20} uniform on [-100, 100], width drawn from {60, 90, 110, 130, 160} (a mix of UNSAT and SAT), seed =
instance id -- already LLL-reduced by the oracle (oracle/lll.py), with the oracle's closest() result for each one
(status, found, niter, nprune, returned vector).  The sequential oracle is the reference's depth-first order, so the
SAT instances pin the batch kernel's order as well.

    python tests/golden/make_batch_fixture.py [count]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
from tests.instances import dense_bilp, reduced_search_input  # noqa: E402

count = int(sys.argv[1]) if len(sys.argv) > 1 else 400
n = 20
widths = [60, 90, 110, 130, 160]
lats, cons, rcs, found, niter, nprune, sols, ws = [], [], [], [], [], [], [], []
for i in range(count):
    w = widths[int(np.random.default_rng(10_000 + i).integers(0, len(widths)))]
    A, C = dense_bilp(n, w, i)
    lat, con, U = reduced_search_input(A, C)
    assert lat.shape == (n, n)
    r = oracle.closest(lat, con)
    lats.append(lat); cons.append(con); rcs.append(r["rc"]); niter.append(r["niter"]); nprune.append(r["nprune"])
    sols.append(r["sol"]); found.append(bool(oracle.check_soln(lat, con, r["sol"]))); ws.append(w)
print(f"{count} instances: {sum(found)} SAT, nodes total {sum(niter) + sum(nprune)}, max niter {max(niter)}")
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "batch20.npz"), lat=np.array(lats), con=np.array(cons),
                    rc=np.array(rcs, dtype=np.int32), found=np.array(found), niter=np.array(niter, dtype=np.uint64),
                    nprune=np.array(nprune, dtype=np.uint64), sol=np.array(sols, dtype=np.int64), w=np.array(ws))
