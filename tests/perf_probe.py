"""Developer probe (not a test): per-warp latency and warp-count scaling of the search kernel on a mid-size instance."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import blt_b200
from tests.instances import dense_bilp, reduced_search_input

n, w = 60, int(sys.argv[1]) if len(sys.argv) > 1 else 85
A, C = dense_bilp(n, w, 0)
lat, con, _ = reduced_search_input(A, C)
blt_b200.closest(lat, con)  # warm-up, allocations
if len(sys.argv) > 2 and sys.argv[2] == 'prof1':
    r = blt_b200.closest(lat, con, timeout_ms=600000, ctas=1, warps_per_cta=1)
    print('prof1 run', r['nodes'], r['kernel_ms'])
    sys.exit(0)
if len(sys.argv) > 2 and sys.argv[2] == 'prof':
    r = blt_b200.closest(lat, con, timeout_ms=600000)
    print('prof run', r['nodes'], r['kernel_ms'])
    sys.exit(0)
big = len(sys.argv) > 2 and sys.argv[2] == 'big'
if big:
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'bench_dense60.npz')); lat, con = z['lat'], z['con']
for opts in ((dict(ctas=148, warps_per_cta=1), dict(ctas=148, warps_per_cta=2), dict(ctas=148, warps_per_cta=4), dict()) if big else ()) or (dict(ctas=1, warps_per_cta=1), dict(ctas=8, warps_per_cta=1), dict(ctas=148, warps_per_cta=1),
             dict(ctas=148, warps_per_cta=2), dict(ctas=148, warps_per_cta=4), dict(ctas=148, warps_per_cta=8), dict()):
    r = blt_b200.closest(lat, con, timeout_ms=120000, **opts)
    print(opts, "rc", r["rc"], "nodes", r["nodes"], "warps", r["warps"], "ms %.2f" % r["kernel_ms"],
          "nodes/s %.3g" % (r["nodes"] / r["kernel_ms"] * 1e3), "per-warp us/node %.1f" % (r["kernel_ms"] * 1e3 * r["warps"] / r["nodes"]),
          "donated", r["donated"], "pivots/node %.2f" % (r["pivots"] / r["nodes"]), "refac", r["refactors"], "busy %.2f" % r["busy_frac"], flush=True)
