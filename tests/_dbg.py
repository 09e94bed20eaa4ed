import sys; sys.path.insert(0,'.')
import numpy as np, oracle, blt_b200
from oracle import bltio, context
for name in ("test_95","test_50","test_50_8"):
    lat=bltio.read_boost(f"tests/golden/{name}.lat"); con=bltio.read_boost(f"tests/golden/{name}.con")
    r=context.solve_raw(lat,con)
    g=blt_b200.closest(r["lat"],con,timeout_ms=20000)
    print(name,'rc',g['rc'],'found',g['found'],'nodes',g['nodes'],'ms',g['kernel_ms'],'refac',g['refactors'],flush=True)
