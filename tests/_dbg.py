import sys; sys.path.insert(0,'.')
import numpy as np, blt_b200
z=np.load("tests/golden/bench_dense60.npz")
for d in (4,6,8):
    r=blt_b200.closest(z["lat"],z["con"],rank=0,world=2,split_depth=d,timeout_ms=60000)
    print(d,'rc',r['rc'],'nodes',r['nodes'],'ms',r['kernel_ms'],'donated',r['donated'],flush=True)
