"""Developer probe (not a test): bltcu_lll_batch on 4000 bases of 20 x 20 (the bench supplement's workload).

    python tests/lll_probe.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import blt_b200
rng = np.random.default_rng(5)
Bs = np.stack([rng.integers(-100, 101, size=(20, 20)).T for _ in range(4000)])
blt_b200.lll_batch(Bs[:64])
for _ in range(2):
    _, _, _, info = blt_b200.lll_batch(Bs)
    print(info)
