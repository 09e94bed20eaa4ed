"""Developer probe (not a test): the size of every subtree of the static split, one search per subtree.

    python tests/subtree_sizes.py [fixture.npz] [out.csv]

Runs rank i of a world as large as the number of subtrees, so that rank i's share is subtree i alone, and writes
index, nodes, kernel ms and the subtree's path (BLTCU_DUMP_KEYS line of the library) to a csv.
"""
import os
import re
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

if len(sys.argv) > 3 and sys.argv[3] == "child":
    import numpy as np
    import blt_b200
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", sys.argv[1]))
    lat, con = z["lat"], z["con"]
    os.environ["BLTCU_DUMP_KEYS"] = "1"
    count = int(sys.argv[2])
    blt_b200.closest(lat, con, rank=count, world=count + 1, no_early_exit=1, timeout_ms=60000)  # warm-up (a rank without a subtree)
    for i in range(count):
        r = blt_b200.closest(lat, con, rank=i, world=count + 1, no_early_exit=1, timeout_ms=60000)
        print(f"ROW {i} {r['nodes']} {r['kernel_ms']:.3f}", flush=True)
    sys.exit(0)

fx = sys.argv[1] if len(sys.argv) > 1 else "bench_dense60_w140.npz"
out = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/subtree_sizes.csv"
import numpy as np
import blt_b200
z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", fx))
os.environ["BLTCU_VERBOSE"] = "1"
# count the subtrees: the verbose line of phase A says how many were published
p = subprocess.run([sys.executable, "-c",
                    "import sys,os; sys.path.insert(0, os.getcwd()); import numpy as np, blt_b200; "
                    f"z=np.load('tests/golden/{fx}'); blt_b200.closest(z['lat'], z['con'], rank=0, world=100000, no_early_exit=1, timeout_ms=60000)"],
                   capture_output=True, text=True)
m = re.search(r"(\d+) subtrees published", p.stderr)
count = int(m.group(1))
print("subtrees", count, flush=True)
env = dict(os.environ)
env.pop("BLTCU_VERBOSE", None)
p = subprocess.run([sys.executable, os.path.abspath(__file__), fx, str(count), "child"], capture_output=True, text=True, env=env)
keys = {}
for line in p.stderr.splitlines():
    m = re.match(r"bltcu: subtree (\d+) of \d+ key (.*)", line)
    if m:
        keys[int(m.group(1))] = m.group(2).strip()
rows = []  # rank 0 also counts the nodes of the static split itself

for line in p.stdout.splitlines():
    t = line.split()
    if t and t[0] == "ROW":
        rows.append((int(t[1]), int(t[2]), float(t[3])))
os.makedirs(os.path.dirname(out) or ".", exist_ok=True)
with open(out, "w") as f:
    f.write("index,nodes,kernel_ms,key\n")
    for i, nodes, ms in rows:
        f.write(f"{i},{nodes},{ms},{keys.get(i, '')}\n")
print("rows", len(rows), "sum of nodes", sum(r[1] for r in rows), "expected", int(z["nodes"]), flush=True)
