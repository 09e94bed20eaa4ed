"""bench.py's reference arm on this machine's cores (no GPU needed): one JSON line with the contract's keys."""
import json
import os
import subprocess
import sys

from tests.conftest import ROOT


def test_reference_arm_prints_the_contract_line():
    env = dict(os.environ, BLT_BENCH_REF_NODES="8000")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                         capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "enum_nodes_per_s" and d["unit"] == "nodes/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["value"] > 0 and d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] == 1
    assert d["e2e"] == {"value": d["value"], "unit": "nodes/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and d["vs_baseline"] is None and d["gpu_launches"] == 0
    # under torchrun the other ranks print nothing
    env2 = dict(env, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out2 = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                           "--warmup", "1"], capture_output=True, text=True, timeout=600, env=env2, cwd=ROOT)
    assert out2.returncode == 0 and out2.stdout.strip() == ""
