"""bench.py's reference arm on the host cores (the one place outside tests/ and smoke() that runs oracle/_ref): the JSON line the driver reads.
The GPU arm needs a device and is exercised on the box; here the CPU contract only."""
import json
import os
import subprocess
import sys

import pytest

from util import ROOT

REF = os.path.join(ROOT, "oracle", "_ref", "ref_bh")


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref not built (oracle/build_ref.sh needs /root/reference)")
def test_reference_arm_prints_the_contract_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["metric"] == "Mpaths/s" and line["unit"] == "Mpaths/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["steps"] == 1 and line["warmup"] == 0 and line["scaling"] == "strong"
    assert "1920x1080 -s 1024" in line["config"]["workload"] and "CBdragon substitute" in line["config"]["workload"]
    cb = line["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] >= 1 and abs(cb["value"] - line["value"]) < 1e-12
    assert "class mix 162101216 : 60825600" in cb["sample"] and "-t 1 vs -t" in cb["sample"]
    assert line["e2e"] == {"value": line["value"], "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_other_ranks_of_the_reference_arm_exit_without_work():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=60, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""
