"""CPU checks of bench.py: the reference arm (the numpy port on host cores, no GPU) prints one JSON line with the keys
the driver reads; the argument parser knows the sub-record switches."""
import json
import subprocess
import sys

from conftest import ROOT


def test_reference_arm_prints_contract_line():
    res = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--grid", "16", "--cpu-fields", "2",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600, cwd=str(ROOT))
    assert res.returncode == 0, res.stderr[-2000:]
    d = json.loads(res.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["unit"] == "fields/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["gpu_launches"] == 0
    assert d["e2e"] == {"value": d["value"], "unit": "fields/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_exit_quietly():
    import os

    res = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--grid", "16", "--gpus", "2"],
                         capture_output=True, text=True, timeout=120, cwd=str(ROOT), env=dict(os.environ, RANK="1", WORLD_SIZE="2"))
    assert res.returncode == 0 and res.stdout.strip() == ""


def test_help_lists_sub_records():
    res = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--help"], capture_output=True, text=True, timeout=120, cwd=str(ROOT))
    assert res.returncode == 0 and "--sub" in res.stdout and "--slab-variants" in res.stdout
