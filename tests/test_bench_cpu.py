"""bench.py pieces that need no GPU: the reference arm (`--impl reference`) runs the unmodified reference on the host
cores and prints one JSON line with the contract's keys; the helper that reads ncu traffic never invents numbers."""
import json
import os
import subprocess
import sys

import pytest

from oracle import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not oracle.have_reference(), reason="oracle/_ref/gt_ref not built")
def test_reference_arm_prints_the_contract_line():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "ecoli",
                        "--steps", "1", "--warmup", "0", "--shrink", "8"], capture_output=True, text=True, cwd=ROOT)
    assert p.returncode == 0, p.stderr
    line = json.loads(p.stdout.strip().splitlines()[-1])
    for key in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert key in line, key
    assert line["impl"] == "reference" and line["unit"] == "suffixes/s" and line["value"] > 1e5
    assert line["cpu_baseline"]["kind"] == "reference" and line["cpu_baseline"]["cores"] == 1
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0


def test_traffic_comes_from_the_profile_file_or_is_null():
    sys.path.insert(0, ROOT)
    import bench
    db = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    assert bench.ncu_traffic("human", "sort_lcp") == pytest.approx(db["human"]["sort_lcp"])
    assert bench.ncu_traffic("human", "no_such_class") is None
    assert bench.ncu_traffic("no_such_workload", "sort_lcp") is None


def test_default_workload_is_the_metrics_configuration():
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert 'ap.add_argument("--workload", default="human"' in src and 'ap.add_argument("--scaling", default="strong"' in src
