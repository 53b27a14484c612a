"""The regression gate on the committed round-1 results: the engine's
JSON records against the NCCL path's, cell by cell.  CPU only — it reads files."""
import json
import os
import sys

import pytest

from conftest import ROOT, run

GATE = os.path.join(ROOT, "scripts", "regression_gate.py")
N8 = os.path.join(ROOT, "profiles", "r01_raw", "n8_run1")
N8B = os.path.join(ROOT, "profiles", "r01_raw", "n8_run2_allpairs")


def gate(*args):
    return run([sys.executable, GATE, *args], check=False)


def test_uni_256mib_matrix_beats_nccl_and_reaches_80_percent_of_link():
    p = gate(os.path.join(N8, "cfg2_engine_dev.json"), os.path.join(N8, "cfg2_nccl_dev.json"),
             "--min-ratio", "1.25", "--min-link-frac", "0.80", "--quiet")
    assert p.returncode == 0, p.stdout
    assert "56 cells compared, 0 below threshold" in p.stdout


def test_bidir_matrix_and_sweep_beat_nccl():
    for e, b in (("cfg3_engine_dev.json", "cfg3_nccl_dev.json"), ("sweep_engine.json", "sweep_nccl.json")):
        p = gate(os.path.join(N8, e), os.path.join(N8, b), "--quiet")
        assert p.returncode == 0, p.stdout


def test_gate_catches_the_regression_it_was_written_for():
    """The first all-pairs default (pushes in a fan-out) lost to NCCL; the gate says so, and passes the fix."""
    bad = gate(os.path.join(N8, "cfg4_engine_dev.json"), os.path.join(N8, "cfg4_nccl_dev.json"))
    assert bad.returncode == 1 and "FAIL" in bad.stdout
    good = gate(os.path.join(N8B, "cfg4_engine_dev.json"), os.path.join(N8B, "cfg4_nccl_dev.json"))
    assert good.returncode == 0 and "FAIL" not in good.stdout


def test_latency_cells_are_lower_is_better(tmp_path):
    e, b = tmp_path / "e.json", tmp_path / "b.json"
    rec = {"mode": "latency", "timing": "dev", "src": 0, "dst": 1, "bytes": 8, "GBps": 0.0}
    e.write_text(json.dumps({**rec, "us": 5.0}) + "\n")
    b.write_text(json.dumps({**rec, "us": 9.0}) + "\n")
    assert gate(str(e), str(b), "--min-ratio", "1.5", "--quiet").returncode == 0
    assert gate(str(b), str(e), "--quiet").returncode == 1
    assert gate(str(e), str(tmp_path / "e.json"), "--quiet").returncode == 0
    (tmp_path / "empty.json").write_text("")
    assert gate(str(e), str(tmp_path / "empty.json")).returncode == 2
