"""The benchmark's report, byte for byte apart from the numbers (reference format:
p2p_matrix.cc:133-140, 178-185, 188-195)."""
import json
import os
import re

import pytest

from conftest import MPIRUN, P2P_MATRIX, ROOT, gpu_count, run

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def mask(text: str) -> str:
    """Numbers in table cells -> '#'; keeps spacing rules: '%6.02f ' cells."""
    return re.sub(r"\d+\.\d\d ", "#.## ", text)


def test_golden_report_text_matches_reference_format_strings():
    golden = open(os.path.join(GOLDEN_DIR, "report_n2_masked.txt")).read()
    assert golden == mask("Evaluating the Uni-Directional NCCL P2P Bandwidth (Gbps)\n"
                          "   D\\D     0      1 \n"
                          "     0   0.00 123.45 \n"
                          "     1 123.45   0.00 \n"
                          "\nEvaluating the Bi-Directional NCCL P2P Bandwidth (Gbps)\n"
                          "   D\\D     0      1 \n"
                          "     0   0.00 123.45 \n"
                          "     1 123.45   0.00 \n")


@pytest.mark.gpu
def test_single_rank_report(built):
    """N = 1: headers and a 0.00 diagonal, exactly as the reference prints for one process."""
    out = run([MPIRUN, "-n", "1", P2P_MATRIX], timeout=120).stdout
    assert out == ("Evaluating the Uni-Directional NCCL P2P Bandwidth (Gbps)\n"
                   "   D\\D     0 \n"
                   "     0   0.00 \n"
                   "\nEvaluating the Bi-Directional NCCL P2P Bandwidth (Gbps)\n"
                   "   D\\D     0 \n"
                   "     0   0.00 \n")


@pytest.mark.gpu
@pytest.mark.skipif(gpu_count() < 2, reason="needs two GPUs")
def test_two_rank_report_matches_golden(built):
    out = run([MPIRUN, "-n", "2", P2P_MATRIX, "--iters", "8"], timeout=300).stdout
    # numbers at B200 rates overflow the %6.02f field; normalise cell width the way printf grew it
    normalised = re.sub(r" *\d+\.\d\d ", " #.## ", out)
    golden = re.sub(r" *#\.## ", " #.## ", open(os.path.join(GOLDEN_DIR, "report_n2_masked.txt")).read())
    assert normalised == golden


@pytest.mark.gpu
def test_local_mode_and_json(built, tmp_path):
    js = tmp_path / "local.json"
    out = run([MPIRUN, "-n", "1", P2P_MATRIX, "--mode", "local", "--sweep-min", "1M", "--sweep-max", "64M", "--iters", "10",
               "--verify", "--json", str(js)], timeout=300).stdout
    assert "verify: OK" in out
    recs = [json.loads(l) for l in js.read_text().splitlines()]
    assert [r["bytes"] for r in recs] == [1 << 20 << i for i in range(7)]
    assert all(r["mode"] == "local" and r["GBps"] > 0 for r in recs)
