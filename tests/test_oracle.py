"""The oracle against its golden vectors (no GPU).

Three restatements must agree with tests/golden/pattern_vectors.json: the C code
(oracle/pattern.c), numpy (oracle/pattern.py) and — in the gpu tests — the CUDA kernels.
Also here: the reference's pure host functions (host hash p2p_matrix.cc:44-61, placement
:63-100, Gbps :177/:258) as restated by the oracle and by the benchmark driver.
"""
import ctypes
import json
import os

import numpy as np
import pytest

from conftest import ROOT, run

GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "pattern_vectors.json")))


def test_c_oracle_matches_golden(oracle):
    lib = oracle.load_c()
    for c in GOLDEN["small"]:
        data = oracle.c_pattern_bytes(c["nbytes"], c["seed"], c["rank"], c["iter"])
        assert data.tobytes().hex() == c["hex"]
        assert oracle.c_digest(data) == c["digest"]
        for i in range(c["nbytes"]):
            assert lib.oracle_pattern_byte(c["seed"], c["rank"], c["iter"], i) == data[i]
    for c in GOLDEN["mix64"]:
        assert lib.oracle_mix64(c["z"]) == c["out"]


def test_numpy_oracle_matches_golden(oracle):
    for c in GOLDEN["small"]:
        data = oracle.pattern_bytes(c["nbytes"], c["seed"], c["rank"], c["iter"])
        assert data.tobytes().hex() == c["hex"]
        assert oracle.digest(data) == c["digest"]
    for c in GOLDEN["mix64"]:
        assert int(oracle.mix64(np.uint64(c["z"]))) == c["out"]


@pytest.mark.parametrize("case", GOLDEN["big"], ids=lambda c: str(c["nbytes"]))
def test_big_vectors(oracle, case):
    n, seed, rank, it = case["nbytes"], case["seed"], case["rank"], case["iter"]
    data = oracle.pattern_bytes(n, seed, rank, it)
    assert data[:32].tobytes().hex() == case["head_hex"]
    assert data[-32:].tobytes().hex() == case["tail_hex"]
    assert oracle.digest(data) == case["digest"] == case["pattern_digest"]
    if n <= (2 << 20):
        assert np.array_equal(oracle.c_pattern_bytes(n, seed, rank, it), data)


def test_transfer_law_and_detection(oracle):
    """expected_recv is the sender's payload; every kind of wrong answer is seen."""
    n, seed = 100_003, 0x5EED0000
    want = oracle.expected_recv(n, seed, 3, 17)
    assert np.array_equal(want, oracle.pattern_bytes(n, seed, 3, 17))
    assert oracle.mismatches(want, seed, 3, 17) == (0, n)
    # wrong peer, stale iteration, shifted copy, dropped tail, single flipped bit
    assert oracle.mismatches(oracle.pattern_bytes(n, seed, 2, 17), seed, 3, 17)[0] > n * 0.99
    assert oracle.mismatches(oracle.pattern_bytes(n, seed, 3, 16), seed, 3, 17)[0] > n * 0.99
    shifted = np.roll(want, 8)
    assert oracle.mismatches(shifted, seed, 3, 17)[0] > n * 0.99
    tail = want.copy()
    tail[-5:] = 0xA5
    bad, first = oracle.mismatches(tail, seed, 3, 17)
    assert 1 <= bad <= 5 and first >= n - 5
    flip = want.copy()
    flip[4242] ^= 0x10
    assert oracle.mismatches(flip, seed, 3, 17) == (1, 4242)
    assert oracle.digest(flip) != oracle.digest(want)
    # digest is position dependent: swapping two different words changes it
    sw = want.copy()
    sw[0:8], sw[8:16] = want[8:16].copy(), want[0:8].copy()
    assert oracle.digest(sw) != oracle.digest(want)
    # C and numpy agree on mismatch counting
    lib = oracle.load_c()
    first_c = ctypes.c_uint64(0)
    p8 = flip.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8))
    assert lib.oracle_pattern_mismatches(p8, n, seed, 3, 17, ctypes.byref(first_c)) == 1
    assert first_c.value == 4242


def test_host_hash_known_answers(oracle, built):
    lib = oracle.load_c()
    for c in GOLDEN["host_hash"]:
        assert oracle.host_hash(c["name"]) == c["hash"]
        assert lib.oracle_host_hash(c["name"].encode()) == c["hash"]
        if c["name"]:
            out = json.loads(run([os.path.join(ROOT, "build", "driver_unit"), "hash", c["name"]]).stdout)
            assert out["hash"] == c["hash"]
    # the reference truncates at the first '.', so FQDN and short name hash alike (p2p_matrix.cc:53-61)
    assert oracle.host_hash("node01.cluster.local") == oracle.host_hash("node01")


PLACEMENTS = [
    ([7, 7], [0, 1]),
    ([7, 7, 7, 7, 7, 7, 7, 7], list(range(8))),
    ([1, 1, 1, 1, 2, 2, 2, 2], [0, 1, 2, 3, 0, 1, 2, 3]),   # two hosts, block placement: accepted
    ([1, 2, 1, 2], [-2] * 4),                                  # round-robin across hosts: rejected (:88-98)
    ([1, 1, 2], [-1] * 3),                                     # uneven: rejected (:83-86)
    ([5], [0]),
]


@pytest.mark.parametrize("hashes,expect", PLACEMENTS)
def test_placement_policy(oracle, built, hashes, expect):
    lib = oracle.load_c()
    arr = (ctypes.c_uint64 * len(hashes))(*hashes)
    for rank, want in enumerate(expect):
        assert oracle.placement(hashes, rank) == want
        assert lib.oracle_placement(arr, len(hashes), rank) == want
        out = json.loads(run([os.path.join(ROOT, "build", "driver_unit"), "placement", str(rank)] + [str(h) for h in hashes]).stdout)
        assert out["device"] == want


def test_gbps_formula(oracle, built):
    lib = oracle.load_c()
    for c in GOLDEN["gbps"]:
        assert oracle.gbps(c["bytes"], c["seconds"], bool(c["bidir"])) == pytest.approx(c["gbps"], rel=1e-12)
        assert lib.oracle_gbps(c["bytes"], c["seconds"], c["bidir"]) == pytest.approx(c["gbps"], rel=1e-12)
        cell = run([os.path.join(ROOT, "build", "driver_unit"), "gbps", str(c["bytes"]), repr(c["seconds"]), str(c["bidir"])]).stdout
        assert cell == "%6.02f \n" % c["gbps"]
    # 900 GB/s is 7200 Gbps: wider than the %6.02f field, columns drift exactly as in the reference
    assert "%6.02f " % oracle.gbps(256 << 20, (256 << 20) / 900e9, False) == "7200.00 "


@pytest.mark.parametrize("text,ok,n", [("8", 1, 8), ("64M", 1, 64 << 20), ("1G", 1, 1 << 30), ("32MiB", 1, 32 << 20),
                                       ("4k", 1, 4096), ("1.5K", 1, 1536), ("x", 0, 0), ("12Q", 0, 0)])
def test_size_parsing(built, text, ok, n):
    out = json.loads(run([os.path.join(ROOT, "build", "driver_unit"), "size", text]).stdout)
    assert out["ok"] == ok
    if ok:
        assert out["bytes"] == n
