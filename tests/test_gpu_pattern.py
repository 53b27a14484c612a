"""The CUDA pattern / check / digest kernels against the golden vectors and the oracle.
They are the third restatement of oracle/pattern.c (after C and numpy) and the tool that
verifies full-size buffers in place, so they must agree bit for bit."""
import json
import os

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "pattern_vectors.json")))


@pytest.fixture(scope="module")
def dev(p2p, oracle):
    import torch
    torch.cuda.set_device(0)
    p2p.init(0, 1, 0)
    buf = p2p.DeviceBuffer((64 << 20) + 64)
    yield p2p, oracle, buf, torch
    torch.cuda.synchronize()
    p2p.finalize()


def test_fill_matches_golden_small(dev):
    p2p, oracle, buf, torch = dev
    for c in GOLDEN["small"]:
        n = c["nbytes"]
        buf.tensor()[: n + 16].fill_(0xEE)
        torch.cuda.synchronize()
        p2p.pattern_fill(buf.ptr, n, c["seed"], c["rank"], c["iter"])
        torch.cuda.synchronize()
        got = buf.tensor()[: n + 16].cpu().numpy()
        assert got[:n].tobytes().hex() == c["hex"]
        assert np.all(got[n:] == 0xEE), "fill wrote past the end"
        bad, first, dig = p2p.pattern_check(buf.ptr, n, c["seed"], c["rank"], c["iter"])
        assert (bad, first, dig) == (0, n, c["digest"])
        assert p2p.digest(buf.ptr, n) == c["digest"]


@pytest.mark.parametrize("case", GOLDEN["big"], ids=lambda c: str(c["nbytes"]))
def test_fill_matches_golden_big(dev, case):
    p2p, oracle, buf, torch = dev
    n = case["nbytes"]
    p2p.pattern_fill(buf.ptr, n, case["seed"], case["rank"], case["iter"])
    bad, first, dig = p2p.pattern_check(buf.ptr, n, case["seed"], case["rank"], case["iter"])
    assert bad == 0 and first == n
    assert dig == case["digest"]
    t = buf.tensor()
    assert t[:32].cpu().numpy().tobytes().hex() == case["head_hex"]
    assert t[n - 32: n].cpu().numpy().tobytes().hex() == case["tail_hex"]


@pytest.mark.parametrize("offset", [0, 1, 3, 4, 7])
@pytest.mark.parametrize("n", [1, 13, 4096, 100_003])
def test_misaligned_and_ragged(dev, offset, n):
    p2p, oracle, buf, torch = dev
    buf.tensor()[: n + 64].fill_(0x11)
    torch.cuda.synchronize()
    p2p.pattern_fill(buf.ptr + offset, n, 7, 2, 5)
    torch.cuda.synchronize()
    got = buf.tensor()[: n + 64].cpu().numpy()
    assert np.array_equal(got[offset: offset + n], oracle.pattern_bytes(n, 7, 2, 5))
    assert np.all(got[:offset] == 0x11) and np.all(got[offset + n:] == 0x11)
    assert p2p.digest(buf.ptr + offset, n) == oracle.pattern_digest(n, 7, 2, 5)


def test_check_counts_and_locates_corruption(dev):
    p2p, oracle, buf, torch = dev
    n = (8 << 20) + 5
    p2p.pattern_fill(buf.ptr, n, 99, 4, 1)
    torch.cuda.synchronize()
    t = buf.tensor()
    clean = p2p.digest(buf.ptr, n)
    for pos in (0, 12345, n - 1):
        t[pos] ^= 0x80
        torch.cuda.synchronize()
        bad, first, dig = p2p.pattern_check(buf.ptr, n, 99, 4, 1)
        assert (bad, first) == (1, pos) and dig != clean
        t[pos] ^= 0x80
    t[1000:1010] ^= 0xFF
    t[n - 3: n] ^= 0x01
    torch.cuda.synchronize()
    bad, first, _ = p2p.pattern_check(buf.ptr, n, 99, 4, 1)
    assert (bad, first) == (13, 1000)
    # a payload from the wrong rank or iteration is (almost) all wrong
    assert p2p.pattern_check(buf.ptr, n, 99, 5, 1)[0] > 0.99 * n
    assert p2p.pattern_check(buf.ptr, n, 99, 4, 2)[0] > 0.99 * n


def test_digest_agrees_with_oracle_on_arbitrary_bytes(dev):
    p2p, oracle, buf, torch = dev
    rng = np.random.default_rng(1234)
    for n in (0, 5, 8, 1000, 1 << 20):
        data = rng.integers(0, 256, size=n, dtype=np.uint8)
        buf.tensor()[:n].copy_(torch.from_numpy(data))
        torch.cuda.synchronize()
        assert p2p.digest(buf.ptr, n) == oracle.digest(data) == oracle.c_digest(data)
