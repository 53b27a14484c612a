"""p2p_digest_noalloc — the one entry point that allocates nothing — against p2p_digest.

The ordinary digest keeps its partial sums and result in memory p2p_init() got from
cudaMalloc.  Processes that cannot allocate (all of HBM already owned by a framework's pool,
a loop that must not trigger cudaMalloc's device-wide synchronisation, a checker that never
starts the engine) call this variant instead: same kernel, scratch inside the library's own
device image.  Because every call reuses that one scratch area, the test makes two calls in a
row on different buffers and sizes and wants both answers right, i.e. nothing left over from
the first call may leak into the second.  Digests are integers: the comparison is exact.

This file sorts last on purpose: a run that stops at the first failure has then already been
through every other test.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def engine(p2p):
    import torch
    torch.cuda.set_device(0)
    p2p.init(0, 1, 0)          # needed by the EXISTING entry point only
    yield p2p, torch
    torch.cuda.synchronize()
    p2p.finalize()


def test_two_calls_in_a_row_match_the_allocating_digest(engine, oracle):
    p2p, torch = engine
    stream = torch.cuda.Stream()
    big, small = (5 << 20) + 123, 70_001
    a = torch.empty(big, dtype=torch.uint8, device="cuda")
    b = torch.empty(small, dtype=torch.uint8, device="cuda")
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    p2p.pattern_fill(a.data_ptr(), big, 11, 0, 1, stream)
    p2p.pattern_fill(b.data_ptr(), small, 12, 3, 2, stream)
    want_a = p2p.digest(a.data_ptr(), big, stream)       # existing entry point (synchronises)
    want_b = p2p.digest(b.data_ptr(), small, stream)
    assert want_a != want_b
    # call 1 uses 21 CTAs' worth of scratch, call 2 a single one, back to back on one stream
    p2p.digest_noalloc(a.data_ptr(), big, out.data_ptr(), stream)
    p2p.digest_noalloc(b.data_ptr(), small, out.data_ptr() + 8, stream)
    stream.synchronize()
    got = out.cpu().numpy().view(np.uint64)
    assert int(got[0]) == want_a
    assert int(got[1]) == want_b
    # and both agree with the CPU oracle's digest of the same payloads
    assert want_a == oracle.pattern_digest(big, 11, 0, 1)
    assert want_b == oracle.pattern_digest(small, 12, 3, 2)
    # a third call, large again, still right (scratch from call 2 is not mistaken for fresh)
    p2p.digest_noalloc(a.data_ptr(), big, out.data_ptr(), stream)
    stream.synchronize()
    assert int(out.cpu().numpy().view(np.uint64)[0]) == want_a


def test_edges_and_limit(engine):
    p2p, torch = engine
    L = p2p.lib()
    out = torch.zeros(1, dtype=torch.int64, device="cuda")
    buf = torch.arange(256, dtype=torch.uint8, device="cuda")
    for n in (0, 1, 255):
        p2p.digest_noalloc(buf.data_ptr(), n, out.data_ptr())
        torch.cuda.synchronize()
        assert int(out.cpu().numpy().view(np.uint64)[0]) == p2p.digest(buf.data_ptr(), n)
    rc = L.p2p_digest_noalloc(ctypes.c_void_p(buf.data_ptr()), p2p.NOALLOC_MAX_BYTES + 1, None, ctypes.c_void_p(out.data_ptr()))
    assert rc == p2p.INVALID_ARGUMENT
    assert "limit" in L.p2p_get_last_error().decode()
    assert L.p2p_digest_noalloc(ctypes.c_void_p(buf.data_ptr()), 16, None, None) == p2p.INVALID_ARGUMENT


def test_exactly_the_largest_buffer_the_scratch_covers(engine):
    """bytes == P2P_NOALLOC_MAX_BYTES: all 4096 partial-sum slots in use (the last CTA covers the last 256 KiB),
    and one byte less (the last slot only partly filled); both must match the allocating digest."""
    p2p, torch = engine
    n = p2p.NOALLOC_MAX_BYTES
    stream = torch.cuda.Stream()
    buf = torch.empty(n, dtype=torch.uint8, device="cuda")
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    p2p.pattern_fill(buf.data_ptr(), n, 77, 1, 5, stream)
    p2p.digest_noalloc(buf.data_ptr(), n, out.data_ptr(), stream)
    p2p.digest_noalloc(buf.data_ptr(), n - 1, out.data_ptr() + 8, stream)
    stream.synchronize()
    got = out.cpu().numpy().view(np.uint64)
    assert int(got[0]) == p2p.digest(buf.data_ptr(), n, stream)
    assert int(got[1]) == p2p.digest(buf.data_ptr(), n - 1, stream)
    assert int(got[0]) != int(got[1])
    del buf


def test_each_device_of_a_process_has_its_own_scratch(p2p, oracle):
    """Two GPUs driven by one process (needs two; skipped otherwise): the scratch array exists once per device, and
    the entry point looks up the current device's copy on its first call there."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    n = (3 << 20) + 5
    data = ((np.arange(n, dtype=np.int64) * 40503) & 0xFF).astype(np.uint8)
    want = oracle.digest(data)
    for dev in (0, 1, 0):
        torch.cuda.set_device(dev)
        buf = torch.from_numpy(data).to(f"cuda:{dev}")
        out = torch.zeros(1, dtype=torch.int64, device=f"cuda:{dev}")
        p2p.digest_noalloc(buf.data_ptr(), n, out.data_ptr())
        torch.cuda.synchronize(dev)
        assert int(out.cpu().numpy().view(np.uint64)[0]) == want, dev
    torch.cuda.set_device(0)


def test_works_in_a_process_that_never_initialised_the_engine(p2p, oracle):
    """Fresh process, no p2p_init, no allocation by the library: the value still matches."""
    from conftest import ROOT, run
    n = (1 << 20) + 7
    code = r'''
import importlib, sys
sys.path.insert(0, %r)
import torch
p2p = importlib.import_module("test-nccl-p2p_b200")
torch.cuda.set_device(0)
data = torch.arange(%d, dtype=torch.int64, device="cuda").mul_(2654435761).to(torch.uint8)
out = torch.zeros(1, dtype=torch.int64, device="cuda")
p2p.digest_noalloc(data.data_ptr(), data.numel(), out.data_ptr())
p2p.digest_noalloc(data.data_ptr(), data.numel(), out.data_ptr())
torch.cuda.synchronize()
import numpy as np
print("digest", int(out.cpu().numpy().view(np.uint64)[0]), flush=True)
''' % (ROOT, n)
    out = run(["python", "-c", code], timeout=180).stdout
    got = int(out.split("digest")[1].split()[0])
    data = ((np.arange(n, dtype=np.int64) * 2654435761) & 0xFF).astype(np.uint8)
    assert got == oracle.digest(data)
