"""Parity tests on one B200: every data path of the engine, called through the C ABI
(ctypes on libb200p2p.so), compared byte for byte with the oracle.

One GPU means one rank, so the transfer is a self send/recv inside a group.  Every test runs
twice: as the engine does it by default (a matched self send/recv is one local copy, no flags)
and with the "loopback" option, which sends the same bytes through the full wire protocol —
eager ring, push with post/done flags, pull with adaptive receive — with the "peer" mapping
pointing at local memory.  Cross-GPU runs are in test_gpu_multi.py.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SEED = 0x5EED0000
CAP = 64 << 20


@pytest.fixture(scope="module", params=["copy", "loopback"])
def eng(request, p2p, oracle):
    import torch
    assert torch.cuda.is_available(), "gpu tests need a GPU"
    torch.cuda.set_device(0)
    p2p.init(0, 1, 0)
    p2p.set_option("loopback", 1 if request.param == "loopback" else 0)
    src = p2p.DeviceBuffer(CAP + 4096)
    dst = p2p.DeviceBuffer(CAP + 4096)
    win = p2p.register(dst.ptr, dst.nbytes)
    p2p.register(src.ptr, src.nbytes)      # lets large messages be pulled
    stream = torch.cuda.Stream()

    class E:
        pass

    e = E()
    e.p2p, e.oracle, e.src, e.dst, e.win, e.stream, e.torch = p2p, oracle, src, dst, win, stream, torch
    e.iter = 0
    e.loopback = request.param == "loopback"
    yield e
    p2p.set_option("path", p2p.PATH_AUTO)
    torch.cuda.synchronize()
    p2p.finalize()


def transfer(e, n, src_off=0, dst_off=0, dtype=None, count=None):
    """One patterned self transfer of n bytes; returns the received bytes plus 32 guard bytes either side."""
    p2p, torch = e.p2p, e.torch
    e.iter += 1
    it = e.iter
    p2p.pattern_fill(e.src.ptr + src_off, n, SEED, 0, it, e.stream)
    e.dst.tensor().fill_(0xA5)
    torch.cuda.synchronize()
    dtype = p2p.INT8 if dtype is None else dtype
    count = n if count is None else count
    with p2p.group():
        p2p.recv(e.dst.ptr + dst_off, count, dtype, 0, e.stream)
        p2p.send(e.src.ptr + src_off, count, dtype, 0, e.stream)
    e.stream.synchronize()
    p2p.raise_async_error()
    lo = max(0, dst_off - 32)
    got = e.dst.tensor()[lo: dst_off + n + 32].cpu().numpy()
    return got, dst_off - lo, it


def assert_exact(e, n, src_off=0, dst_off=0, **kw):
    got, lead, it = transfer(e, n, src_off, dst_off, **kw)
    want = e.oracle.expected_recv(n, SEED, 0, it)
    body = got[lead: lead + n]
    if not np.array_equal(body, want):
        diff = np.nonzero(body != want)[0]
        raise AssertionError(f"{n} bytes (src+{src_off}, dst+{dst_off}): {diff.size} bytes differ, first at {diff[0]}")
    assert np.all(got[:lead] == 0xA5), "wrote before the start of the receive buffer"
    assert np.all(got[lead + n:] == 0xA5), "wrote past the end of the message"


SIZES = [0, 1, 7, 8, 9, 15, 16, 17, 31, 32, 33, 255, 4095, 4096, 4097, 16383, 16384, 16385, 65535, 65536, 65537,
         (512 << 10) - 1, 512 << 10, (512 << 10) + 1, (1 << 20) - 1, 1 << 20, (1 << 20) + 3, (4 << 20) + 16, 16 << 20]


@pytest.mark.parametrize("n", SIZES)
def test_auto_path_sizes(eng, n):
    eng.p2p.set_option("path", eng.p2p.PATH_AUTO)
    assert_exact(eng, n)


@pytest.mark.parametrize("n", [16385, 65536, (1 << 20) + 3, (8 << 20) + 48, 32 << 20])
def test_ldst_push_path(eng, n):
    eng.p2p.set_option("path", eng.p2p.PATH_LDST)
    eng.p2p.set_option("eager_bytes", 16384)     # one rank: the limit may shrink so that small pushes are exercised too
    try:
        assert eng.p2p.query_plan(n, 0)[0] == eng.p2p.PATH_LDST
        assert_exact(eng, n)
    finally:
        eng.p2p.set_option("eager_bytes", 512 << 10)


@pytest.mark.parametrize("n", [4 << 20, (4 << 20) + 1, (8 << 20) + 48, (32 << 20) - 5])
@pytest.mark.parametrize("src_off,dst_off", [(0, 0), (16, 48), (3, 19), (5, 6)])
@pytest.mark.parametrize("tma_from", [-1, 4 << 20], ids=["ldst", "tma"])
def test_pull_path(eng, n, src_off, dst_off, tma_from):
    """At or above the pull threshold (4 MiB) the receiver loads the payload from the sender's
    registered buffer - with the ld/st loop, or through the bulk-copy ring (cp.async.bulk from the peer mapping)
    when both sides share their 16-byte phase.  (Only the loopback run exercises the wire protocol.)"""
    eng.p2p.set_option("path", eng.p2p.PATH_AUTO)
    eng.p2p.set_option("bidir_push", 0)   # sender and receiver are the same rank here; do not take that for a bidirectional exchange
    eng.p2p.set_option("pull_tma_min_bytes", tma_from)
    try:
        assert_exact(eng, n, src_off, dst_off)
        if eng.loopback:
            assert eng.p2p.get_option("stat_last_flavour") == (2 if tma_from >= 0 else 1)
    finally:
        eng.p2p.set_option("bidir_push", 1)
        eng.p2p.set_option("pull_tma_min_bytes", 32 << 20)


@pytest.mark.parametrize("pdl,early", [(0, 0), (1, 0), (2, 0), (3, 0), (3, 1)])
@pytest.mark.parametrize("bidir_push", [0, 1])
def test_pdl_modes_back_to_back(eng, pdl, early, bidir_push):
    """Programmatic dependent launch for every group (trigger at kernel entry, or after the copy loop), with and
    without early posts: six messages of every protocol enqueued back to back on one stream, no host sync in
    between, each into its own part of the receive buffer; every byte compared afterwards."""
    p2p, torch = eng.p2p, eng.torch
    p2p.set_option("path", p2p.PATH_AUTO)
    p2p.set_option("pdl", pdl)
    p2p.set_option("early_post", early)
    p2p.set_option("bidir_push", bidir_push)
    sizes = [(8 << 20) + 16, 2 << 20, 1000, (5 << 20) + 3, 8, 6 << 20]
    try:
        for rep in range(2):
            offs, off = [], 0
            for n in sizes:
                offs.append(off)
                off += (n + 4095) // 4096 * 4096 + 4096
            base_it = 1000 + 10 * rep + 100 * pdl + early
            for k, n in enumerate(sizes):
                p2p.pattern_fill(eng.src.ptr + offs[k], n, SEED, 0, base_it + k, eng.stream)
            eng.dst.tensor().fill_(0xA5)
            torch.cuda.synchronize()
            for k, n in enumerate(sizes):
                with p2p.group():
                    p2p.recv(eng.dst.ptr + offs[k], n, p2p.INT8, 0, eng.stream)
                    p2p.send(eng.src.ptr + offs[k], n, p2p.INT8, 0, eng.stream)
            eng.stream.synchronize()
            p2p.raise_async_error()
            got = eng.dst.tensor()[:off].cpu().numpy()
            for k, n in enumerate(sizes):
                want = eng.oracle.expected_recv(n, SEED, 0, base_it + k)
                assert np.array_equal(got[offs[k]: offs[k] + n], want), (pdl, early, k, n)
                assert np.all(got[offs[k] + n: offs[k] + n + 64] == 0xA5), "wrote past the end of a message"
    finally:
        p2p.set_option("pdl", 3)
        p2p.set_option("early_post", 1)
        p2p.set_option("bidir_push", 1)


def test_large_message_fallbacks(eng):
    """Large message with only one side registered: unregistered sender -> it announces a push;
    unregistered receiver -> pull still works (neither: test_unregistered_buffers_go_through_the_staging_ring)."""
    p2p, torch = eng.p2p, eng.torch
    p2p.set_option("path", p2p.PATH_AUTO)
    p2p.set_option("bidir_push", 0)
    n = (6 << 20) + 11
    loose_src, loose_dst = p2p.DeviceBuffer(n), p2p.DeviceBuffer(n + 64)
    try:
        for src, dst, it in ((loose_src, eng.dst, 31), (eng.src, loose_dst, 32)):
            p2p.pattern_fill(src.ptr, n, SEED, 0, it, eng.stream)
            dst.tensor().fill_(0xA5)
            torch.cuda.synchronize()
            with p2p.group():
                p2p.recv(dst.ptr, n, p2p.INT8, 0, eng.stream)
                p2p.send(src.ptr, n, p2p.INT8, 0, eng.stream)
            eng.stream.synchronize()
            p2p.raise_async_error()
            got = dst.tensor()[: n + 32].cpu().numpy()
            assert np.array_equal(got[:n], eng.oracle.expected_recv(n, SEED, 0, it))
            assert np.all(got[n:] == 0xA5)
    finally:
        p2p.set_option("bidir_push", 1)
        torch.cuda.synchronize()
        loose_src.free()
        loose_dst.free()


@pytest.mark.parametrize("chunk,stages", [(16384, 8), (32768, 4), (8192, 16), (65536, 3)])
@pytest.mark.parametrize("n", [16400, 65536, (1 << 20) + 3, (8 << 20) + 48, 32 << 20])
def test_tma_path(eng, n, chunk, stages):
    p2p = eng.p2p
    p2p.set_option("path", p2p.PATH_TMA)
    p2p.set_option("eager_bytes", 16384)
    p2p.set_option("tma_stages", 2)
    p2p.set_option("tma_chunk", chunk)
    p2p.set_option("tma_stages", stages)
    try:
        assert p2p.query_plan(n, 0)[0] == p2p.PATH_TMA
        assert_exact(eng, n)
    finally:
        p2p.set_option("eager_bytes", 512 << 10)
        p2p.set_option("tma_stages", 2)
        p2p.set_option("tma_chunk", 32768)
        p2p.set_option("tma_stages", 6)


@pytest.mark.parametrize("path", ["ldst", "tma", "auto"])
@pytest.mark.parametrize("src_off,dst_off", [(0, 0), (1, 1), (3, 19), (16, 48), (5, 6), (8, 0), (4, 0), (2, 0), (0, 1), (13, 7)])
def test_misaligned_buffers(eng, path, src_off, dst_off):
    """Same 16-byte phase, and every relative misalignment class (8, 4, 2, 1 bytes)."""
    p2p = eng.p2p
    p2p.set_option("path", {"ldst": p2p.PATH_LDST, "tma": p2p.PATH_TMA, "auto": p2p.PATH_AUTO}[path])
    for n in (100, 70_001, (2 << 20) + 5):
        assert_exact(eng, n, src_off, dst_off)


@pytest.mark.parametrize("dtype_name,size", [("INT8", 1), ("UINT8", 1), ("INT32", 4), ("UINT32", 4), ("INT64", 8),
                                              ("UINT64", 8), ("FLOAT16", 2), ("FLOAT32", 4), ("FLOAT64", 8),
                                              ("BFLOAT16", 2), ("FLOAT8E4M3", 1), ("FLOAT8E5M2", 1)])
def test_datatypes_scale_count(eng, dtype_name, size):
    """dtype only scales count to bytes (ncclDataType_t numbering, nccl.h:278-291)."""
    eng.p2p.set_option("path", eng.p2p.PATH_AUTO)
    count = 100_003
    assert_exact(eng, count * size, dtype=getattr(eng.p2p, dtype_name), count=count)


def test_cfg1_shape_64mib_float(eng):
    """BASELINE config 1's message: 16,777,216 float32 = 64 MiB."""
    eng.p2p.set_option("path", eng.p2p.PATH_AUTO)
    assert_exact(eng, 64 << 20, dtype=eng.p2p.FLOAT32, count=16_777_216)


def test_sequence_rings_wrap(eng):
    """More messages than ring slots on both protocols (64 rendezvous, 4 eager) without a host sync."""
    p2p = eng.p2p
    p2p.set_option("path", p2p.PATH_AUTO)
    for n, reps in ((64, 40), (100_000, 40), ((1 << 20) + 4096, 150)):
        p2p.pattern_fill(eng.src.ptr, n, SEED, 0, 9000 + n, eng.stream)
        eng.dst.tensor().fill_(0xA5)
        eng.torch.cuda.synchronize()
        for _ in range(reps):
            with p2p.group():
                p2p.recv(eng.dst.ptr, n, p2p.INT8, 0, eng.stream)
                p2p.send(eng.src.ptr, n, p2p.INT8, 0, eng.stream)
        eng.stream.synchronize()
        p2p.raise_async_error()
        got = eng.dst.tensor()[:n].cpu().numpy()
        assert np.array_equal(got, eng.oracle.expected_recv(n, SEED, 0, 9000 + n))


def test_group_with_several_transfers(eng):
    """Four sends and four receives of different sizes and protocols in one group (one launch)."""
    p2p = eng.p2p
    p2p.set_option("path", p2p.PATH_AUTO)
    sizes = [8, 5000, 300_000, 3 << 20]
    offs = [0, 1 << 20, 2 << 20, 8 << 20]
    for k, (n, off) in enumerate(zip(sizes, offs)):
        p2p.pattern_fill(eng.src.ptr + off, n, SEED, 0, 500 + k, eng.stream)
    eng.dst.tensor().fill_(0xA5)
    eng.torch.cuda.synchronize()
    with p2p.group():
        for n, off in zip(sizes, offs):
            p2p.recv(eng.dst.ptr + off, n, p2p.INT8, 0, eng.stream)
        for n, off in zip(sizes, offs):
            p2p.send(eng.src.ptr + off, n, p2p.INT8, 0, eng.stream)
    eng.stream.synchronize()
    p2p.raise_async_error()
    for k, (n, off) in enumerate(zip(sizes, offs)):
        got = eng.dst.tensor()[off: off + n + 8].cpu().numpy()
        assert np.array_equal(got[:n], eng.oracle.expected_recv(n, SEED, 0, 500 + k))
        assert np.all(got[n:] == 0xA5)


def test_two_streams_in_one_group(eng):
    """The reference's bidirectional call shape: send on s_0, recv on s_1, grouped (p2p_matrix.cc:212-228)."""
    p2p, torch = eng.p2p, eng.torch
    s1 = torch.cuda.Stream()
    n = (4 << 20) + 7
    p2p.pattern_fill(eng.src.ptr, n, SEED, 0, 77, eng.stream)
    eng.dst.tensor().fill_(0xA5)
    torch.cuda.synchronize()
    with p2p.group():
        p2p.send(eng.src.ptr, n, p2p.INT8, 0, eng.stream)
        p2p.recv(eng.dst.ptr, n, p2p.INT8, 0, s1)
    s1.synchronize()   # completion must be visible on the receive stream alone
    got = eng.dst.tensor()[:n].cpu().numpy()
    eng.stream.synchronize()
    assert np.array_equal(got, eng.oracle.expected_recv(n, SEED, 0, 77))


def test_full_size_256mib_properties(eng, p2p, oracle):
    """BASELINE's 256 MiB message: checked by size-independent properties — device-side byte
    compare against the pattern, digest equal to the oracle's digest of the sender's payload,
    and idempotence (a second transfer changes nothing)."""
    import torch
    n = 256 << 20
    src = p2p.DeviceBuffer(n)
    dst = p2p.DeviceBuffer(n + 64)
    win = p2p.register(dst.ptr, dst.nbytes)
    stream = torch.cuda.Stream()
    try:
        for path in (p2p.PATH_LDST, p2p.PATH_TMA):
            p2p.set_option("path", path)
            p2p.pattern_fill(src.ptr, n, SEED, 0, 1000 + path, stream)
            dst.tensor().fill_(0xA5)
            torch.cuda.synchronize()
            for _ in range(2):
                with p2p.group():
                    p2p.recv(dst.ptr, n, p2p.INT8, 0, stream)
                    p2p.send(src.ptr, n, p2p.INT8, 0, stream)
                bad, first, dig = p2p.pattern_check(dst.ptr, n, SEED, 0, 1000 + path, stream)
                assert bad == 0, f"path {path}: {bad} bytes differ, first at {first}"
                assert dig == oracle.pattern_digest(n, SEED, 0, 1000 + path)
                assert dig == p2p.digest(src.ptr, n, stream)
            assert bytes(dst.tensor()[n: n + 64].cpu().numpy()) == b"\xa5" * 64
            p2p.raise_async_error()
    finally:
        p2p.set_option("path", p2p.PATH_AUTO)
        torch.cuda.synchronize()
        p2p.deregister(win)
        src.free()
        dst.free()


@pytest.mark.parametrize("block", [0, 8, 4097, 600_001, (5 << 20) + 3])
def test_alltoall_one_rank(eng, block):
    """p2p_alltoall with a world of one: its only block is the local one, checked against the
    oracle's all-to-all law (tests/test_gpu_multi.py runs it across GPUs)."""
    p2p, torch = eng.p2p, eng.torch
    p2p.set_option("path", p2p.PATH_AUTO)
    want_send = eng.oracle.alltoall_send(block, SEED, 0, 1)
    eng.src.tensor()[:block].copy_(torch.from_numpy(want_send)) if block else None
    eng.dst.tensor().fill_(0xA5)
    torch.cuda.synchronize()
    p2p.alltoall(eng.src.ptr, eng.dst.ptr, block, p2p.INT8, eng.stream)
    eng.stream.synchronize()
    p2p.raise_async_error()
    got = eng.dst.tensor()[: block + 32].cpu().numpy()
    assert np.array_equal(got[:block], eng.oracle.expected_alltoall_recv(block, SEED, 0, 1))
    assert np.all(got[block:] == 0xA5)


@pytest.mark.parametrize("n", [600_000, (1 << 20) + 3, 2 << 20, (4 << 20) - 1, 4 << 20, (6 << 20) + 11, (21 << 20) + 5])
@pytest.mark.parametrize("src_off,dst_off,send_registered", [(0, 0, False), (3, 5, False), (16, 48, True)])
def test_unregistered_buffers_go_through_the_staging_ring(eng, n, src_off, dst_off, send_registered):
    """The reference never registers anything (plain cudaMalloc, p2p_matrix.cc:127-128).  A receive buffer that is
    in no window cannot be stored into by the sender: the message then goes chunk by chunk through the staging
    ring in the receiver's control region - below the pull threshold always, above it when the send buffer is in
    no window either (otherwise the receiver pulls as ever).  21 MiB is more chunks than the ring has slots."""
    if not eng.loopback:
        pytest.skip("a matched self send/recv is a local copy and needs no window")
    p2p, torch = eng.p2p, eng.torch
    p2p.set_option("path", p2p.PATH_AUTO)
    p2p.set_option("bidir_push", 0)
    loose_src, loose_dst = p2p.DeviceBuffer(n + 64), p2p.DeviceBuffer(n + 128)
    try:
        src = eng.src if send_registered else loose_src
        for rep in range(2):       # twice: the ring's counters carry over from message to message
            it = 7000 + rep
            p2p.pattern_fill(src.ptr + src_off, n, SEED, 0, it, eng.stream)
            loose_dst.tensor().fill_(0xA5)
            torch.cuda.synchronize()
            with p2p.group():
                p2p.recv(loose_dst.ptr + dst_off, n, p2p.INT8, 0, eng.stream)
                p2p.send(src.ptr + src_off, n, p2p.INT8, 0, eng.stream)
            eng.stream.synchronize()
            p2p.raise_async_error()
            got = loose_dst.tensor().cpu().numpy()
            assert np.array_equal(got[dst_off: dst_off + n], eng.oracle.expected_recv(n, SEED, 0, it)), (n, rep)
            assert np.all(got[:dst_off] == 0xA5) and np.all(got[dst_off + n:] == 0xA5), "wrote outside the message"
        # the ordinary paths still work afterwards (flags and counters of the ring do not leak into them)
        assert_exact(eng, 2 << 20)
        assert_exact(eng, (8 << 20) + 16)
    finally:
        p2p.set_option("bidir_push", 1)
        torch.cuda.synchronize()
        loose_src.free()
        loose_dst.free()


def test_argument_errors(eng):
    p2p = eng.p2p
    L = p2p.lib()
    assert L.p2p_send(ctypes.c_void_p(eng.src.ptr), 8, 12, 0, None) == p2p.INVALID_ARGUMENT    # dtype out of range
    assert L.p2p_send(ctypes.c_void_p(eng.src.ptr), 8, -1, 0, None) == p2p.INVALID_ARGUMENT
    assert L.p2p_send(ctypes.c_void_p(eng.src.ptr), 8, 0, 1, None) == p2p.INVALID_ARGUMENT     # peer out of range
    assert L.p2p_recv(ctypes.c_void_p(eng.dst.ptr), 8, 0, -1, None) == p2p.INVALID_ARGUMENT
    assert L.p2p_send(None, 8, 0, 0, None) == p2p.INVALID_ARGUMENT                              # null buffer
    assert L.p2p_send(ctypes.c_void_p(eng.src.ptr), 2**62, 8, 0, None) == p2p.INVALID_ARGUMENT  # overflow
    assert L.p2p_group_end() == p2p.INVALID_USAGE                                               # unbalanced
    assert L.p2p_init(0, 1, 0) == p2p.INVALID_USAGE                                             # twice
    assert L.p2p_set_option(b"nonsense", 1) == p2p.INVALID_ARGUMENT
    assert L.p2p_set_option(b"tma_chunk", 100) == p2p.INVALID_ARGUMENT
    assert p2p.async_error() == p2p.SUCCESS
    assert_exact(eng, 4096)


@pytest.mark.parametrize("nbytes,what", [(2 << 20, "never completed"), (64, "never arrived")])
def test_missing_peer_times_out_instead_of_hanging(eng, nbytes, what):
    """A receive whose sender never comes reports P2P_REMOTE_ERROR after the timeout, and the engine then refuses
    further transfers (its counters and flag rings are no longer consistent) until p2p_finalize.  Run in a process
    of its own for that reason."""
    from conftest import ROOT, run
    code = r'''
import importlib, sys
sys.path.insert(0, %r)
import torch
p2p = importlib.import_module("test-nccl-p2p_b200")
torch.cuda.set_device(0)
p2p.init(0, 1, 0)
buf = p2p.DeviceBuffer(2 << 20)
p2p.register(buf.ptr, buf.nbytes)
p2p.set_option("timeout_ms", 200)
s = torch.cuda.Stream()
p2p.recv(buf.ptr, %d, p2p.INT8, 0, s)      # no matching send
s.synchronize()
assert p2p.async_error() == p2p.REMOTE_ERROR
assert %r in p2p.lib().p2p_get_last_error().decode(), p2p.lib().p2p_get_last_error().decode()
try:
    p2p.recv(buf.ptr, 64, p2p.INT8, 0, s)
    raise SystemExit("a transfer was accepted after a device-side timeout")
except p2p.P2PError as e:
    assert e.code == p2p.REMOTE_ERROR and "error state" in str(e), e
p2p.finalize()
p2p.init(0, 1, 0)                              # a fresh engine works again
assert p2p.async_error() == p2p.SUCCESS
p2p.finalize()
print("timed out cleanly")
''' % (ROOT, nbytes, what)
    assert "timed out cleanly" in run(["python", "-c", code], timeout=120).stdout
