"""Symmetric windows (driver VMM memory shared by POSIX descriptor or fabric handle) and p2p_bcast over the
NVSwitch multicast mapping.

CPU: the descriptor hand-over between processes (Unix-domain socket, SCM_RIGHTS), which is plain host code.
GPU, one rank: a symmetric window over one device, broadcast from a separate send buffer into it — through the
multicast mapping (multimem.st) when the device supports one, and through unicast stores — compared with the
oracle; the window used as an ordinary p2p window.
GPU, several ranks (skipped on a one-GPU box): `p2p_matrix --mode bcast --verify` on every rank count, each way of
sharing the memory, against the NCCL harness's digests.

Law (oracle): after p2p_bcast(root) every rank's receive buffer equals the root's send buffer, byte for byte —
what N-1 ncclSend/ncclRecv pairs in the reference's loop shape (p2p_matrix.cc:141-172) would leave there.
"""
import json
import os

import numpy as np
import pytest

from conftest import MPIRUN, P2P_MATRIX, REF_DIR, ROOT, gpu_count, run

SEED = 0x5EED0010


def test_descriptor_handover_between_processes(built):
    out = run([os.path.join(ROOT, "build", "fdpass_unit")], timeout=60).stdout
    assert "fdpass ok" in out


@pytest.fixture()
def engine(p2p, oracle):
    import torch
    torch.cuda.set_device(0)
    p2p.init(0, 1, 0)
    yield p2p, torch, oracle
    torch.cuda.synchronize()
    p2p.finalize()


@pytest.mark.gpu
@pytest.mark.parametrize("unicast", [False, True], ids=["multicast", "unicast"])
def test_broadcast_into_a_symmetric_window_on_one_gpu(engine, unicast, monkeypatch):
    p2p, torch, oracle = engine
    if unicast:
        monkeypatch.setenv("P2P_BCAST_UNICAST", "1")
    cap = (16 << 20) + 4096
    win = p2p.SymmBuffer(cap)
    src = p2p.DeviceBuffer(cap)
    stream = torch.cuda.Stream()
    try:
        assert win.ptr and win.window >= 0
        if not unicast and not win.multicast_ptr:
            pytest.skip("no multicast object on this device; the unicast run covers the fallback")
        it = 0
        for n, src_off, dst_off in [(8, 0, 0), (1000, 0, 16), ((1 << 20) + 5, 0, 0), (16 << 20, 0, 4096), (70001, 3, 32), (4096, 16, 48)]:
            for rep in range(2):
                it += 1
                p2p.pattern_fill(src.ptr + src_off, n, SEED, 0, it, stream)
                win.tensor().fill_(0xA5)
                torch.cuda.synchronize()
                p2p.bcast(src.ptr + src_off, win.ptr + dst_off, n, p2p.INT8, 0, stream)
                stream.synchronize()
                p2p.raise_async_error()
                got = win.tensor()[: dst_off + n + 64].cpu().numpy()
                assert np.array_equal(got[dst_off: dst_off + n], oracle.expected_recv(n, SEED, 0, it)), (n, src_off, dst_off, rep)
                assert np.all(got[:dst_off] == 0xA5) and np.all(got[dst_off + n:] == 0xA5), "wrote outside the message"
        # typed counts scale like ncclBroadcast's
        p2p.pattern_fill(src.ptr, 4096, SEED, 0, 999, stream)
        p2p.bcast(src.ptr, win.ptr, 1024, p2p.FLOAT32, 0, stream)
        stream.synchronize()
        assert np.array_equal(win.tensor()[:4096].cpu().numpy(), oracle.expected_recv(4096, SEED, 0, 999))
    finally:
        torch.cuda.synchronize()
        win.free()
        src.free()


@pytest.mark.gpu
def test_symmetric_window_serves_send_and_recv(engine):
    """The window is registered with the engine like a cudaIpc window: a rendezvous-sized self transfer through the
    wire protocol (loopback) lands in it."""
    p2p, torch, oracle = engine
    p2p.set_option("loopback", 1)
    p2p.set_option("bidir_push", 0)
    n = (2 << 20) + 7
    a, b = p2p.SymmBuffer(n + 64), p2p.SymmBuffer(n + 64)
    stream = torch.cuda.Stream()
    try:
        p2p.pattern_fill(a.ptr, n, SEED, 0, 5, stream)
        b.tensor().fill_(0xA5)
        torch.cuda.synchronize()
        with p2p.group():
            p2p.recv(b.ptr, n, p2p.INT8, 0, stream)
            p2p.send(a.ptr, n, p2p.INT8, 0, stream)
        stream.synchronize()
        p2p.raise_async_error()
        assert p2p.get_option("stat_last_push_ops") == 1      # pushed into the window, not staged
        got = b.tensor().cpu().numpy()
        assert np.array_equal(got[:n], oracle.expected_recv(n, SEED, 0, 5)) and np.all(got[n:] == 0xA5)
    finally:
        torch.cuda.synchronize()
        a.free()
        b.free()


@pytest.mark.gpu
def test_broadcast_argument_errors(engine):
    p2p, torch, _ = engine
    buf = p2p.DeviceBuffer(4096)
    try:
        with pytest.raises(p2p.P2PError) as ei:
            p2p.bcast(buf.ptr, buf.ptr, 16, p2p.INT8, 0)
        assert ei.value.code == p2p.INVALID_USAGE and "symmetric window" in str(ei.value)
        with pytest.raises(p2p.P2PError):
            p2p.bcast(buf.ptr, buf.ptr, 16, p2p.INT8, 3)       # root out of range
        with pytest.raises(p2p.P2PError):
            p2p.bcast(buf.ptr, buf.ptr, 16, 99, 0)              # datatype
    finally:
        buf.free()


NGPU = gpu_count()
RANKS = sorted({2, min(4, NGPU), min(8, NGPU)} - {0, 1, 3}) or [2]


@pytest.mark.gpu
@pytest.mark.skipif(NGPU < 2, reason="needs two GPUs")
@pytest.mark.parametrize("how", ["multicast", "no_multicast", "sends", "fabric"])
@pytest.mark.parametrize("n", RANKS)
def test_broadcast_across_gpus_verified(built, n, how):
    env = {"multicast": {}, "no_multicast": {"P2P_NO_MULTICAST": "1"}, "sends": {"P2P_BCAST_SENDS": "1"},
           "fabric": {"P2P_SYMM_HANDLE": "fabric"}}[how]
    for size in ("8", "70001", "4M", "33554437"):
        proc = run([MPIRUN, "-n", str(n), P2P_MATRIX, "--mode", "bcast", "--size", size, "--iters", "5", "--warmup", "2", "--timing", "dev", "--verify"],
                   timeout=300, env=env, check=False)
        if how == "fabric" and proc.returncode != 0 and ("fabric" in proc.stdout or "CUresult" in proc.stdout):
            pytest.skip("fabric handles cannot be exported on this box: " + proc.stdout.strip().splitlines()[-1][:200])
        assert proc.returncode == 0 and "verify OK" in proc.stdout and "VERIFY FAILED" not in proc.stdout, proc.stdout[-2000:]


@pytest.mark.gpu
@pytest.mark.skipif(NGPU < 2, reason="needs two GPUs")
def test_broadcast_digests_equal_nccl(built, tmp_path):
    """Every receiver's buffer after p2p_bcast equals what NCCL's N-1 sends from the root leave (digests; NCCL's are
    computed on the CPU by the oracle)."""
    n = RANKS[-1]
    ej, nj = tmp_path / "engine.json", tmp_path / "nccl.json"
    args = ["--mode", "bcast", "--size", "5000011", "--iters", "2", "--warmup", "1", "--timing", "dev", "--verify"]
    run([MPIRUN, "-n", str(n), P2P_MATRIX, *args, "--json", str(ej)], timeout=300)
    run([MPIRUN, "-n", str(n), os.path.join(REF_DIR, "p2p_matrix_nccl"), *args, "--json", str(nj)], timeout=600)

    def digests(path):
        recs = [json.loads(l) for l in open(path)]
        return {(r["dst"], r["iter"]): (r["digest"], r["mismatches"]) for r in recs if r["mode"] == "verify"}

    e, c = digests(ej), digests(nj)
    assert len(e) == 2 * n and e == c and all(bad == 0 for _, bad in e.values())
