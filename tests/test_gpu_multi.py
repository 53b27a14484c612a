"""Cross-GPU parity (needs at least two B200s; skipped on a one-GPU box).

- `mpirun -n N ./p2p_matrix --verify`: every cell's receive buffer compared with the pattern,
  for the uni- and bi-directional loops of the reference (p2p_matrix.cc:154-172, :209-253),
  all-pairs, the ring and the size sweep, on each data path.
- differential test against the reference's library: the NCCL harness (oracle/_ref/
  p2p_matrix_nccl, same loops, payload filled and checked on the CPU by the oracle) must
  leave the same digests in its receive buffers as the engine does.
- the Python binding driven by one process per GPU with a torch.distributed bootstrap.
"""
import json
import os
import sys

import pytest

from conftest import MPIRUN, P2P_MATRIX, REF_DIR, ROOT, gpu_count, run

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(gpu_count() < 2, reason="needs two GPUs")]

NGPU = gpu_count()
RANKS = sorted({2, min(4, NGPU), min(8, NGPU)} - {0, 1, 3}) or [2]


def matrix(n, *args, env=None, timeout=600):
    return run([MPIRUN, "-n", str(n), P2P_MATRIX, *args], timeout=timeout, env=env).stdout


@pytest.mark.parametrize("path", ["auto", "ldst", "tma", "ce_kernels", "ce_memops", "ce_dma", "pdl3_early", "ring_push"])
@pytest.mark.parametrize("n", RANKS)
def test_reference_loops_verified(built, n, path):
    # "ce_*": opt in to the copy-engine path (off by default), flags handled by kernels / stream memory operations / 8-byte DMAs
    if path.startswith("ce_"):
        env = {"P2P_PATH": "auto", "P2P_CE_MIN_BYTES": "4194304", "P2P_CE_WAIT_US": "20000",
               "P2P_CE_FLAGS": {"ce_kernels": "0", "ce_memops": "1", "ce_dma": "2"}[path]}
    elif path == "pdl3_early":
        env = {"P2P_PDL": "3", "P2P_EARLY_POST": "1"}
    elif path == "ring_push":
        env = {"P2P_BIDIR_PUSH": "2"}
    else:
        env = {"P2P_PATH": path}
    out = matrix(n, "--size", "8M", "--iters", "4", "--verify", env=env)
    assert "verify: OK" in out and "VERIFY FAILED" not in out
    assert "Evaluating the Uni-Directional NCCL P2P Bandwidth (Gbps)" in out
    assert "Evaluating the Bi-Directional NCCL P2P Bandwidth (Gbps)" in out


@pytest.mark.parametrize("size", ["8", "1000", "16385", "1M", "1048577", "3M", "4M", "67108864"])
def test_pair_sizes_verified(built, size):
    out = matrix(2, "--size", size, "--iters", "3", "--verify", "--timing", "dev", "--warmup", "2")
    assert "verify: OK" in out


def test_large_message_with_unregistered_send_buffer(built):
    """Only the receive buffer is a window: the sender of a large message announces a push."""
    out = matrix(2, "--size", "8M", "--iters", "3", "--verify", env={"P2P_NO_SEND_WINDOW": "1"})
    assert "verify: OK" in out and "VERIFY FAILED" not in out


@pytest.mark.parametrize("env", [{}, {"P2P_BIDIR_PUSH": "2"}, {"P2P_BIDIR_PUSH": "3"}, {"P2P_PDL": "3", "P2P_EARLY_POST": "1"},
                                 {"P2P_CE_MIN_BYTES": "4194304", "P2P_CE_WAIT_US": "20000"}],
                         ids=["default", "ring_push", "always_push", "pdl3_early", "copy_engine"])
def test_allpairs_and_ring_verified(built, env):
    n = RANKS[-1]
    for mode in ("allpairs", "ring"):
        for size in ("4096", "16M"):
            out = matrix(n, "--mode", mode, "--size", size, "--iters", "3", "--warmup", "1", "--timing", "dev", "--verify", env=env)
            assert "verify: OK" in out, out


@pytest.mark.parametrize("n", RANKS)
def test_alltoall_verified(built, n):
    """p2p_alltoall: every block on every rank compared with the pattern of the rank it came from,
    in each protocol range (eager, push, adaptive)."""
    for size in ("1000", "600K", "2M", "8M"):
        out = matrix(n, "--mode", "alltoall", "--size", size, "--iters", "3", "--warmup", "1", "--timing", "dev", "--verify")
        assert "verify: OK" in out and "VERIFY FAILED" not in out, out


def test_sweep_verified(built):
    out = matrix(2, "--mode", "sweep", "--sweep-max", "16M", "--iters", "20", "--verify")
    assert "verify: OK" in out
    rows = [l.split() for l in out.splitlines() if l.strip() and l.split()[0].isdigit()]
    assert [int(r[0]) for r in rows] == [8 << i for i in range(22)]


def test_digests_equal_nccl(built, tmp_path):
    """Same seeds, same sizes, same loops: the engine and the reference's NCCL path must leave
    identical bytes (compared as digests, NCCL's computed on the CPU by the oracle)."""
    ej, nj = tmp_path / "engine.json", tmp_path / "nccl.json"
    args = ["--size", "5000011", "--iters", "1", "--verify"]
    matrix(2, *args, "--json", str(ej))
    out = run([MPIRUN, "-n", "2", os.path.join(REF_DIR, "p2p_matrix_nccl"), *args, "--json", str(nj)], timeout=600).stdout
    assert "verify: OK" in out

    def digests(path):
        recs = [json.loads(l) for l in open(path)]
        return {(r["dir"], r["src"], r["dst"], r["bytes"]): (r["digest"], r["mismatches"]) for r in recs if r["mode"] == "verify"}

    e, n = digests(ej), digests(nj)
    assert e and e == n
    assert all(bad == 0 for _, bad in e.values())


def test_python_binding_one_process_per_gpu(built, tmp_path):
    """torch.distributed (gloo) carries the bootstrap blobs; ring transfer across two GPUs checked
    against the oracle in each process."""
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    out = run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
               "--master-port", "29611", str(script)], timeout=600).stdout
    assert out.count("rank ok") == 2, out


WORKER = r'''
import importlib, os, sys
sys.path.insert(0, %(root)r)
sys.path.insert(0, os.path.join(%(root)r, "oracle"))
import numpy as np
import torch
import torch.distributed as dist
import pattern as oracle
p2p = importlib.import_module("test-nccl-p2p_b200")
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("gloo", rank=rank, world_size=world)
gather = p2p.torch_allgather()
p2p.init(rank, world, rank, gather)
n = (3 << 20) + 17
src, dst = p2p.DeviceBuffer(n), p2p.DeviceBuffer(n + 64)
p2p.register(dst.ptr, dst.nbytes, gather)
s = torch.cuda.Stream()
right, left = (rank + 1) %% world, (rank - 1) %% world
for it, size in enumerate([8, 4097, n]):
    p2p.pattern_fill(src.ptr, size, 42, rank, it, s)
    dst.tensor().fill_(0xA5)
    torch.cuda.synchronize()
    dist.barrier()
    with p2p.group():
        p2p.recv(dst.ptr, size, p2p.INT8, left, s)
        p2p.send(src.ptr, size, p2p.INT8, right, s)
    s.synchronize()
    p2p.raise_async_error()
    got = dst.tensor()[: size + 64].cpu().numpy()
    assert np.array_equal(got[:size], oracle.expected_recv(size, 42, left, it)), (rank, size)
    assert np.all(got[size:] == 0xA5)
    dist.barrier()
# all-to-all through the binding, against the oracle's law
blk = (1 << 20) + 5
a2a_src, a2a_dst = p2p.DeviceBuffer(blk * world), p2p.DeviceBuffer(blk * world + 64)
p2p.register(a2a_dst.ptr, a2a_dst.nbytes, gather)
p2p.register(a2a_src.ptr, a2a_src.nbytes, gather)
a2a_src.tensor().copy_(torch.from_numpy(oracle.alltoall_send(blk, 7, rank, world)))
a2a_dst.tensor().fill_(0xA5)
torch.cuda.synchronize()
dist.barrier()
p2p.alltoall(a2a_src.ptr, a2a_dst.ptr, blk, p2p.INT8, s)
s.synchronize()
p2p.raise_async_error()
got = a2a_dst.tensor().cpu().numpy()
assert np.array_equal(got[: blk * world], oracle.expected_alltoall_recv(blk, 7, rank, world)), rank
assert np.all(got[blk * world:] == 0xA5)
torch.cuda.synchronize()
p2p.finalize(dist.barrier)
print("rank ok", rank, flush=True)
dist.destroy_process_group()
'''
