"""Host-side logic of the N > 1 path with world_size 2 on CPU (gloo): the bootstrap exchange the
Python binding performs (rank-ordered all-gather of fixed-size blobs), ring neighbours, and the
max-over-ranks reduction bench.py uses.  No GPU, no engine call."""
import os
import sys

import pytest

from conftest import ROOT, run

WORKER = r'''
import importlib, os, sys
sys.path.insert(0, %(root)r)
import torch
import torch.distributed as dist
p2p = importlib.import_module("test-nccl-p2p_b200")
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo", rank=rank, world_size=world)
gather = p2p.torch_allgather()
L = p2p.lib()
n = L.p2p_blob_bytes()
mine = bytes([rank + 1]) * n                       # stands in for p2p_export's blob
blobs = gather(mine)
assert len(blobs) == world and all(len(b) == n for b in blobs)
assert [b[0] for b in blobs] == [r + 1 for r in range(world)], "blobs must come back in rank order"
joined = b"".join(blobs)
assert len(joined) == n * world and joined[rank * n] == rank + 1
# ring neighbours used by bench.py's step
right, left = (rank + 1) %% world, (rank - 1) %% world
peers = gather(bytes([right, left]))
assert all(peers[peers[r][0]][1] == r for r in range(world)), "my right neighbour's left neighbour is me"
# max over ranks of a per-rank time
t = torch.tensor([0.5 + rank], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
assert t.item() == 0.5 + world - 1
# the engine itself refuses to start without a GPU, identically on every rank
try:
    p2p.init(rank, world, 0, gather)
    started = True
except p2p.P2PError as e:
    started = False
    assert e.code == p2p.UNHANDLED_CUDA_ERROR
flags = gather(bytes([int(started)]))
assert len(set(flags)) == 1
print("rank ok", rank, flush=True)
dist.destroy_process_group()
'''


def test_world2_bootstrap_plumbing(built, tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER % {"root": ROOT})
    out = run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
               "--master-port", "29612", str(script)], timeout=300, env={"CUDA_VISIBLE_DEVICES": ""}).stdout
    assert out.count("rank ok") == 2, out
