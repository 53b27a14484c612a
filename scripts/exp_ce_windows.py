"""Experiment (not part of the product): does the copy engine reach its two-way rate (1519 GB/s between two B200s in
one process, scripts/exp_peer_bw.cu) when the destination is another PROCESS's memory - mapped through cudaIpc
(p2p_register_*) or through the driver's virtual memory API (p2p_symm_alloc_*)?  Round 2 found the engine's
copy-engine chain at 1040 GB/s both ways against 1342 for two SM pushes; this separates the mapping from the chain.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/exp_ce_windows.py
"""
import ctypes
import importlib
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
p2p = importlib.import_module("test-nccl-p2p_b200")

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("gloo")
gather = p2p.torch_allgather()
p2p.init(rank, world, rank, gather)
cudart = ctypes.CDLL("libcudart.so.12")
cudart.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
peer = 1 - rank
stream = torch.cuda.Stream()


def timed(dst_ptr, src_ptr, n, iters, active):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    dist.barrier()
    ms = 0.0
    if active:
        for _ in range(3):
            assert cudart.cudaMemcpyAsync(dst_ptr, src_ptr, n, 4, stream.cuda_stream) == 0
        stream.synchronize()
    dist.barrier()
    if active:
        e0.record(stream)
        for _ in range(iters):
            assert cudart.cudaMemcpyAsync(dst_ptr, src_ptr, n, 4, stream.cuda_stream) == 0
        e1.record(stream)
        stream.synchronize()
        ms = e0.elapsed_time(e1)
    t = torch.tensor([ms], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e-3 / iters


for n in (64 << 20, 256 << 20):
    src = p2p.DeviceBuffer(n)
    ipc_dst = p2p.DeviceBuffer(n)
    ipc_win = p2p.register(ipc_dst.ptr, n, gather)
    vmm_dst = p2p.SymmBuffer(n, gather)
    targets = {"cudaIpc mapping": p2p.window_peer_ptr(ipc_win, peer), "VMM mapping": p2p.window_peer_ptr(vmm_dst.window, peer)}
    for name, dst in targets.items():
        one = timed(dst, src.ptr, n, 20, rank == 0)
        both = timed(dst, src.ptr, n, 20, True)
        if rank == 0:
            print(f"{n >> 20:4d} MiB  DMA into the peer's memory through a {name:16s}: one way {n / one / 1e9:7.1f} GB/s   "
                  f"both ways at once {2 * n / both / 1e9:7.1f} GB/s total", flush=True)
    torch.cuda.synchronize()
    dist.barrier()
    vmm_dst.free()
    p2p.deregister(ipc_win)
    src.free()
    ipc_dst.free()
p2p.finalize(dist.barrier)
dist.destroy_process_group()
