#!/usr/bin/env python
"""bench.py — the measurement contract for the transfer engine.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--size BYTES]

A STEP is one pass of the hot path: every GPU sends one message of `size` bytes (default the
256 MiB of BASELINE.json's configs[1]) to the next GPU and receives one from the previous,
as one grouped p2p_send/p2p_recv pair per rank (the reference's call shape,
p2p_matrix.cc:212-228).  With N = 1 the "next GPU" is the GPU itself: a local HBM copy through
the same kernel, flags and C ABI.  N > 1 is launched by the driver as
`python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...`; ranks are
independent (no data-path collective), so scaling is weak.

  value     whole-job payload GB/s, buffers resident in HBM, CUDA events on the launch stream,
            barrier + synchronize either side, max over ranks
  e2e       the same step with the payload starting and ending in pinned HOST memory:
            H2D copy, transfer, D2H copy inside the timed region, through the public C ABI
  roofline  the step IS one kernel (p2p_group_kernel_*): algorithmic bytes per launch divided by
            the event-timed launch duration; N = 1 against measured HBM copy bandwidth
            (read + write bytes), N > 1 against NVLink 5's 900 GB/s per direction
  cpu_baseline   the MPI host-memory send/recv path (oracle/_ref/mpi_host_p2p) on 2 host cores
  --impl reference   the reference's own path — libnccl's ncclSend/ncclRecv driven by the
            reference's loop (oracle/_ref/p2p_matrix_nccl) — on the same step, same timing
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
MPIRUN = os.path.join(ROOT, "build", "mpirun")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
LINK_GBPS = 900.0          # NVLink 5 per direction per GPU (BASELINE.json)
LINK_MEASURED_GBPS = 770.0  # peer-copy figure in /opt/skills/guides/B200_PROFILING.md
METRIC = "P2P payload bandwidth, ring step (each GPU sends one message to the next), whole job"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=400)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--size", type=int, default=256 << 20)
    ap.add_argument("--no-parity", action="store_true", help="skip the engine-vs-NCCL digest comparison at N >= 2")
    ap.add_argument("--no-extras", action="store_true", help="skip the side measurements (latency, all-pairs, NCCL in the same run, CPU baseline); used for profiling runs")
    return ap.parse_args()


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


# ------------------------------------------------------------------ clocks (NVML)

REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
           0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
           0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}


class Clocks:
    """Samples SM clock and throttle reasons of one GPU while a timed region runs."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz, self.ok = [], set(), None, False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = index
            if vis:
                try:
                    phys = int(vis.split(",")[index])
                except Exception:
                    phys = index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:  # no NVML: report that instead of inventing numbers
            self.err = repr(e)

    def sample(self):
        if not self.ok:
            return
        try:
            self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
            fn = getattr(self.nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or self.nv.nvmlDeviceGetCurrentClocksThrottleReasons
            mask = fn(self.h)
            for bit, name in REASONS.items():
                if mask & bit and name != "gpu_idle":
                    self.reasons.add(name)
        except Exception:
            pass

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0,
                    "note": "NVML unavailable" if not self.ok else "no sample landed inside the timed region"}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(s)}


# ------------------------------------------------------------------ baselines run as subprocesses

def run_json(cmd, timeout):
    """Run a benchmark binary that writes JSON lines to --json FILE; return the records."""
    with tempfile.NamedTemporaryFile("r", suffix=".json", delete=False) as tf:
        path = tf.name
    try:
        proc = subprocess.run(cmd + ["--json", path], cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
        if proc.returncode != 0:
            return None, proc.stdout[-1500:]
        with open(path) as f:
            return [json.loads(l) for l in f if l.strip()], proc.stdout
    except Exception as e:
        return None, repr(e)
    finally:
        try:
            os.unlink(path)
        except OSError:
            pass


def cpu_baseline(size: int):
    """MPI host-memory send/recv of the same message between two ranks on the host cores.
    Bounded sample: enough messages for roughly ten seconds."""
    exe = os.path.join(REF_DIR, "mpi_host_p2p")
    if not (os.path.exists(exe) and os.path.exists(MPIRUN)):
        return {"value": None, "unit": "GB/s", "cores": 2, "kind": "port", "sample": "oracle/_ref/mpi_host_p2p not built"}
    probe, _ = run_json([MPIRUN, "-n", "2", exe, "--size", str(size), "--iters", "3", "--mode", "uni", "--first-pair"], 300)
    if not probe:
        return {"value": None, "unit": "GB/s", "cores": 2, "kind": "port", "sample": "mpi_host_p2p failed"}
    per_msg = probe[0]["seconds"]
    iters = max(5, min(400, int(10.0 / max(per_msg, 1e-6))))
    recs, _ = run_json([MPIRUN, "-n", "2", exe, "--size", str(size), "--iters", str(iters), "--warmup", "2", "--mode", "uni", "--first-pair"], 600)
    if not recs:
        return {"value": None, "unit": "GB/s", "cores": 2, "kind": "port", "sample": "mpi_host_p2p failed"}
    r = recs[0]
    return {"value": round(r["GBps"], 3), "unit": "GB/s", "cores": 2, "kind": "port",
            "host_cores": r.get("host_cores"), "transport": r.get("transport"),
            "sample": f"{iters} messages of {size} bytes rank 0 -> rank 1 through host memory (MPI_Send/MPI_Recv, "
                      f"reference loop shape p2p_matrix.cc:141-186), {r['us']:.0f} us each"}


def nccl_run(n: int, mode: str, size: int, iters: int, warmup: int, extra=(), timeout=900):
    exe = os.path.join(REF_DIR, "p2p_matrix_nccl")
    if not (os.path.exists(exe) and os.path.exists(MPIRUN)):
        return None, "oracle/_ref/p2p_matrix_nccl not built"
    return run_json([MPIRUN, "-n", str(n), exe, "--mode", mode, "--size", str(size), "--iters", str(iters), "--warmup", str(warmup),
                     "--timing", "dev", *extra], timeout)


def matrix_run(exe, n: int, args, timeout=900):
    """One run of ./p2p_matrix or of the NCCL harness (same driver, same flags) under the MPI launcher."""
    if not (os.path.exists(exe) and os.path.exists(MPIRUN)):
        return None, f"{exe} or {MPIRUN} not built"
    return run_json([MPIRUN, "-n", str(n), exe, *args], timeout)


def nccl_parity(n: int, size: int):
    """Cross-GPU parity against the reference's library, on this box, in this run: the engine (./p2p_matrix) and
    the reference's ncclSend/ncclRecv loop (oracle/_ref/p2p_matrix_nccl) move the same patterned payloads — the
    ring step at the bench's message size, then every cell of the reference's uni- and bi-directional matrices
    (p2p_matrix.cc:141-186, :196-267) at an odd size — and the digests of what every receiver holds must agree.
    NCCL's buffers are copied to the host and digested by the CPU oracle (oracle/pattern.c); the engine's are
    digested on the device by a kernel that the tests pin to that same oracle."""
    engine_exe, nccl_exe = os.path.join(ROOT, "p2p_matrix"), os.path.join(REF_DIR, "p2p_matrix_nccl")
    legs = [("ring", ["--mode", "ring", "--size", str(size), "--iters", "1", "--warmup", "0", "--timing", "dev", "--verify", "--seed", "0x5EED0002"]),
            ("matrices", ["--mode", "ref", "--size", "5000011", "--iters", "1", "--verify", "--seed", "0x5EED0003"])]
    out = {"nccl_digests_equal": True, "cells": 0, "mismatching_bytes": 0, "what": nccl_parity.__doc__.split(":")[0].strip()}
    for name, args in legs:
        sets = {}
        for who, exe in (("engine", engine_exe), ("nccl", nccl_exe)):
            recs, log = matrix_run(exe, n, args)
            if recs is None:
                return {"nccl_digests_equal": None, "unavailable": f"{who} {name}: {str(log)[-300:]}"}
            sets[who] = {(r["dir"], r["src"], r["dst"], r["bytes"]): (r["digest"], r["mismatches"]) for r in recs if r["mode"] == "verify"}
        e, c = sets["engine"], sets["nccl"]
        equal = bool(e) and e == c
        out[name] = {"cells": len(e), "equal": equal}
        out["cells"] += len(e)
        out["mismatching_bytes"] += sum(bad for _, bad in e.values()) + sum(bad for _, bad in c.values())
        if not equal:
            out["nccl_digests_equal"] = False
            diff = [k for k in set(e) | set(c) if e.get(k) != c.get(k)]
            out[name]["first_difference"] = [list(diff[0]), e.get(diff[0]), c.get(diff[0])] if diff else "one side is empty"
    if out["mismatching_bytes"]:
        out["nccl_digests_equal"] = False
    return out


def nccl_side_numbers(n: int, allpairs_bytes: int):
    """The NCCL figures the north star compares with, taken in the same run: all-pairs aggregate, 8-byte latency."""
    out = {}
    exe = os.path.join(REF_DIR, "p2p_matrix_nccl")
    recs, log = matrix_run(exe, n, ["--mode", "allpairs", "--size", str(allpairs_bytes), "--iters", "30", "--warmup", "3", "--timing", "dev"])
    ap = [r for r in (recs or []) if r["mode"] == "allpairs"]
    out["nccl_allpairs_aggregate_GBps"] = round(ap[0]["GBps"], 2) if ap else None
    recs, log = matrix_run(exe, n, ["--mode", "bcast", "--size", str(allpairs_bytes), "--iters", "30", "--warmup", "3", "--timing", "dev"])
    bc = [r for r in (recs or []) if r["mode"] == "bcast"]
    out["nccl_bcast_grouped_sends_64MiB_GBps_per_receiver"] = round(bc[0]["GBps"] / max(n - 1, 1), 2) if bc else None
    recs, log = matrix_run(exe, n, ["--mode", "latency", "--sweep-min", "8", "--sweep-max", "8", "--iters", "2000"])
    lat = [r for r in (recs or []) if r["mode"] == "latency" and r["bytes"] == 8]
    out["nccl_latency_8B_oneway_us"] = round(lat[0]["us"], 3) if lat else None
    return out


# ------------------------------------------------------------------ the reference arm

def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # rank 0 alone drives all N GPUs through the MPI launcher
    n = args.gpus
    steps, warm = args.steps, max(args.warmup, 3)
    recs, log = nccl_run(n, "ring", args.size, steps, warm)
    if not recs:
        print(json.dumps({"impl": "reference", "unavailable": f"NCCL harness did not run: {str(log)[-300:]}"}))
        return 0
    ring = [r for r in recs if r["mode"] == "ring"][0]
    e2e_steps = max(3, min(steps, 40))
    erecs, _ = nccl_run(n, "ring", args.size, e2e_steps, 3, extra=["--e2e"])
    e2e = [r for r in (erecs or []) if r["mode"] == "ring_e2e"]
    base = cpu_baseline(args.size) if n == 1 else None
    line = {
        "impl": "reference",
        "metric": METRIC, "value": round(ring["GBps"], 3), "unit": "GB/s", "n_gpus": n, "steps": steps, "warmup": warm,
        "ms_per_step": round(ring["seconds"] * 1e3, 6), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": f"ring step, {args.size} bytes per GPU (BASELINE configs[1] message size)", "bytes": args.size,
                   "what": "reference's own path: ncclSend/ncclRecv from libnccl 2.27.3 in the reference's loop "
                           "(oracle/_ref/p2p_matrix_nccl), launched by mpirun on the same GPUs; plain cudaMalloc buffers, "
                           "not registered with NCCL, default NCCL environment - as the reference runs it",
                   "l2": "message larger than L2 (126 MB)" if args.size > (126 << 20) else "message fits L2"},
        "e2e": {"value": round(e2e[0]["GBps"], 3) if e2e else None, "unit": "GB/s",
                "h2d_bytes_per_step": args.size * n if e2e else 0, "d2h_bytes_per_step": args.size * n if e2e else 0,
                "steps": e2e_steps, "note": "same H2D + transfer + D2H step as the engine arm, NCCL doing the transfer"},
        "gpu_launches": 0,
    }
    if base is not None:
        line["cpu_baseline"] = base
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------ the engine arm

def engine_arm(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            # convenience: relaunch ourselves the way the driver would
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.abspath(__file__)] + sys.argv[1:]
            return subprocess.call(cmd)
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU: the engine has no CPU path")
    torch.cuda.set_device(local)
    p2p = importlib.import_module("test-nccl-p2p_b200")
    host_group = None
    if world > 1:
        # torch.distributed is plumbing only (bootstrap blobs, barriers, max over ranks): gloo is enough,
        # the payload never goes through it
        dist.init_process_group("gloo")
        host_group = dist.group.WORLD
    gather = p2p.torch_allgather(host_group) if world > 1 else (lambda b: [b])

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=host_group)

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=host_group)
        return float(t.item())

    B = args.size
    K, W = args.steps, max(args.warmup, 3)
    right, left = (rank + 1) % world, (rank - 1) % world
    p2p.init(rank, world, local, gather)
    src, dst = p2p.DeviceBuffer(B), p2p.DeviceBuffer(B + 256)
    p2p.register(dst.ptr, dst.nbytes, gather)
    p2p.register(src.ptr, src.nbytes, gather)   # large messages may then be pulled by the receiver
    stream = torch.cuda.Stream()
    clocks = Clocks(local)

    def step():
        with p2p.group():
            p2p.recv(dst.ptr, B, p2p.INT8, left, stream)
            p2p.send(src.ptr, B, p2p.INT8, right, stream)

    def verify(it):
        p2p.pattern_fill(src.ptr, B, 0x5EED0001, rank, it, stream)
        dst.tensor().fill_(0xA5)
        barrier()
        step()
        bad, first, dig = p2p.pattern_check(dst.ptr, B, 0x5EED0001, left, it, stream)
        p2p.raise_async_error()
        if bad:
            raise SystemExit(f"rank {rank}: transfer is wrong: {bad} bytes differ, first at {first}")
        return dig

    def timed(fn, steps):
        """barrier+sync, `steps` calls enqueued back to back, events on the launch stream, clocks sampled
        while the GPU works, barrier+sync; returns seconds for all steps (max over ranks)."""
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        while not e1.query():
            clocks.sample()
            time.sleep(0.002)
        stream.synchronize()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1) * 1e-3)

    digest_before = verify(1)
    for _ in range(W):
        step()
    stream.synchronize()
    launches0, ce0 = p2p.get_option("stat_launches"), p2p.get_option("stat_ce_copies")
    total = timed(step, K)
    launches = p2p.get_option("stat_launches") - launches0      # kernels the engine launched in the timed region
    ce_copies = p2p.get_option("stat_ce_copies") - ce0          # payloads it handed to the copy engine instead
    last = {k: p2p.get_option("stat_last_" + k) for k in ("flavour", "grid", "block", "push_ops", "push_ctas", "offer_ops", "recv_ops",
                                                          "recv_ctas", "copy_ctas", "ce_ops", "pdl")}   # the group the timed loop launched
    p2p.raise_async_error()
    digest_after = verify(2)
    per_step = total / K
    value = world * B / per_step / 1e9

    # ---- e2e: payload starts and ends in pinned host memory, copies inside the timed region
    hsrc = torch.empty(B, dtype=torch.uint8).pin_memory()
    hdst = torch.empty(B, dtype=torch.uint8).pin_memory()
    hsrc.fill_(0x3C)
    src_t, dst_t = src.tensor(), dst.tensor()[:B]

    def step_e2e():
        with torch.cuda.stream(stream):
            src_t.copy_(hsrc, non_blocking=True)
        step()
        with torch.cuda.stream(stream):
            hdst.copy_(dst_t, non_blocking=True)

    e2e_steps = max(3, min(K, 40))
    for _ in range(3):
        step_e2e()
    stream.synchronize()
    e2e_total = timed(step_e2e, e2e_steps)
    e2e_ok = bool(hdst[:4096].eq(0x3C).all().item() and hdst[-4096:].eq(0x3C).all().item())
    e2e_value = world * B / (e2e_total / e2e_steps) / 1e9

    # ---- side measurements the north star asks to sit next to the number (not part of `value`)
    extras = {}
    if not args.no_extras:
        extras = side_measurements(p2p, torch, rank, world, stream, src, dst, B, barrier, max_over_ranks, gather)

    peaks, peak_kind = measured_peaks()
    if world == 1:
        roof = {"bound": "hbm", "achieved": round(2 * B / per_step / 1e9, 2), "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "peak_kind": f"{peak_kind} HBM copy (read+write bytes)", "traffic": TRAFFIC_PER_LAUNCH.get("local") if B == (256 << 20) else None}
    else:
        roof = {"bound": "nvlink", "achieved": round(B / per_step / 1e9, 2), "peak": LINK_GBPS, "unit": "GB/s",
                "peak_kind": f"nominal NVLink 5 per direction per GPU; measured peer copy on this pool {LINK_MEASURED_GBPS}",
                "traffic": None}
    roof["frac"] = round(roof["achieved"] / roof["peak"], 4)
    if world > 1 and B == (256 << 20) and not ce_copies:
        mover = "peer_push_ldst" if last["push_ops"] else ("peer_pull_tma" if last["flavour"] == 2 else "peer_pull_ldst")
        roof["traffic"] = TRAFFIC_PER_LAUNCH[mover]
        roof["traffic_from"] = ("profiles/r02_ncu/" + {"peer_push_ldst": "traffic_push_256M.csv", "peer_pull_ldst": "traffic_pull_256M.csv",
                                                       "peer_pull_tma": "traffic_tmapull_256M.csv"}[mover]
                                + " (ncu, single-process two-GPU launch of this mover)")
    if world > 1:
        roof["frac_measured"] = round(roof["achieved"] / LINK_MEASURED_GBPS, 4)
        roof["peak_measured"] = LINK_MEASURED_GBPS
    roof["kernel"] = "p2p_group_kernel_tma" if last["flavour"] == 2 else "p2p_group_kernel_ldst"
    if world == 1:
        how = f"local copy, {'TMA ring' if last['flavour'] == 2 else 'ld/st'}, {last['copy_ctas']} CTAs"
    else:
        # this rank's view of the group it launched: how its send leaves, how its receive is served
        send = (f"send pushed by {last['push_ctas']} CTAs (ld/st stores into the peer's buffer)" if last["push_ops"] else
                "send handed to the copy engine" if last["ce_ops"] else
                "send offered for pulling (the receiver's CTAs load it)" if last["offer_ops"] else "no send")
        recv = f"receive served by {last['recv_ctas']} CTAs (pull, or post-and-wait when the sender pushes)" if last["recv_ops"] else "no receive"
        how = f"{send}; {recv}"
    if ce_copies:
        roof["kernel"] = (f"copy engine (cudaMemcpyAsync to the peer mapping) for {ce_copies} of {K} payloads of this rank; "
                          "receives in p2p_group_kernel_ldst")
    roof["algorithmic_bytes_per_launch"] = (2 if world == 1 else 1) * B
    roof["ctas"] = last["grid"]
    roof["threads_per_cta"] = last["block"]

    torch.cuda.synchronize()
    p2p.finalize((lambda: dist.barrier(group=host_group)) if world > 1 else None)

    parity_failed = False
    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 3), "unit": "GB/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": round(per_step * 1e3, 6), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u8", "data": "synthetic",
            "config": {"workload": f"ring step, {B} bytes per GPU (BASELINE configs[1] message size)"
                                   + ("; N=1: local HBM copy through the same send/recv kernel" if world == 1 else ""),
                       "bytes": B, "l2": "message larger than L2 (126 MB), no flush needed" if B > (126 << 20) else "message fits L2",
                       "path": how, "ctas": last["grid"], "pdl": last["pdl"]},
            "clocks": clocks.summary(),
            "e2e": {"value": round(e2e_value, 3), "unit": "GB/s", "h2d_bytes_per_step": B * world, "d2h_bytes_per_step": B * world,
                    "steps": e2e_steps, "payload_intact": e2e_ok},
            "gpu_launches": launches * world,
            "roofline": roof,
            "parity": {"digest_before": f"{digest_before:#018x}", "digest_after": f"{digest_after:#018x}", "mismatches": 0},
        }
        if world == 1 and not args.no_extras:
            line["cpu_baseline"] = cpu_baseline(B)
        if extras:
            line["extras"] = extras
        if world > 1 and not args.no_parity:
            line["parity"].update(nccl_parity(world, B))
            parity_failed = line["parity"].get("nccl_digests_equal") is False
        if world > 1 and not args.no_extras:
            line["extras"].update(nccl_side_numbers(world, extras.get("allpairs_bytes_per_flow", 64 << 20)))
        if not args.no_extras:
            recs, log = nccl_run(world, "ring", B, min(K, 200), 5)
            ring = [r for r in (recs or []) if r["mode"] == "ring"]
            line["nccl_same_run"] = ({"value": round(ring[0]["GBps"], 3), "unit": "GB/s", "ms_per_step": round(ring[0]["seconds"] * 1e3, 6),
                                      "what": "ncclSend/ncclRecv ring step, libnccl 2.27.3, unregistered buffers (the reference's way), device timing"} if ring
                                     else {"unavailable": str(log)[-200:]})
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(group=host_group)
        dist.destroy_process_group()
    if parity_failed:
        sys.stderr.write("bench.py: the engine's received bytes differ from NCCL's for the same payloads (see the line's parity key)\n")
        return 3
    return 0


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the timed kernel's mover, from committed ncu captures at the
# bench's 256 MiB message size.
#   local:      profiles/r01_ncu/prof_local_tma.raw.csv, this same command at N=1: 268.49 MB read + 212.76 MB written (the
#               rest of the 268 MB written still sits in L2 when the kernel ends).
#   peer_*:     profiles/r02_ncu/traffic_*_256M.csv: the engine's group kernels launched with peer pointers by ONE process
#               that drives two GPUs (scripts/prof_peer.cu; ncu must not wrap a multi-rank command).  A push reads its
#               payload from local DRAM (268.5 MB + 3.2 MB written); a pull writes it to local DRAM (208-209 MB, the rest
#               in L2 at kernel end) and reads nothing locally.  In a ring step the same GPU's DRAM also serves the
#               neighbour's pull (another 268 MB read), which belongs to the neighbour's launch.
TRAFFIC_PER_LAUNCH = {"local": 481247232, "peer_push_ldst": 271659008, "peer_pull_ldst": 208919296, "peer_pull_tma": 207791616}


def side_measurements(p2p, torch, rank, world, stream, src, dst, B, barrier, max_over_ranks, gather):
    """8-byte one-way latency, single-pair uni/bidir bandwidth and all-pairs aggregate (N >= 2)."""
    out = {}
    if world < 2:
        return out

    def events_around(fn, iters, active=True):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        sec = 0.0
        if active:
            e0.record(stream)
            for _ in range(iters):
                fn()
            e1.record(stream)
            stream.synchronize()
            sec = e0.elapsed_time(e1) * 1e-3
        barrier()
        return max_over_ranks(sec)

    pair = rank in (0, 1)
    peer = 1 - rank if pair else 0

    def pingpong(n):
        if rank == 0:
            p2p.send(src.ptr, n, p2p.INT8, 1, stream)
            p2p.recv(dst.ptr, n, p2p.INT8, 1, stream)
        elif rank == 1:
            p2p.recv(dst.ptr, n, p2p.INT8, 0, stream)
            p2p.send(src.ptr, n, p2p.INT8, 0, stream)

    events_around(lambda: pingpong(8), 50, pair)
    t = events_around(lambda: pingpong(8), 2000, pair)
    out["latency_8B_oneway_us"] = round(t / 2000 / 2 * 1e6, 3)

    def uni(n):
        if rank == 0:
            p2p.send(src.ptr, n, p2p.INT8, 1, stream)
        elif rank == 1:
            p2p.recv(dst.ptr, n, p2p.INT8, 0, stream)

    def bidir(n):
        with p2p.group():
            p2p.send(src.ptr, n, p2p.INT8, peer, stream)
            p2p.recv(dst.ptr, n, p2p.INT8, peer, stream)

    for name, n in (("64MiB", 64 << 20), ("256MiB", 256 << 20)):
        if n > B:
            continue
        events_around(lambda: uni(n), 5, pair)
        t = events_around(lambda: uni(n), 50, pair)
        out[f"uni_0to1_{name}_GBps"] = round(n / (t / 50) / 1e9, 2)
        events_around(lambda: bidir(n), 5, pair)
        t = events_around(lambda: bidir(n), 50, pair)
        out[f"bidir_0and1_{name}_GBps"] = round(2 * n / (t / 50) / 1e9, 2)

    # all-pairs: every rank sends 64 MiB to each of the other ranks at once (BASELINE configs[3])
    n = min(64 << 20, B)
    slots = p2p.DeviceBuffer((world - 1) * n)      # one receive slot per peer, as the reference's buffers would need
    p2p.register(slots.ptr, slots.nbytes, gather)

    def allpairs():
        with p2p.group():
            slot = 0
            for p in range(world):
                if p == rank:
                    continue
                p2p.recv(slots.ptr + slot * n, n, p2p.INT8, p, stream)
                slot += 1
            for p in range(world):
                if p != rank:
                    p2p.send(src.ptr, n, p2p.INT8, p, stream)
    events_around(allpairs, 3)
    t = events_around(allpairs, 30)
    out["allpairs_bytes_per_flow"] = n
    out["allpairs_aggregate_GBps"] = round(world * (world - 1) * n / (t / 30) / 1e9, 2)
    p2p.raise_async_error()

    # one-to-many: rank 0's 64 MiB into every rank's buffer through the NVSwitch multicast mapping
    try:
        sym = p2p.SymmBuffer(n + 256, gather)
        def bcast():
            p2p.bcast(src.ptr, sym.ptr, n, p2p.INT8, 0, stream)
        events_around(bcast, 3)
        t = events_around(bcast, 30)
        out["bcast_64MiB_GBps_per_receiver"] = round(n / (t / 30) / 1e9, 2)
        out["bcast_multicast"] = bool(sym.multicast_ptr)
        p2p.raise_async_error()
    except p2p.P2PError as e:
        out["bcast_unavailable"] = str(e)[:200]
    return out


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    return engine_arm(args)


if __name__ == "__main__":
    sys.exit(main())
