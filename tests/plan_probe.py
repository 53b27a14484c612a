"""Helper for tests/test_plan.py: drives the engine in plan-only mode (csrc/p2p_debug.h) in a fresh
process and prints, as JSON, the kernel arguments each group would have been launched with.

    python tests/plan_probe.py '{"rank":0,"nranks":8,"windows":[[addr,bytes],...],
                                 "groups":[[["send"|"recv", ptr, nbytes, peer], ...], ...]}'
"""
import ctypes
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

MAX_OPS = 32
KINDS = {0: "none", 1: "post", 2: "push", 3: "send_ll", 4: "recv_ll", 5: "copy", 6: "recv_adapt", 7: "recv_staged"}
CTL_BASE, CTL_STRIDE = 0x100000000000, 0x1000000000


class OpDesc(ctypes.Structure):
    _fields_ = [(n, ctypes.c_uint64) for n in ("local", "nbytes", "seq", "f_wait", "f_signal", "word", "f_wait2",
                                               "f_signal2", "word2", "aux", "arrive")] + \
               [(n, ctypes.c_uint32) for n in ("kind", "first_block", "nblocks", "peer", "flags", "pad")] + [("stage", ctypes.c_uint64)]


class GroupHdr(ctypes.Structure):
    _fields_ = [(n, ctypes.c_uint32) for n in ("nops", "n_ctrl", "my_token", "tma_chunk", "tma_stages", "pdl")] + \
               [("timeout_ns", ctypes.c_uint64), ("async_err", ctypes.c_uint64), ("tokens", ctypes.c_uint64),
                ("prev_token", ctypes.c_uint32), ("pad", ctypes.c_uint32)]


class GroupArgs(ctypes.Structure):
    _fields_ = [("h", GroupHdr), ("ops", OpDesc * MAX_OPS)]


def owner(addr):
    """Which rank's control region a flag address points into (None: not a control-region address)."""
    if addr < CTL_BASE:
        return None
    r = (addr - CTL_BASE) // CTL_STRIDE
    return int(r) if r < 64 else None


def main():
    arg = sys.argv[1]
    spec = json.load(open(arg[1:])) if arg.startswith("@") else json.loads(arg)   # "@file": spec too long for argv
    p2p = importlib.import_module("test-nccl-p2p_b200")
    L = p2p.lib()
    L.p2p_debug_plan_init.argtypes = [ctypes.c_int] * 5
    L.p2p_debug_plan_window.argtypes = [ctypes.c_uint64, ctypes.c_size_t]
    L.p2p_debug_plan_last.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.POINTER(ctypes.c_uint32),
                                      ctypes.POINTER(ctypes.c_uint32), ctypes.POINTER(ctypes.c_int)]
    rc = L.p2p_debug_plan_init(spec["rank"], spec["nranks"], spec.get("sm_count", 148), spec.get("ctas_ldst", 2), spec.get("ctas_tma", 1))
    assert rc == 0, L.p2p_get_last_error()
    for name, value in spec.get("options", {}).items():
        p2p.set_option(name, value)
    for base, nbytes in spec.get("windows", []):
        assert L.p2p_debug_plan_window(base, nbytes) >= 0
    out = []
    for group in spec["groups"]:
        rec = {"ops": []}
        if not group:            # a rank that takes no part in this phase launches nothing
            rec["rc"] = 0
            out.append(rec)
            continue
        if group[0][0] == "alltoall":
            _, sptr, rptr, count = group[0]
            rc = L.p2p_alltoall(ctypes.c_void_p(sptr), ctypes.c_void_p(rptr), count, p2p.INT8, None)
        else:
            rc = L.p2p_group_start()
            for kind, ptr, nbytes, peer in group:
                fn = L.p2p_send if kind == "send" else L.p2p_recv
                rc = rc or fn(ctypes.c_void_p(ptr), nbytes, p2p.INT8, peer, None)
            end = L.p2p_group_end()
            rc = rc or end
        rec["rc"] = rc
        if rc != 0:
            rec["error"] = L.p2p_get_last_error().decode()
            out.append(rec)
            continue
        args, grid, block, flavour = GroupArgs(), ctypes.c_uint32(), ctypes.c_uint32(), ctypes.c_int()
        size = L.p2p_debug_plan_last(ctypes.byref(args), ctypes.sizeof(args), ctypes.byref(grid), ctypes.byref(block), ctypes.byref(flavour))
        assert size == ctypes.sizeof(args), (size, ctypes.sizeof(args))
        rec.update(grid=grid.value, block=block.value, flavour=flavour.value, n_ctrl=args.h.n_ctrl, pdl=args.h.pdl, nops=args.h.nops,
                   my_token=args.h.my_token, prev_token=args.h.prev_token)
        for i in range(args.h.nops):
            o = args.ops[i]
            rec["ops"].append({
                "kind": KINDS[o.kind], "local": o.local, "nbytes": o.nbytes, "seq": o.seq, "peer": o.peer,
                "first_block": o.first_block, "nblocks": o.nblocks, "flags": o.flags, "word": o.word, "word2": o.word2, "aux": o.aux, "arrive": o.arrive,
                "wait_owner": owner(o.f_wait), "signal_owner": owner(o.f_signal),
                "wait2_owner": owner(o.f_wait2), "signal2_owner": owner(o.f_signal2),
                "f_wait": o.f_wait, "f_signal": o.f_signal, "f_wait2": o.f_wait2, "f_signal2": o.f_signal2,
            })
        out.append(rec)
    p2p.finalize()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
