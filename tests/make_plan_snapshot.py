"""Writes tests/golden/plan_snapshot.json: the complete kernel arguments (every field of every op)
the engine's group builder produces for a fixed set of scenarios, in plan-only mode.
tests/test_plan.py::test_plan_snapshot compares the current build with it, so that any change of
the host logic shows up as a diff that has to be looked at.

    python tests/make_plan_snapshot.py        # regenerate after an INTENDED change of the planner
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROBE = os.path.join(ROOT, "tests", "plan_probe.py")
SEND_WIN, RECV_WIN, LOOSE = 1 << 32, 1 << 33, 1 << 36
WINDOWS = [[SEND_WIN, 1 << 30], [RECV_WIN, 4 << 30]]
MiB = 1 << 20


def allpairs(me, nranks, n):
    g, slot = [], 0
    for p in range(nranks):
        if p != me:
            g += [["recv", RECV_WIN + slot * n, n, p], ["send", SEND_WIN, n, p]]
            slot += 1
    return g


SCENARIOS = {
    "pair_sizes_rank0": dict(rank=0, nranks=2, groups=[[["send", SEND_WIN, n, 1]] for n in
                                                        (0, 8, 4096, 8193, 524288, 524289, 2 * MiB, 4 * MiB - 1, 4 * MiB, 256 * MiB)]),
    "pair_sizes_rank1": dict(rank=1, nranks=2, groups=[[["recv", RECV_WIN + 64, n, 0]] for n in
                                                        (0, 8, 4096, 8193, 524288, 524289, 2 * MiB, 4 * MiB - 1, 4 * MiB, 256 * MiB)]),
    "bidir": dict(rank=0, nranks=2, groups=[[["send", SEND_WIN, n, 1], ["recv", RECV_WIN, n, 1]] for n in (8, 1 * MiB, 32 * MiB, 256 * MiB)]),
    "bidir_unregistered_send": dict(rank=1, nranks=2, groups=[[["send", LOOSE, 8 * MiB, 0], ["recv", RECV_WIN, 8 * MiB, 0]]]),
    "ring8": dict(rank=3, nranks=8, groups=[[["recv", RECV_WIN, 256 * MiB, 2], ["send", SEND_WIN, 256 * MiB, 4]]] * 3),
    "allpairs8_64M": dict(rank=5, nranks=8, groups=[allpairs(5, 8, 64 * MiB)]),
    "allpairs8_1M": dict(rank=5, nranks=8, groups=[allpairs(5, 8, 1 * MiB)]),
    "allpairs8_4K": dict(rank=5, nranks=8, groups=[allpairs(5, 8, 4096)]),
    "allpairs16_512K": dict(rank=0, nranks=16, groups=[allpairs(0, 16, 512 * 1024)]),
    "alltoall8": dict(rank=2, nranks=8, groups=[[["alltoall", SEND_WIN, RECV_WIN, n]] for n in (1000, 1 * MiB, 64 * MiB)]),
    "self": dict(rank=0, nranks=1, groups=[[["recv", RECV_WIN, n, 0], ["send", SEND_WIN, n, 0]] for n in (8, 1 * MiB, 31 * MiB, 32 * MiB, 256 * MiB)]),
    "self_loopback": dict(rank=0, nranks=1, options={"loopback": 1, "bidir_push": 0},
                          groups=[[["recv", RECV_WIN, n, 0], ["send", SEND_WIN, n, 0]] for n in (8, 1 * MiB, 8 * MiB)]),
    "forced_tma": dict(rank=0, nranks=2, options={"path": 2}, groups=[[["send", SEND_WIN, n, 1]] for n in (1 * MiB, 64 * MiB)]),
    "mixed_group": dict(rank=0, nranks=4, groups=[[["send", SEND_WIN, 2 * MiB, 1], ["send", SEND_WIN + 2 * MiB, 8 * MiB, 1],
                                                    ["send", SEND_WIN, 100, 1], ["recv", RECV_WIN, 100, 2],
                                                    ["recv", RECV_WIN + MiB, 3 * MiB, 3], ["send", SEND_WIN, 16 * MiB, 2]]]),
}


def snapshot():
    out = {}
    for name, spec in SCENARIOS.items():
        spec = dict(spec, windows=WINDOWS)
        env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
        for k in list(env):
            if k.startswith("P2P_"):
                del env[k]
        res = subprocess.run([sys.executable, PROBE, json.dumps(spec)], cwd=ROOT, env=env, stdout=subprocess.PIPE,
                             stderr=subprocess.STDOUT, text=True, check=True).stdout
        out[name] = json.loads(res.strip().splitlines()[-1])
    return out


if __name__ == "__main__":
    path = os.path.join(ROOT, "tests", "golden", "plan_snapshot.json")
    with open(path, "w") as f:
        json.dump(snapshot(), f, indent=0, sort_keys=True)
    print("wrote", path)
