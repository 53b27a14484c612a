"""Host logic of the engine on CPU: what one group of p2p_send/p2p_recv calls turns into.

The engine is run in plan-only mode (test-nccl-p2p_b200/csrc/p2p_debug.h): the production code
resolves the group — protocol per message, roles, CTA counts, sequence numbers, flag addresses —
and records the kernel arguments instead of launching.  Addresses are made up but structured, so
a test can tell whose control region every flag lies in.  This is where the multi-GPU decisions
(who moves the bytes in a pair, in a ring, in all-pairs on 8 or 16 ranks) are pinned without GPUs;
the GPU suites then check that the kernels do what these arguments say.
"""
import json
import os
import sys

import pytest

from conftest import ROOT, run

PROBE = os.path.join(ROOT, "tests", "plan_probe.py")
CTL_BASE, CTL_STRIDE = 0x100000000000, 0x1000000000
SECTION = (64 * 4 + 4 + 8) * 8            # sizeof(PeerCtl): post, done, pull_post, pull_done [64], ll_ack[4], pad[8]
OFF = {"post": 0, "done": 512, "pull_post": 1024, "pull_done": 1536, "ll_ack": 2048}
SEND_WIN, RECV_WIN = 1 << 32, 1 << 33     # two fake registered windows: 1 GiB at 4 GiB, 4 GiB at 8 GiB
LOOSE = 1 << 36                           # an address in no window
VALID, PUSH_PLEASE = 1 << 39, 0xFF
MiB = 1 << 20


def flag(region_rank, section_peer, name, seq, ring=64):
    return CTL_BASE + region_rank * CTL_STRIDE + section_peer * SECTION + OFF[name] + (seq % ring) * 8


def plan(rank, nranks, groups, options=None, env=None, **kw):
    spec = {"rank": rank, "nranks": nranks, "groups": groups, "options": options or {},
            "windows": [[SEND_WIN, 1 << 30], [RECV_WIN, 4 << 30]], **kw}
    e = {"CUDA_VISIBLE_DEVICES": ""}
    e.update(env or {})
    out = run([sys.executable, PROBE, json.dumps(spec)], env=e).stdout
    return json.loads(out.strip().splitlines()[-1])


def kinds(group):
    return [o["kind"] for o in group["ops"]]


def test_one_way_large_message_is_pulled_by_the_receiver():
    n = 256 * MiB
    (g,) = plan(0, 2, [[["send", SEND_WIN, n, 1]]])
    (op,) = g["ops"]
    # sender: one control thread publishes where the payload lies, then waits until it has been read
    assert (op["kind"], g["grid"], g["n_ctrl"], g["pdl"]) == ("post", 1, 1, 3)   # PDL, movers trigger after their copy
    assert op["f_signal"] == flag(1, 0, "pull_post", 1) and op["f_wait"] == flag(0, 1, "pull_done", 1)
    assert op["word"] >> 48 == 1 and (op["word"] >> 40) & 0xFF == 0 and op["word"] & VALID and op["word"] & (VALID - 1) == 0
    (g,) = plan(1, 2, [[["recv", RECV_WIN + 4096, n, 0]]])
    (op,) = g["ops"]
    # receiver: 64 CTAs wait for that word in their own region, pull, and tell the sender's region
    assert (op["kind"], op["nblocks"], g["grid"], g["block"]) == ("recv_adapt", 64, 64, 512)
    assert op["f_wait2"] == flag(1, 0, "pull_post", 1) and op["f_signal2"] == flag(0, 1, "pull_done", 1)
    # ... and it is ready to be pushed to instead: its own post word (window 1, offset 4096) and flags
    assert (op["word"] >> 40) & 0xFF == 1 and op["word"] & (VALID - 1) == 4096
    assert op["f_signal"] == flag(0, 1, "post", 1) and op["f_wait"] == flag(1, 0, "done", 1)


def test_unregistered_send_buffer_announces_a_push():
    (g,) = plan(0, 2, [[["send", LOOSE, 8 * MiB, 1]]])
    (op,) = g["ops"]
    assert (op["kind"], op["nblocks"], g["n_ctrl"]) == ("push", 32, 0)
    assert op["f_signal2"] == flag(1, 0, "pull_post", 1) and (op["word2"] >> 40) & 0xFF == PUSH_PLEASE
    assert op["f_wait"] == flag(0, 1, "post", 1) and op["f_signal"] == flag(1, 0, "done", 1)
    # a large receive into an unregistered buffer is accepted (the sender may let us pull) but cannot post
    (g,) = plan(1, 2, [[["recv", LOOSE, 8 * MiB, 0]]])
    assert g["rc"] == 0 and g["ops"][0]["kind"] == "recv_adapt" and g["ops"][0]["word"] == 0


def test_pairwise_exchange_pushes_both_ways():
    n = 256 * MiB
    (g,) = plan(0, 2, [[["send", SEND_WIN, n, 1], ["recv", RECV_WIN, n, 1]]])
    assert kinds(g) == ["recv_adapt", "push"]           # receives are placed first: they start by waiting
    adapt, push = g["ops"]
    assert (adapt["nblocks"], push["nblocks"], g["grid"], g["n_ctrl"]) == (64, 32, 96, 0)
    assert push["signal2_owner"] == 1 and (push["word2"] >> 40) & 0xFF == PUSH_PLEASE
    assert (adapt["first_block"], push["first_block"]) == (0, 64)
    # with the heuristic off, the sender lets the peer pull as in the one-way case
    (g,) = plan(0, 2, [[["send", SEND_WIN, n, 1], ["recv", RECV_WIN, n, 1]]], options={"bidir_push": 0})
    assert kinds(g) == ["post", "recv_adapt"]


def test_ring_step_pulls_from_the_left_while_the_right_pulls_from_us():
    n = 256 * MiB
    (g,) = plan(3, 8, [[["recv", RECV_WIN, n, 2], ["send", SEND_WIN, n, 4]]])
    assert kinds(g) == ["post", "recv_adapt"]
    post, adapt = g["ops"]
    assert post["peer"] == 4 and post["signal_owner"] == 4 and post["wait_owner"] == 3
    # through the bulk-copy ring (flavour 2), with 96 light CTAs: 646-648 GB/s per GPU on 4 and 8 B200s, 64 CTAs 639-640
    assert adapt["peer"] == 2 and adapt["nblocks"] == 96 and adapt["signal2_owner"] == 2
    assert g["grid"] == 97 and g["n_ctrl"] == 1 and g["flavour"] == 2
    (g,) = plan(3, 8, [[["recv", RECV_WIN, n, 2], ["send", SEND_WIN, n, 4]]], options={"pull_tma_min_bytes": -1})
    assert g["flavour"] == 1 and g["grid"] == 65                       # the ld/st loop keeps its 64


@pytest.mark.parametrize("nranks", [4, 8, 16])
def test_all_pairs_large_messages(nranks):
    """Every rank sends 64 MiB to every other and receives as much: pulls that share about 64 CTAs between
    the flows (8 GPUs measured: 8 CTAs per flow 4.83 TB/s, 32 per flow 4.58, pushes 4.28, NCCL 4.6; on 4 GPUs
    8 per flow starved the three flows: 1.73 vs 2.43 TB/s), never fewer than 8 each, everything co-resident."""
    n, me = 64 * MiB, 1
    group = []
    slot = 0
    for p in range(nranks):
        if p != me:
            group += [["recv", RECV_WIN + slot * n, n, p], ["send", SEND_WIN, n, p]]
            slot += 1
    (g,) = plan(me, nranks, [group])
    ks = kinds(g)
    assert ks.count("post") == nranks - 1 and ks.count("recv_adapt") == nranks - 1 and "push" not in ks
    per_flow = max(8, 64 // (nranks - 1))
    assert all(o["nblocks"] == per_flow for o in g["ops"] if o["kind"] == "recv_adapt")
    assert g["n_ctrl"] == nranks - 1 and g["grid"] == 1 + per_flow * (nranks - 1) <= 148 * 2
    # each pull signals the rank it pulls from; sequence numbers are per peer
    for o in g["ops"]:
        assert o["seq"] == 1
        if o["kind"] == "recv_adapt":
            assert o["signal2_owner"] == o["peer"] and o["wait2_owner"] == me
    blocks = sorted((o["first_block"], o["nblocks"]) for o in g["ops"] if o["kind"] != "post")
    assert blocks[0][0] == 1 and all(a[0] + a[1] == b[0] for a, b in zip(blocks, blocks[1:]))   # contiguous, after the control CTA


def test_protocol_by_size_and_eager_launch_shape():
    sizes = [8, 4096, 8193, 512 * 1024, 512 * 1024 + 1, 4 * MiB - 1, 4 * MiB]
    groups = [[["send", SEND_WIN, n, 1]] for n in sizes]
    res = plan(0, 2, groups)
    got = [(r["ops"][0]["kind"], r["ops"][0]["nblocks"], r["block"], r["pdl"]) for r in res]
    assert got == [("send_ll", 1, 128, 1), ("send_ll", 1, 128, 1), ("send_ll", 1, 512, 1), ("send_ll", 16, 512, 1),
                   ("push", 9, 512, 0), ("push", 32, 512, 0), ("post", 0, 512, 3)]
    # eager messages go to the ring slot of (sender, seq) in the RECEIVER's region and poll the ack in ours
    first = res[0]["ops"][0]
    assert first["seq"] == 1 and first["wait_owner"] == 0 and first["f_wait"] == flag(0, 1, "ll_ack", 1, ring=4)
    assert CTL_BASE + CTL_STRIDE <= first["aux"] < CTL_BASE + 2 * CTL_STRIDE
    # eager and rendezvous messages count on separate lanes: 4 eager sends then 3 rendezvous sends
    assert [r["ops"][0]["seq"] for r in res] == [1, 2, 3, 4, 1, 2, 3]


def test_receive_into_an_unregistered_buffer_goes_through_the_staging_ring():
    """The reference's buffers are plain cudaMalloc memory nobody registered (p2p_matrix.cc:127-128).  Between the
    eager limit and the pull threshold the sender cannot store into such a receive buffer: the receiver gets mover
    CTAs of its own and asks for the staging ring instead of posting a window."""
    res = plan(1, 2, [[["recv", LOOSE, 2 * MiB, 0]], [["recv", RECV_WIN, 2 * MiB, 0]], [["recv", LOOSE, 1000, 0]]])
    staged = res[0]["ops"][0]
    assert res[0]["rc"] == 0 and staged["kind"] == "recv_staged" and staged["seq"] == 1 and staged["nblocks"] == 32
    assert staged["signal_owner"] == 0 and res[0]["n_ctrl"] == 0 and res[0]["grid"] == 32
    assert res[1]["rc"] == 0 and res[1]["ops"][0]["kind"] == "post" and res[1]["ops"][0]["seq"] == 2
    assert res[2]["rc"] == 0 and res[2]["ops"][0]["kind"] == "recv_ll"
    # a large message with neither buffer registered: the sender announces a push, the adaptive receive has no
    # window to post and will ask for the ring on the device
    (g,) = plan(1, 2, [[["recv", LOOSE, 8 * MiB, 0]]])
    assert kinds(g) == ["recv_adapt"] and g["ops"][0]["word"] == 0


def test_sequence_numbers_follow_program_order_not_role_order():
    """A pushed and a pulled message to the same peer in one group: the control role is emitted first,
    but the numbers must follow the order of the calls, or the peer would pair them wrongly."""
    (g,) = plan(0, 2, [[["send", SEND_WIN, 2 * MiB, 1], ["send", SEND_WIN + 2 * MiB, 8 * MiB, 1], ["send", SEND_WIN, 100, 1]]])
    by_size = {o["nbytes"]: o for o in g["ops"]}
    assert kinds(g) == ["post", "send_ll", "push"]
    assert (by_size[2 * MiB]["seq"], by_size[8 * MiB]["seq"], by_size[100]["seq"]) == (1, 2, 1)
    # the receiver numbers the matching receives the same way
    (g,) = plan(1, 2, [[["recv", RECV_WIN, 2 * MiB, 0], ["recv", RECV_WIN + 2 * MiB, 8 * MiB, 0], ["recv", RECV_WIN + 16 * MiB, 100, 0]]])
    by_size = {o["nbytes"]: o for o in g["ops"]}
    assert (by_size[2 * MiB]["seq"], by_size[8 * MiB]["seq"], by_size[100]["seq"]) == (1, 2, 1)
    assert (by_size[2 * MiB]["kind"], by_size[8 * MiB]["kind"]) == ("post", "recv_adapt")


def test_send_to_self_is_a_local_copy():
    big, small = 256 * MiB, 1 * MiB
    res = plan(0, 1, [[["recv", RECV_WIN, big, 0], ["send", SEND_WIN, big, 0]],
                      [["recv", RECV_WIN, small, 0], ["send", SEND_WIN, small, 0]],
                      [["recv", RECV_WIN, small, 0], ["send", SEND_WIN, small + 1, 0]]])
    g = res[0]
    (op,) = g["ops"]
    # one flag-free op; >= 32 MiB uses the TMA ring with one CTA per SM, and may overlap launches (PDL)
    assert (op["kind"], op["aux"], op["local"], op["seq"]) == ("copy", RECV_WIN, SEND_WIN, 0)
    assert (g["flavour"], g["grid"], g["pdl"], g["n_ctrl"]) == (2, 148, 1, 0)
    assert (res[1]["flavour"], res[1]["ops"][0]["nblocks"]) == (1, 16)
    assert res[2]["rc"] == 4 and "matched by a receive" in res[2]["error"]         # sizes must agree
    # loopback sends the same pair through the wire protocol instead
    (g,) = plan(0, 1, [[["recv", RECV_WIN, small, 0], ["send", SEND_WIN, small, 0]]], options={"loopback": 1})
    assert kinds(g) == ["post", "push"] and g["ops"][1]["signal_owner"] == 0


def test_alltoall_is_one_group_with_the_local_block_copied():
    blk = 1 * MiB
    (g,) = plan(2, 8, [[["alltoall", SEND_WIN, RECV_WIN, blk]]])
    ks = kinds(g)
    assert ks.count("copy") == 1 and ks.count("post") == 7 and ks.count("push") == 7 and g["flavour"] == 1
    copy = next(o for o in g["ops"] if o["kind"] == "copy")
    assert (copy["local"], copy["aux"]) == (SEND_WIN + 2 * blk, RECV_WIN + 2 * blk)
    for o in g["ops"]:
        if o["kind"] == "push":
            assert o["local"] == SEND_WIN + o["peer"] * blk and o["signal_owner"] == o["peer"]
    assert g["grid"] <= 148 * 2
    # large blocks: a fan-out, so the peers pull - through the bulk-copy ring, and the local block beside them with
    # no more CTAs than one flow could use
    (g,) = plan(2, 8, [[["alltoall", SEND_WIN, RECV_WIN, 64 * MiB]]])
    assert kinds(g).count("recv_adapt") == 7 and kinds(g).count("post") == 7 and g["flavour"] == 2
    assert all(o["nblocks"] == 9 for o in g["ops"] if o["kind"] == "recv_adapt")
    assert next(o for o in g["ops"] if o["kind"] == "copy")["nblocks"] == 32 and g["grid"] == 1 + 63 + 32 <= 148
    (g,) = plan(2, 8, [[["alltoall", SEND_WIN, RECV_WIN, 64 * MiB]]], options={"pull_tma_min_bytes": -1})
    assert g["flavour"] == 1


def test_eager_roles_never_take_more_than_half_the_machine():
    """All-pairs on 16 ranks at 512 KiB: 30 eager roles of 16 CTAs each would be 480 CTAs; everything has to
    be co-resident (roles wait for peers), so they are scaled to at most half of the 296 CTA slots."""
    n, me, nranks = 512 * 1024, 0, 16
    group = []
    for p in range(1, nranks):
        group += [["recv", RECV_WIN + p * n, n, p], ["send", SEND_WIN, n, p]]
    (g,) = plan(me, nranks, [group])
    assert kinds(g) == ["send_ll"] * 15 + ["recv_ll"] * 15       # sends first: they wait for nothing that runs here
    assert sum(o["nblocks"] for o in g["ops"]) <= 148 and g["grid"] <= 148
    assert all(o["nblocks"] >= 1 for o in g["ops"])


def test_group_limits():
    too_many = [["send", SEND_WIN, 8, 1]] * 33
    (g,) = plan(0, 2, [too_many])
    assert g["rc"] == 5 and "at most 32" in g["error"]


def test_plan_snapshot():
    """Every field of every kernel argument, for a fixed set of scenarios, against tests/golden/plan_snapshot.json
    (tests/make_plan_snapshot.py).  A difference is either a regression or an intended change of the planner
    that has to be re-measured on GPUs and re-recorded."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import make_plan_snapshot
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "plan_snapshot.json")))
    now = make_plan_snapshot.snapshot()
    assert sorted(now) == sorted(golden)
    for name in golden:
        assert now[name] == golden[name], f"planner output changed for scenario {name}"


def test_connect_checks_the_peers_blobs():
    """p2p_connect refuses ranks on another host (NVLink peer mappings stop at the host; the reference
    tolerates several hosts, p2p_matrix.cc:78-99), ranks with other protocol limits, blobs out of order, and
    two ranks that drive the same GPU."""
    code = r'''
import ctypes, importlib, struct, sys
sys.path.insert(0, %r)
p2p = importlib.import_module("test-nccl-p2p_b200")
L = p2p.lib()
L.p2p_debug_plan_init.argtypes = [ctypes.c_int] * 5
L.p2p_debug_plan_disconnect.argtypes = [ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_size_t)]
assert L.p2p_debug_plan_init(0, 2, 148, 2, 1) == 0
hh, cb = ctypes.c_uint64(), ctypes.c_size_t()
assert L.p2p_debug_plan_disconnect(ctypes.byref(hh), ctypes.byref(cb)) == 0
n = L.p2p_blob_bytes()

def blob(rank, host_hash=hh.value, eager=p2p.get_option("eager_bytes"), pull=p2p.get_option("pull_min_bytes"), magic=0x42325032, nranks=2, bus=None, stage_chunk=256 << 10):
    # BootBlob: magic, version, rank, nranks, host_hash, pid, device, eager_bytes, pull_min_bytes, ctl_bytes, pci_bus_id[32], ...
    head = struct.pack("<IIiiQiiQqQ", magic, 1, rank, nranks, host_hash, 424242 + rank, rank, eager, pull, cb.value)
    head += (bus if bus is not None else ("0000:%%02x:00.0" %% (rank + 1))).encode().ljust(32, b"\0")
    head += bytes(48 + 64)                                   # hostbox_name, ctl_handle
    head += struct.pack("<II", stage_chunk, 32)              # stage_chunk, stage_lanes
    return head + bytes(n - len(head))

def connect(b0, b1):
    rc = L.p2p_connect(ctypes.create_string_buffer(b0 + b1, 2 * n))
    return rc, L.p2p_get_last_error().decode()

rc, why = connect(blob(0), blob(1, host_hash=hh.value ^ 1))
assert rc == p2p.INVALID_USAGE and "another host" in why, (rc, why)
rc, why = connect(blob(0), blob(1, eager=12345))
assert rc == p2p.INVALID_USAGE and "eager_bytes" in why, (rc, why)
rc, why = connect(blob(0), blob(1, pull=1 << 30))
assert rc == p2p.INVALID_USAGE and "pull_min_bytes" in why, (rc, why)
rc, why = connect(blob(0), blob(1, stage_chunk=1 << 20))
assert rc == p2p.INVALID_USAGE and "staging ring" in why, (rc, why)
rc, why = connect(blob(1), blob(0))
assert rc == p2p.INVALID_ARGUMENT and "rank order" in why, (rc, why)
rc, why = connect(blob(0), blob(1, magic=7))
assert rc == p2p.INVALID_ARGUMENT and "not a p2p bootstrap blob" in why, (rc, why)
rc, why = connect(blob(0), blob(1, nranks=3))
assert rc == p2p.INVALID_ARGUMENT, (rc, why)
# two ranks on one GPU: kernels that wait for each other's flags must not share a device
rc, why = connect(blob(0, bus="0000:1b:00.0"), blob(1, bus="0000:1b:00.0"))
assert rc == p2p.INVALID_USAGE and "both drive GPU 0000:1b:00.0" in why, (rc, why)
print("refused as expected")
''' % ROOT
    out = run([sys.executable, "-c", code], env={"CUDA_VISIBLE_DEVICES": ""}).stdout
    assert "refused as expected" in out


def test_ring_policy_option_makes_the_lone_send_push():
    """bidir_push 2: a lone large send in a group that also receives a large message from ANOTHER peer (ring
    step) pushes instead of offering a pull; 1 (pairwise only) keeps offering; 3 extends it to fan-out groups."""
    n = 256 * MiB
    ring = [[["recv", RECV_WIN, n, 2], ["send", SEND_WIN, n, 4]]]
    (g1,) = plan(3, 8, ring, options={"bidir_push": 1})
    assert kinds(g1) == ["post", "recv_adapt"]
    (g2,) = plan(3, 8, ring, options={"bidir_push": 2})
    assert kinds(g2) == ["recv_adapt", "push"]
    push = g2["ops"][1]
    assert push["peer"] == 4 and push["nblocks"] == 32 and push["word2"] >> 40 & 0xFF == PUSH_PLEASE
    assert push["signal2_owner"] == 4 and push["wait_owner"] == 3
    fan = [[["recv", RECV_WIN, n, 2], ["send", SEND_WIN, n, 4], ["send", SEND_WIN, n, 5]]]
    (g,) = plan(3, 8, fan, options={"bidir_push": 2})
    assert kinds(g).count("push") == 0
    (g,) = plan(3, 8, fan, options={"bidir_push": 3})
    assert kinds(g).count("push") == 2


def test_early_post_only_when_the_launch_ahead_cannot_be_writing_the_payload():
    """With pdl 3 + early_post (the defaults) a send offered for pulling may be published while the engine launch ahead of it
    on the same stream is still running - unless that launch receives into the bytes being sent."""
    n = 16 * MiB
    groups = [[["send", SEND_WIN, n, 1]],                      # nothing ahead: no early post
              [["send", SEND_WIN, n, 1]],                      # ahead: a send (writes nothing)
              [["recv", RECV_WIN, n, 1]],
              [["send", RECV_WIN + 4096, n, 1]],               # forwards what the launch ahead receives: must wait
              [["send", SEND_WIN, n, 1]]]
    res = plan(0, 2, groups)                                   # the default: pdl 3, early_post 1
    flags = [r["ops"][0]["flags"] for r in res]
    assert flags == [0, 1, 0, 0, 1]
    assert [r["pdl"] for r in res] == [3] * 5
    assert [r["prev_token"] for r in res] == [0] + [r["my_token"] for r in res[:-1]]
    for opts in ({"early_post": 0}, {"pdl": 1}, {"pdl": 0}):
        res = plan(0, 2, groups, options=opts)
        assert all(r["ops"][0]["flags"] == 0 for r in res), opts
    assert all(r["pdl"] == 0 for r in res)


def test_options_from_the_environment_are_range_checked():
    """The same checks as p2p_set_option: a bad P2P_* value fails p2p_init instead of dividing by zero later."""
    code = r'''
import ctypes, importlib, sys
sys.path.insert(0, %r)
p2p = importlib.import_module("test-nccl-p2p_b200")
L = p2p.lib()
L.p2p_debug_plan_init.argtypes = [ctypes.c_int] * 5
rc = L.p2p_debug_plan_init(0, 2, 148, 2, 1)
print("rc", rc, L.p2p_get_last_error().decode())
''' % ROOT
    for env, word in (({"P2P_TMA_CHUNK": "0"}, "tma_chunk"), ({"P2P_TMA_STAGES": "0"}, "tma_stages"),
                      ({"P2P_TMA_CHUNK": "4100"}, "multiple of 16"), ({"P2P_TMA_CHUNK": "65536", "P2P_TMA_STAGES": "8"}, "200 KiB"),
                      ({"P2P_NBLOCKS": "-3"}, "nblocks"), ({"P2P_TIMEOUT_MS": "-1"}, "timeout_ms"), ({"P2P_EAGER_BYTES": "2000000"}, "eager_bytes")):
        out = run([sys.executable, "-c", code], env={"CUDA_VISIBLE_DEVICES": "", **env}).stdout
        assert out.startswith("rc 4 ") and word in out, (env, out)
    out = run([sys.executable, "-c", code], env={"CUDA_VISIBLE_DEVICES": "", "P2P_TMA_CHUNK": "16384", "P2P_PATH": "tma"}).stdout
    assert out.startswith("rc 0"), out


def test_pairwise_exchange_pulls_when_only_the_send_buffers_are_windows():
    """A pairwise exchange pushes both ways - unless this rank's own receive buffer is in no window: the peer's then
    most likely is not either, a push would end in the staging ring, and a pull (which needs only the SEND buffer
    registered) goes straight."""
    n = 32 * MiB
    (g,) = plan(0, 2, [[["send", SEND_WIN, n, 1], ["recv", RECV_WIN, n, 1]]])
    assert sorted(kinds(g)) == ["push", "recv_adapt"]
    (g,) = plan(0, 2, [[["send", SEND_WIN, n, 1], ["recv", LOOSE, n, 1]]])
    assert sorted(kinds(g)) == ["post", "recv_adapt"]                      # the send is offered for pulling
    adapt = next(o for o in g["ops"] if o["kind"] == "recv_adapt")
    assert adapt["word"] == 0                                              # nothing to post: it pulls, or asks for the ring
    (g,) = plan(0, 2, [[["send", LOOSE, n, 1], ["recv", LOOSE, n, 1]]])   # nothing registered at all: push into the ring
    assert sorted(kinds(g)) == ["push", "recv_adapt"]
