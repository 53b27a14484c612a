"""The wire protocol across ranks, executed on CPU.

For every rank of a job the production group builder is run in plan-only mode (tests/plan_probe.py),
which yields the exact kernel arguments that rank would launch: roles, sequence numbers, post words
and the ADDRESSES of every flag, in one made-up but job-wide consistent address space.  A small
interpreter then plays the device roles of all ranks against one shared memory, choosing at random
which waiting role gets to run next, and checks that

  * nothing deadlocks (every role of every group finishes, under many interleavings),
  * every receive ends up holding the payload of the send it is paired with (k-th send A->B with
    k-th receive at B from A), through whichever protocol the sizes select,
  * flag rings can be reused (more than 64 rendezvous / 4 eager messages per pair).

The interpreter restates the roles of csrc/p2p_device.cuh (post_and_wait, push_wait_post +
movers_signal, recv_adaptive, ll_send, ll_recv) at message granularity.  It validates the wiring
between ranks and the protocol's design; the kernels themselves are checked on GPUs.
"""
import json
import os
import random
import sys
import tempfile

import pytest

from conftest import ROOT, run

PROBE = os.path.join(ROOT, "tests", "plan_probe.py")
SEND_WIN, RECV_WIN = 1 << 32, 1 << 33
WINDOWS = [[SEND_WIN, 1 << 30], [RECV_WIN, 4 << 30]]
VALID, OFFMASK, PUSH_PLEASE, STAGED = 1 << 39, (1 << 39) - 1, 0xFF, 0xFE
CTL_BASE, CTL_STRIDE = 0x100000000000, 0x1000000000
SECTION = (64 * 4 + 4 + 8) * 8            # sizeof(PeerCtl)
STAGE_CHUNK, STAGE_SLOTS, STAGE_LANES = 256 << 10, 4, 32
MiB = 1 << 20
CLEAR_CONSUMED_SLOTS = True   # what the device code does; the wrap test below also shows what happens without


def stage_flag(region_rank, from_rank, kind, lane):
    """Counter `kind`[lane] in the header of the ring area [region of region_rank][from from_rank] (written by from_rank)."""
    return ("stage", region_rank, from_rank, kind, lane)


def stage_count(seq, chunks):
    return (seq << 24) | chunks


def stage_chunks_of(cta, nctas, nchunks):
    """Chunks a CTA of a staged message serves: lane l belongs to CTA l % nctas; increasing order."""
    return [c for c in range(nchunks) if (c % STAGE_LANES) % nctas == cta]


def interleave(world, ctas):
    """Several CTAs of one role: each step one of those that still run advances, chosen at random."""
    ctas = list(ctas)
    while ctas:
        g = world.rng.choice(ctas)
        try:
            next(g)
        except StopIteration:
            ctas.remove(g)
        yield


def staged_send(world, src, dst, op):
    """Sender side of the staging ring (stage_begin / stage_end of csrc/p2p_device.cuh), CTA by CTA, chunk by chunk."""
    seq, n, nctas = op["seq"], op["nbytes"], max(op["nblocks"], 1)
    nchunks = -(-n // STAGE_CHUNK)
    done = [0]

    def cta(j):
        for c in stage_chunks_of(j, nctas, nchunks):
            lane, k = c % STAGE_LANES, c // STAGE_LANES
            if k >= STAGE_SLOTS:
                while world.flags.get(stage_flag(src, dst, "drained", lane), 0) < stage_count(seq, k + 1 - STAGE_SLOTS):
                    yield
            world.ring[("stage", src, dst, lane, k % STAGE_SLOTS)] = (seq, c, op["local"])
            world.flags[stage_flag(dst, src, "filled", lane)] = stage_count(seq, k + 1)
            yield
        done[0] += 1
        if done[0] == nctas and CLEAR_CONSUMED_SLOTS:
            world.flags[op["f_wait"]] = 0

    yield from interleave(world, [cta(j) for j in range(nctas)])


def staged_recv(world, dst, src, op, consumed=None):
    """Receiver side: take the pair's ring, ask the sender to use it, copy chunk after chunk out of its lanes."""
    seq, n, nctas = op["seq"], op["nbytes"], max(op["nblocks"], 1)
    lock = ("stage_lock", dst, src)
    while world.flags.get(lock, 0):
        yield
    world.flags[lock] = 1
    world.flags[op["f_signal"]] = (seq & 0xFFFF) << 48 | STAGED << 40 | VALID
    nchunks = -(-n // STAGE_CHUNK)
    got_chunks, done = {}, [0]

    def cta(j):
        for c in stage_chunks_of(j, nctas, nchunks):
            lane, k = c % STAGE_LANES, c // STAGE_LANES
            while world.flags.get(stage_flag(dst, src, "filled", lane), 0) < stage_count(seq, k + 1):
                yield
            got = world.ring[("stage", src, dst, lane, k % STAGE_SLOTS)]
            assert got[:2] == (seq, c), f"staging slot holds chunk {got[:2]} where {(seq, c)} was expected"
            got_chunks[c] = got[2]
            world.flags[stage_flag(src, dst, "drained", lane)] = stage_count(seq, k + 1)
            yield
        done[0] += 1
        if done[0] == nctas:
            if consumed is not None and CLEAR_CONSUMED_SLOTS:
                world.flags[consumed] = 0
            assert sorted(got_chunks) == list(range(nchunks)) and len(set(got_chunks.values())) == 1
            world.mem[dst][op["local"]] = ("payload", src, got_chunks[0], n)
            world.flags[lock] = 0

    yield from interleave(world, [cta(j) for j in range(nctas)])


def plans_for(nranks, programs, options=None):
    """programs[r] = list of groups for rank r; returns per rank the list of planned groups."""
    out = []
    for r in range(nranks):
        spec = {"rank": r, "nranks": nranks, "groups": programs[r], "windows": WINDOWS, "options": options or {}}
        text = json.dumps(spec)
        if len(text) > 100_000:
            with tempfile.NamedTemporaryFile("w", suffix=".json", delete=False) as tf:
                tf.write(text)
            try:
                res = run([sys.executable, PROBE, "@" + tf.name], env={"CUDA_VISIBLE_DEVICES": ""}, timeout=900).stdout
            finally:
                os.unlink(tf.name)
        else:
            res = run([sys.executable, PROBE, text], env={"CUDA_VISIBLE_DEVICES": ""}).stdout
        groups = json.loads(res.strip().splitlines()[-1])
        assert all(g["rc"] == 0 for g in groups), groups
        out.append(groups)
    return out


def window_local(word):
    """Local address a post word designates (the windows have the same base on every rank here)."""
    win, off = (word >> 40) & 0xFF, word & OFFMASK
    return WINDOWS[win][0] + off


class World:
    def __init__(self, nranks, rng=None):
        self.rng = rng or random.Random(0)
        self.flags = {}                                  # control words, job-wide addresses
        self.ring = {}                                   # eager slot address -> (seq, payload tag)
        self.mem = [dict() for _ in range(nranks)]       # per rank: buffer address -> payload tag

    def tagged(self, addr, seq):
        v = self.flags.get(addr, 0)
        return v if (v >> 48) == (seq & 0xFFFF) and v & VALID else None


def role(world, rank, op):
    """Generator: yields while it has to wait, returns when the role is complete."""
    k, seq, peer = op["kind"], op["seq"], op["peer"]
    if k == "post":
        world.flags[op["f_signal"]] = op["word"]
        while world.flags.get(op["f_wait"], 0) != seq:
            yield
    elif k == "push":
        if op["f_signal2"]:
            world.flags[op["f_signal2"]] = op["word2"]
        while world.tagged(op["f_wait"], seq) is None:
            yield
        if (world.tagged(op["f_wait"], seq) >> 40) & 0xFF == STAGED:
            yield from staged_send(world, rank, peer, op)
            return
        dst = window_local(world.tagged(op["f_wait"], seq))
        world.mem[peer][dst] = ("payload", rank, op["local"], op["nbytes"])
        if CLEAR_CONSUMED_SLOTS:
            world.flags[op["f_wait"]] = 0      # movers_signal: the last CTA zeroes the slot the post arrived in
        world.flags[op["f_signal"]] = seq
    elif k == "recv_adapt":
        while world.tagged(op["f_wait2"], seq) is None:
            yield
        word = world.tagged(op["f_wait2"], seq)
        if (word >> 40) & 0xFF == PUSH_PLEASE and not op["word"]:
            yield from staged_recv(world, rank, peer, op, consumed=op["f_wait2"])   # neither buffer is in a window
        elif (word >> 40) & 0xFF == PUSH_PLEASE:
            world.flags[op["f_signal"]] = op["word"]
            while world.flags.get(op["f_wait"], 0) != seq:
                yield
            if CLEAR_CONSUMED_SLOTS:
                world.flags[op["f_wait2"]] = 0
        else:
            world.mem[rank][op["local"]] = ("payload", peer, window_local(word), op["nbytes"])
            if CLEAR_CONSUMED_SLOTS:
                world.flags[op["f_wait2"]] = 0
            world.flags[op["f_signal2"]] = seq
    elif k == "recv_staged":
        yield from staged_recv(world, rank, peer, op)
    elif k == "send_ll":
        if seq > 4:
            while world.flags.get(op["f_wait"], 0) != seq - 4:
                yield
        world.ring[op["aux"]] = (seq, ("payload", rank, op["local"], op["nbytes"]))
    elif k == "recv_ll":
        while world.ring.get(op["aux"], (0, None))[0] != seq:
            yield
        world.mem[rank][op["local"]] = world.ring[op["aux"]][1]
        world.flags[op["f_signal"]] = seq
    elif k == "copy":
        world.mem[rank][op["aux"]] = ("payload", rank, op["local"], op["nbytes"])
    else:
        raise AssertionError(k)


def simulate(nranks, plans, seed, on_group_done=None):
    """Each rank runs its groups in order (one stream); roles of the current groups interleave at random."""
    rng = random.Random(seed)
    world = World(nranks, rng)
    cursor = [0] * nranks
    live = [[] for _ in range(nranks)]

    def start(r):
        live[r] = [role(world, r, op) for op in plans[r][cursor[r]]["ops"]] if cursor[r] < len(plans[r]) else []

    def settle(r):
        """Move rank r past every group that has nothing (left) to run."""
        while cursor[r] < len(plans[r]) and not live[r]:
            if on_group_done:
                on_group_done(world, r, cursor[r])
            cursor[r] += 1
            start(r)

    for r in range(nranks):
        start(r)
        settle(r)
    idle_rounds = 0
    while any(cursor[r] < len(plans[r]) for r in range(nranks)):
        candidates = [(r, g) for r in range(nranks) for g in live[r]]
        assert candidates, f"nothing can run, ranks at groups {cursor}"
        r, g = rng.choice(candidates)
        before = (dict(world.flags), dict(world.ring))
        try:
            next(g)
            finished = False
        except StopIteration:
            live[r].remove(g)
            finished = True
        progressed = finished or (world.flags, world.ring) != before
        idle_rounds = 0 if progressed else idle_rounds + 1
        assert idle_rounds < 20000, f"deadlock: ranks stuck at groups {cursor}"
        settle(r)
    return world


def check_phases(nranks, phases, options=None, seeds=(1, 2, 3)):
    """phases: list of lists of (src, dst, nbytes).  Every rank groups its share of a phase; buffers are
    distinct per message so that each receive can be traced to its send."""
    programs = [[] for _ in range(nranks)]
    expect = [[] for _ in range(nranks)]   # per rank and phase: recv buffer -> (src, send buffer, nbytes)
    for phase in phases:
        groups = [[] for _ in range(nranks)]
        exp = [dict() for _ in range(nranks)]
        slot = [0] * nranks
        for src, dst, n in phase:
            sbuf = SEND_WIN + slot[src] * (64 * MiB)
            slot[src] += 1
            rbuf = RECV_WIN + slot[dst] * (64 * MiB)
            slot[dst] += 1
            groups[dst].append(["recv", rbuf, n, src])
            groups[src].append(["send", sbuf, n, dst])
            exp[dst][rbuf] = ("payload", src, sbuf, n)
        for r in range(nranks):
            programs[r].append(groups[r])
            expect[r].append(exp[r])
    plans = plans_for(nranks, programs, options)
    for r in range(nranks):
        assert len(plans[r]) == len(programs[r])

    for seed in seeds:
        def verify(world, r, gi):
            for buf, want in expect[r][gi].items():
                assert world.mem[r].get(buf) == want, (seed, r, gi, hex(buf), world.mem[r].get(buf), want)
        simulate(nranks, plans, seed, on_group_done=verify)


def test_pair_every_protocol_one_way_and_both_ways():
    sizes = [8, 100_000, 512 * 1024, 600_000, 2 * MiB, 4 * MiB, 48 * MiB]
    check_phases(2, [[(0, 1, n)] for n in sizes] + [[(0, 1, n), (1, 0, n)] for n in sizes])


def test_flag_rings_wrap_without_losing_a_message():
    # 70 rendezvous messages and 12 eager ones one way (rings hold 64 and 4), then the same both ways
    check_phases(2, [[(0, 1, 2 * MiB)]] * 70 + [[(0, 1, 64)]] * 12 + [[(0, 1, 8 * MiB), (1, 0, 8 * MiB)]] * 70, seeds=(7,))


@pytest.mark.parametrize("nranks", [3, 8])
def test_ring_and_all_pairs(nranks):
    ring = [[(r, (r + 1) % nranks, 32 * MiB) for r in range(nranks)]] * 3
    allpairs = [[(s, d, n) for s in range(nranks) for d in range(nranks) if s != d] for n in (4096, 1 * MiB, 16 * MiB)]
    check_phases(nranks, ring + allpairs, seeds=(11, 12))


def test_all_pairs_sixteen_ranks_eager_budget():
    check_phases(16, [[(s, d, 512 * 1024) for s in range(16) for d in range(16) if s != d]], seeds=(5,))


def test_random_traffic():
    rng = random.Random(2026)
    nranks = 5
    sizes = [0, 8, 5000, 512 * 1024, 512 * 1024 + 1, 3 * MiB, 4 * MiB, 20 * MiB]
    phases = []
    for _ in range(25):
        phase = []
        for _ in range(rng.randint(1, 8)):
            s, d = rng.sample(range(nranks), 2)
            phase.append((s, d, rng.choice(sizes)))
        phases.append(phase)
    check_phases(nranks, phases, seeds=(21, 22, 23))


def test_unregistered_send_buffers_fall_back_to_push():
    """Large messages from a buffer that is in no window: the sender announces a push, the receiver posts."""
    global SEND_WIN
    keep = SEND_WIN
    try:
        SEND_WIN = 1 << 37          # outside both windows
        check_phases(3, [[(0, 1, 8 * MiB), (1, 2, 8 * MiB), (2, 0, 8 * MiB)], [(0, 2, 32 * MiB)]], seeds=(3,))
    finally:
        SEND_WIN = keep


def test_sequence_tag_wraps_after_65536_messages():
    """Post words carry a 16-bit sequence tag.  Message 1 is pushed (its post lands in slot 1 of the sender's
    post ring), the next 65535 are pulled (they use the other ring and never rewrite that slot), message 65537
    is pushed again into ANOTHER buffer: its tag and slot equal message 1's.  The consumer zeroes a slot once
    the word in it has been used, so the old post cannot be taken for the new one; without that the payload
    goes to message 1's address whenever the sender runs before the receiver has posted."""
    global CLEAR_CONSUMED_SLOTS
    phases = [[(0, 1, 2 * MiB)]] + [[(0, 1, 8 * MiB)]] * 65535 + [[(0, 1, 8), (0, 1, 2 * MiB)]]
    check_phases(2, phases, seeds=(1, 2, 3, 4))
    CLEAR_CONSUMED_SLOTS = False
    try:
        with pytest.raises(AssertionError):
            check_phases(2, phases, seeds=(1, 2, 3, 4))
    finally:
        CLEAR_CONSUMED_SLOTS = True


def test_unregistered_receive_buffers_go_through_the_staging_ring():
    """The reference's own situation: plain buffers, nothing registered.  Every size above the eager limit then goes
    through the staging ring - messages of more chunks than the ring has slots, back to back, both ways at once,
    and a ring of three ranks."""
    global SEND_WIN, RECV_WIN
    keep = SEND_WIN, RECV_WIN
    try:
        SEND_WIN, RECV_WIN = 1 << 37, 1 << 38          # outside both windows
        sizes = [600_000, 1 * MiB, 3 * MiB + 5, 4 * MiB, 9 * MiB, 33 * MiB + 1]   # 33 MiB: every lane wraps its slots
        check_phases(2, [[(0, 1, n)] for n in sizes] + [[(0, 1, n), (1, 0, n)] for n in sizes], seeds=(1, 2, 3))
        check_phases(3, [[(0, 1, 6 * MiB), (1, 2, 6 * MiB), (2, 0, 2 * MiB)]] * 3, seeds=(4, 5))
        # only the receiver unregistered: large messages are pulled as ever, mid-size ones are staged
        SEND_WIN = keep[0]
        check_phases(2, [[(0, 1, n)] for n in sizes] + [[(0, 1, n), (1, 0, n)] for n in sizes], seeds=(6, 7))
    finally:
        SEND_WIN, RECV_WIN = keep
