"""TEST INFRASTRUCTURE — not part of the product (see oracle/pattern.h for the rules).

numpy restatement of oracle/pattern.c, vectorised so that tests can check 256 MiB buffers
in seconds, plus a ctypes loader for the C build (oracle/_ref/libp2poracle.so).

The path under test is ncclSend/ncclRecv as called by the reference (p2p_matrix.cc:157-169,
grouped at :212-227): an identity copy.  `expected_recv` states that law; everything else
exists to make violations visible.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

GAMMA = np.uint64(0x9E3779B97F4A7C15)
_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)
_ITER_SALT = np.uint64(0xD1B54A32D192ED03)


def mix64(z):
    """splitmix64 finaliser on uint64 scalars or arrays (wraps modulo 2**64)."""
    with np.errstate(over="ignore"):
        z = np.asarray(z, dtype=np.uint64)
        z = (z ^ (z >> np.uint64(30))) * _M1
        z = (z ^ (z >> np.uint64(27))) * _M2
        return z ^ (z >> np.uint64(31))


def pattern_key(seed: int, rank: int, it: int) -> np.uint64:
    with np.errstate(over="ignore"):
        a = mix64(np.uint64(seed & 0xFFFFFFFFFFFFFFFF) + GAMMA * np.uint64(rank + 1))
        b = mix64(np.uint64(it & 0xFFFFFFFFFFFFFFFF) + _ITER_SALT)
        return np.uint64(a ^ b)


def pattern_words(nwords: int, seed: int, rank: int, it: int) -> np.ndarray:
    with np.errstate(over="ignore"):
        w = np.arange(1, nwords + 1, dtype=np.uint64)
        return mix64(pattern_key(seed, rank, it) + GAMMA * w)


def pattern_bytes(n: int, seed: int, rank: int, it: int) -> np.ndarray:
    """The payload rank `rank` sends in iteration `it`: n bytes (uint8)."""
    words = pattern_words((n + 7) // 8, seed, rank, it)
    return words.astype("<u8").view(np.uint8)[:n].copy()


def expected_recv(n: int, seed: int, src_rank: int, it: int) -> np.ndarray:
    """Transfer law (p2p_matrix.cc:157-169): dst's receive buffer == src's send buffer."""
    return pattern_bytes(n, seed, src_rank, it)


def alltoall_send(block: int, seed: int, rank: int, nranks: int) -> np.ndarray:
    """What rank `rank` puts in its all-to-all send buffer: block p is the payload (seed, rank, iter=p)."""
    return np.concatenate([pattern_bytes(block, seed, rank, p) for p in range(nranks)]) if block else np.zeros(0, np.uint8)


def expected_alltoall_recv(block: int, seed: int, rank: int, nranks: int) -> np.ndarray:
    """All-to-all law, i.e. the transfer law per peer: block p of rank `rank`'s receive buffer is block
    `rank` of rank p's send buffer.  (What a grouped ncclSend/ncclRecv loop over all peers leaves,
    nccl.h:438-440; p2p_matrix.cc:212-228 is the one-peer case.)"""
    return np.concatenate([pattern_bytes(block, seed, p, rank) for p in range(nranks)]) if block else np.zeros(0, np.uint8)


def mismatches(buf: np.ndarray, seed: int, rank: int, it: int):
    """(number of differing bytes, offset of the first one or len(buf))."""
    buf = np.asarray(buf, dtype=np.uint8).ravel()
    want = pattern_bytes(buf.size, seed, rank, it)
    diff = np.nonzero(buf != want)[0]
    return int(diff.size), int(diff[0]) if diff.size else int(buf.size)


def digest(buf: np.ndarray) -> int:
    buf = np.asarray(buf, dtype=np.uint8).ravel()
    n = buf.size
    nwords = (n + 7) // 8
    padded = np.zeros(nwords * 8, dtype=np.uint8)
    padded[:n] = buf
    words = padded.view("<u8").astype(np.uint64)
    with np.errstate(over="ignore"):
        idx = np.arange(1, nwords + 1, dtype=np.uint64)
        total = np.sum(mix64(words ^ (GAMMA * idx)), dtype=np.uint64) if nwords else np.uint64(0)
        return int(mix64(np.uint64(total) ^ mix64(np.uint64(n))))


def pattern_digest(n: int, seed: int, rank: int, it: int) -> int:
    return digest(pattern_bytes(n, seed, rank, it))


def host_hash(hostname: str) -> int:
    """getHostName + getHostHash, p2p_matrix.cc:44-61."""
    h = 5381
    for ch in hostname.split(".")[0].encode():
        h = (((h << 5) + h) ^ ch) & 0xFFFFFFFFFFFFFFFF
    return h


def placement(hashes, rank: int) -> int:
    """check_process_placement_policy, p2p_matrix.cc:63-100 (-1 uneven, -2 not consecutive)."""
    size = len(hashes)
    num_hosts = len(set(hashes))
    if size % num_hosts:
        return -1
    per = size // num_hosts
    for host in range(num_hosts):
        for g in range(1, per):
            if hashes[host * per + g] != hashes[host * per + g - 1]:
                return -2
    return rank % per


def gbps(msg_bytes: int, seconds_per_message: float, bidirectional: bool) -> float:
    """p2p_matrix.cc:177 and :258."""
    t = msg_bytes * 8.0 / seconds_per_message / 1e9
    return t * 2 if bidirectional else t


# ---------------------------------------------------------------- the C build

_LIB = None


def load_c(path: str | None = None):
    """ctypes handle on oracle/_ref/libp2poracle.so (built by oracle/Makefile)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if path is None:
        path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libp2poracle.so")
    lib = ctypes.CDLL(path)
    u64, i32, sz = ctypes.c_uint64, ctypes.c_int, ctypes.c_size_t
    p8 = ctypes.POINTER(ctypes.c_uint8)
    lib.oracle_mix64.restype = u64
    lib.oracle_mix64.argtypes = [u64]
    lib.oracle_pattern_key.restype = u64
    lib.oracle_pattern_key.argtypes = [u64, i32, u64]
    lib.oracle_pattern_byte.restype = ctypes.c_uint8
    lib.oracle_pattern_byte.argtypes = [u64, i32, u64, u64]
    lib.oracle_pattern_fill.restype = None
    lib.oracle_pattern_fill.argtypes = [p8, sz, u64, i32, u64]
    lib.oracle_expected_recv.restype = None
    lib.oracle_expected_recv.argtypes = [p8, sz, u64, i32, u64]
    lib.oracle_pattern_mismatches.restype = u64
    lib.oracle_pattern_mismatches.argtypes = [p8, sz, u64, i32, u64, ctypes.POINTER(u64)]
    lib.oracle_digest.restype = u64
    lib.oracle_digest.argtypes = [p8, sz]
    lib.oracle_pattern_digest.restype = u64
    lib.oracle_pattern_digest.argtypes = [sz, u64, i32, u64]
    lib.oracle_host_hash.restype = u64
    lib.oracle_host_hash.argtypes = [ctypes.c_char_p]
    lib.oracle_placement.restype = i32
    lib.oracle_placement.argtypes = [ctypes.POINTER(u64), i32, i32]
    lib.oracle_gbps.restype = ctypes.c_double
    lib.oracle_gbps.argtypes = [u64, ctypes.c_double, i32]
    _LIB = lib
    return lib


def c_pattern_bytes(n: int, seed: int, rank: int, it: int) -> np.ndarray:
    lib = load_c()
    out = np.empty(n, dtype=np.uint8)
    lib.oracle_pattern_fill(out.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8)), n, seed, rank, it)
    return out


def c_digest(buf: np.ndarray) -> int:
    lib = load_c()
    buf = np.ascontiguousarray(buf, dtype=np.uint8)
    return int(lib.oracle_digest(buf.ctypes.data_as(ctypes.POINTER(ctypes.c_uint8)), buf.size))
