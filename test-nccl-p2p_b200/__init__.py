"""ctypes binding of libb200p2p.so — the C ABI declared in include/p2p.h.

This is plumbing for bench.py, __graft_entry__.py and the tests: it adds no behaviour of
its own.  Names mirror the C entry points, which mirror the reference's NCCL calls
(ncclSend/ncclRecv/ncclGroupStart/ncclGroupEnd, p2p_matrix.cc:157-169, 212-227).

The directory name carries a hyphen, so import it with
    importlib.import_module("test-nccl-p2p_b200")
with the repository root on sys.path.

There is no fallback: if the shared library has not been built, or no B200 is present,
importing works but every call raises.
"""
from __future__ import annotations

import ctypes
import os
from typing import Callable, List, Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libb200p2p.so")

# ncclDataType_t numbering (nccl.h:278-291)
INT8, UINT8, INT32, UINT32, INT64, UINT64, FLOAT16, FLOAT32, FLOAT64, BFLOAT16, FLOAT8E4M3, FLOAT8E5M2 = range(12)
DTYPE_SIZE = {INT8: 1, UINT8: 1, INT32: 4, UINT32: 4, INT64: 8, UINT64: 8, FLOAT16: 2, FLOAT32: 4,
              FLOAT64: 8, BFLOAT16: 2, FLOAT8E4M3: 1, FLOAT8E5M2: 1}

SUCCESS, UNHANDLED_CUDA_ERROR, SYSTEM_ERROR, INTERNAL_ERROR, INVALID_ARGUMENT, INVALID_USAGE, REMOTE_ERROR, IN_PROGRESS = range(8)

PATH_AUTO, PATH_LDST, PATH_TMA, PATH_CE, PATH_EAGER, PATH_PULL = 0, 1, 2, 3, 4, 5

# every symbol include/p2p.h declares; tests check the library exports exactly these
SYMBOLS = [
    "p2p_init", "p2p_blob_bytes", "p2p_export", "p2p_connect", "p2p_disconnect", "p2p_finalize", "p2p_rank", "p2p_nranks",
    "p2p_register_blob_bytes", "p2p_register_export", "p2p_register_import", "p2p_deregister", "p2p_window_peer_ptr",
    "p2p_mem_alloc", "p2p_mem_free", "p2p_send", "p2p_recv", "p2p_group_start", "p2p_group_end", "p2p_alltoall",
    "p2p_symm_blob_bytes", "p2p_symm_alloc_export", "p2p_symm_alloc_import", "p2p_symm_free", "p2p_symm_multicast_ptr", "p2p_bcast",
    "p2p_get_error_string", "p2p_get_last_error", "p2p_get_async_error", "p2p_set_option", "p2p_get_option",
    "p2p_query_plan", "p2p_pattern_fill", "p2p_pattern_check", "p2p_digest", "p2p_digest_noalloc",
]
NOALLOC_MAX_BYTES = 1 << 30


class P2PError(RuntimeError):
    def __init__(self, code: int, what: str, detail: str):
        super().__init__(f"{what}: {detail}" if detail else what)
        self.code = code


_lib = None


def lib() -> ctypes.CDLL:
    """The loaded shared library.  Fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: run `make` (or __graft_entry__.build()) first. "
            "This engine is CUDA-only; there is no Python or CPU fallback.")
    L = ctypes.CDLL(LIB_PATH)
    vp, sz, i32, u64, ll = ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_uint64, ctypes.c_longlong
    pu64 = ctypes.POINTER(u64)
    sig = {
        "p2p_init": (i32, [i32, i32, i32]),
        "p2p_blob_bytes": (sz, []),
        "p2p_export": (i32, [vp]),
        "p2p_connect": (i32, [vp]),
        "p2p_disconnect": (i32, []),
        "p2p_finalize": (i32, []),
        "p2p_rank": (i32, []),
        "p2p_nranks": (i32, []),
        "p2p_register_blob_bytes": (sz, []),
        "p2p_register_export": (i32, [vp, sz, vp]),
        "p2p_register_import": (i32, [vp, ctypes.POINTER(i32)]),
        "p2p_deregister": (i32, [i32]),
        "p2p_window_peer_ptr": (i32, [i32, i32, ctypes.POINTER(vp)]),
        "p2p_mem_alloc": (i32, [ctypes.POINTER(vp), sz]),
        "p2p_mem_free": (i32, [vp]),
        "p2p_send": (i32, [vp, sz, i32, i32, vp]),
        "p2p_recv": (i32, [vp, sz, i32, i32, vp]),
        "p2p_alltoall": (i32, [vp, vp, sz, i32, vp]),
        "p2p_symm_blob_bytes": (sz, []),
        "p2p_symm_alloc_export": (i32, [sz, vp]),
        "p2p_symm_alloc_import": (i32, [vp, ctypes.POINTER(vp), ctypes.POINTER(i32)]),
        "p2p_symm_free": (i32, [i32]),
        "p2p_symm_multicast_ptr": (i32, [i32, ctypes.POINTER(vp)]),
        "p2p_bcast": (i32, [vp, vp, sz, i32, i32, vp]),
        "p2p_group_start": (i32, []),
        "p2p_group_end": (i32, []),
        "p2p_get_error_string": (ctypes.c_char_p, [i32]),
        "p2p_get_last_error": (ctypes.c_char_p, []),
        "p2p_get_async_error": (i32, [ctypes.POINTER(i32)]),
        "p2p_set_option": (i32, [ctypes.c_char_p, ll]),
        "p2p_get_option": (i32, [ctypes.c_char_p, ctypes.POINTER(ll)]),
        "p2p_query_plan": (i32, [sz, i32, ctypes.POINTER(i32), ctypes.POINTER(i32)]),
        "p2p_pattern_fill": (i32, [vp, sz, u64, i32, u64, vp]),
        "p2p_pattern_check": (i32, [vp, sz, u64, i32, u64, vp, pu64, pu64, pu64]),
        "p2p_digest": (i32, [vp, sz, vp, pu64]),
        "p2p_digest_noalloc": (i32, [vp, sz, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != SUCCESS:
        L = lib()
        raise P2PError(rc, L.p2p_get_error_string(rc).decode(), L.p2p_get_last_error().decode())


Allgather = Callable[[bytes], List[bytes]]


def _single(blob: bytes) -> List[bytes]:
    return [blob]


def torch_allgather(group=None) -> Allgather:
    """An all-gather of byte strings over torch.distributed (any backend: objects go via CPU)."""
    import torch.distributed as dist

    def gather(blob: bytes) -> List[bytes]:
        out: List[Optional[bytes]] = [None] * dist.get_world_size(group)
        dist.all_gather_object(out, blob, group=group)
        return [bytes(b) for b in out]  # type: ignore[arg-type]

    return gather


class DeviceBuffer:
    """Device memory from p2p_mem_alloc; exposes __cuda_array_interface__ so torch can view it."""

    def __init__(self, nbytes: int):
        p = ctypes.c_void_p()
        check(lib().p2p_mem_alloc(ctypes.byref(p), nbytes))
        self.ptr = p.value or 0
        self.nbytes = nbytes
        self._freed = False

    @property
    def __cuda_array_interface__(self):
        return {"shape": (self.nbytes,), "typestr": "|u1", "data": (self.ptr, False), "version": 3, "strides": None}

    def tensor(self):
        import torch
        return torch.as_tensor(self, device="cuda")

    def free(self) -> None:
        if not self._freed and self.ptr:
            lib().p2p_mem_free(ctypes.c_void_p(self.ptr))
            self._freed = True


def init(rank: int, nranks: int, device: int, allgather: Allgather = _single) -> None:
    """p2p_init + blob exchange + p2p_connect (ncclCommInitRank's role, p2p_matrix.cc:115-120)."""
    L = lib()
    check(L.p2p_init(rank, nranks, device))
    n = L.p2p_blob_bytes()
    mine = ctypes.create_string_buffer(n)
    check(L.p2p_export(mine))
    blobs = allgather(mine.raw)
    if len(blobs) != nranks:
        raise P2PError(INVALID_ARGUMENT, "bootstrap", f"allgather returned {len(blobs)} blobs for {nranks} ranks")
    joined = ctypes.create_string_buffer(b"".join(blobs), n * nranks)
    check(L.p2p_connect(joined))


def finalize(barrier: Optional[Callable[[], None]] = None) -> None:
    """p2p_disconnect, a barrier over all ranks (pass one when nranks > 1), p2p_finalize: every rank closes
    its mappings of peer memory before any rank frees what it exported."""
    L = lib()
    rc = L.p2p_disconnect()
    if barrier is not None:
        barrier()
    frc = L.p2p_finalize()
    check(rc)
    check(frc)


def register(ptr: int, nbytes: int, allgather: Allgather = _single) -> int:
    """Collective window registration (ncclCommWindowRegister's role).  Returns the window id."""
    L = lib()
    n = L.p2p_register_blob_bytes()
    mine = ctypes.create_string_buffer(n)
    check(L.p2p_register_export(ctypes.c_void_p(ptr), nbytes, mine))
    blobs = allgather(mine.raw)
    joined = ctypes.create_string_buffer(b"".join(blobs), n * len(blobs))
    win = ctypes.c_int(-1)
    check(L.p2p_register_import(joined, ctypes.byref(win)))
    return win.value


def window_peer_ptr(window: int, peer: int) -> int:
    """Address at which `peer`'s part of the window is mapped in this process (0: that rank contributed nothing)."""
    p = ctypes.c_void_p()
    check(lib().p2p_window_peer_ptr(window, peer, ctypes.byref(p)))
    return p.value or 0


def deregister(window: int) -> None:
    check(lib().p2p_deregister(window))


def _stream_ptr(stream) -> ctypes.c_void_p:
    if stream is None:
        return ctypes.c_void_p(0)
    if hasattr(stream, "cuda_stream"):
        return ctypes.c_void_p(stream.cuda_stream)
    return ctypes.c_void_p(int(stream))


def send(ptr: int, count: int, dtype: int, peer: int, stream=None) -> None:
    check(lib().p2p_send(ctypes.c_void_p(ptr), count, dtype, peer, _stream_ptr(stream)))


def recv(ptr: int, count: int, dtype: int, peer: int, stream=None) -> None:
    check(lib().p2p_recv(ctypes.c_void_p(ptr), count, dtype, peer, _stream_ptr(stream)))


def alltoall(send_ptr: int, recv_ptr: int, count: int, dtype: int, stream=None) -> None:
    """Block p of the send buffer to rank p, block p of the receive buffer from rank p; one fused launch."""
    check(lib().p2p_alltoall(ctypes.c_void_p(send_ptr), ctypes.c_void_p(recv_ptr), count, dtype, _stream_ptr(stream)))


class SymmBuffer:
    """A symmetric window: `nbytes` on every rank from the driver's virtual memory API, peers mapped, bound to the
    NVSwitch multicast object when there is one.  Collective."""

    def __init__(self, nbytes: int, allgather: Allgather = _single):
        L = lib()
        n = L.p2p_symm_blob_bytes()
        mine = ctypes.create_string_buffer(n)
        check(L.p2p_symm_alloc_export(nbytes, mine))
        blobs = allgather(mine.raw)
        joined = ctypes.create_string_buffer(b"".join(blobs), n * len(blobs))
        p, win = ctypes.c_void_p(), ctypes.c_int(-1)
        check(L.p2p_symm_alloc_import(joined, ctypes.byref(p), ctypes.byref(win)))
        self.ptr, self.nbytes, self.window = p.value or 0, nbytes, win.value
        mc = ctypes.c_void_p()
        check(L.p2p_symm_multicast_ptr(self.window, ctypes.byref(mc)))
        self.multicast_ptr = mc.value or 0

    @property
    def __cuda_array_interface__(self):
        return {"shape": (self.nbytes,), "typestr": "|u1", "data": (self.ptr, False), "version": 3, "strides": None}

    def tensor(self):
        import torch
        return torch.as_tensor(self, device="cuda")

    def free(self) -> None:
        if self.window >= 0:
            check(lib().p2p_symm_free(self.window))
            self.window = -1


def bcast(send_ptr: int, recv_ptr: int, count: int, dtype: int, root: int, stream=None) -> None:
    """ncclBroadcast's argument order; recv_ptr lies in a SymmBuffer at the same offset on every rank."""
    check(lib().p2p_bcast(ctypes.c_void_p(send_ptr), ctypes.c_void_p(recv_ptr), count, dtype, root, _stream_ptr(stream)))


def group_start() -> None:
    check(lib().p2p_group_start())


def group_end() -> None:
    check(lib().p2p_group_end())


class group:
    """with p2p.group(): ... — ncclGroupStart/ncclGroupEnd (p2p_matrix.cc:212,227)."""

    def __enter__(self):
        group_start()
        return self

    def __exit__(self, exc_type, exc, tb):
        if exc_type is None:
            group_end()
        else:
            try:
                group_end()
            except P2PError:
                pass
        return False


def set_option(name: str, value: int) -> None:
    check(lib().p2p_set_option(name.encode(), value))


def get_option(name: str) -> int:
    v = ctypes.c_longlong(0)
    check(lib().p2p_get_option(name.encode(), ctypes.byref(v)))
    return v.value


def query_plan(nbytes: int, peer: int):
    path, nb = ctypes.c_int(0), ctypes.c_int(0)
    check(lib().p2p_query_plan(nbytes, peer, ctypes.byref(path), ctypes.byref(nb)))
    return path.value, nb.value


def async_error() -> int:
    r = ctypes.c_int(0)
    check(lib().p2p_get_async_error(ctypes.byref(r)))
    return r.value


def raise_async_error() -> None:
    r = async_error()
    if r != SUCCESS:
        L = lib()
        raise P2PError(r, L.p2p_get_error_string(r).decode(), L.p2p_get_last_error().decode())


def pattern_fill(ptr: int, nbytes: int, seed: int, rank: int, it: int, stream=None) -> None:
    check(lib().p2p_pattern_fill(ctypes.c_void_p(ptr), nbytes, seed, rank, it, _stream_ptr(stream)))


def pattern_check(ptr: int, nbytes: int, seed: int, rank: int, it: int, stream=None):
    """(mismatching bytes, offset of the first one, digest of the buffer); synchronises the stream."""
    bad, first, dig = ctypes.c_uint64(0), ctypes.c_uint64(0), ctypes.c_uint64(0)
    check(lib().p2p_pattern_check(ctypes.c_void_p(ptr), nbytes, seed, rank, it, _stream_ptr(stream),
                                  ctypes.byref(bad), ctypes.byref(first), ctypes.byref(dig)))
    return bad.value, first.value, dig.value


def digest(ptr: int, nbytes: int, stream=None) -> int:
    dig = ctypes.c_uint64(0)
    check(lib().p2p_digest(ctypes.c_void_p(ptr), nbytes, _stream_ptr(stream), ctypes.byref(dig)))
    return dig.value


def digest_noalloc(ptr: int, nbytes: int, out_dev_ptr: int, stream=None) -> None:
    """p2p_digest_noalloc: asynchronous, result (8 bytes) lands in device memory at out_dev_ptr."""
    check(lib().p2p_digest_noalloc(ctypes.c_void_p(ptr), nbytes, _stream_ptr(stream), ctypes.c_void_p(out_dev_ptr)))
