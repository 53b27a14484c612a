"""The C-ABI surface, checked without a GPU: the library loads, exports exactly what
include/p2p.h declares, refuses to work before initialisation, and — on a box without a
B200 — fails loudly instead of falling back to anything."""
import ctypes
import os
import re
import subprocess

import pytest

from conftest import MPIRUN, P2P_MATRIX, ROOT, gpu_count, run

HEADER = os.path.join(ROOT, "include", "p2p.h")


def declared_symbols():
    text = open(HEADER).read()
    return sorted(set(re.findall(r"P2P_API[^;()]*?\b(p2p_[a-z_0-9]+)\s*\(", text)))


def test_header_and_binding_agree(p2p):
    assert declared_symbols() == sorted(p2p.SYMBOLS)


def test_library_exports_every_declared_symbol(p2p):
    out = run(["nm", "-D", "--defined-only", p2p.LIB_PATH]).stdout
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    for sym in declared_symbols():
        assert sym in exported, f"{sym} declared in include/p2p.h but not exported"
    # nothing but the C ABI leaks out (hidden visibility for the C++ internals)
    stray = [s for s in exported if not s.startswith("p2p_") and not s.startswith("_fini") and not s.startswith("_init")]
    assert not stray, f"unexpected exported symbols: {stray[:5]}"
    L = p2p.lib()
    for sym in declared_symbols():
        assert hasattr(L, sym)


def test_every_entry_point_cites_the_reference():
    """Each declaration block in the header names the reference interface it replaces."""
    text = open(HEADER).read()
    assert text.count("p2p_matrix.cc:") >= 10
    assert "nccl.h:" in text


def test_error_strings_and_codes(p2p):
    L = p2p.lib()
    assert L.p2p_get_error_string(0) == b"no error"
    assert L.p2p_get_error_string(4) == b"invalid argument"   # ncclInvalidArgument = 4
    assert L.p2p_get_error_string(5) == b"invalid usage"      # ncclInvalidUsage = 5
    assert L.p2p_get_error_string(99) == b"unknown result code"
    assert L.p2p_blob_bytes() == 256 and L.p2p_register_blob_bytes() == 128


def test_calls_before_init_are_refused():
    """Run in a fresh process so no other test's engine state is visible."""
    code = r'''
import importlib, sys
sys.path.insert(0, %r)
p2p = importlib.import_module("test-nccl-p2p_b200")
L = p2p.lib()
import ctypes
assert L.p2p_send(None, 8, 0, 0, None) == p2p.INVALID_USAGE
assert b"p2p_init" in L.p2p_get_last_error()
assert L.p2p_recv(None, 8, 0, 0, None) == p2p.INVALID_USAGE
assert L.p2p_group_start() == p2p.INVALID_USAGE
assert L.p2p_rank() == -1 and L.p2p_nranks() == 0
assert L.p2p_init(0, 0, 0) == p2p.INVALID_ARGUMENT
assert L.p2p_init(3, 2, 0) == p2p.INVALID_ARGUMENT
assert L.p2p_init(0, 17, 0) == p2p.INVALID_ARGUMENT
print("refused")
''' % ROOT
    out = run(["python", "-c", code]).stdout
    assert "refused" in out


@pytest.mark.skipif(gpu_count() > 0, reason="this box has a GPU")
def test_no_gpu_means_loud_failure_not_fallback(p2p):
    code = r'''
import importlib, sys
sys.path.insert(0, %r)
p2p = importlib.import_module("test-nccl-p2p_b200")
try:
    p2p.init(0, 1, 0)
except p2p.P2PError as e:
    assert e.code == p2p.UNHANDLED_CUDA_ERROR, e.code
    assert "no CPU path" in str(e), str(e)
    print("loud")
''' % ROOT
    assert "loud" in run(["python", "-c", code]).stdout
    # the benchmark binary likewise stops with the reference's CUDACHECK message shape
    proc = run([MPIRUN, "-n", "1", P2P_MATRIX, "--size", "1M", "--iters", "1"], check=False, timeout=60)
    assert proc.returncode != 0
    assert "Failed: Cuda error" in proc.stdout


def test_missing_library_is_an_import_error(tmp_path):
    """The binding must not limp along without the .so."""
    code = r'''
import importlib, sys
sys.path.insert(0, %r)
p2p = importlib.import_module("test-nccl-p2p_b200")
p2p.LIB_PATH = %r
try:
    p2p.lib()
except ImportError as e:
    assert "no Python or CPU fallback" in str(e)
    print("import error")
''' % (ROOT, str(tmp_path / "nope.so"))
    assert "import error" in run(["python", "-c", code]).stdout


def test_product_does_not_link_the_oracle(p2p):
    """oracle/ is test infrastructure: neither the library nor p2p_matrix may contain it."""
    for binary in [p2p.LIB_PATH, P2P_MATRIX]:
        syms = run(["nm", "-C", binary]).stdout
        assert "oracle_" not in syms
    for root, _, files in os.walk(os.path.join(ROOT, "test-nccl-p2p_b200")):
        for f in files:
            if f.endswith((".cu", ".cuh", ".cc", ".hpp", ".h", ".py")):
                text = open(os.path.join(root, f)).read()
                assert '#include "pattern.h"' not in text and "import pattern" not in text, f


@pytest.mark.skipif(gpu_count() > 0, reason="this box has a GPU")
@pytest.mark.parametrize("n", [1, 2, 4])
def test_mpi_glue_fails_on_every_rank_without_hanging(built, n):
    """include/p2p_mpi.h: when the engine cannot start, all ranks learn it and leave the collectives together."""
    exe = os.path.join(ROOT, "build", "mpi_glue_nogpu")
    if not os.path.exists(exe):
        run(["make", "tests"])
    out = run([MPIRUN, "-n", str(n), exe], timeout=60).stdout
    lines = [l for l in out.splitlines() if l.startswith("rank ")]
    assert len(lines) == n and all("init=1 (unhandled cuda error)" in l for l in lines), out


def test_every_option_of_the_engine_is_documented_in_the_header():
    """The option table in csrc/p2p_engine.cu (one entry per name, environment variable and range) against the list
    of tuning knobs in include/p2p.h: an option a caller can set must be described where callers look."""
    import re
    src = open(os.path.join(ROOT, "test-nccl-p2p_b200", "csrc", "p2p_engine.cu")).read()
    table = src[src.index("const OptionSpec kOptionSpecs[]"):]
    table = table[:table.index("};")]
    specs = re.findall(r'\{"(\w+)",\s*"(P2P_\w+)"', table)
    assert len(specs) >= 25
    header = open(os.path.join(ROOT, "include", "p2p.h")).read()
    for name, env in specs:
        assert env == "P2P_" + name.upper(), (name, env)
        assert re.search(r"\b%s\b" % name, header), f"option '{name}' is missing from include/p2p.h"
