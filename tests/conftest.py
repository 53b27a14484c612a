"""Shared fixtures.  `-m "not gpu"` runs everything that needs no GPU (oracle vs golden vectors,
MPI shim, C-ABI surface, host logic, gloo world-2); `-m gpu` runs the parity tests proper,
which call the CUDA path through the C ABI and compare with the oracle."""
from __future__ import annotations

import importlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
ORACLE_DIR = os.path.join(ROOT, "oracle")
if ORACLE_DIR not in sys.path:
    sys.path.insert(0, ORACLE_DIR)

MPIRUN = os.path.join(ROOT, "build", "mpirun")
P2P_MATRIX = os.path.join(ROOT, "p2p_matrix")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def _have(path: str) -> bool:
    return os.path.exists(path)


@pytest.fixture(scope="session")
def built():
    """Make sure the native side exists (builds it when a toolchain is here, e.g. in the CPU container)."""
    needed = [os.path.join(ROOT, "test-nccl-p2p_b200", "lib", "libb200p2p.so"), P2P_MATRIX, MPIRUN,
              os.path.join(REF_DIR, "libp2poracle.so"), os.path.join(REF_DIR, "mpi_host_p2p"),
              os.path.join(ROOT, "build", "minimpi_selftest"), os.path.join(ROOT, "build", "driver_unit"),
              os.path.join(ROOT, "build", "mpi_glue_nogpu"), os.path.join(ROOT, "build", "fdpass_unit")]
    if not all(_have(p) for p in needed):
        sys.path.insert(0, ROOT)
        entry = importlib.import_module("__graft_entry__")
        entry.build()
    missing = [p for p in needed if not _have(p)]
    assert not missing, f"build did not produce {missing}"
    return True


@pytest.fixture(scope="session")
def p2p(built):
    return importlib.import_module("test-nccl-p2p_b200")


@pytest.fixture(scope="session")
def oracle(built):
    import pattern
    pattern.load_c()
    return pattern


def run(cmd, timeout=300, env=None, check=True):
    e = dict(os.environ)
    if env:
        e.update(env)
    proc = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout, env=e)
    if check and proc.returncode != 0:
        raise AssertionError(f"{' '.join(map(str, cmd))} exited {proc.returncode}:\n{proc.stdout[-4000:]}")
    return proc


def gpu_count() -> int:
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0
