"""The in-repo MPI shim and everything launched through it that needs no GPU.

The image has no MPI, yet the reference is an MPI program (`mpirun -n N p2p_matrix`,
reference README.md:5).  These tests run the shim at N = 1, 2, 4, 8, check that a dying rank
takes the job down, that the UNMODIFIED reference links against it with exactly the six MPI
symbols it is known to need, and run the host-memory baseline (the CPU statement of the path).
"""
import json
import os
import subprocess

import pytest

from conftest import MPIRUN, REF_DIR, ROOT, run

SELFTEST = os.path.join(ROOT, "build", "minimpi_selftest")


@pytest.mark.parametrize("n", [1, 2, 4, 8])
def test_selftest(built, n):
    out = run([MPIRUN, "-n", str(n), SELFTEST], timeout=300).stdout
    assert f"OK {n}" in out


def test_singleton_without_launcher(built):
    assert "OK 1" in run([SELFTEST]).stdout


def test_openmpi_style_flags_are_accepted(built):
    out = run([MPIRUN, "--allow-run-as-root", "--oversubscribe", "--bind-to", "none", "-x", "FOO=bar", "-np", "2", SELFTEST]).stdout
    assert "OK 2" in out


def test_failing_rank_stops_the_job(built):
    proc = run([MPIRUN, "-n", "4", SELFTEST, "--fail-rank", "2"], timeout=60, check=False)
    assert proc.returncode == 7
    assert "rank 2 exited with code 7" in proc.stdout


def test_launcher_finds_program_in_cwd(built):
    # `mpirun -n 1 p2p_matrix` (no ./) is how the reference README launches it
    proc = subprocess.run([MPIRUN, "-n", "1", "minimpi_selftest"], cwd=os.path.join(ROOT, "build"),
                          stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
    assert proc.returncode == 0 and "OK 1" in proc.stdout


@pytest.mark.skipif(not os.path.exists(os.path.join(REF_DIR, "p2p_matrix_ref")), reason="reference binary not built (no /root/reference)")
def test_reference_links_with_six_mpi_symbols(built):
    """the unmodified reference needs exactly these from MPI."""
    dis = run(["objdump", "-d", "--no-show-raw-insn", os.path.join(REF_DIR, "p2p_matrix_ref")]).stdout
    # calls made by the reference's own code: main and its (C++-mangled) static helpers
    calls, current = set(), ""
    for line in dis.splitlines():
        if line.endswith(">:"):
            current = line.split("<")[-1].rstrip(">:")
        elif "call" in line and "<MPI_" in line and (current == "main" or current.startswith("_Z")):
            calls.add(line.split("<")[-1].split(">")[0].split("@")[0])
    assert calls == {"MPI_Init_thread", "MPI_Comm_rank", "MPI_Comm_size", "MPI_Allgather", "MPI_Bcast", "MPI_Barrier"}


def test_mpi_host_baseline_runs_and_verifies(built, tmp_path):
    js = tmp_path / "host.json"
    out = run([MPIRUN, "-n", "2", os.path.join(REF_DIR, "mpi_host_p2p"), "--size", str(3 * 1024 * 1024 + 5), "--iters", "5",
               "--verify", "--json", str(js)], timeout=120).stdout
    assert "verify: OK" in out
    assert "Uni-Directional MPI host-memory P2P Bandwidth (Gbps) [minimpi shm" in out
    recs = [json.loads(l) for l in js.read_text().splitlines()]
    assert len(recs) == 4 and {r["mode"] for r in recs} == {"uni", "bidir"}
    assert all(r["GBps"] > 0 and r["transport"] == "minimpi shm" for r in recs)


def test_mpi_host_baseline_four_ranks(built):
    out = run([MPIRUN, "-n", "4", os.path.join(REF_DIR, "mpi_host_p2p"), "--size", "65536", "--iters", "3", "--verify", "--mode", "uni"],
              timeout=120).stdout
    assert "verify: OK" in out
    rows = [l for l in out.splitlines() if l[:7].strip().isdigit()]
    assert len(rows) == 4 and all(len(r.split()) == 5 for r in rows)
