"""The Makefile prefers a real MPI when one is installed (first half).
No MPI exists in this image, so the test plants an `mpicxx` that answers the OpenMPI wrapper
queries (`--showme:compile`, `--showme:link`) and checks that a fresh build picks it up instead
of the in-repo shim.  (The planted wrapper points at the shim's own header and library, which is
all there is to link against here.)"""
import os
import shutil
import stat
import subprocess

import pytest

from conftest import ROOT


@pytest.mark.skipif(shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"), reason="no nvcc")
def test_makefile_uses_mpicxx_when_present(tmp_path):
    clone = tmp_path / "repo"
    subprocess.run(["git", "clone", "-q", ROOT, str(clone)], check=True)
    # a pre-built shim somewhere else plays the part of the installed MPI
    mpi_home = tmp_path / "fakempi"
    shutil.copytree(os.path.join(ROOT, "third_party", "minimpi"), mpi_home)
    subprocess.run(["make", "-C", str(mpi_home)], check=True, stdout=subprocess.DEVNULL)
    bindir = tmp_path / "bin"
    bindir.mkdir()
    wrapper = bindir / "mpicxx"
    wrapper.write_text(f"""#!/bin/sh
case "$1" in
  --showme:compile) echo "-I{mpi_home}" ;;
  --showme:link) echo "-L{mpi_home} -lmpi -lpthread" ;;
  *) exec g++ "$@" ;;
esac
""")
    wrapper.chmod(wrapper.stat().st_mode | stat.S_IEXEC)
    shutil.copy(mpi_home / "bin" / "mpirun", bindir / "mpirun")
    env = dict(os.environ, PATH=f"{bindir}:{os.environ['PATH']}")
    proc = subprocess.run(["make", "-j4", "all"], cwd=clone, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert proc.returncode == 0, proc.stdout[-3000:]
    assert f"built with: system MPI ({wrapper})" in proc.stdout
    assert f"-I{mpi_home}" in proc.stdout and "-Ithird_party/minimpi" not in proc.stdout
    assert not (clone / "third_party" / "minimpi" / "libmpi.a").exists(), "the shim must not be built when an MPI is installed"
    assert (clone / "p2p_matrix").exists()
    assert os.path.realpath(clone / "build" / "mpirun") == os.path.realpath(bindir / "mpirun")
    # and the program starts under that launcher (it stops at the missing GPU, as it should here)
    run = subprocess.run([str(clone / "build" / "mpirun"), "-n", "2", str(clone / "p2p_matrix"), "--help"], env=env,
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
    assert run.returncode == 0 and "usage: mpirun -n N" in run.stdout
