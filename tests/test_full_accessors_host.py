"""The full-metric accessors of somar_b200/csrc/stencil_full.cuh are __host__ __device__: tests/cuda/full_accessors_host.cu
runs them on the CPU.  ExView (ex_slow) is the accessor whose values were established against the oracle on the GPU; the
kind-dispatched ones the rim kernels and the persistent bottom kernel use (FastView on CELL_FAST cells, DirView<D> on
CELL_DIR<D> cells, TabView on any cell) and the statements built on them must give the same bits, cell by cell, over box
layouts with interfaces in every direction, boxes of one and two cells, Neumann / Dirichlet / periodic mixes and the
y-slab of a middle or last rank.  No GPU needed: this is host code compiled by nvcc."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _nvcc():
    return shutil.which("nvcc") or ("/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else None)


@pytest.mark.skipif(_nvcc() is None, reason="nvcc not found")
def test_kind_dispatched_accessors_equal_exview_on_the_host(tmp_path):
    exe = tmp_path / "full_accessors_host"
    cmd = [_nvcc(), "-O1", "-std=c++17", "-fmad=false", "-gencode", "arch=compute_100a,code=sm_100a",
           "-I", os.path.join(ROOT, "somar_b200", "csrc"), "-o", str(exe),
           os.path.join(ROOT, "tests", "cuda", "full_accessors_host.cu")]
    b = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert b.returncode == 0, b.stdout + b.stderr
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    last = r.stdout.strip().splitlines()[-1]
    n, bad = int(last.split()[0]), int(last.split()[2])
    assert n > 300000 and bad == 0, last
    # every kind occurs somewhere in the set
    tot = [0, 0, 0, 0, 0]
    for line in r.stdout.splitlines()[:-1]:
        f = line.split("fast")[1].replace("dir", " ").replace("general", " ").replace("/", " ").split()
        for q in range(5):
            tot[q] += int(f[q])
    assert all(t > 0 for t in tot), tot
