"""The C++ host side above the C ABI (include/somar_b200/adapter.hpp + somar_b200/host/pressure_demo.cpp):
mirrors AMRPressureSolver::levelDefine/levelSolve and the MappedAMRPoissonOp method set.  CPU: it builds with plain
g++ and fails loudly without a GPU (no fallback).  GPU: it solves a discrete manufactured problem held in host
"LevelData" FABs (4 boxes with ghost cells) and recovers the solution."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "somar_b200", "host")
DEMO = os.path.join(HOST, "pressure_demo")


def _build():
    r = subprocess.run(["make", "-C", HOST], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert os.path.exists(DEMO)


def test_adapter_builds_and_has_no_cpu_fallback():
    _build()
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present: the fail-loudly path cannot be observed")
    r = subprocess.run([DEMO, "16", "16", "8"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 2, r.stdout + r.stderr          # somar_b200::Error, not a silent CPU solve
    assert "DEMO-OK" not in r.stdout
    assert "error" in r.stderr.lower()


@pytest.mark.gpu
@pytest.mark.parametrize("args", [["64", "64", "32", "1"], ["64", "64", "32", "3"], ["48", "96", "16", "1"],
                                  ["64", "64", "16", "1", "leptic"]])      # adapter's LevelLepticSolver on a thin domain
def test_pressure_demo_recovers_manufactured_solution(args):
    if not os.path.exists(DEMO):
        _build()
    r = subprocess.run([DEMO] + args, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "DEMO-OK" in r.stdout, r.stdout + r.stderr
