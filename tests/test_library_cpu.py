"""CPU checks of the product library: it loads, exports every symbol of include/somar_b200.h, and refuses to
compute without a GPU (no CPU fallback)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_every_declared_symbol(somar):
    hdr = open(os.path.join(ROOT, "include", "somar_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(somar_[a-z0-9_]+)\s*\(", hdr)))
    assert declared, "no declarations found"
    L = somar.lib()
    missing = [s for s in declared if not hasattr(L, s)]
    assert not missing, missing
    assert sorted(somar.SYMBOLS) == declared


def test_default_params_match_reference_defaults(somar):
    p = somar.default_params()   # src/utils/ProblemContext.cpp:1147-1216
    assert (p.eps, p.num_smooth_down, p.num_smooth_up, p.num_smooth_bottom, p.numMG, p.imin, p.imax) == \
        (1e-6, 2, 2, 2, 1, 5, 20)
    assert (p.relaxMode, p.bottom_imax, p.bottom_numRestarts, p.bottom_eps, p.bottom_reps) == (1, 80, 5, 1e-6, 1e-12)


def test_no_cpu_fallback(somar):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(somar.SomarError):
        somar.Level((16, 16, 16), (1.0, 1.0, 1.0))


def test_product_never_imports_oracle():
    bad = []
    for dp, _, fs in os.walk(os.path.join(ROOT, "somar_b200")):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", ".cpp")):
                src = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"(from|import)\s+oracle|#include\s+\"[^\"]*oracle", src):
                    bad.append(f)
    assert not bad, bad
