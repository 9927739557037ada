"""Writes tests/golden/refkernels_{2,3}d.npz: the outputs of the reference's own Chombo-Fortran kernels (executed through
oracle/chf2py.py from /root/reference) for every case of tests/refkernels.py.  Inputs are regenerated from seeds, so only
outputs are stored.  Run here (where /root/reference exists); the fixtures travel."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import refkernels as rk  # noqa: E402

for D in (2, 3):
    ref = rk.reference_module(D)
    assert ref is not None, "needs /root/reference"
    out = {}
    for name, fn in rk.CASES.items():
        _, out_r = fn(D, ref)
        for k, v in out_r.items():
            out[name + "/" + k] = np.asarray(v)
    path = os.path.join(ROOT, "tests", "golden", "refkernels_%dd.npz" % D)
    np.savez_compressed(path, **out)
    print(path, len(out), "arrays,", os.path.getsize(path) // 1024, "KiB")
