"""full-metric MAC gradient / divergence on y-slab boxes: GPU (one rank) vs oracle"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from common import Problem
from somar_b200 import capi
from oracle import pyoracle as po
for parts in [(1, 8, 1), (1, 4, 1), (1, 2, 1), (1, 1, 1)]:
    pb = Problem(n=(32, 64, 16), map_name="terrain", aniso=4.0, parts=parts)
    orc = pb.oracle(po); lev = pb.level(capi)
    rng = np.random.default_rng(3)
    phi = rng.uniform(-1, 1, size=pb.shape)
    fp = lev.field(0, phi)
    gr = [lev.field(0) for _ in range(3)]
    lev.mac_gradient(gr, fp, homogeneous=False)
    fd = lev.field(0)
    lev.mac_divergence(fd, gr, divide_by=0.5)
    g_o = orc.mac_gradient(phi, homogeneous=False)
    d_o = orc.mac_divergence(g_o) / 0.5
    d = fd.get()
    err = np.abs(d - d_o)
    kk, jj, ii = np.unravel_index(err.argmax(), err.shape)
    print(parts, "max |div - oracle|", err.max(), "at (i,j,k)", (ii, jj, kk), "apply vs oracle apply:",
          np.abs(d_o * 0.5 - orc.apply_op(phi, homogeneous=False)).max())
    for dd in range(3):
        e = np.abs(gr[dd].get_face(dd) - g_o[dd])
        print("   grad", dd, e.max(), np.unravel_index(e.argmax(), e.shape)[::-1])
    lev.destroy()
