"""Shared problem builders for the parity tests: the SAME seeded inputs go to the CPU oracle and to the CUDA path."""
from __future__ import annotations

import numpy as np

from somar_b200 import synthetic as syn


class Problem:
    def __init__(self, n, *, dim=3, map_name="stretched", aniso=4.0, periodic=(0, 0, 0), parts=(1, 1, 1), bc_type=None,
                 bc_value=None, alpha=0.0, beta=1.0, map_params=None, seed=7, L=None):
        self.n = tuple(n)
        self.dim = dim
        self.periodic = tuple(int(x) for x in periodic)
        if L is not None:   # explicit computational extent (the shipped decks' amr.length)
            L = tuple(float(x) for x in L) + (1.0,) * (3 - len(L))
        elif dim == 3:
            L = (float(n[0]), float(n[1]), float(n[2]) / aniso)
        else:
            L = (float(n[0]), float(n[1]) / aniso, 1.0)
        mp = {"sigma": 1.2}
        mp.update(map_params or {})
        self.spec = syn.GridSpec(n=self.n, L=L, dim=dim, periodic=tuple(bool(x) for x in periodic), map_name=map_name,
                                 map_params=mp)
        self.is_diagonal = self.spec.is_diagonal
        m = syn.build_metric(self.spec)
        self.jgup = [j.numpy() for j in m.jgup]
        if m.ncomp == 1:
            self.jgup = [np.ascontiguousarray(j[0]) for j in self.jgup]
        self.jinv = m.jinv.numpy()
        self.ncomp = m.ncomp
        self.boxes = syn.split_boxes(self.n, parts)
        self.bc_type = bc_type
        self.bc_value = bc_value
        self.alpha, self.beta = alpha, beta
        self.shape = (self.n[2], self.n[1], self.n[0])
        self.rng = np.random.default_rng(seed)

    def rand(self, shape=None):
        return self.rng.uniform(-1.0, 1.0, size=shape or self.shape)

    def rhs(self):
        return syn.random_rhs(self.spec, syn.build_metric(self.spec).jinv).numpy()

    def oracle(self, po, **params):
        p = po.default_params(**params)
        return po.Oracle(self.n, self.spec.dx, self.jgup, self.jinv, dim=self.dim, periodic=self.periodic,
                         boxes=self.boxes, alpha=self.alpha, beta=self.beta, is_diagonal=self.is_diagonal,
                         bc_type=self.bc_type, bc_value=self.bc_value, params=p)

    def level(self, capi, box_rank=None, rank=0, nranks=1, comm_id=None, device=0, **params):
        p = capi.default_params(**params)
        lev = capi.Level(self.n, self.spec.dx, dim=self.dim, periodic=self.periodic, boxes=self.boxes,
                         box_rank=box_rank, alpha=self.alpha, beta=self.beta, is_diagonal=self.is_diagonal,
                         bc_type=self.bc_type, bc_value=self.bc_value, params=p, device=device, rank=rank,
                         nranks=nranks, comm_id=comm_id)
        lev.set_metric_global(self.jgup, self.jinv)
        lev.finalize()
        return lev


def rel_err(a, b):
    a = np.asarray(a)
    b = np.asarray(b)
    den = max(np.abs(b).max(), 1e-300)
    return float(np.abs(a - b).max() / den)
