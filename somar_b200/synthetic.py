"""Synthetic inputs for the mapped-Poisson hot path: coordinate maps -> J g^{ij}, 1/J, rhs.

Input synthesis only -- nothing here is on the timed path.  Everything is float64 torch (CPU or CUDA);
array convention is the Chombo/Fortran one seen from C: shape (nz, ny, nx), x fastest.  Face-centred
arrays of direction d have one extra point along d.

The metric is built the way the reference builds it (src/geometry/GeoSourceInterface.cpp:65-120,417-455):
the Jacobian matrix dx^mu/dXi^nu at a cell or face centre is a centred difference of the map x(Xi) over
+-1/2 dXi^nu; J = det, dXi/dx = inverse, J g^{mu nu} = J sum_rho dXi^mu/dx^rho dXi^nu/dx^rho.

Maps:
  * ``stretched``  -- family S of SURVEY.md 8(d): x=xi, y=eta, z = H s(zeta/H),
                      s(f) = 1 - tanh(sigma (1-f)) / tanh(sigma)  (profile sketched at
                      src/geometry/maps/BathymetricBaseMapF.ChF:38-42).  Diagonal metric.
  * ``terrain``    -- family T: z = d(x,y) + (H-d) zeta/H (BathymetricBaseMapF.ChF:89-120) over a Gaussian
                      seamount d = 0.3 H exp(-r^2/(0.1 Lx)^2).  Full metric.
  * ``twisted``    -- src/geometry/maps/TwistedMapF.ChF:28-95 (TaylorGreen deck, pert = 0.05).
  * ``cartesian``  -- identity (src/geometry/maps/CartesianMapF.ChF).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Callable, List, Optional, Sequence, Tuple

import torch

DT = torch.float64


@dataclass
class GridSpec:
    """A single-level problem description (domain, computational spacing, map, BCs)."""
    n: Tuple[int, int, int]                 # cells (nx, ny, nz); nz == 1 for 2-D
    L: Tuple[float, float, float]           # computational extent; dXi = L/n
    dim: int = 3
    periodic: Tuple[bool, bool, bool] = (False, False, False)
    map_name: str = "stretched"
    map_params: dict = field(default_factory=dict)

    @property
    def dx(self) -> Tuple[float, float, float]:
        return tuple(self.L[d] / self.n[d] if d < self.dim else 1.0 for d in range(3))

    @property
    def is_diagonal(self) -> bool:
        return self.map_name in ("stretched", "cartesian")


def headline_grid(n=(1024, 1024, 256), aniso=16.0, periodic=(False, False, False), map_name="stretched") -> GridSpec:
    """L = (nx*D, ny*D, nz*D/A) with D = 1, anisotropy A = dx/dz."""
    nx, ny, nz = n
    return GridSpec(n=n, L=(float(nx), float(ny), float(nz) / aniso), dim=3, periodic=periodic, map_name=map_name,
                    map_params={"sigma": 1.2})


# ------------------------------------------------------------------------------------------------
# Maps: (X0, X1, X2 tensors of computational coordinates) -> tuple of physical coordinates
# ------------------------------------------------------------------------------------------------
def _map_fn(spec: GridSpec) -> Callable:
    L = spec.L
    vd = spec.dim - 1  # vertical direction
    H = L[vd]
    if spec.map_name == "cartesian":
        return lambda X: tuple(X)
    if spec.map_name == "stretched":
        sigma = float(spec.map_params.get("sigma", 1.2))
        th = math.tanh(sigma)

        def f(X):
            out = list(X)
            out[vd] = H * (1.0 - torch.tanh(sigma * (1.0 - X[vd] / H)) / th)
            return tuple(out)
        return f
    if spec.map_name == "terrain":
        amp = float(spec.map_params.get("amp", 0.3))
        width = float(spec.map_params.get("width", 0.1))

        def f(X):
            out = list(X)
            r2 = (X[0] - 0.5 * L[0]) ** 2
            if spec.dim == 3:
                r2 = r2 + (X[1] - 0.5 * L[1]) ** 2
            d = amp * H * torch.exp(-r2 / (width * L[0]) ** 2)
            out[vd] = d + (H - d) * X[vd] / H
            return tuple(out)
        return f
    if spec.map_name == "twisted":
        pert = float(spec.map_params.get("pert", 0.05))
        tp = 2.0 * math.pi

        def f(X):
            s = [torch.sin(tp * X[d] / L[d]) if d < spec.dim else None for d in range(3)]
            if spec.dim == 2:
                return (X[0] + pert * s[1], X[1] + pert * s[0], X[2])
            return (X[0] + pert * s[1] * s[2], X[1] + pert * s[2] * s[0], X[2] + pert * s[0] * s[1])
        return f
    raise ValueError(f"unknown map {spec.map_name}")


def _coords(spec: GridSpec, stagger: Sequence[int], device, yrange=None) -> List[torch.Tensor]:
    """1-D computational coordinates of the points of a box with centring `stagger` (1 = node).
    yrange = (j0, j1) restricts direction 1 to the cells j0..j1-1 (a rank-local y-slab)."""
    out = []
    for d in range(3):
        lo, hi = (0, spec.n[d]) if (d != 1 or yrange is None) else yrange
        n = hi - lo + (1 if stagger[d] else 0)
        off = 0.0 if stagger[d] else 0.5
        out.append((torch.arange(n, dtype=DT, device=device) + (lo + off)) * spec.dx[d])
    return out


def _jacobian(spec: GridSpec, stagger: Sequence[int], device, yrange=None):
    """dx^mu/dXi^nu (list of lists of broadcastable tensors, shape (nz,ny,nx)-compatible) at the given centring."""
    fmap = _map_fn(spec)
    c = _coords(spec, stagger, device, yrange)
    X = [c[0].view(1, 1, -1), c[1].view(1, -1, 1), c[2].view(-1, 1, 1)]
    dim = spec.dim
    Jm = [[None] * dim for _ in range(dim)]
    for nu in range(dim):
        h = 0.5 * spec.dx[nu]
        Xp = list(X)
        Xm = list(X)
        Xp[nu] = X[nu] + h
        Xm[nu] = X[nu] - h
        xp = fmap(Xp)
        xm = fmap(Xm)
        for mu in range(dim):
            d = (xp[mu] - xm[mu]) / spec.dx[nu]
            Jm[mu][nu] = d
    return Jm


def _det_inv(Jm, dim):
    if dim == 2:
        a, b, c, d = Jm[0][0], Jm[0][1], Jm[1][0], Jm[1][1]
        det = a * d - b * c
        inv = [[d / det, -b / det], [-c / det, a / det]]
        return det, inv
    (a, b, c), (d, e, f), (g, h, i) = Jm
    A = e * i - f * h
    B = -(d * i - f * g)
    C = d * h - e * g
    det = a * A + b * B + c * C
    inv = [[A / det, -(b * i - c * h) / det, (b * f - c * e) / det],
           [B / det, (a * i - c * g) / det, -(a * f - c * d) / det],
           [C / det, -(a * h - b * g) / det, (a * e - b * d) / det]]
    return det, inv


@dataclass
class Metric:
    """jgup[d]: tensor (ncomp, nz(+1), ny(+1), nx(+1)) face-centred in d; ncomp = dim (full) or 1 (J g^{dd} only)."""
    jgup: List[torch.Tensor]
    jinv: torch.Tensor
    ncomp: int


def build_metric(spec: GridSpec, device="cpu", diag_only: Optional[bool] = None, yrange=None) -> Metric:
    """J g^{ij} on faces and 1/J at cell centres.  diag_only stores only J g^{dd} (diagonal metrics).
    yrange = (j0, j1): only the y-slab of cells j0..j1-1 (what one rank owns)."""
    dim = spec.dim
    if diag_only is None:
        diag_only = spec.is_diagonal
    ny = spec.n[1] if yrange is None else yrange[1] - yrange[0]
    shape_c = (spec.n[2], ny, spec.n[0])
    det, _ = _det_inv(_jacobian(spec, (0, 0, 0), device, yrange), dim)
    jinv = (1.0 / det).expand(shape_c).contiguous()
    jg = []
    for d in range(dim):
        st = [0, 0, 0]
        st[d] = 1
        shape_f = (spec.n[2] + st[2], ny + st[1], spec.n[0] + st[0])
        det, inv = _det_inv(_jacobian(spec, st, device, yrange), dim)
        comps = []
        for c in range(dim):
            if diag_only and c != d:
                continue
            g = None
            for rho in range(dim):
                t = inv[d][rho] * inv[c][rho]
                g = t if g is None else g + t
            comps.append((det * g).expand(shape_f))
        jg.append(torch.stack(comps, 0).contiguous())
    return Metric(jgup=jg, jinv=jinv, ncomp=(1 if diag_only else dim))


def random_rhs(spec: GridSpec, jinv: torch.Tensor, seed=12345, device="cpu") -> torch.Tensor:
    """Seeded uniform(-1,1) field projected to zero dV-weighted mean (dV = prod(dx)/Jinv): solvable for
    all-Neumann / periodic problems."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    r = torch.rand(jinv.shape, dtype=DT, device=device, generator=g) * 2.0 - 1.0
    w = 1.0 / jinv
    r -= (r * w).sum() / w.sum()
    return r


def manufactured_phi(spec: GridSpec, modes=(1, 1, 1), device="cpu") -> torch.Tensor:
    """phi* = prod_d cos(pi m_d Xi_d / L_d) at cell centres (Neumann-compatible in computational space)."""
    c = _coords(spec, (0, 0, 0), device)
    X = [c[0].view(1, 1, -1), c[1].view(1, -1, 1), c[2].view(-1, 1, 1)]
    out = torch.ones((spec.n[2], spec.n[1], spec.n[0]), dtype=DT, device=device)
    for d in range(spec.dim):
        k = 2.0 * modes[d] if spec.periodic[d] else float(modes[d])
        out = out * torch.cos(math.pi * k * X[d] / spec.L[d])
    return out


def split_boxes(n: Sequence[int], parts: Sequence[int]) -> List[Tuple[Tuple[int, int, int], Tuple[int, int, int]]]:
    """Cut the domain [0,n) into parts[0] x parts[1] x parts[2] equal boxes (lo, hi inclusive), x fastest."""
    boxes = []
    for kz in range(parts[2]):
        for ky in range(parts[1]):
            for kx in range(parts[0]):
                idx = (kx, ky, kz)
                lo, hi = [], []
                for d in range(3):
                    assert n[d] % parts[d] == 0, "domain not divisible by the requested partition"
                    w = n[d] // parts[d]
                    lo.append(idx[d] * w)
                    hi.append((idx[d] + 1) * w - 1)
                boxes.append((tuple(lo), tuple(hi)))
    return boxes
