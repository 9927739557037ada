// projector.cu -- the calculus around the pressure solve, so that a whole
// BaseProjector::project (src/projection/BaseProjectorI.H:176-300, a_velIsFlux = true) stays on the device:
//   k_mac_div    Divergence::levelDivergenceMAC -> MAPPEDFLUXDIVERGENCE{2,3}D
//                src/calculus/DivCurlGrad/Divergence.cpp:44-127, DivCurlGradF.ChF:1034-1214   (+ the "/= dt" pass)
//   k_mac_grad   Gradient::levelGradientMAC (normal components) -> singleBoxMacGrad -> MAPPEDMACGRAD[ORTHO]
//                Gradient.cpp:85-170, 946-1100, DivCurlGradF.ChF:87-262; BC ghosts + BC fluxes as the operator's
//                optionally fused with LevelMACProjector::applyCorrection (LevelMACProjector.cpp:221-240)
//
// Face fields use the brick of the cell fields (geom.cuh): face (i,j,k) of direction d is the LOW face of cell
// (i,j,k); the high face of the last cell of a row/column sits in the first ghost slot.
//
// J g^{dj} d_j phi on a face is exactly the flux MAPPEDGETFLUX[ORTHO] forms inside applyOp (same extrapolated
// copy for the transverse terms, same BC ghosts, same boundary-flux overwrite; with beta = 1 the statements of
// MAPPEDMACGRAD and MAPPEDGETFLUX differ only by the order of one commutative product), so the kernel reuses
// face_flux_ortho / face_flux_full: div(grad phi) == L phi bit for bit, which tests/ check on both sides.
// HBM-bound: gradient 8 (phi) + 8 ND (J g^{dd}; 8 ND^2 full) + 8 ND (out) B/cell, +8 ND when it accumulates into
// the velocity; divergence 8 ND + 8 (Jinv) + 8 (out) B/cell.
#include "level.hpp"
#include "stencil_ortho.cuh"
#include "stencil_full.cuh"

namespace somar {

#define PBX 64
#define PBY 4

struct FaceOut {
    double* p[3];
};

// store: out = v, or (accumulate) out = out + v * scale   [FArrayBox::plus(src, scale)]
__device__ __forceinline__ void face_store(double* __restrict__ out, long long f, double v, double scale, int acc) {
    if (acc) out[f] = out[f] + v * scale;
    else out[f] = v;
}

template <int ND, bool FULL>
__global__ void __launch_bounds__(PBX* PBY) k_mac_grad(const Geom g, const Metric m, FaceOut o,
                                                       const double* __restrict__ phi, bool homog, double scale,
                                                       int acc, int kz) {
    const int i = blockIdx.x * PBX + threadIdx.x;
    const int j = blockIdx.y * PBY + threadIdx.y;
    if (i >= g.n[0] || j >= g.n[1]) return;
    const int k0 = blockIdx.z * kz;
    const int k1 = min(k0 + kz, g.n[2]);
    for (int k = k0; k < k1; ++k) {
        const long long c = g.idx(i, j, k);
        const unsigned sm = side_mask(g, i, j, k);
        const double pc = phi[c];
        if (FULL) {
            const ExView<ND> v(g, phi, i, j, k, EXMODE_OP);
            face_store(o.p[0], c, face_flux_full<ND, 0, 0>(g, m, v, phi, i, j, k, c, sm & 1u, homog, pc), scale, acc);
            if (i == g.n[0] - 1)
                face_store(o.p[0], c + 1, face_flux_full<ND, 0, 1>(g, m, v, phi, i, j, k, c, sm & 2u, homog, pc), scale, acc);
            face_store(o.p[1], c, face_flux_full<ND, 1, 0>(g, m, v, phi, i, j, k, c, sm & 4u, homog, pc), scale, acc);
            if (j == g.n[1] - 1)
                face_store(o.p[1], c + g.sy, face_flux_full<ND, 1, 1>(g, m, v, phi, i, j, k, c, sm & 8u, homog, pc), scale, acc);
            if (ND > 2) {
                face_store(o.p[2], c, face_flux_full<ND, (ND > 2 ? 2 : 0), 0>(g, m, v, phi, i, j, k, c, sm & 16u, homog, pc), scale, acc);
                if (k == g.n[2] - 1)
                    face_store(o.p[2], c + g.sz, face_flux_full<ND, (ND > 2 ? 2 : 0), 1>(g, m, v, phi, i, j, k, c, sm & 32u, homog, pc), scale, acc);
            }
        } else {
            face_store(o.p[0], c, face_flux_ortho<0>(g, phi, m.jg[0][0], c, 1, 0, sm & 1u, homog, pc), scale, acc);
            if (i == g.n[0] - 1)
                face_store(o.p[0], c + 1, face_flux_ortho<1>(g, phi, m.jg[0][0], c, 1, 0, sm & 2u, homog, pc), scale, acc);
            face_store(o.p[1], c, face_flux_ortho<0>(g, phi, m.jg[1][1], c, g.sy, 1, sm & 4u, homog, pc), scale, acc);
            if (j == g.n[1] - 1)
                face_store(o.p[1], c + g.sy, face_flux_ortho<1>(g, phi, m.jg[1][1], c, g.sy, 1, sm & 8u, homog, pc), scale, acc);
            if (ND > 2) {
                face_store(o.p[2], c, face_flux_ortho<0>(g, phi, m.jg[2][2], c, g.sz, 2, sm & 16u, homog, pc), scale, acc);
                if (k == g.n[2] - 1)
                    face_store(o.p[2], c + g.sz, face_flux_ortho<1>(g, phi, m.jg[2][2], c, g.sz, 2, sm & 32u, homog, pc), scale, acc);
            }
        }
    }
}

// The caller has exchanged phi.  scale/acc: see face_store.
void launch_mac_grad(Ctx& cx, const Geom& g0, const Metric& m, double* const out[3], const double* phi, bool homog,
                     double scale, bool accumulate) {
    Geom g = g0;
    g.alpha = 0.0;
    g.beta = 1.0;   // the gradient carries no beta (x * 1.0 is exact)
    FaceOut o;
    for (int d = 0; d < 3; ++d) o.p[d] = out[d];
    const int kz = (g.n[2] >= 64) ? 8 : 1;
    dim3 grid((g.n[0] + PBX - 1) / PBX, (g.n[1] + PBY - 1) / PBY, (g.n[2] + kz - 1) / kz), block(PBX, PBY);
    const int acc = accumulate ? 1 : 0;
    if (m.diagonal) {
        if (g.nd == 3) k_mac_grad<3, false><<<grid, block, 0, cx.stream>>>(g, m, o, phi, homog, scale, acc, kz);
        else k_mac_grad<2, false><<<grid, block, 0, cx.stream>>>(g, m, o, phi, homog, scale, acc, kz);
    } else {
        if (g.nd == 3) k_mac_grad<3, true><<<grid, block, 0, cx.stream>>>(g, m, o, phi, homog, scale, acc, kz);
        else k_mac_grad<2, true><<<grid, block, 0, cx.stream>>>(g, m, o, phi, homog, scale, acc, kz);
    }
    cx.launches++;
}

// div = Jinv * sum_d (F_d(+) - F_d(-)) / dx_d, then "/ divideBy" when divideBy != 0 (BaseProjectorI.H:251-258).
template <int ND>
__global__ void __launch_bounds__(PBX* PBY) k_mac_div(const Geom g, const double* __restrict__ jinv,
                                                      double* __restrict__ div, const double* __restrict__ f0,
                                                      const double* __restrict__ f1, const double* __restrict__ f2,
                                                      double divideBy, int kz) {
    const int i = blockIdx.x * PBX + threadIdx.x;
    const int j = blockIdx.y * PBY + threadIdx.y;
    if (i >= g.n[0] || j >= g.n[1]) return;
    const int k0 = blockIdx.z * kz;
    const int k1 = min(k0 + kz, g.n[2]);
    for (int k = k0; k < k1; ++k) {
        const long long c = g.idx(i, j, k);
        double s = (f0[c + 1] - f0[c]) * g.rdx[0];
        s = s + (f1[c + g.sy] - f1[c]) * g.rdx[1];
        if (ND > 2) s = s + (f2[c + g.sz] - f2[c]) * g.rdx[2];
        double r = jinv[c] * s;
        if (divideBy != 0.0) r = r / divideBy;
        div[c] = r;
    }
}

void launch_mac_div(Ctx& cx, const Geom& g, const Metric& m, double* div, const double* const flux[3], double divideBy) {
    const int kz = (g.n[2] >= 64) ? 8 : 1;
    dim3 grid((g.n[0] + PBX - 1) / PBX, (g.n[1] + PBY - 1) / PBY, (g.n[2] + kz - 1) / kz), block(PBX, PBY);
    if (g.nd == 3) k_mac_div<3><<<grid, block, 0, cx.stream>>>(g, m.jinv, div, flux[0], flux[1], flux[2], divideBy, kz);
    else k_mac_div<2><<<grid, block, 0, cx.stream>>>(g, m.jinv, div, flux[0], flux[1], nullptr, divideBy, kz);
    cx.launches++;
}

// ---------------------------------------------------------------------------------------------------------
// Host side (level stream).
// ---------------------------------------------------------------------------------------------------------
// LevelMACProjector::computeGrad -- LevelMACProjector.cpp:186-218 (single level: no coarse-fine interpolation).
int op_mac_gradient(SomarLevel* L, SomarField* const grad[3], SomarField* phi, bool homog, double scale, bool accumulate) {
    Depth& dp = L->depths[0];
    int rc = exchange(L, 0, phi->p, 1);                 // Gradient.cpp:120-121
    if (rc != SOMAR_OK) return rc;
    double* out[3] = {grad[0]->p, grad[1]->p, (dp.g.nd > 2 && grad[2]) ? grad[2]->p : nullptr};
    launch_mac_grad(L->cx, dp.g, dp.m, out, phi->p, homog, scale, accumulate);
    return SOMAR_OK;
}

// LevelMACProjector::computeDiv -- LevelMACProjector.cpp:156-181 (fluxes carry their boundary values).
int op_mac_divergence(SomarLevel* L, SomarField* div, const SomarField* const flux[3], double divideBy) {
    Depth& dp = L->depths[0];
    const double* f[3] = {flux[0]->p, flux[1]->p, (dp.g.nd > 2 && flux[2]) ? flux[2]->p : nullptr};
    launch_mac_div(L->cx, dp.g, dp.m, div->p, f, divideBy);
    return SOMAR_OK;
}

// BaseProjector::project, a_velIsFlux = true, one level -- BaseProjectorI.H:176-300.  The gradient is never
// stored: k_mac_grad accumulates  vel += dtScale * grad  face by face (same statement as FArrayBox::plus).
int project(SomarLevel* L, SomarField* const vel[3], SomarField* pressure, double dt, bool zeroPressure, bool forceHomog,
            SomarSolveStats* st) {
    int rc;
    if (!L->pwork && (rc = field_alloc(L, 0, &L->pwork)) != SOMAR_OK) return rc;
    if ((rc = op_mac_divergence(L, L->pwork, vel, dt)) != SOMAR_OK) return rc;        // :244-258
    if ((rc = mg_solve(L, pressure, L->pwork, zeroPressure, forceHomog, st)) != SOMAR_OK) return rc;   // :261-266
    const double dtScale = (dt == 0.0) ? -1.0 : -dt;                                   // LevelMACProjector.cpp:227
    return op_mac_gradient(L, vel, pressure, false, dtScale, true);                    // :281-282
}

}  // namespace somar
