// projector.cu -- the calculus around the pressure solve, so that a whole
// BaseProjector::project (src/projection/BaseProjectorI.H:176-300, a_velIsFlux = true) stays on the device:
//   k_mac_div    Divergence::levelDivergenceMAC -> MAPPEDFLUXDIVERGENCE{2,3}D
//                src/calculus/DivCurlGrad/Divergence.cpp:44-127, DivCurlGradF.ChF:1034-1214   (+ the "/= dt" pass)
//   k_mac_grad   Gradient::levelGradientMAC (normal components) -> singleBoxMacGrad -> MAPPEDMACGRAD[ORTHO]
//                Gradient.cpp:85-170, 946-1100, DivCurlGradF.ChF:87-262; BC ghosts + BC fluxes as the operator's
//                optionally fused with LevelMACProjector::applyCorrection (LevelMACProjector.cpp:221-240)
//
//   k_cc_div, k_cc_grad   LevelCCProjector::computeDiv / computeGrad / applyCorrection (LevelCCProjector.cpp:156-255):
//                the cell-centred ("approximate") projection of the reference's time step, the second of its two
//                projections (AMRNavierStokes::doCCProjection, AMRNavierStokesAdvancePPM.cpp:1513-1573).  CellToEdge /
//                EdgeToCell, multByJ / divByJ and the two BC holders are folded into the two passes (below).
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
static Geom grad_geom(const Geom& g0, int gradBC);
void launch_mac_grad(Ctx& cx, const Geom& g0, const Metric& m, double* const out[3], const double* phi, bool homog,
                     double scale, bool accumulate, int gradBC) {
    const Geom g = grad_geom(g0, gradBC);
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
// The reference projectors' own BC holders (PhysBCUtil.cpp; LevelMACProjector.cpp:99-101, LevelCCProjector.cpp:103-105)
//   m_gradBC = basicGradPressureFuncBC: EllipticExtrapBCGhostClass(2) -- every non-periodic domain side gets an
//              order-2 extrapolated ghost and there is no flux method, so a wall face carries a one-sided difference;
//   m_divBC  = uStarFuncBC: on solid walls the NORMAL face value is set to zero, in place (BasicVelocityBCGhostClass ->
//              setSideDiriBC, face-centred branch, EllipticBCUtils.cpp:41-118).
// ---------------------------------------------------------------------------------------------------------
static Geom grad_geom(const Geom& g0, int gradBC) {
    Geom g = g0;
    g.alpha = 0.0;
    g.beta = 1.0;   // the gradient carries no beta (x * 1.0 is exact)
    if (gradBC == 1)
        for (int d = 0; d < 3; ++d)
            for (int s = 0; s < 2; ++s) {
                if (g.bc[d][s] != BC_NONE) g.bc[d][s] = BC_EXTRAP2;
                if (g.bct[d][s] != BC_PERIODIC && g.bct[d][s] != BC_NONE) g.bct[d][s] = BC_EXTRAP2;
            }
    return g;
}

// boundary faces of direction d of f_d <- 0 on the non-periodic domain sides this rank touches
template <int ND>
__global__ void k_wall_flux(const Geom g, FaceOut o) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y * blockDim.y + threadIdx.y;
    const int d = blockIdx.z >> 1, s = blockIdx.z & 1;
    if (d >= ND || g.bc[d][s] == BC_NONE) return;
    const int da = (d == 0) ? 1 : 0, db = (d == 2) ? 1 : 2;   // the two directions spanning the side
    if (a >= g.n[da] || b >= g.n[db]) return;
    int q[3];
    q[d] = s ? g.n[d] : 0;   // the high wall face is the low face of the first ghost cell
    q[da] = a;
    q[db] = b;
    o.p[d][g.idx(q[0], q[1], q[2])] = 0.0;
}

void launch_wall_flux(Ctx& cx, const Geom& g, double* const flux[3]) {
    FaceOut o;
    for (int d = 0; d < 3; ++d) o.p[d] = flux[d];
    int na = 1, nb = 1;
    for (int d = 0; d < g.nd; ++d) {
        const int da = (d == 0) ? 1 : 0, db = (d == 2) ? 1 : 2;
        na = max(na, g.n[da]);
        nb = max(nb, g.n[db]);
    }
    dim3 block(64, 4), grid((na + 63) / 64, (nb + 3) / 4, 2 * g.nd);
    if (g.nd == 3) k_wall_flux<3><<<grid, block, 0, cx.stream>>>(g, o);
    else k_wall_flux<2><<<grid, block, 0, cx.stream>>>(g, o);
    cx.launches++;
}

// ---- cell-centred projector ------------------------------------------------------------------------------
// Divergence::levelDivergenceCC (Divergence.cpp:359-396; USE_SIMPLE_STENCIL is off) in one pass, nothing stored on faces:
//   [u_d * J]                                multByJ with the cached cell-centred J (LevelGeometryUtil.cpp:302-340)
//   e_d(i) = half (u_d(i) + u_d(i - e_d))    CellToEdge (Chombo AMRTools CellToEdgeF.ChF)
//   wall faces <- 0                          m_divBC
//   Jinv sum_d (e_d(+) - e_d(-)) / dx_d      MAPPEDFLUXDIVERGENCE (the statements of k_mac_div), then "/ divideBy"
// 8 ND (u) + 8 (J) + 8 (Jinv) + 8 (out) B/cell.
struct CellIn {
    const double* p[3];
};
template <int ND>
__global__ void __launch_bounds__(PBX* PBY) k_cc_div(const Geom g, const double* __restrict__ jinv,
                                                     const double* __restrict__ J, CellIn u, double* __restrict__ div,
                                                     double divideBy, int kz) {
    const int i = blockIdx.x * PBX + threadIdx.x;
    const int j = blockIdx.y * PBY + threadIdx.y;
    if (i >= g.n[0] || j >= g.n[1]) return;
    const int k0 = blockIdx.z * kz;
    const int k1 = min(k0 + kz, g.n[2]);
    const long long st[3] = {1, g.sy, g.sz};
    for (int k = k0; k < k1; ++k) {
        const long long c = g.idx(i, j, k);
        const unsigned sm = side_mask(g, i, j, k);
        const double Jc = J ? J[c] : 1.0;
        double df[3] = {0.0, 0.0, 0.0};
#pragma unroll
        for (int d = 0; d < ND; ++d) {
            const double* __restrict__ ud = u.p[d];
            const double uc = J ? ud[c] * Jc : ud[c];
            double lo = 0.0, hi = 0.0;
            if (!(sm & (1u << (2 * d)))) {
                const double um = J ? ud[c - st[d]] * J[c - st[d]] : ud[c - st[d]];
                lo = 0.5 * (uc + um);
            }
            if (!(sm & (2u << (2 * d)))) {
                const double up = J ? ud[c + st[d]] * J[c + st[d]] : ud[c + st[d]];
                hi = 0.5 * (up + uc);
            }
            df[d] = hi - lo;
        }
        double s = df[0] * g.rdx[0];
        s = s + df[1] * g.rdx[1];
        if (ND > 2) s = s + df[2] * g.rdx[2];
        double r = jinv[c] * s;
        if (divideBy != 0.0) r = r / divideBy;
        div[c] = r;
    }
}

// Gradient::levelGradientCC (Gradient.cpp:469-494): levelGradientMAC with m_gradBC, then EdgeToCell
// (cell_d = half (face_d(i) + face_d(i + e_d)), Chombo AMRTools EdgeToCellF.ChF); the face gradients are formed twice
// (once per adjacent cell) and never stored.  mode 0: out_d = gradient.  mode 1 = the rest of the projection,
// LevelCCProjector::applyCorrection (:236-254) between multByJ and divByJ:  u_d <- ((u_d J) + grad_d dtScale) Jinv.
struct CellOut {
    double* p[3];
};
template <int ND, bool FULL>
__global__ void __launch_bounds__(PBX* PBY) k_cc_grad(const Geom g, const Metric m, const double* __restrict__ phi,
                                                      const double* __restrict__ J, CellOut o, double dtScale,
                                                      int mode, int kz) {
    const int i = blockIdx.x * PBX + threadIdx.x;
    const int j = blockIdx.y * PBY + threadIdx.y;
    if (i >= g.n[0] || j >= g.n[1]) return;
    const int k0 = blockIdx.z * kz;
    const int k1 = min(k0 + kz, g.n[2]);
    for (int k = k0; k < k1; ++k) {
        const long long c = g.idx(i, j, k);
        const unsigned sm = side_mask(g, i, j, k);
        const double pc = phi[c];
        double lo[3] = {0.0, 0.0, 0.0}, hi[3] = {0.0, 0.0, 0.0};
        if (FULL) {
            const ExView<ND> v(g, phi, i, j, k, EXMODE_OP);
            lo[0] = face_flux_full<ND, 0, 0>(g, m, v, phi, i, j, k, c, sm & 1u, false, pc);
            hi[0] = face_flux_full<ND, 0, 1>(g, m, v, phi, i, j, k, c, sm & 2u, false, pc);
            lo[1] = face_flux_full<ND, 1, 0>(g, m, v, phi, i, j, k, c, sm & 4u, false, pc);
            hi[1] = face_flux_full<ND, 1, 1>(g, m, v, phi, i, j, k, c, sm & 8u, false, pc);
            if (ND > 2) {
                lo[2] = face_flux_full<ND, (ND > 2 ? 2 : 0), 0>(g, m, v, phi, i, j, k, c, sm & 16u, false, pc);
                hi[2] = face_flux_full<ND, (ND > 2 ? 2 : 0), 1>(g, m, v, phi, i, j, k, c, sm & 32u, false, pc);
            }
        } else {
            lo[0] = face_flux_ortho<0>(g, phi, m.jg[0][0], c, 1, 0, sm & 1u, false, pc);
            hi[0] = face_flux_ortho<1>(g, phi, m.jg[0][0], c, 1, 0, sm & 2u, false, pc);
            lo[1] = face_flux_ortho<0>(g, phi, m.jg[1][1], c, g.sy, 1, sm & 4u, false, pc);
            hi[1] = face_flux_ortho<1>(g, phi, m.jg[1][1], c, g.sy, 1, sm & 8u, false, pc);
            if (ND > 2) {
                lo[2] = face_flux_ortho<0>(g, phi, m.jg[2][2], c, g.sz, 2, sm & 16u, false, pc);
                hi[2] = face_flux_ortho<1>(g, phi, m.jg[2][2], c, g.sz, 2, sm & 32u, false, pc);
            }
        }
#pragma unroll
        for (int d = 0; d < ND; ++d) {
            const double gc = 0.5 * (lo[d] + hi[d]);
            if (mode == 0) {
                o.p[d][c] = gc;
            } else {
                double v = o.p[d][c];
                if (J) v = v * J[c];
                v = v + gc * dtScale;
                if (J) v = v * m.jinv[c];
                o.p[d][c] = v;
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// Host side (level stream).
// ---------------------------------------------------------------------------------------------------------
// LevelMACProjector::computeGrad -- LevelMACProjector.cpp:186-218 (single level: no coarse-fine interpolation).
static int check_extrap_bc(const Geom& g) {
    for (int d = 0; d < g.nd; ++d)
        if (!g.periodic[d] && g.n[d] < 3) { set_error("gradient BC 1 (order-2 extrapolated ghosts) needs 3 cells across a non-periodic direction"); return SOMAR_ERR_ARG; }
    return SOMAR_OK;
}
int op_mac_gradient(SomarLevel* L, SomarField* const grad[3], SomarField* phi, bool homog, double scale, bool accumulate,
                    int gradBC) {
    Depth& dp = L->depths[0];
    int rc;
    if (gradBC == 1 && (rc = check_extrap_bc(dp.g)) != SOMAR_OK) return rc;
    rc = exchange(L, 0, phi->p, 1);                     // Gradient.cpp:120-121
    if (rc != SOMAR_OK) return rc;
    double* out[3] = {grad[0]->p, grad[1]->p, (dp.g.nd > 2 && grad[2]) ? grad[2]->p : nullptr};
    launch_mac_grad(L->cx, dp.g, dp.m, out, phi->p, homog, scale, accumulate, gradBC);
    return SOMAR_OK;
}

// m_divBC of the projectors, in place (Divergence.cpp:79-100 casts the const away and writes the caller's fluxes)
int op_wall_flux_bc(SomarLevel* L, SomarField* const flux[3]) {
    Depth& dp = L->depths[0];
    double* f[3] = {flux[0]->p, flux[1]->p, (dp.g.nd > 2 && flux[2]) ? flux[2]->p : nullptr};
    launch_wall_flux(L->cx, dp.g, f);
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
// L->projector_bc (somar_level_set_projector_bc): 0 = the fluxes carry their wall values and the gradient uses the
// operator's BC holder (round 1); 1 = the reference projector's own m_divBC / m_gradBC (above).
int project(SomarLevel* L, SomarField* const vel[3], SomarField* pressure, double dt, bool zeroPressure, bool forceHomog,
            SomarSolveStats* st) {
    int rc;
    if (!L->pwork && (rc = field_alloc(L, 0, &L->pwork)) != SOMAR_OK) return rc;
    if (L->projector_bc == 1 && (rc = op_wall_flux_bc(L, vel)) != SOMAR_OK) return rc;
    if ((rc = op_mac_divergence(L, L->pwork, vel, dt)) != SOMAR_OK) return rc;        // :244-258
    if ((rc = mg_solve(L, pressure, L->pwork, zeroPressure, forceHomog, st)) != SOMAR_OK) return rc;   // :261-266
    const double dtScale = (dt == 0.0) ? -1.0 : -dt;                                   // LevelMACProjector.cpp:227
    return op_mac_gradient(L, vel, pressure, false, dtScale, true, L->projector_bc);   // :281-282
}

// ---- LevelCCProjector ------------------------------------------------------------------------------------
static void launch_cc_div(Ctx& cx, const Geom& g, const Metric& m, const double* J, const double* const u[3], double* div,
                          double divideBy) {
    CellIn in;
    for (int d = 0; d < 3; ++d) in.p[d] = u[d];
    const int kz = (g.n[2] >= 64) ? 8 : 1;
    dim3 grid((g.n[0] + PBX - 1) / PBX, (g.n[1] + PBY - 1) / PBY, (g.n[2] + kz - 1) / kz), block(PBX, PBY);
    if (g.nd == 3) k_cc_div<3><<<grid, block, 0, cx.stream>>>(g, m.jinv, J, in, div, divideBy, kz);
    else k_cc_div<2><<<grid, block, 0, cx.stream>>>(g, m.jinv, J, in, div, divideBy, kz);
    cx.launches++;
}
static void launch_cc_grad(Ctx& cx, const Geom& g0, const Metric& m, const double* phi, const double* J,
                           double* const out[3], double dtScale, int mode) {
    const Geom g = grad_geom(g0, 1);   // m_gradBC = gradPiFuncBC, LevelCCProjector.cpp:105
    CellOut o;
    for (int d = 0; d < 3; ++d) o.p[d] = out[d];
    const int kz = (g.n[2] >= 64) ? 8 : 1;
    dim3 grid((g.n[0] + PBX - 1) / PBX, (g.n[1] + PBY - 1) / PBY, (g.n[2] + kz - 1) / kz), block(PBX, PBY);
    if (m.diagonal) {
        if (g.nd == 3) k_cc_grad<3, false><<<grid, block, 0, cx.stream>>>(g, m, phi, J, o, dtScale, mode, kz);
        else k_cc_grad<2, false><<<grid, block, 0, cx.stream>>>(g, m, phi, J, o, dtScale, mode, kz);
    } else {
        if (g.nd == 3) k_cc_grad<3, true><<<grid, block, 0, cx.stream>>>(g, m, phi, J, o, dtScale, mode, kz);
        else k_cc_grad<2, true><<<grid, block, 0, cx.stream>>>(g, m, phi, J, o, dtScale, mode, kz);
    }
    cx.launches++;
}

// LevelCCProjector::computeDiv (:156-187).  The ghosts CellToEdge reads across box interfaces, rank boundaries and
// periodic sides are the neighbours' cells: one halo exchange per component (and of J); ghosts beyond a solid wall are
// never used (m_divBC overwrites the wall faces).
int op_cc_divergence(SomarLevel* L, SomarField* div, SomarField* const vel[3], SomarField* J, double divideBy) {
    Depth& dp = L->depths[0];
    int rc;
    for (int d = 0; d < dp.g.nd; ++d)
        if ((rc = exchange(L, 0, vel[d]->p, 1)) != SOMAR_OK) return rc;
    if (J && (rc = exchange(L, 0, J->p, 1)) != SOMAR_OK) return rc;
    const double* u[3] = {vel[0]->p, vel[1]->p, (dp.g.nd > 2 && vel[2]) ? vel[2]->p : nullptr};
    launch_cc_div(L->cx, dp.g, dp.m, J ? J->p : nullptr, u, div->p, divideBy);
    return SOMAR_OK;
}

// LevelCCProjector::computeGrad (:199-231)
int op_cc_gradient(SomarLevel* L, SomarField* const grad[3], SomarField* phi) {
    Depth& dp = L->depths[0];
    int rc;
    if ((rc = check_extrap_bc(dp.g)) != SOMAR_OK) return rc;
    if ((rc = exchange(L, 0, phi->p, 1)) != SOMAR_OK) return rc;
    double* out[3] = {grad[0]->p, grad[1]->p, (dp.g.nd > 2 && grad[2]) ? grad[2]->p : nullptr};
    launch_cc_grad(L->cx, dp.g, dp.m, phi->p, nullptr, out, 0.0, 0);
    return SOMAR_OK;
}

// BaseProjector<FArrayBox>::project on one level (BaseProjectorI.H:176-300) with LevelCCProjector's virtuals.
// J == nullptr: a_velIsFlux = true.  Two sweeps around the solve; multByJ / divByJ never touch memory on their own.
int project_cc(SomarLevel* L, SomarField* const vel[3], SomarField* J, SomarField* pressure, double dt, bool zeroPressure,
               bool forceHomog, SomarSolveStats* st) {
    Depth& dp = L->depths[0];
    int rc;
    if ((rc = check_extrap_bc(dp.g)) != SOMAR_OK) return rc;
    if (!L->pwork && (rc = field_alloc(L, 0, &L->pwork)) != SOMAR_OK) return rc;
    if ((rc = op_cc_divergence(L, L->pwork, vel, J, dt)) != SOMAR_OK) return rc;       // :244-258
    if ((rc = mg_solve(L, pressure, L->pwork, zeroPressure, forceHomog, st)) != SOMAR_OK) return rc;
    if ((rc = exchange(L, 0, pressure->p, 1)) != SOMAR_OK) return rc;
    const double dtScale = (dt == 0.0) ? -1.0 : -dt;                                    // LevelCCProjector.cpp:242
    double* out[3] = {vel[0]->p, vel[1]->p, (dp.g.nd > 2 && vel[2]) ? vel[2]->p : nullptr};
    launch_cc_grad(L->cx, dp.g, dp.m, pressure->p, J ? J->p : nullptr, out, dtScale, 1);
    return SOMAR_OK;
}

}  // namespace somar
