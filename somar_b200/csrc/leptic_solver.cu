// leptic_solver.cu -- the leptic ITERATION (SURVEY.md 8(a) "LevelLepticSolver" row, 8(f)-2), device-resident:
//   somar_leptic_create   LevelLepticSolver::define   src/calculus/LepticSolver/LevelLepticSolver.cpp:147-437
//   somar_leptic_solve    LevelLepticSolver::solve    :646-956
//       k_horiz_rhs       ::computeHorizRHS           :981-1097  (MAPPEDMACGRADORTHO, DivCurlGradF.ChF:221-262 ->
//                                                      UNMAPPEDVERTINTEGRAL, SubspaceF.ChF:33-58 -> LEPTICACCUMDIV,
//                                                      LevelLepticSolverF.ChF:291-331, fused: the face gradients and their
//                                                      vertical averages are never stored)
//       k_vert_avg_dense  LevelGeometry::createVertAvgFCJgupPtr  src/geometry/LevelGeometryBasics.cpp:499-571
//       setZeroAvg        :1665-1709 (k_reduce kind 5 + k_sub_scalar)
// The sweeps it sequences (vertical solves, excess, extrusion) are leptic.cu's; the operator, norms and the two
// multigrid solvers are the library's own (level.cu / solver.cu) on two derived levels:
//   vert   the level's J g^{ij} with Jinv = 1 (:197-205): residuals/norms of the J-scaled equation and the 3-D multigrid
//          of the fall-back (:251-300);
//   horiz  the flattened level with vertically averaged J g^{ab}, Jinv = 1, zero-Neumann sides (:355-407).
// Flat fields (excess, boundary data, horizontal right-hand side) are SomarFields of the vertical level that own only
// the planes the flat kernels touch.  Each level has its own stream; the driver is host-sequenced (it reads a norm
// back every order, as the reference does), so hand-overs between levels are stream synchronisations.
// Scope: one level; diagonal metric on one rank or y-slab ranks (the vertical level borrows the level's communicator; the
// small horizontal level is replicated on every rank, its rows all-gathered); full (terrain-following) metric on one
// rank: k_horiz_rhs_full, k_vert_horiz_grad, a 2-D full-metric horizontal level, a horizontal solve at every order; see
// include/somar_b200.h.
#include "level.hpp"
#include "cells.cuh"
#include "stencil_full.cuh"
#include <cmath>
#include <cstring>
#include <algorithm>
#include <utility>
#include <vector>

struct SomarLeptic {
    SomarLevel* orig = nullptr;    // not owned
    SomarLevel* vert = nullptr;
    SomarLevel* horiz = nullptr;
    SomarLepticParams prm;
    bool doHoriz = false;
    int vertBC[2] = {0, 0};
    // vertical level: full fields
    SomarField *phiTotal = nullptr, *vertPhi = nullptr, *rhsA = nullptr, *rhsB = nullptr;
    // flat fields (short allocations on the vertical level)
    SomarField *bdryLo = nullptr, *bdryHi = nullptr, *excess = nullptr, *flatRhs = nullptr, *flatPhi = nullptr;
    size_t flatDoubles = 0;
    // original level
    SomarField* res = nullptr;
    // horizontal level
    SomarField *hPhi = nullptr, *hRhs = nullptr;
    // several ranks: the horizontal level is REPLICATED (every rank holds the whole flattened domain, like the
    // gathered bottom depth of bottom_gather.cu); rows travel with one ncclAllGather on the vertical level's stream
    std::vector<int> row0, rows;       // first row / number of rows of each rank at depth 0
    int maxRows = 0;
    double *gsend = nullptr, *gall = nullptr, *gdense = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

namespace somar {

#define QTPB 128

// out = a / b on the valid cells (FArrayBox::divide, LevelLepticSolver.cpp:712)
__global__ void __launch_bounds__(QTPB) k_div_cells(const Geom g, double* __restrict__ out, const double* __restrict__ a,
                                                    const double* __restrict__ b, int kz) {
    const int i = blockIdx.x * QTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= g.n[0]) return;
    const int k0 = blockIdx.z * kz, k1 = min(k0 + kz, g.n[2]);
    for (int k = k0; k < k1; ++k) {
        const long long c = g.idx(i, j, k);
        out[c] = a[c] / b[c];
    }
}

// flat y = y + x * a (FArrayBox::plus(src, scale) on a flat holder: BoundaryData::vertPlus, :766,787; :806)
__global__ void __launch_bounds__(QTPB) k_flat_axpy(const Geom g, double* __restrict__ y, const double* __restrict__ x, double a) {
    const int i = blockIdx.x * QTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= g.n[0]) return;
    const long long c = g.idx(i, j, 0);
    y[c] = y[c] + x[c] * a;
}

// computeHorizRHS, diagonal metric, one thread per column:
//   F_d(face) = sum_k [ (1/dx_d * J g^{dd}) * (phi(i) - phi(i - e_d)) ] * (1/Nz)   on interior faces, k ascending,
//             = 0 on domain-boundary faces (the BoundaryData values, zero for a diagonal metric, :1063-1069)
//   rhs = 0 + (F_0(i+1) - F_0(i)) * (-1/dx_0) + (F_1(j+1) - F_1(j)) * (-1/dx_1)  [+ excess * exScale, :803-808]
__global__ void __launch_bounds__(QTPB) k_horiz_rhs(const Geom g, double* __restrict__ rhs, const double* __restrict__ phi,
                                                    const double* __restrict__ jg0, const double* __restrict__ jg1,
                                                    const double* __restrict__ excess, double exScale, double dzScale) {
    const int i = blockIdx.x * QTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= g.n[0]) return;
    const int gi = g.g0[0] + i, gj = g.g0[1] + j;
    const bool xl = gi >= 1, xh = gi + 1 <= g.gn[0] - 1, yl = gj >= 1, yh = gj + 1 <= g.gn[1] - 1;
    const double rx = g.rdx[0], ry = g.rdx[1];
    double fxl = 0.0, fxh = 0.0, fyl = 0.0, fyh = 0.0;
    const long long c0 = g.idx(i, j, 0);
    const int nz = g.n[2];
#pragma unroll 2
    for (int k = 0; k < nz; ++k) {
        const long long c = c0 + g.sz * k;
        const double p = phi[c];
        if (xl) fxl = fxl + (rx * jg0[c]) * (p - phi[c - 1]) * dzScale;
        if (xh) fxh = fxh + (rx * jg0[c + 1]) * (phi[c + 1] - p) * dzScale;
        if (yl) fyl = fyl + (ry * jg1[c]) * (p - phi[c - g.sy]) * dzScale;
        if (yh) fyh = fyh + (ry * jg1[c + g.sy]) * (phi[c + g.sy] - p) * dzScale;
    }
    double r = 0.0;
    r = r + (fxh - fxl) * (-1.0 / g.dx[0]);
    r = r + (fyh - fyl) * (-1.0 / g.dx[1]);
    if (excess) r = r + excess[c0] * exScale;
    rhs[c0] = r;
}

// ---- full (terrain-following) metric ------------------------------------------------------------------------
// phi as computeHorizRHS sees it from the box of the column being processed (:1004-1014): extrapAllGhosts
// (ExtrapolationUtils.cpp:388-421) extrapolates EVERY ghost of every box with order 2, direction by direction on a
// growing box (x innermost), BEFORE the exchange; the exchange (faces, then corners) overwrites the ghosts that lie
// inside the domain with the neighbours' cells; what lies outside the domain keeps the per-box extrapolation -- in
// every direction in which it is outside the BOX, so an edge ghost next to a box interface depends on the box layout.
template <int ND>
struct LepticView {
    const Geom& g;
    const double* __restrict__ phi;
    BoxOf b;
    __device__ LepticView(const Geom& g_, const double* __restrict__ phi_, int i, int j, int k)
        : g(g_), phi(phi_), b(box_of<ND>(g_, i, j, k)) {}
    __device__ double ex(int p0, int p1, int p2) const {
        const int p[3] = {p0, p1, p2};
        int o[3] = {0, 0, 0}, ord[3] = {0, 0, 0};
        bool outBox = false, outDom = false;
#pragma unroll
        for (int d = 0; d < ND; ++d) {
            if (p[d] < b.lo[d]) o[d] = -1;
            else if (p[d] > b.hi[d]) o[d] = 1;
            if (o[d] != 0) {
                outBox = true;
                ord[d] = 2;
            }
            const int gq = p[d] + g.g0[d];
            if (gq < 0 || gq >= g.gn[d]) outDom = true;
        }
        if (!outBox || !outDom) return phi[g.idx(p0, p1, p2)];
        return ExRec<ND - 1>::eval(g, phi, p0, p1, p2, o, ord);
    }
};

// computeHorizRHS, full metric, one thread per column: the statements of k_horiz_rhs with MAPPEDMACGRAD
// (DivCurlGradF.ChF:87-147: normal difference + the four-point transverse differences in the other two directions) for the
// face gradients.  Both faces of a column are formed with the column's own box view, as the reference forms every face
// once per box.
__global__ void __launch_bounds__(QTPB) k_horiz_rhs_full(const Geom g, const Metric m, double* __restrict__ rhs,
                                                         const double* __restrict__ phi,
                                                         const double* __restrict__ excess, double exScale,
                                                         double dzScale) {
    const int i = blockIdx.x * QTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= g.n[0]) return;
    const int gi = g.g0[0] + i, gj = g.g0[1] + j;
    const bool xl = gi >= 1, xh = gi + 1 <= g.gn[0] - 1, yl = gj >= 1, yh = gj + 1 <= g.gn[1] - 1;
    double fxl = 0.0, fxh = 0.0, fyl = 0.0, fyh = 0.0;
    const long long c0 = g.idx(i, j, 0);
    const int nz = g.n[2];
    const LepticView<3> v(g, phi, i, j, 0);
    for (int k = 0; k < nz; ++k) {
        const long long c = c0 + g.sz * k;
        const double p = phi[c];
        if (xl) fxl = fxl + face_flux_full<3, 0, 0>(g, m, v, phi, i, j, k, c, false, true, p) * dzScale;
        if (xh) fxh = fxh + face_flux_full<3, 0, 1>(g, m, v, phi, i, j, k, c, false, true, p) * dzScale;
        if (yl) fyl = fyl + face_flux_full<3, 1, 0>(g, m, v, phi, i, j, k, c, false, true, p) * dzScale;
        if (yh) fyh = fyh + face_flux_full<3, 1, 1>(g, m, v, phi, i, j, k, c, false, true, p) * dzScale;
    }
    double r = 0.0;
    r = r + (fxh - fxl) * (-1.0 / g.dx[0]);
    r = r + (fyh - fyl) * (-1.0 / g.dx[1]);
    if (excess) r = r + excess[c0] * exScale;
    rhs[c0] = r;
}

// levelVertHorizGradient (:1107-1176) + LEPTICVERTHORIZGRAD (LevelLepticSolverF.ChF:74-122): on the lower and the upper
// domain face of every column,  scale * sum_m J g^{z m} d_m phi  from the four-point differences of the ghost-side and
// the valid-side plane.  ExtrapolateFaceAndCopy(phi, phi, FAB box & domain, z, side, 2) has just rewritten the ghost-side
// plane (order 2 from the column) and, in BOTH planes, the cells beyond the horizontal domain sides (order 2 along
// the plane); cells across box interfaces are the neighbours' (the residual evaluation that closed the previous order
// exchanged phi).  One thread per column, both sides.
__global__ void __launch_bounds__(QTPB) k_vert_horiz_grad(const Geom g, const double* __restrict__ jgz0,
                                                          const double* __restrict__ jgz1, const double* __restrict__ phi,
                                                          double* __restrict__ flatLo, double* __restrict__ flatHi,
                                                          double scale) {
    const int i = blockIdx.x * QTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= g.n[0]) return;
    const double dxinv0 = scale * 0.25 / g.dx[0];
    const double dxinv1 = scale * 0.25 / g.dx[1];
    for (int side = 0; side < 2; ++side) {
        const int k0 = side ? g.n[2] - 1 : 0, ks = side ? -1 : 1;
        // plane value at an in-domain column: which = 0 the ghost-side plane (extrapolated), 1 the valid-side plane
        auto plane = [&](int ii, int jj, int which) {
            const long long c = g.idx(ii, jj, k0);
            if (which) return phi[c];
            return 3.0 * (phi[c] - phi[c + ks * g.sz]) + phi[c + 2 * ks * g.sz];
        };
        // the same, one column beyond a horizontal domain side: order 2 along the plane
        auto at = [&](int ii, int jj, int which) {
            const int gi = ii + g.g0[0], gj = jj + g.g0[1];
            if (gi < 0) return 3.0 * (plane(ii + 1, jj, which) - plane(ii + 2, jj, which)) + plane(ii + 3, jj, which);
            if (gi >= g.gn[0]) return 3.0 * (plane(ii - 1, jj, which) - plane(ii - 2, jj, which)) + plane(ii - 3, jj, which);
            if (gj < 0) return 3.0 * (plane(ii, jj + 1, which) - plane(ii, jj + 2, which)) + plane(ii, jj + 3, which);
            if (gj >= g.gn[1]) return 3.0 * (plane(ii, jj - 1, which) - plane(ii, jj - 2, which)) + plane(ii, jj - 3, which);
            return plane(ii, jj, which);
        };
        const long long f = g.idx(i, j, side ? g.n[2] : 0);     // the domain face (the high one sits in the first ghost slot)
        const double v = jgz0[f] * dxinv0 * (at(i + 1, j, 0) - at(i - 1, j, 0) + at(i + 1, j, 1) - at(i - 1, j, 1)) +
                         jgz1[f] * dxinv1 * (at(i, j + 1, 0) - at(i, j - 1, 0) + at(i, j + 1, 1) - at(i, j - 1, 1));
        (side ? flatHi : flatLo)[g.idx(i, j, 0)] = v;
    }
}

// dense out(i, j) = 0 + sum_k jg(i, j, k) * scale over (nx + ex) x (ny + ey) faces
__global__ void __launch_bounds__(QTPB) k_vert_avg_dense(const Geom g, double* __restrict__ out, const double* __restrict__ jg,
                                                         int ex, int ey, double scale) {
    const int i = blockIdx.x * QTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= g.n[0] + ex || j >= g.n[1] + ey) return;
    const long long c0 = g.idx(i, j, 0);
    double s = 0.0;
    for (int k = 0; k < g.n[2]; ++k) s = s + jg[c0 + g.sz * k] * scale;
    out[i + (long long)(g.n[0] + ex) * j] = s;
}

__global__ void k_fill_dense(double* __restrict__ x, long long n, double v) {
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) x[t] = v;
}

// x = x - a on the valid cells of a 2-D level (FArrayBox::operator-=, :1707)
__global__ void __launch_bounds__(QTPB) k_sub_scalar(const Geom g, double* __restrict__ x, double avg) {
    const int i = blockIdx.x * QTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= g.n[0]) return;
    const long long c = g.idx(i, j, 0);
    x[c] = x[c] - avg;
}

static inline dim3 plane_grid(const Geom& g, int ex = 0, int ey = 0, int nz = 1) {
    return dim3((g.n[0] + ex + QTPB - 1) / QTPB, g.n[1] + ey, nz);
}

static int flat_alloc(SomarLeptic* S, SomarField** out) {
    SomarField* f = new SomarField{S->vert, 0, nullptr};
    if (cudaMalloc(&f->p, S->flatDoubles * sizeof(double)) != cudaSuccess) {
        delete f;
        set_error("leptic: out of device memory (flat field)");
        return SOMAR_ERR_CUDA;
    }
    *out = f;
    return SOMAR_OK;
}
static inline int flat_zero(SomarLeptic* S, SomarField* f) {
    SOMAR_CUDA(cudaMemsetAsync(f->p, 0, S->flatDoubles * sizeof(double), S->vert->cx.stream));
    return SOMAR_OK;
}
static void field_free(SomarField* f) {
    if (!f) return;
    cudaFree(f->p);
    delete f;
}
static inline int sync(SomarLevel* L) {
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    return SOMAR_OK;
}
static int norm_of(SomarLevel* L, const SomarField* x, int ord, double* out) {
    int rc = reduce_global(L, x->depth, x->p, nullptr, ord, out);
    if (rc) return rc;
    if (ord == 2) *out = std::sqrt(*out);
    return SOMAR_OK;
}

// Every rank's rows [0, rows_r) (+ one more row from the last rank when extraRow) of a row-pitched local array
// -> the dense global array dst ((ny + extraRow) x width) on every rank.  Stream: the vertical level's.
static int gather_rows(SomarLeptic* S, const double* src, size_t srcPitch, int width, int extraRow, double* dst) {
    SomarLevel* V = S->vert;
    const int me = V->desc.rank, nr = V->desc.nranks;
    cudaStream_t st = V->cx.stream;
    const size_t chunk = (size_t)(S->maxRows + 1) * (size_t)(S->orig->depths[0].g.n[0] + 1);
    SOMAR_CUDA(cudaMemcpy2DAsync(S->gsend, width * sizeof(double), src, srcPitch * sizeof(double), width * sizeof(double),
                                 S->rows[me] + extraRow, cudaMemcpyDeviceToDevice, st));
    SOMAR_NCCL(ncclAllGather(S->gsend, S->gall, chunk, ncclDouble, V->comm, st));
    for (int r = 0; r < nr; ++r) {
        const bool last = S->row0[r] + S->rows[r] == V->depths[0].g.gn[1];
        const size_t nrow = (size_t)S->rows[r] + ((extraRow && last) ? 1 : 0);
        SOMAR_CUDA(cudaMemcpyAsync(dst + (size_t)S->row0[r] * width, S->gall + r * chunk, nrow * width * sizeof(double),
                                   cudaMemcpyDeviceToDevice, st));
    }
    return SOMAR_OK;
}

}  // namespace somar

using namespace somar;

extern "C" {

int somar_leptic_default_params(SomarLepticParams* p) {
    if (!p) { set_error("null argument"); return SOMAR_ERR_ARG; }
    // LevelLepticSolver::setDefaultParameters, LevelLepticSolver.cpp:463-514
    p->max_order = 4; p->eps = 1.0e-15; p->hang = 1.0e-15; p->norm_type = 0; p->horiz_rhs_tol = 1.0e-14;
    somar_default_params(&p->horiz);
    SomarSolverParams& h = p->horiz;
    h.imin = 5; h.imax = 20; h.num_smooth_down = 4; h.num_smooth_bottom = 4; h.num_smooth_up = 4; h.num_smooth_precond = 2;
    h.relaxMode = SOMAR_RELAX_LEVEL_GSRB; h.precondMode = 1; h.maxDepth = -1; h.eps = 1.0e-12; h.hang = 1.0e-15;
    h.normThresh = 1.0e-30; h.numMG = 1;   // "hardcoding this because 1 is all that seems to work" (:565)
    h.bottom_imax = 80; h.bottom_eps = 1.0e-12; h.bottom_numRestarts = 5; h.bottom_hang = 1.0e-15;
    somar_default_params(&p->full);
    SomarSolverParams& f = p->full;
    f.imin = 5; f.imax = 20; f.num_smooth_down = 4; f.num_smooth_bottom = 4; f.num_smooth_up = 4; f.num_smooth_precond = 4;
    f.relaxMode = SOMAR_RELAX_LINE_GSRB; f.precondMode = 2; f.maxDepth = -1; f.eps = 1.0e-6; f.hang = 1.0e-15;
    f.normThresh = 1.0e-30; f.numMG = 1;
    f.bottom_imax = 80; f.bottom_numRestarts = 5;   // eps / hang of this BiCGStab are never set (:269-274)
    return SOMAR_OK;
}

int somar_leptic_destroy(SomarLeptic* S) {
    if (!S) return SOMAR_OK;
    if (S->vert) { cudaSetDevice(S->vert->device); cudaStreamSynchronize(S->vert->cx.stream); }
    if (S->horiz) cudaStreamSynchronize(S->horiz->cx.stream);
    for (SomarField* f : {S->phiTotal, S->vertPhi, S->rhsA, S->rhsB, S->bdryLo, S->bdryHi, S->excess, S->flatRhs,
                          S->flatPhi, S->res, S->hPhi, S->hRhs})
        field_free(f);
    cudaFree(S->gsend); cudaFree(S->gall); cudaFree(S->gdense);
    if (S->ev0) cudaEventDestroy(S->ev0);
    if (S->ev1) cudaEventDestroy(S->ev1);
    somar_level_destroy(S->horiz);
    somar_level_destroy(S->vert);
    delete S;
    return SOMAR_OK;
}

int somar_leptic_create(SomarLevel* L, const SomarLepticParams* prm, SomarLeptic** out) {
    if (!L || !prm || !out) { set_error("null argument"); return SOMAR_ERR_ARG; }
    if (!L->finalized) { set_error("leptic: the level is not finalized"); return SOMAR_ERR_ARG; }
    const SomarLevelDesc& D = L->desc;
    const int nd = D.dim, vd = nd - 1;
    const bool fullMetric = !D.is_diagonal;
    if (fullMetric && (nd != 3 || D.nranks > 1)) {
        set_error("leptic solver: a full metric is built for 3-D levels on one rank");
        return SOMAR_ERR_ARG;
    }
    if (fullMetric && (D.n[0] < 3 || D.n[1] < 3 || D.n[2] < 3)) {
        set_error("leptic solver: a full metric needs 3 cells in every direction (order-2 extrapolated ghosts)");
        return SOMAR_ERR_ARG;
    }
    if (D.periodic[vd]) { set_error("leptic solver: the vertical direction is periodic (LevelLepticSolver.cpp:1316)"); return SOMAR_ERR_ARG; }
    if (prm->max_order < 0 || prm->max_order > 14) { set_error("leptic solver: max_order must be in 0..14"); return SOMAR_ERR_ARG; }
    cudaSetDevice(L->device);
    SomarLeptic* S = new SomarLeptic;
    S->orig = L;
    S->prm = *prm;
    // gatherVerticalBCTypes (:1522-1644): one box column per (i,j), both ends on the physical boundary
    for (int s = 0; s < 2; ++s) {
        S->vertBC[s] = D.bc_type[2 * vd + s];
        if (S->vertBC[s] != 0 && S->vertBC[s] != 1) {
            set_error("leptic solver: vertical sides must be Neumann (0) or Dirichlet (1)");
            delete S;
            return SOMAR_ERR_ARG;
        }
    }
    S->doHoriz = (S->vertBC[0] == 0 && S->vertBC[1] == 0);
    if (S->doHoriz && (nd != 3 || D.periodic[0] || D.periodic[1])) {
        set_error("leptic solver: a horizontal solve needs a 3-D level with non-periodic horizontal directions "
                  "(computeHorizRHS leaves periodic boundary faces unset, LevelLepticSolver.cpp:997-1001,1064-1069)");
        delete S;
        return SOMAR_ERR_ARG;
    }
    int rc;
    auto fail = [&](int code) { somar_leptic_destroy(S); return code; };
    // ---- vertical level: metric copied, Jinv = 1 (:197-205); multigrid of the fall-back with the `full` parameters
    {
        SomarLevelDesc dv = D;
        dv.alpha = 0.0;
        dv.beta = 1.0;
        SomarSolverParams pv = prm->full;
        pv.bottom_normType = prm->norm_type;   // :274
        if ((rc = level_create_impl(&dv, &pv, L->device, nullptr, L->comm, &S->vert))) return fail(rc);
        Depth& src = L->depths[0];
        Depth& dst = S->vert->depths[0];
        if (src.g.total != dst.g.total) { set_error("leptic: vertical level layout differs"); return fail(SOMAR_ERR_ARG); }
        for (int d = 0; d < nd; ++d)
            for (int e = 0; e < nd; ++e) {
                if (e != d && !fullMetric) continue;
                if (cudaMemcpyAsync(dst.jg[d][e], src.jg[d][e], sizeof(double) * src.g.total, cudaMemcpyDeviceToDevice,
                                    S->vert->cx.stream) != cudaSuccess) { set_error("leptic: metric copy failed"); return fail(SOMAR_ERR_CUDA); }
            }
        if ((rc = sync(L))) return fail(rc);
        launch_fill(S->vert->cx, dst.g, dst.jinv, 1.0);
        if ((rc = level_finalize(S->vert))) return fail(rc);
        if ((rc = sync(S->vert))) return fail(rc);
    }
    const Geom& g = S->vert->depths[0].g;
    S->flatDoubles = (size_t)(g.off0 + g.sz);
    for (SomarField** f : {&S->phiTotal, &S->vertPhi, &S->rhsA, &S->rhsB})
        if ((rc = field_alloc(S->vert, 0, f))) return fail(rc);
    if ((rc = field_alloc(L, 0, &S->res))) return fail(rc);
    for (SomarField** f : {&S->bdryLo, &S->bdryHi})
        if ((rc = flat_alloc(S, f)) || (rc = flat_zero(S, *f))) return fail(rc);
    // ---- horizontal level (:303-434)
    if (S->doHoriz) {
        for (SomarField** f : {&S->excess, &S->flatRhs, &S->flatPhi})
            if ((rc = flat_alloc(S, f)) || (rc = flat_zero(S, *f))) return fail(rc);
        SomarLevelDesc dh = D;
        dh.dim = 2;
        dh.n[2] = 1;
        dh.periodic[2] = 0;
        dh.dx[2] = 1.0;
        dh.alpha = 0.0;
        dh.beta = 1.0;
        dh.is_diagonal = fullMetric ? 0 : 1;   // J g^{ab}, a, b horizontal (LevelGeometryBasics.cpp:545-568)
        for (int s = 0; s < 6; ++s) { dh.bc_type[s] = 0; dh.bc_value[s] = 0.0; }   // EllipticConstNeumBCFluxClass(0,0) (:363-368)
        // flattened boxes of the level, each footprint once (LepticBoxUtils::FlattenTransform; the reference then
        // re-balances them, createHorizontalSolverGrids -- the layout only enters through the multigrid schedule)
        std::vector<int> lo, hi;
        for (int b = 0; b < D.nboxes; ++b) {
            const int* bl = &L->box_lo[3 * b];
            const int* bh = &L->box_hi[3 * b];
            bool seen = false;
            for (size_t q = 0; q < lo.size() / 3; ++q)
                if (lo[3 * q] == bl[0] && lo[3 * q + 1] == bl[1] && hi[3 * q] == bh[0] && hi[3 * q + 1] == bh[1]) seen = true;
            if (seen) continue;
            lo.insert(lo.end(), {bl[0], bl[1], 0});
            hi.insert(hi.end(), {bh[0], bh[1], 0});
        }
        dh.nboxes = (int)(lo.size() / 3);
        dh.box_lo = lo.data();
        dh.box_hi = hi.data();
        dh.box_rank = nullptr;                 // the whole flattened domain on this rank (replicated when nranks > 1)
        dh.rank = 0;
        dh.nranks = 1;
        SomarSolverParams ph = prm->horiz;
        ph.bottom_normType = prm->norm_type;   // :387
        if ((rc = level_create_impl(&dh, &ph, L->device, nullptr, nullptr, &S->horiz))) return fail(rc);
        const int nxg = g.gn[0], nyg = g.gn[1];
        if (D.nranks > 1) {
            S->row0.assign(D.nranks, 1 << 30);
            S->rows.assign(D.nranks, 0);
            std::vector<int> top(D.nranks, -1);
            for (int b = 0; b < D.nboxes; ++b) {
                const int r = L->box_rank[b];
                S->row0[r] = std::min(S->row0[r], L->box_lo[3 * b + 1]);
                top[r] = std::max(top[r], L->box_hi[3 * b + 1]);
            }
            for (int r = 0; r < D.nranks; ++r) {
                if (top[r] < 0) { set_error("leptic: a rank owns no box"); return fail(SOMAR_ERR_ARG); }
                S->rows[r] = top[r] - S->row0[r] + 1;
                S->maxRows = std::max(S->maxRows, S->rows[r]);
            }
            const size_t chunk = (size_t)(S->maxRows + 1) * (size_t)(nxg + 1);
            if (cudaMalloc(&S->gsend, chunk * sizeof(double)) != cudaSuccess ||
                cudaMalloc(&S->gall, chunk * D.nranks * sizeof(double)) != cudaSuccess ||
                cudaMalloc(&S->gdense, (size_t)(nyg + 1) * (nxg + 1) * sizeof(double)) != cudaSuccess) {
                set_error("leptic: out of device memory (gather buffers)");
                return fail(SOMAR_ERR_CUDA);
            }
        }
        // vertically averaged J g^{00}, J g^{11}: scale = 1 / Nz (LevelGeometryBasics.cpp:545-568), on this rank's rows
        const int nc = fullMetric ? 2 : 1;     // components per face direction, component outermost
        const long long n0 = (long long)(g.n[0] + 1) * g.n[1] * nc, n1 = (long long)g.n[0] * (g.n[1] + 1) * nc;
        const long long G0 = (long long)(nxg + 1) * nyg * nc, G1 = (long long)nxg * (nyg + 1) * nc, GC = (long long)nxg * nyg;
        double *dense = nullptr, *glob = nullptr;
        if (cudaMalloc(&dense, sizeof(double) * (n0 + n1)) != cudaSuccess || cudaMalloc(&glob, sizeof(double) * (G0 + G1 + GC)) != cudaSuccess) {
            cudaFree(dense);
            set_error("leptic: out of device memory");
            return fail(SOMAR_ERR_CUDA);
        }
        cudaStream_t st = S->vert->cx.stream;
        const double scale = 1.0 / (double)g.n[2];
        const Depth& dv = S->vert->depths[0];
        if (fullMetric) {
            for (int e = 0; e < 2; ++e) {
                k_vert_avg_dense<<<plane_grid(g, 1, 0), QTPB, 0, st>>>(g, dense + e * (n0 / 2), dv.jg[0][e], 1, 0, scale);
                k_vert_avg_dense<<<plane_grid(g, 0, 1), QTPB, 0, st>>>(g, dense + n0 + e * (n1 / 2), dv.jg[1][e], 0, 1, scale);
            }
            S->vert->cx.launches += 2;
        } else {
            k_vert_avg_dense<<<plane_grid(g, 1, 0), QTPB, 0, st>>>(g, dense, dv.jg[0][0], 1, 0, scale);
            k_vert_avg_dense<<<plane_grid(g, 0, 1), QTPB, 0, st>>>(g, dense + n0, dv.jg[1][1], 0, 1, scale);
        }
        k_fill_dense<<<64, 256, 0, st>>>(glob + G0 + G1, GC, 1.0);   // "because we are solving L[phi] = J*rhs" (:428-430)
        S->vert->cx.launches += 3;
        rc = SOMAR_OK;
        if (D.nranks > 1) {
            rc = gather_rows(S, dense, g.n[0] + 1, nxg + 1, 0, glob);
            if (!rc) rc = gather_rows(S, dense + n0, g.n[0], nxg, 1, glob + G0);
        } else {
            cudaMemcpyAsync(glob, dense, sizeof(double) * (n0 + n1), cudaMemcpyDeviceToDevice, st);
        }
        cudaError_t ce = cudaStreamSynchronize(st);
        if (!rc && ce == cudaSuccess) {
            rc = somar_level_set_metric_device(S->horiz, nc, glob, glob + G0, nullptr, glob + G0 + G1);
            if (!rc) rc = somar_level_finalize(S->horiz);
        }
        cudaFree(dense);
        cudaFree(glob);
        if (ce != cudaSuccess) { set_error(std::string("leptic: ") + cudaGetErrorString(ce)); return fail(SOMAR_ERR_CUDA); }
        if (rc) return fail(rc);
        if ((rc = field_alloc(S->horiz, 0, &S->hPhi)) || (rc = field_alloc(S->horiz, 0, &S->hRhs))) return fail(rc);
        launch_fill(S->horiz->cx, S->horiz->depths[0].g, S->hPhi->p, 0.0);
        launch_fill(S->horiz->cx, S->horiz->depths[0].g, S->hRhs->p, 0.0);
    }
    if (cudaEventCreate(&S->ev0) != cudaSuccess || cudaEventCreate(&S->ev1) != cudaSuccess) { set_error("leptic: event"); return fail(SOMAR_ERR_CUDA); }
    if ((rc = sync(S->vert))) return fail(rc);
    *out = S;
    return SOMAR_OK;
}

int somar_leptic_level(SomarLeptic* S, int which, SomarLevel** out) {
    if (!S || !out || which < 0 || which > 1) { set_error("leptic_level: bad argument"); return SOMAR_ERR_ARG; }
    *out = which == 0 ? S->vert : S->horiz;
    return SOMAR_OK;
}

int somar_leptic_solve(SomarLeptic* S, SomarField* phi, const SomarField* rhs, int homogeneous, SomarLepticStats* st) {
    if (!S || !phi || !rhs || !st) { set_error("null argument"); return SOMAR_ERR_ARG; }
    SomarLevel *L = S->orig, *V = S->vert, *Hz = S->horiz;
    if (phi->lev != L || rhs->lev != L || phi->depth != 0 || rhs->depth != 0 || phi == rhs) {
        set_error("leptic solve: phi and rhs must be distinct depth-0 fields of the level handed to somar_leptic_create");
        return SOMAR_ERR_ARG;
    }
    cudaSetDevice(L->device);
    const SomarLepticParams& P = S->prm;
    const Geom& g = V->depths[0].g;
    const Depth& dv = V->depths[0];
    const int vd = g.nd - 1;
    const double dz = g.dx[vd];
    const double H = dz * (double)g.gn[vd];       // LevelGeometry::getDomainLength(SpaceDim-1) (:661-662)
    const int maxOrder = P.max_order;
    const int kz = 8;
    int rc;
    std::memset(st, 0, sizeof(*st));
    st->exit_status = -1;
    const long long launches0 = L->cx.launches + V->cx.launches + (Hz ? Hz->cx.launches : 0);
    if ((rc = sync(L))) return rc;
    SOMAR_CUDA(cudaEventRecord(S->ev0, V->cx.stream));

    // J * residual to the vertical grids (:698-716); the bricks of the two levels have the same layout
    if ((rc = op_apply(L, 0, S->res->p, phi, rhs->p, 1, homogeneous != 0))) return rc;
    k_div_cells<<<plane_grid(g, 0, 0, (g.n[2] + kz - 1) / kz), QTPB, 0, L->cx.stream>>>(g, S->rhsA->p, S->res->p,
                                                                                      L->depths[0].jinv, kz);
    L->cx.launches++;
    if ((rc = sync(L))) return rc;
    SomarField *rhsP = S->rhsA, *tmpP = S->rhsB;

    double resNorm = 0.0;
    if ((rc = norm_of(V, rhsP, P.norm_type, &resNorm))) return rc;
    st->res_norms[st->num_res_norms++] = resNorm;
    const double res0 = resNorm;
    if ((rc = flat_zero(S, S->bdryLo)) || (rc = flat_zero(S, S->bdryHi))) return rc;      // bdryData.setVal(0) (:733)
    launch_fill(V->cx, g, S->phiTotal->p, 0.0);                                            // :746
    launch_fill(V->cx, g, S->vertPhi->p, 0.0);
    bool useExcess = S->doHoriz, useHoriz = S->doHoriz;
    cudaStream_t vs = V->cx.stream;

    for (int order = 0; order <= maxOrder; ++order) {
        st->orders = order + 1;
        if (S->doHoriz) {
            if (order >= 1 && dv.m.diagonal) {
                // levelVertHorizGradient: zero for a diagonal metric (:1122-1125)
                if ((rc = flat_zero(S, S->bdryLo)) || (rc = flat_zero(S, S->bdryHi))) return rc;
            } else if (order >= 1) {
                // -Grad[phi_{order-1}]^z on the vertical boundary faces (:760-762)
                k_vert_horiz_grad<<<plane_grid(g), QTPB, 0, vs>>>(g, dv.jg[2][0], dv.jg[2][1], S->vertPhi->p, S->bdryLo->p,
                                                                  S->bdryHi->p, -1.0);
                V->cx.launches++;
            }
            if (order >= 1 && useExcess) {                                                 // vertPlus(excess, 1, Hi) (:766)
                k_flat_axpy<<<plane_grid(g), QTPB, 0, vs>>>(g, S->bdryHi->p, S->excess->p, 1.0);
                V->cx.launches++;
            }
            if (useExcess) {
                if ((rc = leptic_vertical_excess(V, S->excess, rhsP, S->bdryLo, S->bdryHi))) return rc;
                if (order == 1) useExcess = false;
            }
            if (order == 0 && useExcess) {                                                 // :786-788
                k_flat_axpy<<<plane_grid(g), QTPB, 0, vs>>>(g, S->bdryHi->p, S->excess->p, -1.0);
                V->cx.launches++;
            }
        }
        // vertical problem (:795); dzCrse is only read on coarse-fine sides, which one level does not have
        if ((rc = leptic_vertical_solve(V, S->vertPhi, rhsP, S->vertBC[0], S->vertBC[1], dz, S->bdryLo, S->bdryHi))) return rc;

        if (useHoriz) {
            // flatRhs = computeHorizRHS [- excess / H] (:800-808); "this exchanges vertPhi" (:1006)
            if ((rc = exchange(V, 0, S->vertPhi->p, 1))) return rc;
            if (dv.m.diagonal)
                k_horiz_rhs<<<plane_grid(g), QTPB, 0, vs>>>(g, S->flatRhs->p, S->vertPhi->p, dv.jg[0][0], dv.jg[1][1],
                                                            useExcess ? S->excess->p : nullptr, -1.0 / H, 1.0 / (double)g.n[2]);
            else
                k_horiz_rhs_full<<<plane_grid(g), QTPB, 0, vs>>>(g, dv.m, S->flatRhs->p, S->vertPhi->p,
                                                                 useExcess ? S->excess->p : nullptr, -1.0 / H,
                                                                 1.0 / (double)g.n[2]);
            V->cx.launches++;
            // to the horizontal grids (:811-812)
            const Geom& gh = Hz->depths[0].g;
            if (V->desc.nranks > 1) {
                if ((rc = gather_rows(S, S->flatRhs->p + g.idx(0, 0, 0), g.sy, gh.n[0], 0, S->gdense))) return rc;
                SOMAR_CUDA(cudaMemcpy2DAsync(S->hRhs->p + gh.idx(0, 0, 0), gh.sy * sizeof(double), S->gdense,
                                             gh.n[0] * sizeof(double), gh.n[0] * sizeof(double), gh.n[1], cudaMemcpyDeviceToDevice, vs));
            } else {
                SOMAR_CUDA(cudaMemcpy2DAsync(S->hRhs->p + gh.idx(0, 0, 0), gh.sy * sizeof(double), S->flatRhs->p + g.idx(0, 0, 0),
                                             g.sy * sizeof(double), g.n[0] * sizeof(double), g.n[1], cudaMemcpyDeviceToDevice, vs));
            }
            if ((rc = sync(V))) return rc;
            double hnorm = 0.0;
            if ((rc = norm_of(Hz, S->hRhs, P.norm_type, &hnorm))) return rc;
            if (P.horiz_rhs_tol * res0 > hnorm) useHoriz = false;                          // :820-825
        }
        if (useHoriz) {
            // horizontalSolver (:1427-1462) + addHorizontalCorrection (:1489-1515)
            const Geom& gh = Hz->depths[0].g;
            SomarSolveStats hs;
            if ((rc = mg_solve(Hz, S->hPhi, S->hRhs, true, true, &hs))) return rc;
            st->horiz_iters_total += hs.iters;
            st->horiz_solves++;
            // setZeroAvg: the flattened boxes cover the horizontal domain (:410-433)
            double hsum = 0.0;
            if ((rc = reduce_global(Hz, 0, S->hPhi->p, nullptr, 5, &hsum))) return rc;   // (the whole horizontal domain is here)
            k_sub_scalar<<<plane_grid(gh), QTPB, 0, Hz->cx.stream>>>(gh, S->hPhi->p, hsum / ((double)gh.gn[0] * (double)gh.gn[1]));
            Hz->cx.launches++;
            SOMAR_CUDA(cudaMemcpy2DAsync(S->flatPhi->p + g.idx(0, 0, 0), g.sy * sizeof(double), S->hPhi->p + gh.idx(0, g.g0[1], 0),
                                         gh.sy * sizeof(double), g.n[0] * sizeof(double), g.n[1], cudaMemcpyDeviceToDevice,
                                         Hz->cx.stream));
            if ((rc = sync(Hz))) return rc;
            if ((rc = leptic_add_extrusion(V, S->vertPhi, S->flatPhi))) return rc;
        }

        // finalize the order (:837-927)
        if ((rc = op_apply(V, 0, tmpP->p, S->vertPhi, rhsP->p, 1, true))) return rc;
        if ((rc = norm_of(V, tmpP, P.norm_type, &resNorm))) return rc;
        double relResNorm = resNorm / res0;
        const double prevRel = st->res_norms[st->num_res_norms - 1] / res0;
        double redu = prevRel - relResNorm;
        if (redu <= P.hang && order == maxOrder) {
            // not converging: the 3-D multigrid solver, zero initial guess, homogeneous (:851-875)
            SomarSolveStats fs;
            if ((rc = mg_solve(V, S->vertPhi, rhsP, true, true, &fs))) return rc;
            if ((rc = op_apply(V, 0, tmpP->p, S->vertPhi, rhsP->p, 1, true))) return rc;
            if ((rc = norm_of(V, tmpP, P.norm_type, &resNorm))) return rc;
            relResNorm = resNorm / res0;
            st->krylov_used = 1;
        }
        std::swap(rhsP, tmpP);                                                             // :878
        st->res_norms[st->num_res_norms++] = resNorm;
        redu = prevRel - relResNorm;
        if (redu > P.hang || order < maxOrder) {
            launch_axpy(V->cx, g, S->phiTotal->p, S->vertPhi->p, 1.0);                     // addVerticalCorrection (:885)
            st->exit_status = (order < maxOrder - 1) ? 0 : 1;
        } else if (-redu > P.hang) {
            st->exit_status = (order == 0) ? 4 : 3;
            break;
        } else {
            st->exit_status = (order == 0) ? 4 : 2;
            break;
        }
        if (dv.m.diagonal) useHoriz = false;                                               // LevelGeometry::isDiagonal() (:924-926)
    }

    if ((rc = sync(V))) return rc;
    if (st->exit_status != 4) {                                                            // :932-939
        launch_axpy(L->cx, L->depths[0].g, phi->p, S->phiTotal->p, 1.0);
        if ((rc = sync(L))) return rc;
    }
    SOMAR_CUDA(cudaEventRecord(S->ev1, V->cx.stream));
    SOMAR_CUDA(cudaEventSynchronize(S->ev1));
    float ms = 0.f;
    SOMAR_CUDA(cudaEventElapsedTime(&ms, S->ev0, S->ev1));
    st->solve_ms = ms;
    st->launches = L->cx.launches + V->cx.launches + (Hz ? Hz->cx.launches : 0) - launches0;
    return SOMAR_OK;
}

}  // extern "C"
