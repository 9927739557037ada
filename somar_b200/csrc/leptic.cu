// leptic.cu -- the sweeps of SOMAR's leptic solver (SURVEY.md 8(a) "LevelLepticSolver" row / 8(f)-2) as device kernels:
//   k_leptic_nn / k_leptic_ldl   LevelLepticSolver::verticalLineSolver   src/calculus/LepticSolver/LevelLepticSolver.cpp:1247-1425
//        Neum-Neum columns:      TRIDIAGPOISSONNN1DFAB                   src/utils/TridiagUtilsF.ChF:85-164
//        every other BC pair:    LEPTICLAPACKVERTICALSOLVER -> dptsv     LevelLepticSolverF.ChF:161-289 (LAPACK dpttrf + dptts2,
//                                not in /root/reference: restated from the published unblocked algorithm, as oracle/leptic.cpp)
//   k_vert_integral              UNMAPPEDVERTINTEGRAL                    src/utils/SubspaceF.ChF:33-58
//   k_add_extrusion              ADDEXTRUSION                            src/utils/SubspaceF.ChF:66-110
//   k_accum_div                  LEPTICACCUMDIV                          LevelLepticSolverF.ChF:291-331
//   k_vertical_excess            LevelLepticSolver::computeVerticalExcess LevelLepticSolver.cpp:1183-1240
// The leptic ITERATION (LevelLepticSolver::solve, :646-935: excess bookkeeping, horizontal multigrid solves on the
// flattened grid, Krylov fall-back) is host logic that is not built yet; these are its per-cell passes.
//
// One thread owns one vertical column; consecutive threads own consecutive x, so every load and store of a plane is
// a full-sector coalesced access.  The matrix comes from J g^{zz} on the fly (no D/DL/DU work arrays as in the
// Fortran); the only scratch is one field (the pivots / the gam coefficients), which the backward sweep re-reads.
// "Flat" (vertically collapsed) fields are ordinary fields whose vertical index 0 holds the data -- the
// reference's VertShifter does the same (LevelLepticSolver.cpp:1213-1226).
// HBM traffic per cell: NN 80 B (rhs, Jgzz | x, gam written; x, gam read, x written; x read, phi written),
// LDL^T 72 B.  Arithmetic is statement-for-statement the reference's (-fmad=false): bit-identical to the oracle.
#include "level.hpp"

namespace somar {

#define LTPB 128

struct Col {
    int nv;             // cells per column
    long long sv;       // vertical stride
    int nx, ny;         // horizontal extent (ny = 1 for 2-D problems)
};

__host__ __device__ inline Col col_of(const Geom& g) {
    Col c;
    if (g.nd > 2) { c.nv = g.n[2]; c.sv = g.sz; c.nx = g.n[0]; c.ny = g.n[1]; }
    else { c.nv = g.n[1]; c.sv = g.sy; c.nx = g.n[0]; c.ny = 1; }
    return c;
}
// address of vertical index 0 of the column of thread (i, j)
__device__ __forceinline__ long long col_base(const Geom& g, int i, int j) {
    return (g.nd > 2) ? g.idx(i, j, 0) : g.idx(i, 0, 0);
}

struct Roll {
    const double* lo;   // flat Neumann values (may be null)
    const double* hi;
    double sLo, sHi;    // -isign / dz
};

// rhs with the Neumann boundary values rolled in (LevelLepticSolver.cpp:1318-1352), and what the reference
// leaves in rhs afterwards: (rhs + r) + r * -1.0 (:1419-1420).
__device__ __forceinline__ double rolled(double r, bool on, double roll) { return on ? r + roll * 1.0 : r; }

// TRIDIAGPOISSONNN1DFAB: sigma = J g^zz, a(r) = sigma(r), c(r) = sigma(r+1), b = -(a + c) with the boundary face dropped.
__global__ void __launch_bounds__(LTPB) k_leptic_nn(const Geom g, const double* __restrict__ jgzz,
                                                    double* __restrict__ phi, double* __restrict__ rhs,
                                                    double* __restrict__ gam, double dz, Roll rl) {
    const Col cl = col_of(g);
    const int i = blockIdx.x * LTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= cl.nx) return;
    const long long c0 = col_base(g, i, j);
    const int nv = cl.nv;
    const long long sv = cl.sv;
    const double dxsq = dz * dz;
    const bool onLo = rl.lo != nullptr, onHi = rl.hi != nullptr;
    const double rLo = onLo ? 0.0 + rl.lo[c0] * rl.sLo : 0.0;
    const double rHi = onHi ? 0.0 + rl.hi[c0] * rl.sHi : 0.0;
    // decomposition and forward substitution
    double cr = jgzz[c0 + sv];                       // c(0)
    double bet = -cr;
    double r0 = rolled(rhs[c0], onLo, rLo);
    if (nv == 1) r0 = rolled(r0, onHi, rHi);
    double x = r0 * dxsq / bet;
    double gm = cr / bet;
    phi[c0] = x;
    gam[c0] = gm;
    double aPrev = 0.0;                              // a(r-1) for the quirk of the last row
    int r;
    for (r = 1; r <= nv - 2; ++r) {
        const long long c = c0 + sv * r;
        const double ar = cr;                        // a(r) = sigma(r) = c(r-1)
        cr = jgzz[c + sv];
        const double b = -(ar + cr);
        bet = b - ar * gm;
        x = (rhs[c] * dxsq - ar * x) / bet;
        gm = cr / bet;
        phi[c] = x;
        gam[c] = gm;
        aPrev = ar;
    }
    if (nv >= 2) {
        // "Last index is a special case": a(r-1) and the unmodified diagonal b(r) = -a(Nx-1)   (TridiagUtilsF.ChF:141-143)
        const long long c = c0 + sv * r;
        const double aLast = cr;                     // a(Nx-1) = sigma(Nx-1)
        const double am1 = (nv == 2) ? 1.2345e10 : aPrev;   // a(0) is the Fortran's filler value
        const double rr = rolled(rhs[c], onHi, rHi);
        x = (rr * dxsq - am1 * x) / (-aLast);
        phi[c] = x;
    }
    // back substitution
    double avg = x;
    for (r = nv - 2; r >= 0; --r) {
        const long long c = c0 + sv * r;
        x = phi[c] - gam[c] * x;
        phi[c] = x;
        avg = avg + x;
    }
    avg = avg / (double)nv;
    for (r = 0; r < nv; ++r) {
        const long long c = c0 + sv * r;
        phi[c] = phi[c] - avg;
    }
    // what the roll-in / unroll pair leaves in rhs
    if (onLo) rhs[c0] = (rhs[c0] + rLo * 1.0) + rLo * -1.0;
    if (onHi) { const long long c = c0 + sv * (nv - 1); rhs[c] = (rhs[c] + rHi * 1.0) + rHi * -1.0; }
}

// LEPTICLAPACKVERTICALSOLVER: D(k) = (J(k) + J(k+1)) / dz^2 with the BC rows, E(k) = -J(k+1) / dz^2, B = -rhs; dptsv.
// flag[0] receives the LAPACK INFO of a column that broke down before its last pivot (the reference aborts).
__global__ void __launch_bounds__(LTPB) k_leptic_ldl(const Geom g, const double* __restrict__ jgzz,
                                                     double* __restrict__ phi, double* __restrict__ rhs,
                                                     double* __restrict__ piv, double dz, double dzCrse, int loBC,
                                                     int hiBC, Roll rl, int* __restrict__ flag) {
    const Col cl = col_of(g);
    const int i = blockIdx.x * LTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= cl.nx) return;
    const long long c0 = col_base(g, i, j);
    const int nv = cl.nv;
    const long long sv = cl.sv;
    const double invdzsq = 1.0 / (dz * dz);
    const bool onLo = rl.lo != nullptr && loBC == BC_NEUM, onHi = rl.hi != nullptr && hiBC == BC_NEUM;
    const double rLo = onLo ? 0.0 + rl.lo[c0] * rl.sLo : 0.0;
    const double rHi = onHi ? 0.0 + rl.hi[c0] * rl.sHi : 0.0;
    const double alpha = 1.0 - 2.0 * dz / (dzCrse + dz);
    auto rhsAt = [&](int k) {
        double v = rhs[c0 + sv * k];
        if (k == 0) v = rolled(v, onLo, rLo);
        if (k == nv - 1) v = rolled(v, onHi, rHi);
        return v;
    };
    auto diag = [&](int k, double jl, double jh) {       // jl = J(k), jh = J(k+1)
        if (k == 0) {
            if (loBC == BC_NEUM) return jh * invdzsq;
            if (loBC == BC_DIRI) return (2.0 * jl + jh) * invdzsq;
            return ((1.0 - alpha) * jl + jh) * invdzsq;
        }
        if (k == nv - 1) {
            if (hiBC == BC_NEUM) return jl * invdzsq;
            if (hiBC == BC_DIRI) return (jl + 2.0 * jh) * invdzsq;
            return (jl + (1.0 - alpha) * jh) * invdzsq;
        }
        return (jl + jh) * invdzsq;
    };
    // dpttrf fused with the forward substitution of dptts2 (row k needs only e(k-1), which is final)
    double jl = jgzz[c0], jh = jgzz[c0 + sv];
    double d = diag(0, jl, jh);
    double b = -rhsAt(0);
    int info = 0;
    piv[c0] = d;
    phi[c0] = b;
    for (int k = 1; k < nv; ++k) {
        if (d <= 0.0) { info = k; break; }
        const long long c = c0 + sv * k;
        const double ei = -jh * invdzsq;                 // E(k) = DL(k) = -J(k) / dz^2
        const double e = ei / d;
        jl = jh;
        jh = jgzz[c + sv];
        d = diag(k, jl, jh) - e * ei;
        b = -rhsAt(k) - b * e;
        piv[c] = d;
        phi[c] = b;
    }
    if (info == 0 && d <= 0.0) info = nv;
    if (info == 0) {
        double x = b / d;
        phi[c0 + sv * (nv - 1)] = x;
        for (int k = nv - 2; k >= 0; --k) {
            const long long c = c0 + sv * k;
            const double dk = piv[c];
            const double e = (-jgzz[c + sv] * invdzsq) / dk;
            x = phi[c] / dk - x * e;
            phi[c] = x;
        }
    } else {
        // the solve is skipped: B still holds -rhs (INFO == N is tolerated, LevelLepticSolverF.ChF:277-286)
        for (int k = 0; k < nv; ++k) phi[c0 + sv * k] = -rhsAt(k);
        if (info != nv) atomicCAS(flag, 0, info);
    }
    if (onLo) rhs[c0] = (rhs[c0] + rLo * 1.0) + rLo * -1.0;
    if (onHi) { const long long c = c0 + sv * (nv - 1); rhs[c] = (rhs[c] + rHi * 1.0) + rHi * -1.0; }
}

// flat(i,j) = flat(i,j) + func(i,j,k) * dz, k ascending
__global__ void __launch_bounds__(LTPB) k_vert_integral(const Geom g, double* __restrict__ flat,
                                                        const double* __restrict__ func, double dz) {
    const Col cl = col_of(g);
    const int i = blockIdx.x * LTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= cl.nx) return;
    const long long c0 = col_base(g, i, j);
    double s = flat[c0];
    for (int k = 0; k < cl.nv; ++k) s = s + func[c0 + cl.sv * k] * dz;
    flat[c0] = s;
}

// excess = hiBC + loBC * -1 + sum_k rhs * (-dz)
__global__ void __launch_bounds__(LTPB) k_vertical_excess(const Geom g, double* __restrict__ excess,
                                                          const double* __restrict__ rhs, const double* __restrict__ blo,
                                                          const double* __restrict__ bhi, double dzScale) {
    const Col cl = col_of(g);
    const int i = blockIdx.x * LTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= cl.nx) return;
    const long long c0 = col_base(g, i, j);
    double s = bhi[c0];
    s = s + blo[c0] * -1.0;
    for (int k = 0; k < cl.nv; ++k) s = s + rhs[c0 + cl.sv * k] * dzScale;
    excess[c0] = s;
}

// dest(i,j,k) = dest(i,j,k) + flat(i,j)
__global__ void __launch_bounds__(LTPB) k_add_extrusion(const Geom g, double* __restrict__ dest,
                                                        const double* __restrict__ flat, int kz) {
    const Col cl = col_of(g);
    const int i = blockIdx.x * LTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= cl.nx) return;
    const long long c0 = col_base(g, i, j);
    const double v = flat[c0];
    const int k0 = blockIdx.z * kz, k1 = min(k0 + kz, cl.nv);
    for (int k = k0; k < k1; ++k) dest[c0 + cl.sv * k] = dest[c0 + cl.sv * k] + v;
}

// flat rhs(i,j) = rhs + (flux(i + e_dir) - flux(i)) * dxScale
__global__ void __launch_bounds__(LTPB) k_accum_div(const Geom g, double* __restrict__ rhs,
                                                    const double* __restrict__ flux, long long st, double dxScale) {
    const Col cl = col_of(g);
    const int i = blockIdx.x * LTPB + threadIdx.x;
    const int j = blockIdx.y;
    if (i >= cl.nx) return;
    const long long c0 = col_base(g, i, j);
    rhs[c0] = rhs[c0] + (flux[c0 + st] - flux[c0]) * dxScale;
}

static inline dim3 col_grid(const Geom& g, int nz = 1) {
    const Col cl = col_of(g);
    return dim3((cl.nx + LTPB - 1) / LTPB, cl.ny, nz);
}

// ---------------------------------------------------------------------------------------------------------
int leptic_vertical_solve(SomarLevel* L, SomarField* phi, SomarField* rhs, int loBC, int hiBC, double dzCrse,
                          const SomarField* bdryLo, const SomarField* bdryHi) {
    Depth& dp = L->depths[0];
    const Geom& g = dp.g;
    const int vd = g.nd - 1;
    if (g.periodic[vd]) { set_error("leptic vertical solve: the vertical direction is periodic (LevelLepticSolver.cpp:1316)"); return SOMAR_ERR_ARG; }
    auto okBC = [](int t) { return t == 0 || t == 1 || t == 3; };
    if (!okBC(loBC) || !okBC(hiBC)) { set_error("leptic vertical solve: unforeseen vertical BC type"); return SOMAR_ERR_ARG; }
    const double dz = g.dx[vd];
    Roll rl;
    rl.lo = (loBC == 0 && bdryLo) ? bdryLo->p : nullptr;
    rl.hi = (hiBC == 0 && bdryHi) ? bdryHi->p : nullptr;
    rl.sLo = 1.0 / dz;      // -isign / dz, isign = -1
    rl.sHi = -1.0 / dz;
    const double* jgzz = dp.jg[vd][vd];
    if (loBC == 0 && hiBC == 0) {
        k_leptic_nn<<<col_grid(g), LTPB, 0, L->cx.stream>>>(g, jgzz, phi->p, rhs->p, dp.tmp, dz, rl);
        L->cx.launches++;
        return SOMAR_OK;
    }
    int* flag = reinterpret_cast<int*>(L->d_scal + 6);
    SOMAR_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), L->cx.stream));
    k_leptic_ldl<<<col_grid(g), LTPB, 0, L->cx.stream>>>(g, jgzz, phi->p, rhs->p, dp.tmp, dz, dzCrse, loBC, hiBC, rl, flag);
    L->cx.launches++;
    int info = 0;
    SOMAR_CUDA(cudaMemcpyAsync(&info, flag, sizeof(int), cudaMemcpyDeviceToHost, L->cx.stream));
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    if (info != 0) {
        set_error("LepticLapackVerticalSolver: INFO = " + std::to_string(info));
        return SOMAR_ERR_ARG;
    }
    return SOMAR_OK;
}

int leptic_vert_integral(SomarLevel* L, SomarField* flat, const SomarField* func, double dzScale) {
    const Geom& g = L->depths[0].g;
    k_vert_integral<<<col_grid(g), LTPB, 0, L->cx.stream>>>(g, flat->p, func->p, dzScale);
    L->cx.launches++;
    return SOMAR_OK;
}

int leptic_vertical_excess(SomarLevel* L, SomarField* excess, const SomarField* rhs, const SomarField* bdryLo,
                           const SomarField* bdryHi) {
    const Geom& g = L->depths[0].g;
    const double dzScale = -1.0 * g.dx[g.nd - 1];
    k_vertical_excess<<<col_grid(g), LTPB, 0, L->cx.stream>>>(g, excess->p, rhs->p, bdryLo->p, bdryHi->p, dzScale);
    L->cx.launches++;
    return SOMAR_OK;
}

int leptic_add_extrusion(SomarLevel* L, SomarField* dest, const SomarField* flat) {
    const Geom& g = L->depths[0].g;
    const Col cl = col_of(g);
    const int kz = 8;
    k_add_extrusion<<<col_grid(g, (cl.nv + kz - 1) / kz), LTPB, 0, L->cx.stream>>>(g, dest->p, flat->p, kz);
    L->cx.launches++;
    return SOMAR_OK;
}

int leptic_accum_div(SomarLevel* L, SomarField* flatRhs, const SomarField* flatFlux, int dir, double dxScale) {
    const Geom& g = L->depths[0].g;
    if (dir < 0 || dir >= g.nd - 1) { set_error("LepticAccumDiv: FCdir must be a horizontal direction"); return SOMAR_ERR_ARG; }
    const long long st = (dir == 0) ? 1 : g.sy;
    k_accum_div<<<col_grid(g), LTPB, 0, L->cx.stream>>>(g, flatRhs->p, flatFlux->p, st, dxScale);
    L->cx.launches++;
    return SOMAR_OK;
}

}  // namespace somar
