// oracle/leptic.cpp -- TEST INFRASTRUCTURE ONLY (CPU oracle; never the product path).
//
// Plain C++ restatement of the sweeps of SOMAR's leptic solver (SURVEY.md 8(a) "LevelLepticSolver" row, 8(f)-2):
//   orc_leptic_vertical_solve   LevelLepticSolver::verticalLineSolver      src/calculus/LepticSolver/LevelLepticSolver.cpp:1247-1425
//                               -> TRIDIAGPOISSONNN1DFAB                    src/utils/TridiagUtilsF.ChF:85-164   (Neum-Neum columns)
//                               -> LEPTICLAPACKVERTICALSOLVER               src/calculus/LepticSolver/LevelLepticSolverF.ChF:161-289
//                                  (LAPACK dptsv = dpttrf + dpttrs; LAPACK is NOT in /root/reference and its version is
//                                  unpinned -- restated from the published reference algorithm, unblocked order)
//   orc_leptic_vert_integral    UNMAPPEDVERTINTEGRAL                        src/utils/SubspaceF.ChF:33-58
//   orc_leptic_add_extrusion    ADDEXTRUSION                                src/utils/SubspaceF.ChF:66-110
//   orc_leptic_accum_div        LEPTICACCUMDIV                              LevelLepticSolverF.ChF:291-331
//   orc_leptic_vertical_excess  LevelLepticSolver::computeVerticalExcess    LevelLepticSolver.cpp:1183-1240
// PINNED to the reference's own Fortran for TRIDIAGPOISSONNN1DFAB, UNMAPPEDVERTINTEGRAL, ADDEXTRUSION and LEPTICACCUMDIV
// (executed through oracle/chf2py.py; tests/test_reference_kernels.py case "leptic", identical bits); the LAPACK branch
// (dptsv) is pinned by tests/test_leptic.py to the exact solution of the tridiagonal systems.
//
// Arrays are global, x fastest: cell fields (nv, [ny,] nx), J g^{zz} on vertical faces (nv+1, [ny,] nx), flat fields
// ([ny,] nx).  `nh` = number of columns (nx*ny), `nv` = cells per column; the vertical stride is nh.
#include <cmath>
#include <cstdio>
#include <vector>

extern "C" {

enum { BCT_NEUM = 0, BCT_DIRI = 1, BCT_CF = 3 };

// Returns 0, or the LAPACK INFO of the first column whose factorisation broke down before the last pivot
// (the reference aborts there, LevelLepticSolverF.ChF:277-280).  bdry_lo / bdry_hi (flat, may be NULL): Neumann
// boundary values rolled into the right-hand side, rhs -/+ bc/dz (LevelLepticSolver.cpp:1318-1352); rhs is left as
// the reference leaves it: (rhs + r) - r.
int orc_leptic_vertical_solve(double* phi, double* rhs, const double* jgzz, long nh, int nv, double dz, double dzCrse,
                              int loBC, int hiBC, const double* bdry_lo, const double* bdry_hi) {
    int bad = 0;
    const double invdzsq = 1.0 / (dz * dz);
#pragma omp parallel for schedule(static)
    for (long c = 0; c < nh; ++c) {
        auto J = [&](int k) { return jgzz[c + nh * (long)k]; };
        auto R = [&](int k) -> double& { return rhs[c + nh * (long)k]; };
        auto P = [&](int k) -> double& { return phi[c + nh * (long)k]; };
        // roll in Neumann values: scale = -isign/dz (:1343)
        double rl = 0.0, rh = 0.0;
        const bool rollLo = (loBC == BCT_NEUM) && bdry_lo, rollHi = (hiBC == BCT_NEUM) && bdry_hi;
        if (rollLo) { rl = 0.0 + bdry_lo[c] * (1.0 / dz); R(0) = R(0) + rl * 1.0; }
        if (rollHi) { rh = 0.0 + bdry_hi[c] * (-1.0 / dz); R(nv - 1) = R(nv - 1) + rh * 1.0; }
        if (loBC == BCT_NEUM && hiBC == BCT_NEUM) {
            // TRIDIAGPOISSONNN1DFAB -- TridiagUtilsF.ChF:85-164 (sigma = J g^zz, Nx = nv, dx = dz)
            const double dxsq = dz * dz;
            std::vector<double> x(nv), gam(nv > 1 ? nv - 1 : 1);
            auto a = [&](int r) { return r == 0 ? 1.2345e10 : J(r); };   // a(0) is the Fortran's filler (:113)
            auto cc = [&](int r) { return J(r + 1); };
            double bet = -cc(0);
            x[0] = R(0) * dxsq / bet;
            gam[0] = cc(0) / bet;
            int r;
            for (r = 1; r <= nv - 2; ++r) {
                const double b = -(a(r) + cc(r));
                bet = b - a(r) * gam[r - 1];
                x[r] = (R(r) * dxsq - a(r) * x[r - 1]) / bet;
                gam[r] = cc(r) / bet;
            }
            // "Last index is a special case" (:141-143): a(r-1), and the unmodified diagonal b(r) = -a(Nx-1)
            x[r] = (R(r) * dxsq - a(r - 1) * x[r - 1]) / (-a(nv - 1));
            double avg = x[r];
            for (r = nv - 2; r >= 0; --r) {
                x[r] = x[r] - gam[r] * x[r + 1];
                avg = avg + x[r];
            }
            avg = avg / (double)nv;
            for (r = 0; r < nv; ++r) P(r) = x[r] - avg;
        } else {
            // LEPTICLAPACKVERTICALSOLVER -- LevelLepticSolverF.ChF:161-289; D/DL/B are 1-based there
            std::vector<double> D(nv), E(nv > 1 ? nv - 1 : 1), B(nv);
            D[0] = (J(0) + J(1)) * invdzsq;
            if (nv > 1) E[0] = -J(1) * invdzsq;
            B[0] = -R(0);
            for (int k = 2; k <= nv - 1; ++k) {
                D[k - 1] = (J(k - 1) + J(k)) * invdzsq;
                E[k - 1] = -J(k) * invdzsq;
                B[k - 1] = -R(k - 1);
            }
            D[nv - 1] = (J(nv - 1) + J(nv)) * invdzsq;
            B[nv - 1] = -R(nv - 1);
            const double alpha = 1.0 - 2.0 * dz / (dzCrse + dz);
            if (loBC == BCT_NEUM) D[0] = J(1) * invdzsq;
            else if (loBC == BCT_DIRI) D[0] = (2.0 * J(0) + J(1)) * invdzsq;
            else D[0] = ((1.0 - alpha) * J(0) + J(1)) * invdzsq;
            if (hiBC == BCT_NEUM) D[nv - 1] = J(nv - 1) * invdzsq;
            else if (hiBC == BCT_DIRI) D[nv - 1] = (J(nv - 1) + 2.0 * J(nv)) * invdzsq;
            else D[nv - 1] = (J(nv - 1) + (1.0 - alpha) * J(nv)) * invdzsq;
            // dpttrf (L D L^T of an SPD tridiagonal matrix)
            int info = 0;
            for (int i = 0; i < nv - 1; ++i) {
                if (D[i] <= 0.0) { info = i + 1; break; }
                const double ei = E[i];
                E[i] = ei / D[i];
                D[i + 1] = D[i + 1] - E[i] * ei;
            }
            if (info == 0 && D[nv - 1] <= 0.0) info = nv;
            if (info == 0) {
                // dpttrs -> dptts2
                for (int i = 1; i < nv; ++i) B[i] = B[i] - B[i - 1] * E[i - 1];
                B[nv - 1] = B[nv - 1] / D[nv - 1];
                for (int i = nv - 2; i >= 0; --i) B[i] = B[i] / D[i] - B[i + 1] * E[i];
            } else if (info != nv) {
#pragma omp critical
                if (!bad) bad = info;
            }
            for (int k = 0; k < nv; ++k) P(k) = B[k];   // INFO == N is tolerated: B is still -rhs (:277-286)
        }
        // unroll (:1419-1420)
        if (rollLo) R(0) = R(0) + rl * -1.0;
        if (rollHi) R(nv - 1) = R(nv - 1) + rh * -1.0;
    }
    return bad;
}

// levelSum(i,j) += func(i,j,k) * dz, k ascending -- SubspaceF.ChF:52-55
void orc_leptic_vert_integral(double* flat, const double* func, long nh, int nv, double dz) {
    for (int k = 0; k < nv; ++k)
        for (long c = 0; c < nh; ++c) flat[c] = flat[c] + func[c + nh * (long)k] * dz;
}

// dest(i,j,k) += src(i,j) -- SubspaceF.ChF:66-110
void orc_leptic_add_extrusion(double* dest, const double* flat, long nh, int nv) {
    for (int k = 0; k < nv; ++k)
        for (long c = 0; c < nh; ++c) dest[c + nh * (long)k] = dest[c + nh * (long)k] + flat[c];
}

// rhs(i) += (flux(i + e_dir) - flux(i)) * dxScale on a flat grid -- LevelLepticSolverF.ChF:291-331.
// flux has nx+1 (dir 0) or ny+1 (dir 1) points in its own direction.
void orc_leptic_accum_div(double* rhs, const double* flux, int nx, int ny, int dir, double dxScale) {
    const long fx = nx + (dir == 0);
    for (int j = 0; j < ny; ++j)
        for (int i = 0; i < nx; ++i) {
            const long f = i + fx * (long)j;
            const long fp = (dir == 0) ? f + 1 : f + fx;
            rhs[i + (long)nx * j] = rhs[i + (long)nx * j] + (flux[fp] - flux[f]) * dxScale;
        }
}

// excess = hiBC - loBC - Integral[rhs] dz -- LevelLepticSolver.cpp:1206-1238
void orc_leptic_vertical_excess(double* excess, const double* rhs, const double* bdry_lo, const double* bdry_hi,
                                long nh, int nv, double dz) {
    for (long c = 0; c < nh; ++c) excess[c] = bdry_hi[c];
    for (long c = 0; c < nh; ++c) excess[c] = excess[c] + bdry_lo[c] * -1.0;
    orc_leptic_vert_integral(excess, rhs, nh, nv, -1.0 * dz);
}

}  // extern "C"
