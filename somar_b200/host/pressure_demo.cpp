// pressure_demo.cpp -- the C++ host side driving the library the way SOMAR's projector drives AMRPressureSolver
// (BaseProjector::project -> AMRPressureSolver::levelSolve, src/projection/BaseProjectorI.H:176-300): host
// "LevelData" FABs (one per box, with ghost cells, Fortran order), metric handed over box by box, solve, read back.
// Everything numerical happens on the GPU behind include/somar_b200.h; this file contains no arithmetic of the
// solver.  It checks the result against a discrete manufactured solution (rhs = L[phi*] computed by the device
// operator itself), so it needs no oracle.   Usage: pressure_demo [nx ny nz] [relax_mode]
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "somar_b200/adapter.hpp"

using namespace somar_b200;

namespace {

const double kPi = 3.14159265358979323846;
const double kSigma = 1.2;   // vertical stretching z = H s(zeta/H), s(f) = 1 - tanh(sigma (1-f)) / tanh(sigma)

double dzdzeta(double f) {
    const double c = std::cosh(kSigma * (1.0 - f));
    return kSigma / (c * c) / std::tanh(kSigma);
}

struct BoxFab {
    Int3 lo, hi;
    int ghost;
    std::vector<double> v;
    BoxFab(const Int3& l, const Int3& h, int g) : lo(l), hi(h), ghost(g) {
        size_t n = 1;
        for (int d = 0; d < 3; ++d) n *= (size_t)(h[d] - l[d] + 1 + 2 * g);
        v.assign(n, 0.0);
    }
    double& at(int i, int j, int k) {
        const int nx = hi[0] - lo[0] + 1 + 2 * ghost, ny = hi[1] - lo[1] + 1 + 2 * ghost;
        return v[(size_t)(i - lo[0] + ghost) + (size_t)nx * ((j - lo[1] + ghost) + (size_t)ny * (k - lo[2] + ghost))];
    }
    HostFab view() {
        HostFab f;
        f.data = v.data();
        for (int d = 0; d < 3; ++d) {
            f.lo[d] = lo[d] - ghost;
            f.hi[d] = hi[d] + ghost;
            f.vlo[d] = lo[d];
            f.vhi[d] = hi[d];
        }
        return f;
    }
};

}  // namespace

int main(int argc, char** argv) {
    Int3 n{{64, 64, 32}};
    if (argc >= 4)
        for (int d = 0; d < 3; ++d) n[d] = std::atoi(argv[1 + d]);
    const int relaxMode = (argc >= 5) ? std::atoi(argv[4]) : SOMAR_RELAX_LEVEL_GSRB;
    // "leptic": a thin domain (dx / H = 4) with a Dirichlet top, solved by LevelLepticSolver instead of multigrid
    const bool leptic = argc >= 6 && std::string(argv[5]) == "leptic";
    const double aniso = leptic ? 4.0 * n[2] : 4.0;
    try {
        // ---- geometry: 2 x 2 x 1 boxes, all on this rank -------------------------------------------------------
        LevelGeometryDesc geo;
        geo.n = n;
        geo.dx = {{1.0 / n[0], 1.0 / n[0], 1.0 / (aniso * n[0])}};
        const double Hc = geo.dx[2] * n[2];   // computational depth
        if (leptic) geo.bcType[5] = SOMAR_BC_DIRI;
        for (int by = 0; by < 2; ++by)
            for (int bx = 0; bx < 2; ++bx) {
                geo.boxLo.push_back({{bx * n[0] / 2, by * n[1] / 2, 0}});
                geo.boxHi.push_back({{(bx + 1) * n[0] / 2 - 1, (by + 1) * n[1] / 2 - 1, n[2] - 1}});
            }
        PressureSolver solver;
        //                        imin imax eps   maxDepth precond down up bottom precondMode relaxMode numMG hang   thresh verb
        solver.setAMRMGParameters(0, 30, 1e-10, -1, 2, 4, 4, 2, 1, relaxMode, 1, 1e-15, 1e-30, 0);
        Level& lev = solver.levelDefine(geo);

        // ---- metric of each box: J = dz/dzeta; J g^xx = J g^yy = J (at x/y faces), J g^zz = 1/J (at z faces) -----
        for (size_t b = 0; b < geo.boxLo.size(); ++b) {
            const Int3 lo = geo.boxLo[b], hi = geo.boxHi[b];
            const int bx = hi[0] - lo[0] + 1, by = hi[1] - lo[1] + 1, bz = hi[2] - lo[2] + 1;
            std::vector<double> j0((size_t)(bx + 1) * by * bz), j1((size_t)bx * (by + 1) * bz), j2((size_t)bx * by * (bz + 1)),
                ji((size_t)bx * by * bz);
            for (int k = 0; k <= bz; ++k) {
                const double Jc = dzdzeta((lo[2] + k + 0.5) * geo.dx[2] / Hc), Jf = dzdzeta((lo[2] + k) * geo.dx[2] / Hc);
                for (int j = 0; j <= by; ++j)
                    for (int i = 0; i <= bx; ++i) {
                        if (k < bz && j < by) j0[i + (size_t)(bx + 1) * (j + (size_t)by * k)] = Jc;
                        if (k < bz && i < bx) j1[i + (size_t)bx * (j + (size_t)(by + 1) * k)] = Jc;
                        if (i < bx && j < by) j2[i + (size_t)bx * (j + (size_t)by * k)] = 1.0 / Jf;
                        if (i < bx && j < by && k < bz) ji[i + (size_t)bx * (j + (size_t)by * k)] = 1.0 / Jc;
                    }
            }
            lev.setMetric(lo, hi, 1, j0.data(), j1.data(), j2.data(), ji.data());
        }
        solver.ready();

        // ---- manufactured discrete problem: rhs = L[phi*] by the device operator ---------------------------------
        std::vector<BoxFab> phiStar, phi, rhs;
        LevelFabs vStar, vPhi, vRhs;
        for (size_t b = 0; b < geo.boxLo.size(); ++b) {
            phiStar.emplace_back(geo.boxLo[b], geo.boxHi[b], 1);
            phi.emplace_back(geo.boxLo[b], geo.boxHi[b], 1);
            rhs.emplace_back(geo.boxLo[b], geo.boxHi[b], 0);
        }
        for (size_t b = 0; b < geo.boxLo.size(); ++b) {
            BoxFab& f = phiStar[b];
            for (int k = f.lo[2]; k <= f.hi[2]; ++k)
                for (int j = f.lo[1]; j <= f.hi[1]; ++j)
                    for (int i = f.lo[0]; i <= f.hi[0]; ++i)
                        f.at(i, j, k) = std::cos(kPi * (i + 0.5) / n[0]) * std::cos(2 * kPi * (j + 0.5) / n[1]) *
                                        std::cos(kPi * (k + 0.5) / n[2]);
            vStar.push_back(phiStar[b].view());
            vPhi.push_back(phi[b].view());
            vRhs.push_back(rhs[b].view());
        }
        PoissonOp op(lev, 0);
        Field dStar(lev, 0), dRhs(lev, 0);
        dStar.copyFrom(vStar);
        op.applyOp(dRhs, dStar, true);
        dRhs.copyTo(vRhs);

        if (leptic) {
            // ---- LevelLepticSolver::define(op) + solve(phi, rhs) on device fields (LevelLepticSolver.cpp:147,646) ----
            LevelLepticSolver ls;
            ls.define(lev, true);
            Field dPhi(lev, 0);
            op.setToZero(dPhi);
            const SomarLepticStats ls_st = ls.solve(dPhi, dRhs);
            const std::vector<double> nr = ls.resNorms();
            bool falling = nr.size() >= 2;
            for (size_t q = 1; q < nr.size(); ++q) falling = falling && nr[q] < 0.5 * nr[q - 1];
            op.incr(dPhi, dStar, -1.0);
            const double err = op.norm(dPhi, 0), ref = op.norm(dStar, 0);
            std::printf("leptic solver %dx%dx%d: %d orders, exit status %s, |r|/|r0| = %.3e, max |phi - phi*| / max |phi*| = %.3e, "
                        "%.2f ms, %lld launches\n", n[0], n[1], n[2], ls_st.orders, ls.exitStatusStr(), nr.back() / nr.front(),
                        err / ref, ls_st.solve_ms, ls_st.launches);
            const bool ok = falling && ls_st.exit_status != 4 && nr.back() <= 1e-3 * nr.front() && err <= 1e-2 * ref;
            std::printf(ok ? "DEMO-OK\n" : "DEMO-FAILED\n");
            return ok ? 0 : 1;
        }
        // ---- the projector's call ---------------------------------------------------------------------------
        const SomarSolveStats st = solver.levelSolve(vPhi, vRhs, true, false);

        // phi is determined up to a constant (all-Neumann): compare after removing the means
        double mS = 0, mP = 0, cnt = 0;
        for (size_t b = 0; b < phi.size(); ++b)
            for (int k = phi[b].lo[2]; k <= phi[b].hi[2]; ++k)
                for (int j = phi[b].lo[1]; j <= phi[b].hi[1]; ++j)
                    for (int i = phi[b].lo[0]; i <= phi[b].hi[0]; ++i) {
                        mS += phiStar[b].at(i, j, k);
                        mP += phi[b].at(i, j, k);
                        cnt += 1;
                    }
        mS /= cnt;
        mP /= cnt;
        double err = 0;
        for (size_t b = 0; b < phi.size(); ++b)
            for (int k = phi[b].lo[2]; k <= phi[b].hi[2]; ++k)
                for (int j = phi[b].lo[1]; j <= phi[b].hi[1]; ++j)
                    for (int i = phi[b].lo[0]; i <= phi[b].hi[0]; ++i)
                        err = std::fmax(err, std::fabs((phi[b].at(i, j, k) - mP) - (phiStar[b].at(i, j, k) - mS)));
        const double rel = st.final_rnorm / st.initial_rnorm;
        std::printf("pressure_demo %dx%dx%d relax_mode %d: %d V-cycles, |r|/|r0| = %.3e, exit_status %d, "
                    "max |phi - phi*| = %.3e, %.2f ms, %lld launches, %d depths\n",
                    n[0], n[1], n[2], relaxMode, st.iters, rel, st.exitStatus, err, st.solve_ms, st.launches,
                    lev.numDepths());
        // ---- a whole projection on the device: v = J g grad(phi*) is a pure gradient, so projecting it (dt = 1)
        //      must return the pressure phi* (up to a constant) and leave a divergence-free remainder ------------
        MACProjector proj(solver);
        Field ux(lev, 0), uy(lev, 0), uz(lev, 0), pres(lev, 0), div(lev, 0);
        proj.computeGrad(ux, uy, &uz, dStar);
        proj.computeDiv(div, ux, uy, &uz);
        const double div0 = op.norm(div, 0);
        const SomarSolveStats ps = proj.project(ux, uy, &uz, pres, 1.0, true, false);
        proj.computeDiv(div, ux, uy, &uz);
        const double div1 = op.norm(div, 0);
        std::printf("projection: %d V-cycles, max|div| %.3e -> %.3e\n", ps.iters, div0, div1);
        // ---- and the cell-centred projection of the time step (LevelCCProjector): the cell-centred gradient of phi* as
        //      the velocity flux; the projection is approximate (Div_cc Grad_cc is the 2h-wide Laplacian), so the
        //      cell-centred divergence falls but does not vanish ------------------------------------------------------
        CCProjector ccProj(solver);
        Field cx(lev, 0), cy(lev, 0), cz(lev, 0), cpres(lev, 0);
        ccProj.computeGrad(cx, cy, &cz, dStar);
        ccProj.computeDiv(div, cx, cy, &cz);
        const double cdiv0 = op.norm(div, 0);
        const SomarSolveStats cs = ccProj.project(cx, cy, &cz, cpres, 1.0, nullptr, true, false);
        ccProj.computeDiv(div, cx, cy, &cz);
        const double cdiv1 = op.norm(div, 0);
        std::printf("cc projection: %d V-cycles, max|div_cc| %.3e -> %.3e\n", cs.iters, cdiv0, cdiv1);
        const bool ok = rel <= 1e-9 && err <= 1e-7 && div1 <= 1e-8 * div0 && cdiv1 < cdiv0;
        std::printf(ok ? "DEMO-OK\n" : "DEMO-FAILED\n");
        return ok ? 0 : 1;
    } catch (const Error& e) {
        std::fprintf(stderr, "pressure_demo: error %d: %s\n", e.code, e.what());
        return 2;
    }
}
