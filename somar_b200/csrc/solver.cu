// solver.cu -- multigrid drivers on the level stream:
//   mg_cycle      MappedMultiGrid<T>::cycle              src/calculus/AMRElliptic/MappedMultiGrid.H:555-653
//   bottom_solve  Chombo 3.1 BiCGStabSolver<T>::solve     -> bicgstab.cu
//   mg_solve      MappedAMRMultiGrid<T>::solveNoInitResid src/calculus/AMRElliptic/MappedAMRMultiGrid.H:979-1186
//                 (l_base == l_max: AMRVCycle :1510-1513 -> oneCycle; postVCycleOps :1189-1215)
#include "level.hpp"
#include <cmath>
#include <cstdio>

namespace somar {

static inline Geom& G(SomarLevel* L, int d) { return L->depths[d].g; }

int mg_cycle(SomarLevel* L, int depth, SomarField* e, const SomarField* r, bool eIsZero) {
    const SomarSolverParams& p = L->prm;
    int rc;
    if (depth == L->ndepths() - 1) {
        const Geom& g = G(L, depth);
        const long long bottomCells = (long long)g.gn[0] * g.gn[1] * g.gn[2];
        ProfScope ps(L, PH_BOTTOM);
        if (eIsZero) launch_fill(L->cx, g, e->p, 0.0);
        if (bottomCells == 1) return op_relax(L, depth, e, r, 1);                       // :566-568
        if (bottom_persistent_usable(L)) return bottom_persistent(L, e, r);             // both steps, one launch
        if (bottom_gathered_usable(L)) return bottom_gathered(L, e, r);                 // the same on the replicated depth
        if ((rc = op_relax(L, depth, e, r, p.num_smooth_bottom)) != SOMAR_OK) return rc;
        int it = 0;
        if ((rc = bottom_solve(L, e, r, &it)) != SOMAR_OK) return rc;
        L->bottom_iters_total += it;
        return SOMAR_OK;
    }
    Depth& nx = L->depths[depth + 1];
    int cycles = p.numMG;
    if (cycles < 0) {   // F-cycle branch :578-619
        cycles = -cycles;
        if (eIsZero) launch_fill(L->cx, G(L, depth), e->p, 0.0);
        if ((rc = op_restrict(L, depth, nx.res, e, r))) return rc;
        launch_fill(L->cx, nx.g, nx.cor->p, 0.0);
        if ((rc = mg_cycle(L, depth + 1, nx.cor, nx.res))) return rc;
        if ((rc = op_prolong(L, depth, e, nx.cor))) return rc;
        if ((rc = op_relax(L, depth, e, r, p.num_smooth_down))) return rc;
        for (int img = 0; img < cycles; ++img) {
            if ((rc = op_restrict(L, depth, nx.res, e, r))) return rc;
            launch_fill(L->cx, nx.g, nx.cor->p, 0.0);
            const int save = L->prm.numMG;
            L->prm.numMG = 1;
            rc = mg_cycle(L, depth + 1, nx.cor, nx.res);
            L->prm.numMG = save;
            if (rc) return rc;
            if ((rc = op_prolong(L, depth, e, nx.cor))) return rc;
        }
        return op_relax(L, depth, e, r, p.num_smooth_up);
    }
    if ((rc = op_relax(L, depth, e, r, p.num_smooth_down, -100, eIsZero))) return rc;  // :622
    {
        ProfScope ps(L, PH_RESTRICT);
        if ((rc = op_restrict(L, depth, nx.res, e, r))) return rc;                     // :628
    }
    // "set the coarse correction to zero" (:629) is handed down: the first consumer clears it, or -- the pipelined
    // smoother -- never reads it
    for (int img = 0; img < cycles; ++img)
        if ((rc = mg_cycle(L, depth + 1, nx.cor, nx.res, img == 0))) return rc;         // :631-633
    {
        ProfScope ps(L, PH_PROLONG);
        if ((rc = op_prolong(L, depth, e, nx.cor))) return rc;                         // :638
    }
    return op_relax(L, depth, e, r, p.num_smooth_up);                                   // :641
}

// ---------------------------------------------------------------------------------------------------------
int mg_solve(SomarLevel* L, SomarField* phi, const SomarField* rhs, bool zeroPhi, bool forceHomog,
             SomarSolveStats* st) {
    const SomarSolverParams& p = L->prm;
    const Geom& g = G(L, 0);
    Ctx& cx = L->cx;
    int rc;
    if (L->swork.empty()) {
        L->swork.resize(4, nullptr);
        for (auto& f : L->swork)
            if ((rc = field_alloc(L, 0, &f))) return rc;
    }
    SomarField *corr = L->swork[0], *resid = L->swork[1], *best = L->swork[2], *spare = L->swork[3];
    // Best-phi bookkeeping (MappedAMRMultiGrid.H:1105-1117) without copies: phi += corr is written out of place into
    // `spare` and the buffers trade places, so the previous iterate survives for free.  While the newest iterate is
    // the best one nothing is kept; when an iterate gets worse the previous one (still in `spare`) becomes `best`.
    bool bestIsCurrent = true;
    const long long launches0 = cx.launches;
    SOMAR_CUDA(cudaEventRecord(L->ev0, cx.stream));
    // (the correction's "set to zero" before every V-cycle, MappedAMRMultiGrid.H:1064,1215, is handed to mg_cycle)
    if (zeroPhi) launch_fill(cx, g, phi->p, 0.0);
    auto computeResidual = [&](double* out) {
        return residual_maxnorm(L, resid, phi, rhs, forceHomog, out);
    };
    double initial_rnorm;
    if ((rc = computeResidual(&initial_rnorm))) return rc;
    if (L->convergenceMetric != 0.0) initial_rnorm = L->convergenceMetric;   // m_convergenceMetric :1036-1038
    double rnorm = initial_rnorm, norm_last = 2 * initial_rnorm, best_rnorm = rnorm;
    bool useBest = false;
    bool somethingConverged = false;       // stays false while no V-cycle improves on what we were given :1043
    L->bottomMetric = initial_rnorm;       // setConvergenceMetrics(initial_rnorm, cushion * eps) :1047
    L->bottomEps = 1.0 * p.eps;
    if ((rc = bottom_collect_iters(L))) return rc;    // counts left by direct mg_cycle calls
    L->bottom_iters_total = 0;
    int hl = 0;
    st->history[hl++] = rnorm;
    int iter = 0;
    bool goNorm = rnorm > p.normThresh;
    bool goRedu = rnorm > p.eps * initial_rnorm;
    bool goIter = iter < p.imax;
    bool goHang = iter < p.imin || rnorm < (1 - p.hang) * norm_last;
    while (goIter && goRedu && goHang && goNorm) {
        norm_last = rnorm;
        if ((rc = mg_cycle(L, 0, corr, resid, true))) return rc; // AMRVCycle, single level; corr = 0 on entry
        {
            ProfScope ps(L, PH_OUTER);
            launch_axby(cx, g, spare->p, phi->p, corr->p, 1.0, 1.0);   // postVCycleOps: phi + corr (1.0 * x is exact)
            std::swap(spare->p, phi->p);                               // phi = the new iterate, spare = the previous one
            if ((rc = computeResidual(&rnorm))) return rc;
            ++iter;
            if (hl < 128) st->history[hl++] = rnorm;
            if (rnorm <= best_rnorm) {
                best_rnorm = rnorm;
                bestIsCurrent = true;
                useBest = false;
                somethingConverged = true;
            } else {
                if (bestIsCurrent) std::swap(best->p, spare->p);   // the previous iterate was the best one: keep it
                bestIsCurrent = false;
                useBest = true;
            }
        }
        if (p.verbosity >= 3)
            fprintf(stderr, "   %s somar_b200 MG iteration = %d, residual norm = %.6e, rate = %.3f\n", useBest ? "[D]" : "   ",
                    iter, rnorm, rnorm > 0 ? norm_last / rnorm : 0.0);
        goNorm = rnorm > p.normThresh;
        goRedu = rnorm > p.eps * initial_rnorm;
        goIter = iter < p.imax;
        goHang = iter < p.imin || rnorm < (1 - p.hang) * norm_last;
    }
    if (useBest) {
        rnorm = best_rnorm;
        std::swap(phi->p, best->p);
    }
    SOMAR_CUDA(cudaEventRecord(L->ev1, cx.stream));
    SOMAR_CUDA(cudaEventSynchronize(L->ev1));
    float ms = 0.f;
    SOMAR_CUDA(cudaEventElapsedTime(&ms, L->ev0, L->ev1));
    if ((rc = bottom_collect_iters(L))) return rc;    // iterations the persistent bottom kernel counted on the device
    st->iters = iter;
    st->initial_rnorm = initial_rnorm;
    st->final_rnorm = rnorm;
    st->exitStatus = int(!goRedu) + int(!goIter) * 2 + int(!goHang) * 4 + int(!goNorm) * 8;   // :1148
    st->bottom_iters_total = L->bottom_iters_total;
    st->history_len = hl;
    st->solve_ms = ms;
    st->launches = cx.launches - launches0;
    if (rnorm > 10.0 * initial_rnorm && rnorm > 10.0 * p.eps) {   // :1133-1138 "kaboom"
        set_error("solver diverged (residual > 10 x initial)");
        return SOMAR_ERR_DIVERGED;
    }
    if (!somethingConverged && rnorm >= initial_rnorm && rnorm >= p.eps) {   // :1140-1144 "solver blew up"
        set_error("MappedAMRMultiGrid solver blew up (no V-cycle improved on the initial guess)");
        return SOMAR_ERR_DIVERGED;
    }
    return SOMAR_OK;
}

}  // namespace somar
