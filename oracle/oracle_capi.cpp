// oracle/oracle_capi.cpp -- TEST INFRASTRUCTURE ONLY.  C interface (ctypes) to the CPU oracle.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
//
// Fields cross this interface as GLOBAL arrays over the level's domain without ghost cells, x fastest
// (numpy shape (nz, ny, nx), C order).  Face-centred metric arrays have +1 point in their own direction
// and `jgup_ncomp` components outermost (component c of direction d is J g^{d c}).
#include "somar_oracle.hpp"
#include <omp.h>

using namespace orc;

extern "C" {

struct OrcDesc {
    int dim;             // 2 or 3 (2-D problems: n[2] == 1)
    int n[3];
    int periodic[3];
    int nboxes;
    const int* box_lo;   // nboxes x 3
    const int* box_hi;   // nboxes x 3 (inclusive)
    double dx[3];
    double alpha, beta;
    int is_diagonal;
    int bc_type[6];      // [dir*2 + side], BCType codes
    double bc_value[6];
    int jgup_ncomp;      // dim (full FluxBox) or 1 (only J g^{dd})
    const double* jgup[3];
    const double* jinv;
};

struct OrcParams {
    double eps;
    int num_smooth_down, num_smooth_up, num_smooth_bottom, num_smooth_precond;
    int numMG, imin, imax;
    double hang, normThresh;
    int maxDepth, relaxMode, precondMode;
    double bottom_eps, bottom_reps, bottom_hang, bottom_small;
    int bottom_imax, bottom_numRestarts, bottom_normType;
    int verbosity;
};

struct OrcStats {
    int iters, exitStatus;
    double initial_rnorm, final_rnorm;
    int bottom_iters_total;
    int history_len;
    double history[128];
};

struct OrcHandle {
    std::unique_ptr<MG> mg;
    int dim;
};

static void scatter(const Op& op, LevelData& ld, const double* g) {
    const Box& db = op.domain.box;
    const long nx = db.size(0), ny = db.size(1);
    for (size_t b = 0; b < op.grids.size(); ++b) {
        FAB& f = ld[b];
        ORC_LOOP(op.grids[b], i, j, k) f(i, j, k) = g[(i - db.lo[0]) + nx * ((j - db.lo[1]) + ny * (long)(k - db.lo[2]))];
    }
}
static void gather(const Op& op, const LevelData& ld, double* g) {
    const Box& db = op.domain.box;
    const long nx = db.size(0), ny = db.size(1);
    for (size_t b = 0; b < op.grids.size(); ++b) {
        const FAB& f = ld[b];
        ORC_LOOP(op.grids[b], i, j, k) g[(i - db.lo[0]) + nx * ((j - db.lo[1]) + ny * (long)(k - db.lo[2]))] = f(i, j, k);
    }
}

void orc_default_params(OrcParams* p) {
    SolverParams s;
    p->eps = s.eps; p->num_smooth_down = s.num_smooth_down; p->num_smooth_up = s.num_smooth_up;
    p->num_smooth_bottom = s.num_smooth_bottom; p->num_smooth_precond = s.num_smooth_precond;
    p->numMG = s.numMG; p->imin = s.imin; p->imax = s.imax; p->hang = s.hang; p->normThresh = s.normThresh;
    p->maxDepth = s.maxDepth; p->relaxMode = s.relaxMode; p->precondMode = s.precondMode;
    p->bottom_eps = s.bottom_eps; p->bottom_reps = s.bottom_reps; p->bottom_hang = s.bottom_hang;
    p->bottom_small = s.bottom_small; p->bottom_imax = s.bottom_imax; p->bottom_numRestarts = s.bottom_numRestarts;
    p->bottom_normType = s.bottom_normType; p->verbosity = s.verbosity;
}

static SolverParams toParams(const OrcParams* p) {
    SolverParams s;
    s.eps = p->eps; s.num_smooth_down = p->num_smooth_down; s.num_smooth_up = p->num_smooth_up;
    s.num_smooth_bottom = p->num_smooth_bottom; s.num_smooth_precond = p->num_smooth_precond;
    s.numMG = p->numMG; s.imin = p->imin; s.imax = p->imax; s.hang = p->hang; s.normThresh = p->normThresh;
    s.maxDepth = p->maxDepth; s.relaxMode = p->relaxMode; s.precondMode = p->precondMode;
    s.bottom_eps = p->bottom_eps; s.bottom_reps = p->bottom_reps; s.bottom_hang = p->bottom_hang;
    s.bottom_small = p->bottom_small; s.bottom_imax = p->bottom_imax; s.bottom_numRestarts = p->bottom_numRestarts;
    s.bottom_normType = p->bottom_normType; s.verbosity = p->verbosity;
    return s;
}

void* orc_create(const OrcDesc* d, const OrcParams* p) {
    const int nd = d->dim;
    Domain dom;
    dom.box = Box(IntVect(0, 0, 0), IntVect(d->n[0] - 1, d->n[1] - 1, d->n[2] - 1));
    for (int i = 0; i < 3; ++i) dom.periodic[i] = (i < nd) && d->periodic[i];
    std::vector<Box> grids(d->nboxes);
    for (int b = 0; b < d->nboxes; ++b)
        grids[b] = Box(IntVect(d->box_lo[3 * b], d->box_lo[3 * b + 1], d->box_lo[3 * b + 2]),
                       IntVect(d->box_hi[3 * b], d->box_hi[3 * b + 1], d->box_hi[3 * b + 2]));
    BCSpec bc;
    for (int dir = 0; dir < 3; ++dir)
        for (int s = 0; s < 2; ++s) { bc.type[dir][s] = d->bc_type[2 * dir + s]; bc.value[dir][s] = d->bc_value[2 * dir + s]; }
    FluxLevelData Jg;
    Jg.define(grids, nd, nd);
    LevelData Jinv;
    Jinv.define(grids, 1, IntVect::Zero());
    const long nx = d->n[0], ny = d->n[1], nz = d->n[2];
    for (size_t b = 0; b < grids.size(); ++b) {
        ORC_LOOP(grids[b], i, j, k) Jinv[b](i, j, k) = d->jinv[i + nx * (j + ny * (long)k)];
        for (int dir = 0; dir < nd; ++dir) {
            const long fx = nx + (dir == 0), fy = ny + (dir == 1), fz = nz + (dir == 2);
            FAB& f = Jg.fabs[b][dir];
            for (int c = 0; c < nd; ++c) {
                ORC_LOOP(f.box(), i, j, k) {
                    double v = 0.0;
                    if (d->jgup_ncomp == 1) { if (c == dir) v = d->jgup[dir][i + fx * (j + fy * (long)k)]; }
                    else v = d->jgup[dir][i + fx * (j + fy * ((long)k + fz * c))];
                    f(i, j, k, c) = v;
                }
            }
        }
    }
    OrcHandle* h = new OrcHandle;
    h->dim = nd;
    h->mg = buildMG(dom, grids, d->dx, nd, d->alpha, d->beta, d->is_diagonal != 0, bc, toParams(p), std::move(Jg), std::move(Jinv));
    return h;
}

void orc_destroy(void* hv) { delete (OrcHandle*)hv; }
void orc_set_threads(int n) { omp_set_num_threads(n); }
int orc_max_threads() { return omp_get_max_threads(); }
int orc_num_depths(void* hv) { return ((OrcHandle*)hv)->mg->depth(); }

void orc_depth_info(void* hv, int depth, int* n, int* ratio, double* dx, int* zeroavg, int* nboundary) {
    const Op& op = *((OrcHandle*)hv)->mg->ops[depth];
    for (int d = 0; d < 3; ++d) { n[d] = op.domain.box.size(d); ratio[d] = op.mgCrseRefRatio[d]; dx[d] = op.dx[d]; }
    *zeroavg = op.zeroAvgProlong;
    *nboundary = (int)op.boundaryBoxData.size();
}

// Copies the depth's metric fields into global arrays (jg[d]: nd comps outermost).  Any pointer may be NULL.
void orc_get_metric(void* hv, int depth, double* jg0, double* jg1, double* jg2, double* jinv, double* lapdiag) {
    OrcHandle* h = (OrcHandle*)hv;
    const Op& op = *h->mg->ops[depth];
    double* jg[3] = {jg0, jg1, jg2};
    const long nx = op.domain.box.size(0), ny = op.domain.box.size(1), nz = op.domain.box.size(2);
    for (size_t b = 0; b < op.grids.size(); ++b) {
        for (int dir = 0; dir < op.nd; ++dir) {
            if (!jg[dir]) continue;
            const long fx = nx + (dir == 0), fy = ny + (dir == 1), fz = nz + (dir == 2);
            const FAB& f = op.Jgup.fabs[b][dir];
            for (int c = 0; c < op.nd; ++c) ORC_LOOP(f.box(), i, j, k) jg[dir][i + fx * (j + fy * ((long)k + fz * c))] = f(i, j, k, c);
        }
    }
    if (jinv) gather(op, op.Jinv, jinv);
    if (lapdiag) gather(op, op.lapDiag, lapdiag);
}

// Boundary-slab decomposition (collectBoundaryData) for inspection: writes up to `cap` entries of
// (box index, lo[3], hi[3], stencil[6]) = 13 ints each; returns the number of entries.
int orc_boundary_boxes(void* hv, int depth, int simple, int* out, int cap) {
    const Op& op = *((OrcHandle*)hv)->mg->ops[depth];
    const auto& v = simple ? op.simpleBoundaryBoxData : op.boundaryBoxData;
    int n = 0;
    for (const auto& d : v) {
        if (n >= cap) break;
        int* o = out + 13 * n++;
        o[0] = d.index;
        for (int i = 0; i < 3; ++i) { o[1 + i] = d.validBdry.lo[i]; o[4 + i] = d.validBdry.hi[i]; }
        for (int i = 0; i < 3; ++i) { o[7 + 2 * i] = d.stencil[i][0]; o[8 + 2 * i] = d.stencil[i][1]; }
    }
    return (int)v.size();
}

void orc_apply_op(void* hv, int depth, double* lhs, const double* phi, int homog) {
    Op& op = *((OrcHandle*)hv)->mg->ops[depth];
    LevelData p, l;
    op.create(p, 1); op.create(l, 0);
    scatter(op, p, phi);
    op.applyOp(l, p, homog != 0);
    gather(op, l, lhs);
}

void orc_residual(void* hv, int depth, double* res, const double* phi, const double* rhs, int homog) {
    Op& op = *((OrcHandle*)hv)->mg->ops[depth];
    LevelData p, l, r;
    op.create(p, 1); op.create(l, 0); op.create(r, 0);
    scatter(op, p, phi); scatter(op, r, rhs);
    op.residual(l, p, r, homog != 0);
    gather(op, l, res);
}

void orc_relax(void* hv, int depth, double* phi, const double* rhs, int iters) {
    Op& op = *((OrcHandle*)hv)->mg->ops[depth];
    LevelData p, r;
    op.create(p, 1); op.create(r, 0);
    scatter(op, p, phi); scatter(op, r, rhs);
    op.relax(p, r, iters);
    gather(op, p, phi);
}

void orc_precond(void* hv, int depth, double* phi, const double* rhs) {
    Op& op = *((OrcHandle*)hv)->mg->ops[depth];
    LevelData p, r;
    op.create(p, 1); op.create(r, 0);
    scatter(op, r, rhs);
    op.preCond(p, r);
    gather(op, p, phi);
}

void orc_restrict_residual(void* hv, int depth, double* crse, const double* phi, const double* rhs) {
    MG& mg = *((OrcHandle*)hv)->mg;
    Op& op = *mg.ops[depth];
    Op& cop = *mg.ops[depth + 1];
    LevelData p, r, c;
    op.create(p, 1); op.create(r, 0); cop.create(c, 0);
    scatter(op, p, phi); scatter(op, r, rhs);
    op.restrictResidual(c, p, r, cop);
    gather(cop, c, crse);
}

void orc_prolong_increment(void* hv, int depth, double* phi, const double* crse) {
    MG& mg = *((OrcHandle*)hv)->mg;
    Op& op = *mg.ops[depth];
    Op& cop = *mg.ops[depth + 1];
    LevelData p, c;
    op.create(p, 1); cop.create(c, 1);
    scatter(op, p, phi); scatter(cop, c, crse);
    op.prolongIncrement(p, c);
    gather(op, p, phi);
}

double orc_norm(void* hv, int depth, const double* x, int ord) {
    Op& op = *((OrcHandle*)hv)->mg->ops[depth];
    LevelData p;
    op.create(p, 0);
    scatter(op, p, x);
    return op.norm(p, ord);
}

// One multigrid cycle from depth `depth` down (MappedMultiGrid::cycle); e is in/out.
void orc_cycle(void* hv, int depth, double* e, const double* r) {
    MG& mg = *((OrcHandle*)hv)->mg;
    Op& op = *mg.ops[depth];
    LevelData le, lr;
    op.create(le, 1); op.create(lr, 0);
    scatter(op, le, e); scatter(op, lr, r);
    mg.cycle(depth, le, lr);
    gather(op, le, e);
}

// Bottom solve alone (on the coarsest depth); bottom metric/eps as after a solve() call or the defaults.
int orc_bottom_solve(void* hv, double* e, const double* r, double metric, double eps) {
    MG& mg = *((OrcHandle*)hv)->mg;
    Op& op = *mg.ops.back();
    mg.bottomMetric = metric; mg.bottomEps = eps;
    LevelData le, lr;
    op.create(le, 1); op.create(lr, 0);
    scatter(op, le, e); scatter(op, lr, r);
    const int it = mg.bottomSolve(le, lr);
    gather(op, le, e);
    return it;
}

void orc_solve(void* hv, double* phi, const double* rhs, int zeroPhi, int forceHomogeneous, OrcStats* out) {
    MG& mg = *((OrcHandle*)hv)->mg;
    Op& op = *mg.ops[0];
    LevelData p, r;
    op.create(p, 1); op.create(r, 0);
    scatter(op, p, phi); scatter(op, r, rhs);
    SolveStats st;
    mg.solve(p, r, zeroPhi != 0, forceHomogeneous != 0, st);
    gather(op, p, phi);
    out->iters = st.iters; out->exitStatus = st.exitStatus;
    out->initial_rnorm = st.initial_rnorm; out->final_rnorm = st.final_rnorm;
    out->bottom_iters_total = st.bottom_iters_total;
    out->history_len = (int)std::min<size_t>(128, st.history.size());
    for (int i = 0; i < out->history_len; ++i) out->history[i] = st.history[i];
}

// m_convergenceMetric of the next solves (0: the initial residual norm); did the last solve hit the "blew up" abort?
void orc_set_convergence_metric(void* hv, double metric) { ((OrcHandle*)hv)->mg->convergenceMetric = metric; }
int orc_blew_up(void* hv) { return ((OrcHandle*)hv)->mg->blewUp ? 1 : 0; }

// Solve used by the CPU-baseline timer: keeps fields inside the handle so that the timed region is the
// solve only.  Returns seconds spent in mg.solve().
double orc_timed_solve(void* hv, const double* rhs, int ncycles_cap, OrcStats* out) {
    MG& mg = *((OrcHandle*)hv)->mg;
    Op& op = *mg.ops[0];
    LevelData p, r;
    op.create(p, 1); op.create(r, 0);
    scatter(op, r, rhs);
    const int save = mg.p.imax;
    if (ncycles_cap > 0) mg.p.imax = ncycles_cap;
    SolveStats st;
    const double t0 = omp_get_wtime();
    mg.solve(p, r, true, true, st);
    const double t1 = omp_get_wtime();
    mg.p.imax = save;
    out->iters = st.iters; out->exitStatus = st.exitStatus;
    out->initial_rnorm = st.initial_rnorm; out->final_rnorm = st.final_rnorm;
    out->bottom_iters_total = st.bottom_iters_total;
    out->history_len = (int)std::min<size_t>(128, st.history.size());
    for (int i = 0; i < out->history_len; ++i) out->history[i] = st.history[i];
    return t1 - t0;
}

// ---- projector calculus.  Face arrays: global, +1 point in their own direction, one component. ----
static void scatter_faces(const Op& op, FluxLevelData& fl, const double* const* g) {
    const long nx = op.domain.box.size(0), ny = op.domain.box.size(1);
    for (size_t b = 0; b < op.grids.size(); ++b)
        for (int d = 0; d < op.nd; ++d) {
            const long fx = nx + (d == 0), fy = ny + (d == 1);
            FAB& f = fl.fabs[b][d];
            ORC_LOOP(f.box(), i, j, k) f(i, j, k) = g[d][i + fx * (j + fy * (long)k)];
        }
}
static void gather_faces(const Op& op, const FluxLevelData& fl, double* const* g) {
    const long nx = op.domain.box.size(0), ny = op.domain.box.size(1);
    for (size_t b = 0; b < op.grids.size(); ++b)     // ascending boxes: an interface face ends up as its high-side box wrote it
        for (int d = 0; d < op.nd; ++d) {
            const long fx = nx + (d == 0), fy = ny + (d == 1);
            const FAB& f = fl.fabs[b][d];
            ORC_LOOP(f.box(), i, j, k) g[d][i + fx * (j + fy * (long)k)] = f(i, j, k);
        }
}

void orc_mac_gradient(void* hv, int depth, double* g0, double* g1, double* g2, const double* phi, int homog) {
    Op& op = *((OrcHandle*)hv)->mg->ops[depth];
    LevelData p;
    op.create(p, 1);
    scatter(op, p, phi);
    FluxLevelData gr;
    gr.define(op.grids, 1, op.nd);
    op.macGradient(gr, p, homog != 0);
    double* g[3] = {g0, g1, g2};
    gather_faces(op, gr, g);
}

// the same with the BC holder named: 0 the operator's, 1 order-2 extrapolated ghosts (the projectors' m_gradBC)
void orc_mac_gradient_bc(void* hv, int depth, double* g0, double* g1, double* g2, const double* phi, int homog, int gradBC) {
    Op& op = *((OrcHandle*)hv)->mg->ops[depth];
    LevelData p;
    op.create(p, 1);
    scatter(op, p, phi);
    FluxLevelData gr;
    gr.define(op.grids, 1, op.nd);
    op.macGradient(gr, p, homog != 0, gradBC);
    double* g[3] = {g0, g1, g2};
    gather_faces(op, gr, g);
}

// Divergence::levelDivergenceCC (Divergence.cpp:359-396) with the projector's wall BC, and Gradient::levelGradientCC
// (Gradient.cpp:469-494) with the projector's gradient BC, on depth 0.  u*, c*: cell arrays.
void orc_cc_divergence(void* hv, double* div, const double* u0, const double* u1, const double* u2) {
    Op& op = *((OrcHandle*)hv)->mg->ops[0];
    const double* u[3] = {u0, u1, u2};
    std::array<LevelData, 3> vel;
    for (int d = 0; d < op.nd; ++d) {
        op.create(vel[d], 1);
        scatter(op, vel[d], u[d]);
        op.exchange(vel[d], false);
    }
    FluxLevelData edge;
    edge.define(op.grids, 1, op.nd);
    op.cellToEdge(edge, vel);
    op.wallFluxBC(edge);
    LevelData d;
    op.create(d, 0);
    op.macDivergence(d, edge);
    gather(op, d, div);
}

void orc_cc_gradient(void* hv, double* c0, double* c1, double* c2, const double* phi) {
    Op& op = *((OrcHandle*)hv)->mg->ops[0];
    LevelData p;
    op.create(p, 1);
    scatter(op, p, phi);
    FluxLevelData eg;
    eg.define(op.grids, 1, op.nd);
    op.macGradient(eg, p, false, 1);
    std::array<LevelData, 3> cc;
    for (int d = 0; d < op.nd; ++d) op.create(cc[d], 0);
    op.edgeToCell(cc, eg);
    double* c[3] = {c0, c1, c2};
    for (int d = 0; d < op.nd; ++d) gather(op, cc[d], c[d]);
}

// BaseProjector<FArrayBox>::project via LevelCCProjector on depth 0.  J may be NULL (a_velIsFlux = true).
void orc_project_cc(void* hv, double* u0, double* u1, double* u2, const double* J, double* pressure, double dt,
                    int zeroPressure, int forceHomogeneous, OrcStats* out) {
    MG& mg = *((OrcHandle*)hv)->mg;
    Op& op = *mg.ops[0];
    double* u[3] = {u0, u1, u2};
    std::array<LevelData, 3> vel;
    for (int d = 0; d < op.nd; ++d) {
        op.create(vel[d], 1);
        scatter(op, vel[d], u[d]);
    }
    LevelData Jl;
    if (J) {
        op.create(Jl, 0);
        scatter(op, Jl, J);
    }
    LevelData p;
    op.create(p, 1);
    scatter(op, p, pressure);
    SolveStats st;
    mg.projectCC(vel, J ? &Jl : nullptr, p, dt, zeroPressure != 0, forceHomogeneous != 0, st);
    gather(op, p, pressure);
    for (int d = 0; d < op.nd; ++d) gather(op, vel[d], u[d]);
    out->iters = st.iters; out->exitStatus = st.exitStatus;
    out->initial_rnorm = st.initial_rnorm; out->final_rnorm = st.final_rnorm;
    out->bottom_iters_total = st.bottom_iters_total;
    out->history_len = (int)std::min<size_t>(128, st.history.size());
    for (int i = 0; i < out->history_len; ++i) out->history[i] = st.history[i];
}

// LevelLepticSolver::computeHorizRHS / ::levelVertHorizGradient on depth 0 (flat outputs: one value per column)
void orc_leptic_horiz_rhs(void* hv, double* flatRhs, const double* phi) {
    Op& op = *((OrcHandle*)hv)->mg->ops[0];
    LevelData p;
    op.create(p, 1);
    scatter(op, p, phi);
    op.lepticHorizRHS(flatRhs, p);
}

void orc_leptic_vert_horiz_grad(void* hv, double* flatLo, double* flatHi, const double* phi, double scale) {
    Op& op = *((OrcHandle*)hv)->mg->ops[0];
    LevelData p;
    op.create(p, 1);
    scatter(op, p, phi);
    // the reference makes no exchange here: it finds the ghosts as the residual evaluation that closed the previous
    // order left them (LevelLepticSolver.cpp:839: exchange + homogeneous BC ghosts) -- reproduce that state
    LevelData tmp;
    op.create(tmp, 0);
    op.applyOp(tmp, p, true);
    op.lepticVertHorizGradient(flatLo, flatHi, p, scale);
}

void orc_mac_divergence(void* hv, int depth, double* div, const double* f0, const double* f1, const double* f2) {
    Op& op = *((OrcHandle*)hv)->mg->ops[depth];
    FluxLevelData fl;
    fl.define(op.grids, 1, op.nd);
    const double* f[3] = {f0, f1, f2};
    scatter_faces(op, fl, f);
    LevelData d;
    op.create(d, 0);
    op.macDivergence(d, fl);
    gather(op, d, div);
}

void orc_project(void* hv, double* v0, double* v1, double* v2, double* pressure, double dt, int zeroPressure,
                 int forceHomogeneous, OrcStats* out) {
    MG& mg = *((OrcHandle*)hv)->mg;
    Op& op = *mg.ops[0];
    FluxLevelData vel;
    vel.define(op.grids, 1, op.nd);
    double* v[3] = {v0, v1, v2};
    scatter_faces(op, vel, v);
    LevelData p;
    op.create(p, 1);
    scatter(op, p, pressure);
    SolveStats st;
    mg.project(vel, p, dt, zeroPressure != 0, forceHomogeneous != 0, st);
    gather(op, p, pressure);
    gather_faces(op, vel, v);
    out->iters = st.iters; out->exitStatus = st.exitStatus;
    out->initial_rnorm = st.initial_rnorm; out->final_rnorm = st.final_rnorm;
    out->bottom_iters_total = st.bottom_iters_total;
    out->history_len = (int)std::min<size_t>(128, st.history.size());
    for (int i = 0; i < out->history_len; ++i) out->history[i] = st.history[i];
}

}  // extern "C"
