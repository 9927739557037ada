// bicgstab.cu -- the bottom solver of the V-cycle with its scalar recurrences RESIDENT ON THE DEVICE.
//
// Chombo 3.1 BiCGStabSolver<T>::solve (not under the reference tree; configured at
// src/projection/AMRPressureSolver.cpp:253-262, called from MappedMultiGrid<T>::cycle,
// src/calculus/AMRElliptic/MappedMultiGrid.H:566-573) takes six global reductions per iteration and branches on
// every one of them.  The bottom depth is tiny (a few hundred thousand cells, L2 resident), so a host that waits
// for each scalar spends the iteration in launch + synchronise latency.  Here rho/alpha/beta/omega, the norms and
// the branch flags live in a BicgState in HBM:
//   * reductions land in state.red[] (all-reduced in place by NCCL when there are several ranks);
//   * one-thread "control" kernels (k_bicg_c1..c5) do the scalar arithmetic of the algorithm in the same order,
//     in IEEE fp64 (-fmad=false), so every value equals the host-sequenced version bit for bit;
//   * the vector kernels read their coefficient and their branch flag from the state;
//   * once the state says "finished" (or "restart requested") the kernels that touch persistent data (e, the
//     state) do nothing, so the host may enqueue iterations ahead and looks at the state once per chunk.
// The (rare) restart branch is done by the host between chunks with the ordinary operator calls.
// bottom_solve_host() is the host-sequenced version (SOMAR_B200_BOTTOM=0), kept as the A/B partner of the tests.
#include "bicg_persistent.cuh"
#include "level.hpp"
#include <cmath>
#include <cstdlib>
#include <cstring>

namespace somar {

static inline Geom& G(SomarLevel* L, int d) { return L->depths[d].g; }

__global__ void k_bicg_c1(BicgState* st) { if (st->status == 0) ctl1(st); }
__global__ void k_bicg_c2(BicgState* st) { if (st->status == 0) ctl2(st); }
__global__ void k_bicg_c3(BicgState* st) { if (st->status == 0) ctl3(st); }
__global__ void k_bicg_c4(BicgState* st) { if (st->status == 0) ctl4(st); }
__global__ void k_bicg_c5(BicgState* st) { if (st->status == 0) ctl5(st); }

// p = r (first iteration after a start)   or   p = ((beta p) + (-beta omega) v) + r
__global__ void k_bicg_p(const Geom g, double* __restrict__ pv, const double* __restrict__ v,
                         const double* __restrict__ r, const BicgState* __restrict__ st) {
    const bool init = st->init != 0;
    const double b = st->bet[1], cv = st->c_pv;
    double pp[4], vv[4], rr[4];
    FOR_VALID_CELLS4(g, c, u, { rr[u] = r[c]; if (!init) { pp[u] = pv[c]; vv[u] = v[c]; } }, {
        if (init) {
            pv[c] = rr[u];
        } else {
            const double t1 = pp[u] * b;
            const double t2 = t1 + cv * vv[u];
            pv[c] = t2 + 1.0 * rr[u];
        }
    })
}
// r += coef * x with the norm of the new r in the same pass (flag set);  flag clear: r = 0 (elseZero) or untouched.
__global__ void k_bicg_r(const Geom g, double* __restrict__ r, const double* __restrict__ x,
                         const BicgState* __restrict__ st, const double* __restrict__ coef,
                         const int* __restrict__ flag, bool elseZero, double* __restrict__ scratch,
                         unsigned int* __restrict__ tickets, double* __restrict__ out) {
    const bool on = *flag != 0;
    if (!on && !elseZero) return;
    const int kind = st->normKind;
    const bool isMax = kind == 0;
    const double a = *coef;
    double acc = 0.0;
    double xv[4], rv[4];
    FOR_VALID_CELLS4(g, c, u, { if (on) { xv[u] = x[c]; rv[u] = r[c]; } }, {
        double nv = 0.0;
        if (on) nv = rv[u] + a * xv[u];
        r[c] = nv;
        if (kind == 0) acc = fmax(acc, fabs(nv));
        else if (kind == 1) acc += fabs(nv);
        else acc += nv * nv;
    })
    acc = block_red(acc, isMax, 0.0);
    finish_reduce(acc, isMax, 0.0, scratch, out, tickets, 0, false);
}
// e += coef * x     (only while iterating and when the branch flag is set)
__global__ void k_bicg_e(const Geom g, double* __restrict__ e, const double* __restrict__ x,
                         const BicgState* __restrict__ st, const double* __restrict__ coef,
                         const int* __restrict__ flag) {
    if (st->status != 0 || *flag == 0) return;
    const double a = *coef;
    double xv[4], ev[4];
    FOR_VALID_CELLS4(g, c, u, { xv[u] = x[c]; ev[u] = e[c]; }, { e[c] = ev[u] + a * xv[u]; })
}
// red[0] = <t, r>, red[1] = <t, t> in one pass (each sum in the order k_reduce would take)
__global__ void k_bicg_dot2(const Geom g, const double* __restrict__ t, const double* __restrict__ r,
                            double* __restrict__ scratch, unsigned int* __restrict__ tickets,
                            double* __restrict__ out) {
    double a0 = 0.0, a1 = 0.0;
    double tv[4], rv[4];
    FOR_VALID_CELLS4(g, c, u, { tv[u] = t[c]; rv[u] = r[c]; }, {
        a0 += tv[u] * rv[u];
        a1 += tv[u] * tv[u];
    })
    a0 = block_red(a0, false, 0.0);
    a1 = block_red(a1, false, 0.0);
    finish_reduce(a0, false, 0.0, scratch, out, tickets, 1, false);
    finish_reduce(a1, false, 0.0, scratch + MAX_RED_BLOCKS, out + 1, tickets, 2, false);
}

static int allreduce_state(SomarLevel* L, double* p, int n, bool isMax) {
    if (L->desc.nranks > 1)
        SOMAR_NCCL(ncclAllReduce(p, p, n, ncclDouble, isMax ? ncclMax : ncclSum, L->comm, L->cx.stream));
    return SOMAR_OK;
}

struct BicgDev;
static int ensure_state(SomarLevel* L);
static int bottom_solve_device(SomarLevel* L, SomarField* a_phi, const SomarField* a_rhs, int* iters) {
    const SomarSolverParams& p = L->prm;
    const int bd = L->ndepths() - 1;
    const Geom& g = G(L, bd);
    Ctx& cx = L->cx;
    int rc;
    if (L->bwork.empty()) {
        L->bwork.resize(8, nullptr);
        for (auto& f : L->bwork)
            if ((rc = field_alloc(L, bd, &f))) return rc;
    }
    if ((rc = ensure_state(L))) return rc;
    BicgState* ds = static_cast<BicgState*>(L->d_bicg);   // first member of BicgDev
    BicgState* hs = static_cast<BicgState*>(L->h_bicg);
    SomarField *r = L->bwork[0], *r_tilde = L->bwork[1], *e = L->bwork[2], *pv = L->bwork[3], *p_tilde = L->bwork[4],
               *s_tilde = L->bwork[5], *t = L->bwork[6], *v = L->bwork[7];
    const int kind = (p.bottom_normType == 0) ? 0 : (p.bottom_normType == 1 ? 1 : 2);
    const bool normMax = kind == 0;
    auto norm_host = [&](SomarField* x, double* out) {
        int rc2 = reduce_global(L, bd, x->p, nullptr, kind, out);
        if (kind == 2) *out = std::sqrt(*out);
        return rc2;
    };
    const int nb = row_blocks(g, cx.sms);

    if ((rc = op_apply(L, bd, r->p, a_phi, a_rhs->p, 1, true))) return rc;
    launch_copy(cx, g, r_tilde->p, r->p);
    launch_fill(cx, g, e->p, 0.0);
    launch_fill(cx, g, p_tilde->p, 0.0);
    launch_fill(cx, g, s_tilde->p, 0.0);
    double n0;
    if ((rc = norm_host(r, &n0))) return rc;

    BicgState s0 = {};
    s0.nrm[0] = s0.nrm[1] = n0;
    s0.initial_norm = (L->bottomMetric > 0) ? L->bottomMetric : n0;
    s0.initial_rnorm = n0;
    s0.eps = L->bottomEps;
    s0.reps = p.bottom_reps;
    s0.small = p.bottom_small;
    s0.one_minus_hang = 1.0 - p.bottom_hang;
    s0.imax = p.bottom_imax;
    s0.init = 1;
    s0.normKind = kind;
    int restarts = 0;
    auto loop_top = [](const BicgState& s) { return (s.i < s.imax && s.nrm[0] > s.eps * s.nrm[1]) && (s.nrm[1] > 0); };
    *hs = s0;
    hs->status = loop_top(*hs) ? 0 : 1;

    const int CHUNK = 8;
    // One iteration is ~40 launches of kernels that run 3-30 us each on the bottom grid: sequenced launch by launch the host
    // cannot keep the stream fed (0.8 ms per iteration against 0.4 ms of kernel time, profiles/r2_ncu_k_gsrb_full.txt).
    // The iterations never read the host (every scalar lives in *ds), so a full chunk is captured once as a CUDA graph
    // and replayed.  The graph holds pointers and by-value geometry: it is rebuilt when anything it captured changes
    // (the key below).  One rank and full metric only -- the diagonal metric has the resident kernel, several ranks
    // have NCCL calls in the loop.
    static const bool graphOn = []() { const char* e = getenv("SOMAR_B200_BOTTOM_GRAPH"); return !(e && e[0] == '0'); }();
    const Depth& bdp = L->depths[bd];
    const bool graphable = graphOn && L->desc.nranks == 1 && !bdp.m.diagonal && p.relaxMode == SOMAR_RELAX_LEVEL_GSRB &&
                           p.precondMode != 2;
    struct GraphKey {
        double alpha, beta;
        const void* ptr[12];
        int relaxMode, precondMode, nsp, kind, nb, n[3];
    } key;
    std::memset(&key, 0, sizeof(key));
    static_assert(sizeof(GraphKey) <= sizeof(L->bicg_graph_key), "bicg_graph_key too small");
    if (graphable) {
        key.alpha = g.alpha; key.beta = g.beta;
        for (int q = 0; q < 8; ++q) key.ptr[q] = L->bwork[q]->p;
        key.ptr[8] = bdp.tmp; key.ptr[9] = ds; key.ptr[10] = cx.scratch; key.ptr[11] = bdp.m.jinv;
        key.relaxMode = p.relaxMode; key.precondMode = p.precondMode; key.nsp = p.num_smooth_precond; key.kind = kind;
        key.nb = nb; key.n[0] = g.n[0]; key.n[1] = g.n[1]; key.n[2] = g.n[2];
    }
    auto iterate = [&](int todo) -> int {
        for (int it = 0; it < todo; ++it) {
            launch_reduce(cx, g, r_tilde->p, r->p, 3, &ds->red[0]);
            if ((rc = allreduce_state(L, &ds->red[0], 1, false))) return rc;
            k_bicg_c1<<<1, 1, 0, cx.stream>>>(ds);
            k_bicg_p<<<nb, TPB, 0, cx.stream>>>(g, pv->p, v->p, r->p, ds);
            cx.launches += 2;
            if ((rc = op_precond(L, bd, p_tilde, pv))) return rc;
            if ((rc = op_apply(L, bd, v->p, p_tilde, nullptr, 0, true))) return rc;
            launch_reduce(cx, g, r_tilde->p, v->p, 3, &ds->red[0]);
            if ((rc = allreduce_state(L, &ds->red[0], 1, false))) return rc;
            k_bicg_c2<<<1, 1, 0, cx.stream>>>(ds);
            k_bicg_r<<<nb, TPB, 0, cx.stream>>>(g, r->p, v->p, ds, &ds->neg_alpha, &ds->doA, true, cx.scratch, cx.tickets,
                                                &ds->red[0]);
            if ((rc = allreduce_state(L, &ds->red[0], 1, normMax))) return rc;
            k_bicg_c3<<<1, 1, 0, cx.stream>>>(ds);
            k_bicg_e<<<nb, TPB, 0, cx.stream>>>(g, e->p, p_tilde->p, ds, &ds->alph[0], &ds->doA);
            cx.launches += 4;
            if ((rc = op_precond(L, bd, s_tilde, r))) return rc;
            if ((rc = op_apply(L, bd, t->p, s_tilde, nullptr, 0, true))) return rc;
            k_bicg_dot2<<<nb, TPB, 0, cx.stream>>>(g, t->p, r->p, cx.scratch, cx.tickets, &ds->red[0]);
            if ((rc = allreduce_state(L, &ds->red[0], 2, false))) return rc;
            k_bicg_c4<<<1, 1, 0, cx.stream>>>(ds);
            k_bicg_e<<<nb, TPB, 0, cx.stream>>>(g, e->p, s_tilde->p, ds, &ds->omega[0], &ds->doB);
            k_bicg_r<<<nb, TPB, 0, cx.stream>>>(g, r->p, t->p, ds, &ds->neg_omega, &ds->doB, false, cx.scratch, cx.tickets,
                                                &ds->red[0]);
            if ((rc = allreduce_state(L, &ds->red[0], 1, normMax))) return rc;
            k_bicg_c5<<<1, 1, 0, cx.stream>>>(ds);
            cx.launches += 5;
        }
        return SOMAR_OK;
    };
    while (hs->status == 0) {
        SOMAR_CUDA(cudaMemcpyAsync(ds, hs, sizeof(BicgState), cudaMemcpyHostToDevice, cx.stream));
        int todo = hs->imax - hs->i;
        if (todo > CHUNK) todo = CHUNK;
        if (graphable && todo == CHUNK) {
            if (!L->bicg_graph || std::memcmp(&key, L->bicg_graph_key, sizeof(key)) != 0) {
                if (L->bicg_graph) { cudaGraphExecDestroy(L->bicg_graph); L->bicg_graph = nullptr; }
                const long long l0 = cx.launches;
                SOMAR_CUDA(cudaStreamBeginCapture(cx.stream, cudaStreamCaptureModeThreadLocal));
                rc = iterate(CHUNK);
                cudaGraph_t graph = nullptr;
                const cudaError_t ce = cudaStreamEndCapture(cx.stream, &graph);
                const int captured = (int)(cx.launches - l0);
                cx.launches = l0;
                if (rc != SOMAR_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
                SOMAR_CUDA(ce);
                const cudaError_t ie = cudaGraphInstantiate(&L->bicg_graph, graph, 0);
                cudaGraphDestroy(graph);
                SOMAR_CUDA(ie);
                L->bicg_graph_launches = captured;
                std::memcpy(L->bicg_graph_key, &key, sizeof(key));
            }
            SOMAR_CUDA(cudaGraphLaunch(L->bicg_graph, cx.stream));
            cx.launches += L->bicg_graph_launches;
        } else if ((rc = iterate(todo))) {
            return rc;
        }
        SOMAR_CUDA(cudaMemcpyAsync(hs, ds, sizeof(BicgState), cudaMemcpyDeviceToHost, cx.stream));
        SOMAR_CUDA(cudaStreamSynchronize(cx.stream));
        if (hs->status == 2) {
            // restart: fold the correction into phi and begin again from the true residual
            launch_axpy(cx, g, a_phi->p, e->p, 1.0);
            launch_fill(cx, g, e->p, 0.0);
            if (restarts == p.bottom_numRestarts) { hs->status = 1; break; }
            if ((rc = op_apply(L, bd, r->p, a_phi, a_rhs->p, 1, true))) return rc;
            if ((rc = norm_host(r, &hs->nrm[0]))) return rc;
            for (int k = 0; k < 4; ++k) hs->rho[k] = 0.0;
            hs->alph[0] = hs->alph[1] = hs->bet[0] = hs->bet[1] = hs->omega[0] = hs->omega[1] = 0.0;
            ++restarts;
            hs->init = 1;
            launch_copy(cx, g, r_tilde->p, r->p);
            hs->status = loop_top(*hs) ? 0 : 1;
        }
    }
    launch_axpy(cx, g, a_phi->p, e->p, 1.0);
    *iters = hs->i;
    return SOMAR_OK;
}

// ---------------------------------------------------------------------------------------------------------
// The same algorithm sequenced by the host: every scalar is read back (six synchronisations per iteration).
static int bottom_solve_host(SomarLevel* L, SomarField* a_phi, const SomarField* a_rhs, int* iters) {
    const SomarSolverParams& p = L->prm;
    const int bd = L->ndepths() - 1;
    const Geom& g = G(L, bd);
    int rc;
    if (L->bwork.empty()) {
        L->bwork.resize(8, nullptr);
        for (auto& f : L->bwork)
            if ((rc = field_alloc(L, bd, &f))) return rc;
    }
    SomarField *r = L->bwork[0], *r_tilde = L->bwork[1], *e = L->bwork[2], *pv = L->bwork[3], *p_tilde = L->bwork[4],
               *s_tilde = L->bwork[5], *t = L->bwork[6], *v = L->bwork[7];
    Ctx& cx = L->cx;
    auto norm = [&](SomarField* x, double* out) {
        int kind = (p.bottom_normType == 0) ? 0 : (p.bottom_normType == 1 ? 1 : 2);
        int rc2 = reduce_global(L, bd, x->p, nullptr, kind, out);
        if (kind == 2) *out = std::sqrt(*out);
        return rc2;
    };
    auto dot = [&](SomarField* a, SomarField* b, double* out) { return reduce_global(L, bd, a->p, b->p, 3, out); };

    int recount = 0;
    if ((rc = op_apply(L, bd, r->p, a_phi, a_rhs->p, 1, true))) return rc;
    launch_copy(cx, g, r_tilde->p, r->p);
    launch_fill(cx, g, e->p, 0.0);
    launch_fill(cx, g, p_tilde->p, 0.0);
    launch_fill(cx, g, s_tilde->p, 0.0);
    int i = 0;
    double rho[4] = {0, 0, 0, 0}, nrm[2];
    if ((rc = norm(r, &nrm[0]))) return rc;
    double initial_norm = nrm[0];
    const double initial_rnorm = nrm[0];
    nrm[1] = nrm[0];
    double alph[2] = {0, 0}, bet[2] = {0, 0}, omega[2] = {0, 0};
    bool init = true;
    int restarts = 0;
    if (L->bottomMetric > 0) initial_norm = L->bottomMetric;
    const double eps = L->bottomEps;
    while ((i < p.bottom_imax && nrm[0] > eps * nrm[1]) && (nrm[1] > 0)) {
        ++i;
        nrm[1] = nrm[0];
        alph[1] = alph[0]; bet[1] = bet[0]; omega[1] = omega[0];
        rho[3] = rho[2]; rho[2] = rho[1];
        if ((rc = dot(r_tilde, r, &rho[1]))) return rc;
        if (rho[1] == 0.0) break;
        if (init) {
            launch_copy(cx, g, pv->p, r->p);
            init = false;
        } else {
            bet[1] = (rho[1] / rho[2]) * (alph[1] / omega[1]);
            launch_scale(cx, g, pv->p, bet[1]);
            launch_axpy(cx, g, pv->p, v->p, -bet[1] * omega[1]);
            launch_axpy(cx, g, pv->p, r->p, 1.0);
        }
        if ((rc = op_precond(L, bd, p_tilde, pv))) return rc;
        if ((rc = op_apply(L, bd, v->p, p_tilde, nullptr, 0, true))) return rc;
        double m;
        if ((rc = dot(r_tilde, v, &m))) return rc;
        alph[0] = rho[1] / m;
        if (std::fabs(m) > p.bottom_small * std::fabs(rho[1])) {
            launch_axpy(cx, g, r->p, v->p, -alph[0]);
            if ((rc = norm(r, &nrm[0]))) return rc;
            launch_axpy(cx, g, e->p, p_tilde->p, alph[0]);
        } else {
            launch_fill(cx, g, r->p, 0.0);
            nrm[0] = 0.0;
        }
        if (nrm[0] > eps * initial_norm && nrm[0] > p.bottom_reps * initial_rnorm) {
            if ((rc = op_precond(L, bd, s_tilde, r))) return rc;
            if ((rc = op_apply(L, bd, t->p, s_tilde, nullptr, 0, true))) return rc;
            double tr, tt;
            if ((rc = dot(t, r, &tr))) return rc;
            if ((rc = dot(t, t, &tt))) return rc;
            omega[0] = tr / tt;
            launch_axpy(cx, g, e->p, s_tilde->p, omega[0]);
            launch_axpy(cx, g, r->p, t->p, -omega[0]);
            if ((rc = norm(r, &nrm[0]))) return rc;
        }
        if (nrm[0] <= eps * initial_norm || nrm[0] <= p.bottom_reps * initial_rnorm) break;
        if (omega[0] == 0.0 || nrm[0] > (1.0 - p.bottom_hang) * nrm[1]) {
            if (recount == 0) {
                recount = 1;
            } else {
                recount = 0;
                launch_axpy(cx, g, a_phi->p, e->p, 1.0);
                if (restarts == p.bottom_numRestarts) { launch_fill(cx, g, e->p, 0.0); break; }
                if ((rc = op_apply(L, bd, r->p, a_phi, a_rhs->p, 1, true))) return rc;
                if ((rc = norm(r, &nrm[0]))) return rc;
                rho[0] = rho[1] = rho[2] = rho[3] = 0.0;
                alph[0] = alph[1] = bet[0] = bet[1] = omega[0] = omega[1] = 0.0;
                launch_fill(cx, g, e->p, 0.0);
                ++restarts;
                init = true;
                launch_copy(cx, g, r_tilde->p, r->p);
            }
        }
    }
    launch_axpy(cx, g, a_phi->p, e->p, 1.0);
    *iters = i;
    return SOMAR_OK;
}

int bottom_solve(SomarLevel* L, SomarField* a_phi, const SomarField* a_rhs, int* iters) {
    return L->bottom_mode ? bottom_solve_device(L, a_phi, a_rhs, iters) : bottom_solve_host(L, a_phi, a_rhs, iters);
}

// ---- persistent kernel: eligibility, launch -------------------------------------------------------------
struct BicgDev {          // device allocation behind SomarLevel::d_bicg
    BicgState st;
    int iters;
    int pad;
    double part[4 * MAX_RED_BLOCKS];
    unsigned long long bar[BP_BAR_WORDS];   // tree_barrier counters of the persistent kernel
};

static int ensure_state(SomarLevel* L) {
    if (L->d_bicg) return SOMAR_OK;
    SOMAR_CUDA(cudaMalloc(&L->d_bicg, sizeof(BicgDev)));
    SOMAR_CUDA(cudaMemsetAsync(L->d_bicg, 0, sizeof(BicgDev), L->cx.stream));
    SOMAR_CUDA(cudaMallocHost(&L->h_bicg, sizeof(BicgState)));
    return SOMAR_OK;
}

bool bottom_persistent_usable(SomarLevel* L) {
    if (L->bottom_mode < 2) return false;
    const Depth& dp = L->depths[L->ndepths() - 1];
    const Geom& g = dp.g;
    if (L->desc.nranks > 1 || g.nbr[0] >= 0 || g.nbr[1] >= 0) return false;
    if (!dp.m.diagonal) {   // the FULL instantiation (SOMAR_B200_BOTTOM_FULL_PERSISTENT=0: the launch-per-step graph instead)
        static const bool on = []() { const char* e = getenv("SOMAR_B200_BOTTOM_FULL_PERSISTENT"); return !(e && e[0] == '0'); }();
        if (!on) return false;
    }
    return bottom_persistent_params_ok(L, g, !dp.m.diagonal);
}

// solver parameters and size the persistent kernel supports, for a single-rank diagonal-metric geometry g
bool bottom_persistent_params_ok(SomarLevel* L, const Geom& g, bool full) {
    const SomarSolverParams& p = L->prm;
    if (g.periodic[0] || g.periodic[1] || (g.nd > 2 && g.periodic[2])) return false;
    if (p.relaxMode != SOMAR_RELAX_LEVEL_GSRB) return false;
    const bool precond = !(p.num_smooth_precond == 0 || p.precondMode == 0);
    if (precond && p.precondMode == 2) return false;
    if ((long long)g.gn[0] * g.gn[1] * g.gn[2] == 1) return false;
    if (L->bottom_coop_blocks < 0) {   // blocks of the kernel that can be resident at once
        int dev = 0, coop = 0, perSm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, dev);
        // (a level has one metric kind, so one cached count per level serves)
        if (full && g.nd == 3) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_bicg_persistent<3, true>, BP_TPB, 0);
        else if (full) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_bicg_persistent<2, true>, BP_TPB, 0);
        else if (g.nd == 3) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_bicg_persistent<3>, BP_TPB, 0);
        else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, k_bicg_persistent<2>, BP_TPB, 0);
        int cap = coop ? perSm * L->cx.sms : 0;
        L->bottom_coop_blocks = cap > MAX_RED_BLOCKS ? MAX_RED_BLOCKS : cap;
    }
    return row_blocks(g, L->cx.sms) <= L->bottom_coop_blocks && g.cells() <= (1ll << 24);
}

// blocks of the persistent kernel: one 32-cell item per warp and round, the rounds evenly filled
static int persistent_blocks(const Geom& g, int cap) {
    const long long items = (long long)g.n[1] * g.n[2] * ((g.n[0] + 31) / 32);
    const long long want = (items + BP_TPB / 32 - 1) / (BP_TPB / 32);
    const long long rounds = (want + cap - 1) / cap;
    const long long nb = (want + rounds - 1) / rounds;
    return (int)(nb < 1 ? 1 : nb);   // (the caller raises it to the canonical block count if that is larger)
}

// bottom of the V-cycle: num_smooth_bottom relaxations + BiCGStab (MappedMultiGrid.H:566-573) in one launch
int bottom_persistent(SomarLevel* L, SomarField* e_mg, const SomarField* r_mg) {
    const int bd = L->ndepths() - 1;
    Depth& dp = L->depths[bd];
    int rc;
    if (L->bwork.empty()) {
        L->bwork.resize(8, nullptr);
        for (auto& f : L->bwork)
            if ((rc = field_alloc(L, bd, &f))) return rc;
    }
    double* work[8];
    for (int q = 0; q < 8; ++q) work[q] = L->bwork[q]->p;
    return launch_bottom_persistent(L, dp.g, dp.m, e_mg->p, r_mg->p, work);
}

// the launch, on any single-rank geometry (the rank's own bottom depth, or the gathered one of bottom_gather.cu)
int launch_bottom_persistent(SomarLevel* L, const Geom& g_, const Metric& m_, double* phi, double* rhs,
                             double* const work[8]) {
    const SomarSolverParams& p = L->prm;
    int rc;
    if ((rc = ensure_state(L))) return rc;
    BicgDev* dv = static_cast<BicgDev*>(L->d_bicg);
    if (bottom_resident_usable(L, g_, m_))   // mode 3: the whole bottom problem resident on chip (bicg_resident.cu)
        return launch_bottom_resident(L, g_, m_, phi, rhs, work, dv->bar, dv->part, &dv->iters);
    BicgFields f;
    f.phi = phi; f.rhs = rhs;
    f.r = work[0]; f.rt = work[1]; f.e = work[2]; f.pv = work[3]; f.pt = work[4];
    f.stl = work[5]; f.t = work[6]; f.v = work[7];
    BicgArgs a;
    a.presmooth = p.num_smooth_bottom;
    a.nprecond = (p.num_smooth_precond == 0 || p.precondMode == 0) ? 0 : p.num_smooth_precond;
    a.numRestarts = p.bottom_numRestarts;
    a.bottomMetric = L->bottomMetric;
    a.eps = L->bottomEps; a.reps = p.bottom_reps; a.small = p.bottom_small; a.one_minus_hang = 1.0 - p.bottom_hang;
    a.imax = p.bottom_imax;
    a.normKind = (p.bottom_normType == 0) ? 0 : (p.bottom_normType == 1 ? 1 : 2);
    Geom g = g_;
    Metric m = m_;
    // tile of the fused preconditioner + operator pass (bp_tile_relax_apply): 16 x 8 core, shrunk until the tile with
    // its halo of 2*nprecond+1 cells and every z-plane fits; SOMAR_B200_BOTTOM_TILED=0 keeps one pass per barrier
    a.tcx = a.tcy = 0;
    {
        static const bool tiledOn = []() { const char* e = getenv("SOMAR_B200_BOTTOM_TILED"); return !(e && e[0] == '0'); }();
        if (tiledOn && m.diagonal && g.nd == 3 && a.nprecond > 0) {
            const int h = 2 * a.nprecond + 1;
            int cx = 16, cy = 8;
            while ((cx + 2 * h) * (cy + 2 * h) * g.n[2] > BP_TILE_CAP && (cx > 4 || cy > 4)) {
                if (cx > cy) cx /= 2; else cy /= 2;
            }
            if ((cx + 2 * h) * (cy + 2 * h) * g.n[2] <= BP_TILE_CAP) { a.tcx = cx; a.tcy = cy; }
        }
    }
    double* part = dv->part;
    int* itersAccum = &dv->iters;
    int cb = row_blocks(g, L->cx.sms);   // the blocks k_reduce would use: the canonical summation order
    int cap = L->bottom_coop_blocks;
    {   // SOMAR_B200_BOTTOM_BPS = blocks per SM of the persistent kernel (fewer blocks: cheaper grid barriers,
        // more cells per thread); default: everything that is co-resident
        static const int bps = []() { const char* e = getenv("SOMAR_B200_BOTTOM_BPS"); return e ? atoi(e) : 0; }();
        if (bps > 0 && bps * L->cx.sms < cap) cap = bps * L->cx.sms;
    }
    int nb = persistent_blocks(g, cap);
    if (nb < cb) nb = cb;
    // SOMAR_B200_GBAR=1 swaps cooperative groups' grid.sync() for the two-level counter barrier (tree_barrier).
    // Measured on B200 (headline, 380 BiCGStab iterations per solve): bottom 59.5 ms with it, 55.1 ms without --
    // grid.sync() is not limited by its single arrival word, so it stays the default.
    static const bool treeBar = []() { const char* e = getenv("SOMAR_B200_GBAR"); return e && e[0] == '1'; }();
    unsigned long long* bar = treeBar ? dv->bar : nullptr;
    if (bar) SOMAR_CUDA(cudaMemsetAsync(bar, 0, sizeof(dv->bar), L->cx.stream));
    void* args[] = {&g, &m, &f, &a, &part, &itersAccum, &cb, &bar};
    const dim3 grid(nb), block(BP_TPB);
    const void* fn = m.diagonal ? ((g.nd == 3) ? (const void*)k_bicg_persistent<3> : (const void*)k_bicg_persistent<2>)
                                : ((g.nd == 3) ? (const void*)k_bicg_persistent<3, true> : (const void*)k_bicg_persistent<2, true>);
    SOMAR_CUDA(cudaLaunchCooperativeKernel(fn, grid, block, args, 0, L->cx.stream));
    L->cx.launches++;
    L->bottom_iters_pending = true;
    return SOMAR_OK;
}

// iterations counted on the device since the last call (one 4-byte copy; called once per solve / cycle)
int bottom_collect_iters(SomarLevel* L) {
    if (!L->bottom_iters_pending) return SOMAR_OK;
    BicgDev* dv = static_cast<BicgDev*>(L->d_bicg);
    int n = 0;
    SOMAR_CUDA(cudaMemcpyAsync(L->h_bicg, &dv->iters, sizeof(int), cudaMemcpyDeviceToHost, L->cx.stream));
    SOMAR_CUDA(cudaMemsetAsync(&dv->iters, 0, sizeof(int), L->cx.stream));
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    n = *static_cast<int*>(L->h_bicg);
    L->bottom_iters_total += n;
    L->bottom_iters_pending = false;
    return SOMAR_OK;
}

}  // namespace somar
