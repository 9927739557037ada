// level.cu -- level construction (semicoarsening schedule, HBM allocation, coarse metrics) and the per-depth
// operator methods that the C ABI exposes.
#include "level.hpp"
#include <algorithm>
#include <array>
#include <cmath>
#include <cstring>

namespace somar {

static thread_local std::string g_err;
void set_error(const std::string& s) { g_err = s; }
const char* last_error() { return g_err.c_str(); }

// ---------------------------------------------------------------------------------------------------------
// Semicoarsening schedule: MappedAMRPoissonOpFactory::MGnewOp, MappedAMRPoissonOpFactory.cpp:362-553
// (single AMR level, unforced ratios).  Computational dx only; s_maxCoarse = 4 (MappedAMRPoissonOp.cpp:55).
// ---------------------------------------------------------------------------------------------------------
static bool coarsenable(const SomarLevel* L, const int r[3]) {
    const int nb = L->desc.nboxes;
    for (int b = 0; b < nb; ++b)
        for (int d = 0; d < 3; ++d) {
            if (L->box_lo[3 * b + d] % r[d] != 0) return false;
            if ((L->box_hi[3 * b + d] + 1) % r[d] != 0) return false;
        }
    return true;
}

// Returns false where the factory would return NULL.  `ratios` holds the ratios of the depths above.
static bool schedule_depth(const SomarLevel* L, int depth, std::vector<std::array<int, 3>>& ratios, int coarsening[3],
                           double dx[3]) {
    const int nd = L->desc.dim;
    if (L->prm.maxDepth >= 0 && depth > L->prm.maxDepth) return false;
    const int maxCoarse[3] = {4, nd > 1 ? 4 : 1, nd > 2 ? 4 : 1};
    for (int d = 0; d < 3; ++d) { dx[d] = L->desc.dx[d]; coarsening[d] = 1; }
    int r[3] = {1, 1, 1};
    for (int i = 0; i < depth; ++i) {
        if (i < depth - 1) {
            for (int d = 0; d < 3; ++d) r[d] = ratios[i][d];
        } else {
            double maxDx = 0.0;
            for (int d = 0; d < nd; ++d) maxDx = std::max(maxDx, dx[d]);
            r[0] = r[1] = r[2] = 1;
            for (int d = 0; d < nd; ++d)
                if (dx[d] <= maxDx / 2.0) r[d] = 2;
            if (r[0] * r[1] * r[2] == 1) { r[0] = 2; r[1] = nd > 1 ? 2 : 1; r[2] = nd > 2 ? 2 : 1; }
        }
        for (int d = 0; d < 3; ++d) { dx[d] *= r[d]; coarsening[d] *= r[d]; }
    }
    auto test = [&](const int extra[3]) {
        int t[3];
        for (int d = 0; d < 3; ++d) t[d] = coarsening[d] * maxCoarse[d] * (extra ? extra[d] : 1);
        return coarsenable(L, t);
    };
    if (coarsening[0] * coarsening[1] * coarsening[2] > 1 && !test(nullptr)) {
        for (int d = 0; d < 3; ++d) { dx[d] /= r[d]; coarsening[d] /= r[d]; }
        r[0] = r[1] = r[2] = 1;
        for (int d = 0; d < nd; ++d) {
            r[d] = 2;
            if (!test(r)) r[d] = 1;
        }
        if (r[0] * r[1] * r[2] == 1) return false;
        int refDir = 0;
        for (refDir = 0; refDir < nd; ++refDir)
            if (r[refDir] == 1) break;
        if (refDir >= nd) refDir = nd - 1;   // see oracle/somar_oracle.cpp: reference reads dx[dims] here
        for (int d = 0; d < nd; ++d)
            if (r[d] > 1 && dx[d] / dx[refDir] > 0.5) r[d] = 1;
        if (r[0] * r[1] * r[2] == 1) return false;
        for (int d = 0; d < 3; ++d) { dx[d] *= r[d]; coarsening[d] *= r[d]; }
        if (coarsening[0] * coarsening[1] * coarsening[2] > 1 && !test(nullptr)) return false;
    }
    if (depth > 0) ratios.push_back({r[0], r[1], r[2]});
    return true;
}

static void make_geom(const SomarLevel* L, const int coarsening[3], const double dx[3], Geom& g) {
    const SomarLevelDesc& D = L->desc;
    std::memset(&g, 0, sizeof(g));
    g.nd = D.dim;
    for (int d = 0; d < 3; ++d) {
        g.gn[d] = D.n[d] / coarsening[d];
        g.g0[d] = L->local_lo[d] / coarsening[d];
        g.n[d] = (L->local_hi[d] + 1) / coarsening[d] - g.g0[d];
        g.dx[d] = dx[d];
        g.rdx[d] = 1.0 / dx[d];
        g.rdx2[d] = 1.0 / (dx[d] * dx[d]);
        g.periodic[d] = (d < D.dim) && D.periodic[d];
        g.bs[d] = (L->box_size[d] > 0) ? L->box_size[d] / coarsening[d] : g.gn[d];
        if (g.bs[d] < 1) g.bs[d] = 1;
    }
    g.gh[0] = SOMAR_XO;
    g.gh[1] = SOMAR_GHOST;
    g.gh[2] = (D.dim > 2) ? SOMAR_GHOST : 0;
    const long long pitch = ((g.n[0] + 2 * SOMAR_XO + 15) / 16) * 16;
    g.sy = pitch;
    g.sz = pitch * (g.n[1] + 2 * g.gh[1]);
    g.total = g.sz * (g.n[2] + 2 * g.gh[2]);
    g.off0 = SOMAR_XO + g.sy * g.gh[1] + g.sz * g.gh[2];
    for (int d = 0; d < 3; ++d)
        for (int s = 0; s < 2; ++s) {
            const bool active = d < D.dim;
            g.bct[d][s] = !active ? BC_NONE : (g.periodic[d] ? BC_PERIODIC : D.bc_type[2 * d + s]);
            const bool atDom = (s == 0) ? (g.g0[d] == 0) : (g.g0[d] + g.n[d] == g.gn[d]);
            g.bc[d][s] = (active && !g.periodic[d] && atDom) ? D.bc_type[2 * d + s] : BC_NONE;
            g.bcval[d][s] = D.bc_value[2 * d + s];
        }
    g.alpha = D.alpha;
    g.beta = D.beta;
    g.diriOrder = 1;   // EllipticConstDiriBCGhostClass: EllipticBCUtils.cpp:423
    // y neighbours
    g.nbr[0] = g.nbr[1] = -1;
    if (D.nranks > 1) {
        if (g.g0[1] > 0) g.nbr[0] = D.rank - 1;
        else if (g.periodic[1]) g.nbr[0] = D.nranks - 1;
        if (g.g0[1] + g.n[1] < g.gn[1]) g.nbr[1] = D.rank + 1;
        else if (g.periodic[1]) g.nbr[1] = 0;
    }
}

static int dalloc(double** p, long long n) {
    SOMAR_CUDA(cudaMalloc(p, sizeof(double) * (size_t)n));
    SOMAR_CUDA(cudaMemset(*p, 0, sizeof(double) * (size_t)n));
    // cudaMemset of device memory runs on the legacy default stream, asynchronously to the host; the level's stream is
    // non-blocking and would not wait for it: a kernel launched next on that stream could be overtaken by the zeros
    SOMAR_CUDA(cudaStreamSynchronize(nullptr));
    return SOMAR_OK;
}

int field_alloc(SomarLevel* L, int depth, SomarField** out) {
    if (depth < 0 || depth >= L->ndepths()) { set_error("field_alloc: bad depth"); return SOMAR_ERR_ARG; }
    SomarField* f = new SomarField{L, depth, nullptr};
    int rc = dalloc(&f->p, L->depths[depth].g.total);
    if (rc != SOMAR_OK) { delete f; return rc; }
    *out = f;
    return SOMAR_OK;
}

int level_build(SomarLevel* L) {
    const SomarLevelDesc& D = L->desc;
    const int nd = D.dim;
    // rank-local region = bounding box of the rank's boxes; must be tiled exactly and span x and z
    long long cells = 0;
    for (int d = 0; d < 3; ++d) { L->local_lo[d] = 1 << 30; L->local_hi[d] = -1; }
    int mine = 0;
    for (int b = 0; b < D.nboxes; ++b) {
        if (L->box_rank[b] != D.rank) continue;
        ++mine;
        long long c = 1;
        for (int d = 0; d < 3; ++d) {
            L->local_lo[d] = std::min(L->local_lo[d], L->box_lo[3 * b + d]);
            L->local_hi[d] = std::max(L->local_hi[d], L->box_hi[3 * b + d]);
            c *= (L->box_hi[3 * b + d] - L->box_lo[3 * b + d] + 1);
        }
        cells += c;
    }
    if (mine == 0) { set_error("this rank owns no box"); return SOMAR_ERR_LAYOUT; }
    long long bb = 1;
    for (int d = 0; d < 3; ++d) bb *= (L->local_hi[d] - L->local_lo[d] + 1);
    if (bb != cells) { set_error("boxes of this rank do not tile a rectangular region"); return SOMAR_ERR_LAYOUT; }
    if (D.nranks > 1 && nd == 2) { set_error("multi-rank runs need dim == 3 (partition in y)"); return SOMAR_ERR_LAYOUT; }
    for (int d = 0; d < 3; ++d) {
        if (d == 1 && nd == 3) continue;   // y is the partitioned direction
        if (L->local_lo[d] != 0 || L->local_hi[d] != D.n[d] - 1) {
            set_error("rank-local region must span the domain in x and z");
            return SOMAR_ERR_LAYOUT;
        }
    }
    // uniform tiling? (needed only by the full-metric kernels)
    for (int d = 0; d < 3; ++d) L->box_size[d] = L->box_hi[d] - L->box_lo[d] + 1;
    for (int b = 0; b < D.nboxes && L->box_size[0] > 0; ++b)
        for (int d = 0; d < 3; ++d) {
            const int s = L->box_hi[3 * b + d] - L->box_lo[3 * b + d] + 1;
            if (s != L->box_size[d] || L->box_lo[3 * b + d] % s != 0) { L->box_size[0] = L->box_size[1] = L->box_size[2] = 0; break; }
        }
    if (!D.is_diagonal && L->box_size[0] == 0) {
        set_error("full-metric operator needs a uniform box tiling (box-relative extrapolation, stencil_full.cuh)");
        return SOMAR_ERR_LAYOUT;
    }
    // depths
    std::vector<std::array<int, 3>> ratios;
    for (int depth = 0;; ++depth) {
        int coarsening[3];
        double dx[3];
        if (!schedule_depth(L, depth, ratios, coarsening, dx)) break;
        Depth dp;
        make_geom(L, coarsening, dx, dp.g);
        for (int d = 0; d < 3; ++d) dp.coarsening[d] = coarsening[d];
        dp.dxProduct = dx[0] * (nd > 1 ? dx[1] : 1.0) * (nd > 2 ? dx[2] : 1.0);
        if (depth > 0)
            for (int d = 0; d < 3; ++d) L->depths.back().ratio[d] = ratios.back()[d];
        L->depths.push_back(dp);
    }
    if (L->depths.empty()) { set_error("no multigrid depth could be built"); return SOMAR_ERR_ARG; }
    // allocation
    for (int k = 0; k < L->ndepths(); ++k) {
        Depth& dp = L->depths[k];
        const long long n = dp.g.total;
        int rc;
        for (int d = 0; d < nd; ++d)
            for (int c = 0; c < nd; ++c) {
                if (D.is_diagonal && c != d) continue;
                if ((rc = dalloc(&dp.jg[d][c], n)) != SOMAR_OK) return rc;
            }
        if ((rc = dalloc(&dp.jinv, n)) != SOMAR_OK) return rc;
        if ((rc = dalloc(&dp.lapdiag, n)) != SOMAR_OK) return rc;
        if ((rc = dalloc(&dp.tmp, n)) != SOMAR_OK) return rc;
        for (int d = 0; d < 3; ++d)
            for (int c = 0; c < 3; ++c) dp.m.jg[d][c] = dp.jg[d][c];
        dp.m.jinv = dp.jinv;
        dp.m.lapdiag = dp.lapdiag;
        dp.m.diagonal = D.is_diagonal;
        if (k > 0) {
            if ((rc = field_alloc(L, k, &dp.res)) != SOMAR_OK) return rc;
            if ((rc = field_alloc(L, k, &dp.cor)) != SOMAR_OK) return rc;
        }
        if (D.nranks > 1) {
            const long long hb = (long long)SOMAR_GHOST * (dp.g.n[0] + 2 * SOMAR_GHOST) * (dp.g.n[2] + 2 * dp.g.gh[2]);
            for (int s = 0; s < 2; ++s) {
                if ((rc = dalloc(&dp.sendbuf[s], hb)) != SOMAR_OK) return rc;
                if ((rc = dalloc(&dp.recvbuf[s], hb)) != SOMAR_OK) return rc;
            }
        }
    }
    return SOMAR_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Halo exchange: replaces LevelData::exchange(Copier) (GSRB.cpp:81,175; MappedAMRPoissonOp.cpp:2222-2238).
// y neighbours through NCCL send/recv of packed rows (NVLink), then periodic images in x and z inside the
// brick; sweeping y -> x -> z over the allocated extent also fills edges and corners.
// ---------------------------------------------------------------------------------------------------------
int exchange(SomarLevel* L, int depth, double* x, int width) {
    Depth& dp = L->depths[depth];
    const Geom& g = dp.g;
    int p2p = 0;
    if (L->desc.nranks > 1 && (g.nbr[0] >= 0 || g.nbr[1] >= 0)) {
        p2p = halo_p2p_exchange(L, depth, x, width);   // peer-memory path (two launches), when the arenas are mapped
        if (p2p < 0) return SOMAR_ERR_COMM;
    }
    if (p2p == 1) {
        // done
    } else if (L->desc.nranks > 1 && (g.nbr[0] >= 0 || g.nbr[1] >= 0)) {
        const size_t cnt = (size_t)width * (g.n[0] + 2 * SOMAR_GHOST) * (g.n[2] + 2 * g.gh[2]);
        for (int s = 0; s < 2; ++s)
            if (g.nbr[s] >= 0) launch_pack_y(L->cx, g, x, dp.sendbuf[s], s, width, false);
        SOMAR_NCCL(ncclGroupStart());
        // Messages between one pair of ranks match in posting order.  With two ranks and periodic y both
        // neighbours are the SAME peer, so post by travel direction: first the rows going down (my low rows ->
        // the peer's high ghosts), then the rows going up.
        for (int s = 0; s < 2; ++s) {
            if (g.nbr[s] >= 0)
                SOMAR_NCCL(ncclSend(dp.sendbuf[s], cnt, ncclDouble, g.nbr[s], L->comm, L->cx.stream));
            if (g.nbr[1 - s] >= 0)
                SOMAR_NCCL(ncclRecv(dp.recvbuf[1 - s], cnt, ncclDouble, g.nbr[1 - s], L->comm, L->cx.stream));
        }
        SOMAR_NCCL(ncclGroupEnd());
        for (int s = 0; s < 2; ++s)
            if (g.nbr[s] >= 0) launch_pack_y(L->cx, g, x, dp.recvbuf[s], s, width, true);
    } else if (g.periodic[1]) {
        launch_periodic_fill(L->cx, g, x, 1, width);
    }
    if (g.periodic[0]) launch_periodic_fill(L->cx, g, x, 0, width);
    if (g.nd > 2 && g.periodic[2]) launch_periodic_fill(L->cx, g, x, 2, width);
    return SOMAR_OK;
}

// Global scalar: local two-stage reduction, NCCL all-reduce across ranks (replaces the MPI_Allreduce calls of
// MappedAMRMultiGrid.H:825, computeMappedNorm.cpp, LevelDataOps), one 8-byte D2H copy.
int reduce_global(SomarLevel* L, int depth, const double* x, const double* y, int kind, double* out) {
    const Geom& g = L->depths[depth].g;
    launch_reduce(L->cx, g, x, y, kind, L->d_scal);
    if (L->desc.nranks > 1) {
        const bool isMax = (kind == 0 || kind == 4);
        SOMAR_NCCL(ncclAllReduce(L->d_scal, L->d_scal, 1, ncclDouble, isMax ? ncclMax : ncclSum, L->comm, L->cx.stream));
    }
    SOMAR_CUDA(cudaMemcpyAsync(L->h_scal, L->d_scal, sizeof(double), cudaMemcpyDeviceToHost, L->cx.stream));
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    *out = L->h_scal[0];
    return SOMAR_OK;
}

// ---------------------------------------------------------------------------------------------------------
// Operator methods
// ---------------------------------------------------------------------------------------------------------
// applyOpI / residualI -- MappedAMRPoissonOp.cpp:772-898, 646-677.  mode 0: out = L phi; 1: out = rhs - L phi.
int op_apply(SomarLevel* L, int depth, double* out, SomarField* phi, const double* rhs, int mode, bool homog) {
    Depth& dp = L->depths[depth];
    if (out == phi->p) { set_error("applyOp/residual: lhs must not alias phi (MappedAMRPoissonOp.cpp:653)"); return SOMAR_ERR_ARG; }
    int rc = exchange(L, depth, phi->p, 1);
    if (rc != SOMAR_OK) return rc;
    launch_apply(L->cx, dp.g, dp.m, out, phi->p, rhs, mode, homog);
    return SOMAR_OK;
}

// residual at depth 0 and its max-norm over all ranks (MappedAMRMultiGrid::computeAMRResidual + the norm of
// solveNoInitResid, MappedAMRMultiGrid.H:793-839): one pass where k_march runs, else residual then k_reduce
int residual_maxnorm(SomarLevel* L, SomarField* res, SomarField* phi, const SomarField* rhs, bool homog, double* out) {
    Depth& dp = L->depths[0];
    if (res->p == phi->p) { set_error("residual: lhs must not alias phi (MappedAMRPoissonOp.cpp:653)"); return SOMAR_ERR_ARG; }
    int rc = exchange(L, 0, phi->p, 1);
    if (rc != SOMAR_OK) return rc;
    SOMAR_CUDA(cudaMemsetAsync(L->d_scal, 0, sizeof(double), L->cx.stream));
    if (!launch_residual_maxnorm(L->cx, dp.g, dp.m, res->p, phi->p, rhs->p, homog, L->d_scal)) {
        launch_apply(L->cx, dp.g, dp.m, res->p, phi->p, rhs->p, 1, homog);
        return reduce_global(L, 0, res->p, nullptr, 0, out);
    }
    if (L->desc.nranks > 1)
        SOMAR_NCCL(ncclAllReduce(L->d_scal, L->d_scal, 1, ncclDouble, ncclMax, L->comm, L->cx.stream));
    SOMAR_CUDA(cudaMemcpyAsync(L->h_scal, L->d_scal, sizeof(double), cudaMemcpyDeviceToHost, L->cx.stream));
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    *out = L->h_scal[0];
    return SOMAR_OK;
}

static int ensure_line(SomarLevel* L, Depth& dp) {
    if (dp.line_valid) return SOMAR_OK;
    int rc;
    if (!dp.line_fact && (rc = dalloc(&dp.line_fact, dp.g.total)) != SOMAR_OK) return rc;
    if (!dp.line_dmod && (rc = dalloc(&dp.line_dmod, dp.g.total)) != SOMAR_OK) return rc;
    launch_line_factor(L->cx, dp.g, dp.m, dp.line_fact, dp.line_dmod);
    dp.line_valid = true;
    return SOMAR_OK;
}

// MappedAMRPoissonOp::relax -- MappedAMRPoissonOp.cpp:1733-1743 -> RelaxationMethod::relax
int op_relax(SomarLevel* L, int depth, SomarField* phi, const SomarField* rhs, int iters, int modeOverride, bool phiIsZero) {
    Depth& dp = L->depths[depth];
    const int mode = (modeOverride != -100) ? modeOverride : L->prm.relaxMode;
    int rc;
    if (mode == SOMAR_RELAX_LINE_GSRB && dp.g.bs[dp.g.nd - 1] != dp.g.gn[dp.g.nd - 1]) {
        // LineGSRB::relax solves one tridiagonal system per BOX column (GSRB.cpp:177-330); a box side inside the domain gets
        // the low-order closure (coefficient 0, GSRBF.ChF:1590-1604), so vertically split boxes give a different smoother.
        // The kernels here solve whole columns, which is that smoother only when every box spans the vertical.
        set_error("relax: LineGSRB needs boxes that span the vertical (per-box line solves on vertically split boxes are not built)");
        return SOMAR_ERR_LAYOUT;
    }
    // phiIsZero: the caller skipped "set the correction to zero".  The pipelined smoother then takes every input plane
    // of its first iteration from the level's zero plane (no DRAM read of phi, no fill pass, no halo exchange: every
    // rank's phi is zero); any other smoother gets the fill it would have had.
    bool zeroIn = false;
    if (phiIsZero) {
        zeroIn = iters > 0 && mode == SOMAR_RELAX_LEVEL_GSRB && L->use_fused == 2 && gsrb_async_usable(dp.g, dp.m);
        if (zeroIn && !L->zeroPlane) {
            const Geom& g0 = L->depths[0].g;
            const size_t nz = (size_t)(g0.off0 + g0.sz);
            if (cudaMalloc(&L->zeroPlane, nz * sizeof(double)) != cudaSuccess ||
                cudaMemsetAsync(L->zeroPlane, 0, nz * sizeof(double), L->cx.stream) != cudaSuccess) {
                (void)cudaGetLastError();
                cudaFree(L->zeroPlane);
                L->zeroPlane = nullptr;
                zeroIn = false;
            }
        }
        if (!zeroIn) launch_fill(L->cx, dp.g, phi->p, 0.0);
    }
    const bool coarse = depth > 0 && depth < L->ndepths() - 1;
    ProfScope psCoarse(L, coarse ? PH_SMOOTH_COARSE : -1);
    for (int it = 0; it < iters; ++it) {
        if (mode == SOMAR_RELAX_LEVEL_GSRB && L->use_fused == 2 && gsrb_async_usable(dp.g, dp.m)) {
            // Both colour passes in one sweep.  With y-neighbour ranks the sweep also relaxes the red cells of the
            // first ghost row on each side (same inputs, same arithmetic as their owner), so ONE exchange of two
            // ghost rows of phi per iteration replaces the reference's exchange before each colour pass; the
            // metric's ghost rows were filled at finalize, the right-hand side's are filled here once.
            const bool ranks = dp.g.nbr[0] >= 0 || dp.g.nbr[1] >= 0 || dp.g.periodic[1];
            if (ranks && it == 0 && (rc = exchange(L, depth, const_cast<double*>(rhs->p), 1)) != SOMAR_OK) return rc;
            const bool fromZero = zeroIn && it == 0;
            if (ranks && !fromZero && (rc = exchange(L, depth, phi->p, 2)) != SOMAR_OK) return rc;
            ProfScope psK(L, depth == 0 ? PH_SMOOTH0_KERNEL : -1);
            const double* zp = fromZero ? L->zeroPlane : nullptr;
            const bool launched = gsrb_tma_usable(dp.g, dp.m) ? launch_gsrb_tma(L->cx, dp.g, dp.m, dp.tmp, phi->p, rhs->p, zp)
                                                              : launch_gsrb_async(L->cx, dp.g, dp.m, dp.tmp, phi->p, rhs->p, zp);
            if (!launched) {
                set_error("relax: k_gsrb_async could not reserve its shared memory");
                return SOMAR_ERR_CUDA;
            }
            std::swap(dp.tmp, phi->p);
        } else if (mode == SOMAR_RELAX_LEVEL_GSRB && L->use_fused && gsrb_fused_usable(dp.g, dp.m)) {
            // register-pipelined variant (single rank only)
            ProfScope psK(L, depth == 0 ? PH_SMOOTH0_KERNEL : -1);
            launch_gsrb_fused(L->cx, dp.g, dp.m, dp.tmp, phi->p, rhs->p);
            std::swap(dp.tmp, phi->p);
        } else if (mode == SOMAR_RELAX_LEVEL_GSRB) {   // LevelGSRB::relax, GSRB.cpp:58-98
            for (int pass = 0; pass < 2; ++pass) {
                if ((rc = exchange(L, depth, phi->p, 1)) != SOMAR_OK) return rc;
                ProfScope psK(L, depth == 0 ? PH_SMOOTH0_KERNEL : -1);
                if (dp.m.diagonal) {
                    launch_gsrb(L->cx, dp.g, dp.m, phi->p, rhs->p, pass);
                } else {
                    launch_gsrb_full(L->cx, dp.g, dp.m, dp.tmp, phi->p, rhs->p, pass);
                    std::swap(dp.tmp, phi->p);
                }
            }
        } else if (mode == SOMAR_RELAX_LINE_GSRB && line_compact_usable(L, dp)) {
            // all the iterations of the call in the colour-split layout (line_compact.cu): same statements, same bits
            const bool stale = !dp.line_valid;
            if ((rc = ensure_line(L, dp)) != SOMAR_OK) return rc;
            if ((rc = line_compact_relax(L, dp, phi->p, rhs->p, iters, stale, depth == 0 ? PH_SMOOTH0_KERNEL : -1)) != SOMAR_OK)
                return rc;
            break;
        } else if (mode == SOMAR_RELAX_LINE_GSRB) {    // LineGSRB::relax, GSRB.cpp:148-330
            if ((rc = ensure_line(L, dp)) != SOMAR_OK) return rc;
            for (int pass = 0; pass < 2; ++pass) {
                if ((rc = exchange(L, depth, phi->p, 1)) != SOMAR_OK) return rc;
                ProfScope psK(L, depth == 0 ? PH_SMOOTH0_KERNEL : -1);
                if (dp.m.diagonal) {
                    launch_line_gsrb(L->cx, dp.g, dp.m, phi->p, phi->p, rhs->p, dp.line_fact, dp.line_dmod, pass);
                } else {
                    launch_line_gsrb(L->cx, dp.g, dp.m, dp.tmp, phi->p, rhs->p, dp.line_fact, dp.line_dmod, pass);
                    std::swap(dp.tmp, phi->p);
                }
            }
        } else if (mode == SOMAR_RELAX_JACOBI) {       // Jacobi::relax, Jacobi.cpp:50-90
            if ((rc = op_apply(L, depth, dp.tmp, phi, rhs->p, 1, true)) != SOMAR_OK) return rc;
            launch_jacobi(L->cx, dp.g, dp.m, phi->p, dp.tmp);
        } else if (mode == SOMAR_RELAX_NONE) {
            return SOMAR_OK;
        } else if (mode == SOMAR_RELAX_LOOSE_GSRB) {   // LooseGSRB::relax, GSRB.cpp:104-141 (see k_gsrb_loose)
            if (!dp.m.diagonal) {
                set_error("relax: LooseGSRB with a full metric is not built (its extrap copy carries box-interface ghosts "
                          "of the previous iteration)");
                return SOMAR_ERR_ARG;
            }
            if (L->box_size[0] == 0) { set_error("relax: LooseGSRB needs a uniform box tiling"); return SOMAR_ERR_LAYOUT; }
            launch_gsrb_loose(L->cx, dp.g, dp.m, phi->p, nullptr, rhs->p, 0, 0);
            launch_gsrb_loose(L->cx, dp.g, dp.m, phi->p, nullptr, rhs->p, 1, 0);
            if ((rc = exchange(L, depth, phi->p, 1)) != SOMAR_OK) return rc;      // exchangeEnd
            launch_copy(L->cx, dp.g, dp.tmp, phi->p);                               // what the box ghosts hold from here on
            launch_gsrb_loose(L->cx, dp.g, dp.m, phi->p, dp.tmp, rhs->p, 0, 1);
            launch_gsrb_loose(L->cx, dp.g, dp.m, phi->p, dp.tmp, rhs->p, 1, 1);
        } else {
            set_error("relax: unknown relax_mode");
            return SOMAR_ERR_ARG;
        }
    }
    return SOMAR_OK;
}

// MappedAMRPoissonOp::preCond -- MappedAMRPoissonOp.cpp:684-734
int op_precond(SomarLevel* L, int depth, SomarField* cor, const SomarField* res) {
    Depth& dp = L->depths[depth];
    if (L->prm.num_smooth_precond == 0 || L->prm.precondMode == 0) {
        launch_copy(L->cx, dp.g, cor->p, res->p);
        return SOMAR_OK;
    }
    launch_diag_precond(L->cx, dp.g, dp.m, cor->p, res->p);
    const int mode = (L->prm.precondMode == 2) ? SOMAR_RELAX_LINE_GSRB : L->prm.relaxMode;
    return op_relax(L, depth, cor, res, L->prm.num_smooth_precond, mode);
}

// restrictResidual -- MappedAMRPoissonOp.cpp:1281-1304 (homogeneous residual, then J-weighted average)
int op_restrict(SomarLevel* L, int depth, SomarField* crse, SomarField* phi, const SomarField* rhs) {
    if (depth + 1 >= L->ndepths()) { set_error("restrictResidual: no coarser depth"); return SOMAR_ERR_ARG; }
    Depth& dp = L->depths[depth];
    int rc = exchange(L, depth, phi->p, 1);
    if (rc != SOMAR_OK) return rc;
    launch_restrict_residual(L->cx, dp.g, dp.m, L->depths[depth + 1].g, crse->p, phi->p, rhs->p, dp.ratio, dp.tmp);
    return SOMAR_OK;
}

// prolongIncrement -- ConstInterpPS / ZeroAvgConstInterpPS, ProlongationStrategy.cpp:49-165
int op_prolong(SomarLevel* L, int depth, SomarField* phi, const SomarField* crse) {
    if (depth + 1 >= L->ndepths()) { set_error("prolongIncrement: no coarser depth"); return SOMAR_ERR_ARG; }
    Depth& dp = L->depths[depth];
    double* sums = dp.zeroAvg ? (L->d_scal + 2) : nullptr;
    launch_prolong(L->cx, dp.g, dp.m, L->depths[depth + 1].g, phi->p, crse->p, dp.ratio, dp.dxProduct, sums);
    if (dp.zeroAvg) {
        if (L->desc.nranks > 1)   // the two MPI_Allreduce(SUM) of ProlongationStrategy.cpp:141,148 in one call
            SOMAR_NCCL(ncclAllReduce(sums, sums, 2, ncclDouble, ncclSum, L->comm, L->cx.stream));
        launch_prolong_finish(L->cx, dp.g, L->depths[depth + 1].g, phi->p, crse->p, dp.ratio, sums);
    }
    return SOMAR_OK;
}

// ---------------------------------------------------------------------------------------------------------
// finalize: coarse metrics (fill_MGfields, MappedAMRPoissonOpFactory.cpp:1164-1235), lapDiag, and the
// null-space probe that selects the prolongation strategy (:659-693).
// ---------------------------------------------------------------------------------------------------------
int level_finalize(SomarLevel* L) {
    const int nd = L->desc.dim;
    int rc;
    for (int k = 0; k < L->ndepths(); ++k) {
        Depth& dp = L->depths[k];
        if (k > 0) {
            Depth& fn = L->depths[k - 1];
            for (int d = 0; d < nd; ++d)
                for (int c = 0; c < nd; ++c)
                    if (dp.jg[d][c]) launch_avg_face(L->cx, fn.g, dp.g, dp.jg[d][c], fn.jg[d][c], d, fn.ratio);
            launch_avg_harmonic(L->cx, fn.g, dp.g, dp.jinv, fn.jinv, fn.ratio);
        }
        // The operator needs no metric halo: a cell only reads its own faces, and the high face of the last cell of
        // a row/column is supplied by the caller's FluxBox (stored in the first ghost slot).
        launch_lapdiag(L->cx, dp.g, dp.m, dp.lapdiag);
        dp.line_valid = false;
        // The fused smoother also relaxes the red cells of the first ghost row towards a neighbour rank
        // (op_relax), so the diagonal metric gets two ghost rows from the neighbours, once.
        if ((L->desc.nranks > 1 || dp.g.periodic[1]) && dp.m.diagonal) {
            double* arr[5] = {dp.jg[0][0], dp.jg[1][1], nd > 2 ? dp.jg[2][2] : nullptr, dp.jinv, dp.lapdiag};
            for (double* a : arr)
                if (a && (rc = exchange(L, k, a, SOMAR_GHOST)) != SOMAR_OK) return rc;
        }
    }
    // null-space probe per depth: rhs = L[0]; res = rhs - L[1]; zeroAvg iff |max(res)| < 0.01 * eps
    for (int k = 0; k < L->ndepths(); ++k) {
        Depth& dp = L->depths[k];
        SomarField *phi = nullptr, *rhs = nullptr, *res = nullptr;
        if ((rc = field_alloc(L, k, &phi)) || (rc = field_alloc(L, k, &rhs)) || (rc = field_alloc(L, k, &res))) return rc;
        if ((rc = op_apply(L, k, rhs->p, phi, nullptr, 0, true)) != SOMAR_OK) return rc;
        launch_fill(L->cx, dp.g, phi->p, 1.0);
        if ((rc = op_apply(L, k, res->p, phi, rhs->p, 1, true)) != SOMAR_OK) return rc;
        double mx = 0.0;
        if ((rc = reduce_global(L, k, res->p, nullptr, 4, &mx)) != SOMAR_OK) return rc;
        dp.zeroAvg = (std::fabs(mx) < 0.01 * L->prm.eps);
        cudaFree(phi->p); cudaFree(rhs->p); cudaFree(res->p);
        delete phi; delete rhs; delete res;
    }
    if ((rc = gathered_bottom_build(L)) != SOMAR_OK) return rc;   // multi-rank: replicate the bottom depth
    L->finalized = true;
    return SOMAR_OK;
}

}  // namespace somar
