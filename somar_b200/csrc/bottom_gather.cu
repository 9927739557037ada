// bottom_gather.cu -- the coarse-level gather: with several ranks the bottom depth of the V-cycle is replicated.
//
// The reference keeps the bottom depth distributed over all MPI ranks and pays, per BiCGStab iteration, five
// MPI_Allreduce calls and six ghost exchanges on a grid of a few hundred thousand cells
// (MappedMultiGrid.H:566-573 -> BiCGStabSolver; LevelData::exchange in every relaxation and operator call).
// On NVLink that is pure latency: measured at 8 GPUs the distributed bottom took 26 ms per V-cycle against 5 ms
// for the whole bottom visit on ONE GPU.  So every rank keeps a full-domain copy of the bottom depth's metric
// (gathered once at finalize), and a bottom visit is
//     all-gather of the ranks' slabs of the residual and the correction (a few MB over NVLink, one NCCL call)
//  -> the persistent single-GPU bottom kernel (bicg_persistent.cuh) on the full domain, redundantly on every rank
//  -> each rank keeps its own rows of the correction.
// Every rank computes the same bits, and they are the bits a single-rank run computes.
#include "level.hpp"
#include <algorithm>

namespace somar {

static int dalloc0(double** p, long long n) {
    SOMAR_CUDA(cudaMalloc(p, sizeof(double) * (size_t)n));
    SOMAR_CUDA(cudaMemset(*p, 0, sizeof(double) * (size_t)n));
    SOMAR_CUDA(cudaStreamSynchronize(nullptr));   // the level's stream is non-blocking: see dalloc in level.cu
    return SOMAR_OK;
}

void gathered_bottom_free(SomarLevel* L) {
    GatheredBottom* gb = L->gb;
    if (!gb) return;
    for (double* p : {gb->jg[0], gb->jg[1], gb->jg[2], gb->jinv, gb->lapdiag, gb->phi, gb->rhs, gb->sendbuf, gb->allbuf})
        if (p) cudaFree(p);
    for (double* p : gb->work)
        if (p) cudaFree(p);
    delete gb;
    L->gb = nullptr;
}

// rows [first valid row, +rows) of every z plane of a brick  <->  dense [planes][rows][pitch]
static int slab_copy(SomarLevel* L, const Geom& g, double* brick, int row0, int rows, double* dense, bool toBrick) {
    const int planes = g.n[2] + 2 * g.gh[2];
    double* b = brick + g.sy * (g.gh[1] + row0);
    const size_t width = sizeof(double) * (size_t)rows * g.sy;
    if (toBrick)
        SOMAR_CUDA(cudaMemcpy2DAsync(b, sizeof(double) * g.sz, dense, width, width, planes, cudaMemcpyDeviceToDevice,
                                     L->cx.stream));
    else
        SOMAR_CUDA(cudaMemcpy2DAsync(dense, width, b, sizeof(double) * g.sz, width, planes, cudaMemcpyDeviceToDevice,
                                     L->cx.stream));
    return SOMAR_OK;
}

// all-gather `narr` arrays of the local bottom depth (rows 0 .. n1-1+extra of every plane) into full-domain bricks
static int gather_arrays(SomarLevel* L, double* const* local, double* const* full, int narr, int extra) {
    GatheredBottom* gb = L->gb;
    const Geom& gl = L->depths[L->ndepths() - 1].g;
    const int planes = gl.n[2] + 2 * gl.gh[2];
    const size_t per = (size_t)planes * (gb->maxRows + 1) * gl.sy;   // one array of one rank (padded to the largest)
    int rc;
    for (int a = 0; a < narr; ++a)
        if ((rc = slab_copy(L, gl, local[a], 0, gl.n[1] + extra, gb->sendbuf + a * per, false))) return rc;
    SOMAR_NCCL(ncclAllGather(gb->sendbuf, gb->allbuf, per * narr, ncclDouble, L->comm, L->cx.stream));
    for (int r = 0; r < L->desc.nranks; ++r)   // in rank order: a rank's extra row is the next rank's row 0
        for (int a = 0; a < narr; ++a)
            if ((rc = slab_copy(L, gb->g, full[a], gb->row0[r], gb->rows[r] + extra,
                                gb->allbuf + ((size_t)r * narr + a) * per, true)))
                return rc;
    return SOMAR_OK;
}

// called at the end of level_finalize (metric of every depth complete, halos exchanged)
int gathered_bottom_build(SomarLevel* L) {
    gathered_bottom_free(L);
    const SomarLevelDesc& D = L->desc;
    if (D.nranks <= 1 || L->bottom_mode < 2 || L->ndepths() < 1) return SOMAR_OK;
    const int bd = L->ndepths() - 1;
    const Depth& dp = L->depths[bd];
    const Geom& gl = dp.g;
    if (!dp.m.diagonal || gl.periodic[0] || gl.periodic[1] || (gl.nd > 2 && gl.periodic[2])) return SOMAR_OK;
    if ((long long)gl.gn[0] * gl.gn[1] * gl.gn[2] > (1ll << 22)) return SOMAR_OK;   // replicate small bottoms only
    GatheredBottom* gb = new GatheredBottom();
    L->gb = gb;
    // every rank's rows at the bottom depth, from the box layout
    gb->row0.assign(D.nranks, 1 << 30);
    gb->rows.assign(D.nranks, 0);
    std::vector<int> hi(D.nranks, -1);
    for (int b = 0; b < D.nboxes; ++b) {
        const int r = L->box_rank[b];
        gb->row0[r] = std::min(gb->row0[r], L->box_lo[3 * b + 1]);
        hi[r] = std::max(hi[r], L->box_hi[3 * b + 1]);
    }
    gb->maxRows = 0;
    for (int r = 0; r < D.nranks; ++r) {
        const int lo = gb->row0[r] / dp.coarsening[1];
        gb->rows[r] = (hi[r] + 1) / dp.coarsening[1] - lo;
        gb->row0[r] = lo;
        gb->maxRows = std::max(gb->maxRows, gb->rows[r]);
    }
    // full-domain geometry: the local one with all rows, no neighbours, the domain's boundary conditions in y
    Geom g = gl;
    g.n[1] = g.gn[1];
    g.g0[1] = 0;
    g.sz = g.sy * (g.n[1] + 2 * g.gh[1]);
    g.total = g.sz * (g.n[2] + 2 * g.gh[2]);
    g.off0 = SOMAR_XO + g.sy * g.gh[1] + g.sz * g.gh[2];
    for (int s = 0; s < 2; ++s) g.bc[1][s] = D.bc_type[2 + s];
    g.nbr[0] = g.nbr[1] = -1;
    gb->g = g;
    int rc;
    for (int d = 0; d < g.nd; ++d)
        if ((rc = dalloc0(&gb->jg[d], g.total))) return rc;
    if ((rc = dalloc0(&gb->jinv, g.total)) || (rc = dalloc0(&gb->lapdiag, g.total)) || (rc = dalloc0(&gb->phi, g.total)) ||
        (rc = dalloc0(&gb->rhs, g.total)))
        return rc;
    for (double*& w : gb->work)
        if ((rc = dalloc0(&w, g.total))) return rc;
    const int planes = g.n[2] + 2 * g.gh[2];
    const size_t per = (size_t)planes * (gb->maxRows + 1) * g.sy;
    const int narrMax = 5;
    if ((rc = dalloc0(&gb->sendbuf, per * narrMax)) || (rc = dalloc0(&gb->allbuf, per * narrMax * D.nranks))) return rc;
    gb->m = dp.m;
    for (int d = 0; d < 3; ++d)
        for (int c = 0; c < 3; ++c) gb->m.jg[d][c] = (d == c) ? gb->jg[d] : nullptr;
    gb->m.jinv = gb->jinv;
    gb->m.lapdiag = gb->lapdiag;
    // metric: rows 0 .. n1 (the extra row holds the high y-face of the last cell row)
    double* loc[5] = {dp.jg[0][0], dp.jg[1][1], dp.jinv, dp.lapdiag, g.nd > 2 ? dp.jg[2][2] : nullptr};
    double* ful[5] = {gb->jg[0], gb->jg[1], gb->jinv, gb->lapdiag, g.nd > 2 ? gb->jg[2] : nullptr};
    return gather_arrays(L, loc, ful, g.nd > 2 ? 5 : 4, 1);
}

bool bottom_gathered_usable(SomarLevel* L) {
    if (!L->gb || L->bottom_mode < 2) return false;
    return bottom_persistent_params_ok(L, L->gb->g);
}

// one bottom visit on the replicated bottom depth
int bottom_gathered(SomarLevel* L, SomarField* e_mg, const SomarField* r_mg) {
    GatheredBottom* gb = L->gb;
    const Geom& gl = L->depths[L->ndepths() - 1].g;
    int rc;
    double* loc[2] = {r_mg->p, e_mg->p};
    double* ful[2] = {gb->rhs, gb->phi};
    if ((rc = gather_arrays(L, loc, ful, 2, 0))) return rc;
    if ((rc = launch_bottom_persistent(L, gb->g, gb->m, gb->phi, gb->rhs, gb->work))) return rc;
    // my rows of the correction: full brick -> local brick, plane by plane
    const int planes = gl.n[2] + 2 * gl.gh[2];
    const size_t width = sizeof(double) * (size_t)gl.n[1] * gl.sy;
    SOMAR_CUDA(cudaMemcpy2DAsync(e_mg->p + gl.sy * gl.gh[1], sizeof(double) * gl.sz,
                                 gb->phi + gb->g.sy * (gb->g.gh[1] + gl.g0[1]), sizeof(double) * gb->g.sz, width, planes,
                                 cudaMemcpyDeviceToDevice, L->cx.stream));
    return SOMAR_OK;
}

}  // namespace somar
