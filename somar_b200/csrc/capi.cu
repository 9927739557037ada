// capi.cu -- the extern "C" surface declared in include/somar_b200.h.
#include "level.hpp"
#include <cstdlib>
#include <cstring>

namespace somar { const char* last_error(); }
using namespace somar;

#define CHECK_LEVEL(L) do { if (!(L)) { set_error("null level"); return SOMAR_ERR_ARG; } } while (0)
#define CHECK_FINAL(L) do { CHECK_LEVEL(L); if (!(L)->finalized) { set_error("level not finalized"); return SOMAR_ERR_ARG; } \
                            cudaSetDevice((L)->device); } while (0)
#define SAME_DEPTH(a, b) do { if ((a)->depth != (b)->depth || (a)->lev != (b)->lev) { set_error("fields of different depth/level"); return SOMAR_ERR_ARG; } } while (0)

// 3-D copy between a host Fortran-ordered array over [lo,hi] (global indices) and the brick, region [rlo,rhi].
static int copy3d(SomarLevel* L, const Geom& g, double* brick, double* host, const int lo[3], const int hi[3],
                  const int rlo[3], const int rhi[3], bool toDevice) {
    for (int d = 0; d < 3; ++d)
        if (rlo[d] < lo[d] || rhi[d] > hi[d] || rhi[d] < rlo[d]) { set_error("copy3d: region outside the host array"); return SOMAR_ERR_ARG; }
    const size_t fnx = hi[0] - lo[0] + 1, fny = hi[1] - lo[1] + 1;
    cudaMemcpy3DParms p;
    std::memset(&p, 0, sizeof(p));
    cudaPitchedPtr hp = make_cudaPitchedPtr((void*)host, fnx * sizeof(double), fnx * sizeof(double), fny);
    cudaPitchedPtr dp = make_cudaPitchedPtr((void*)brick, g.sy * sizeof(double), g.sy * sizeof(double), g.n[1] + 2 * g.gh[1]);
    cudaPos hpos = make_cudaPos((size_t)(rlo[0] - lo[0]) * sizeof(double), rlo[1] - lo[1], rlo[2] - lo[2]);
    const int li = rlo[0] - g.g0[0], lj = rlo[1] - g.g0[1], lk = rlo[2] - g.g0[2];
    if (li < -g.gh[0] || lj < -g.gh[1] || lk < -g.gh[2] || rhi[0] - g.g0[0] >= g.n[0] + g.gh[0] ||
        rhi[1] - g.g0[1] >= g.n[1] + g.gh[1] || rhi[2] - g.g0[2] >= g.n[2] + (g.gh[2] ? g.gh[2] : 1)) {
        set_error("copy3d: region outside this rank's brick");
        return SOMAR_ERR_LAYOUT;
    }
    cudaPos dpos = make_cudaPos((size_t)(SOMAR_XO + li) * sizeof(double), g.gh[1] + lj, g.gh[2] + lk);
    p.extent = make_cudaExtent((size_t)(rhi[0] - rlo[0] + 1) * sizeof(double), rhi[1] - rlo[1] + 1, rhi[2] - rlo[2] + 1);
    if (toDevice) { p.srcPtr = hp; p.srcPos = hpos; p.dstPtr = dp; p.dstPos = dpos; p.kind = cudaMemcpyHostToDevice; }
    else { p.srcPtr = dp; p.srcPos = dpos; p.dstPtr = hp; p.dstPos = hpos; p.kind = cudaMemcpyDeviceToHost; }
    SOMAR_CUDA(cudaMemcpy3DAsync(&p, L->cx.stream));
    return SOMAR_OK;
}

extern "C" {

const char* somar_last_error(void) { return last_error(); }
const char* somar_version(void) { return "somar_b200 0.1 (sm_100a)"; }

int somar_default_params(SomarSolverParams* p) {
    if (!p) return SOMAR_ERR_ARG;
    // src/utils/ProblemContext.cpp:1147-1216
    p->eps = 1e-6; p->num_smooth_down = 2; p->num_smooth_up = 2; p->num_smooth_bottom = 2; p->num_smooth_precond = 2;
    p->numMG = 1; p->imin = 5; p->imax = 20; p->hang = 1e-15; p->normThresh = 1e-30; p->maxDepth = -1;
    p->relaxMode = SOMAR_RELAX_LEVEL_GSRB; p->precondMode = 1;
    p->bottom_eps = 1e-6; p->bottom_reps = 1e-12; p->bottom_hang = 1e-7; p->bottom_small = 1e-30;
    p->bottom_imax = 80; p->bottom_numRestarts = 5; p->bottom_normType = 2; p->verbosity = 0;
    return SOMAR_OK;
}

int somar_device_count(int* count) {
    if (!count) return SOMAR_ERR_ARG;
    SOMAR_CUDA(cudaGetDeviceCount(count));
    return SOMAR_OK;
}

int somar_comm_unique_id(unsigned char id[128]) {
    ncclUniqueId u;
    SOMAR_NCCL(ncclGetUniqueId(&u));
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
    std::memcpy(id, &u, 128);
    return SOMAR_OK;
}

int somar_level_create(const SomarLevelDesc* desc, const SomarSolverParams* params, int device,
                       const unsigned char* comm_id, SomarLevel** out) {
    return level_create_impl(desc, params, device, comm_id, nullptr, out);
}

}  // extern "C"

// `borrowed`: a communicator owned by another level of the same ranks (the leptic solver's derived levels); the new
// level uses it and does not destroy it.  The caller keeps the levels' NCCL work from overlapping (stream syncs).
int somar::level_create_impl(const SomarLevelDesc* desc, const SomarSolverParams* params, int device,
                             const unsigned char* comm_id, ncclComm_t borrowed, SomarLevel** out) {
    if (!desc || !params || !out) { set_error("null argument"); return SOMAR_ERR_ARG; }
    if (desc->dim != 2 && desc->dim != 3) { set_error("dim must be 2 or 3"); return SOMAR_ERR_ARG; }
    if (desc->dim == 2 && desc->n[2] != 1) { set_error("2-D problems need n[2] == 1"); return SOMAR_ERR_ARG; }
    if (desc->nboxes < 1 || !desc->box_lo || !desc->box_hi) { set_error("no boxes"); return SOMAR_ERR_ARG; }
    if (desc->nranks > 1 && !comm_id && !borrowed) { set_error("nranks > 1 needs a communicator id"); return SOMAR_ERR_ARG; }
    int ndev = 0;
    SOMAR_CUDA(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) { set_error("no such CUDA device (there is no CPU fallback)"); return SOMAR_ERR_CUDA; }
    SOMAR_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    SOMAR_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) { set_error("libsomar_b200 is built for sm_100a only"); return SOMAR_ERR_CUDA; }
    SomarLevel* L = new SomarLevel;
    L->desc = *desc;
    L->prm = *params;
    L->device = device;
    L->box_lo.assign(desc->box_lo, desc->box_lo + 3 * desc->nboxes);
    L->box_hi.assign(desc->box_hi, desc->box_hi + 3 * desc->nboxes);
    if (desc->box_rank) L->box_rank.assign(desc->box_rank, desc->box_rank + desc->nboxes);
    else L->box_rank.assign(desc->nboxes, 0);
    L->desc.box_lo = L->box_lo.data();
    L->desc.box_hi = L->box_hi.data();
    L->desc.box_rank = L->box_rank.data();
    L->cx.sms = prop.multiProcessorCount;
    if (const char* e = getenv("SOMAR_B200_FUSED")) L->use_fused = (e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 2;
    if (const char* e = getenv("SOMAR_B200_BOTTOM")) L->bottom_mode = (e[0] >= '0' && e[0] <= '3') ? e[0] - '0' : 3;
    int rc;
    cudaError_t ce;
    if ((ce = cudaStreamCreateWithFlags(&L->cx.stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (ce = cudaMalloc(&L->cx.scratch, sizeof(double) * 2 * SOMAR_MAX_RED_BLOCKS)) != cudaSuccess ||
        (ce = cudaMalloc(&L->cx.tickets, sizeof(unsigned) * 8)) != cudaSuccess ||
        (ce = cudaMemset(L->cx.tickets, 0, sizeof(unsigned) * 8)) != cudaSuccess ||
        (ce = cudaMalloc(&L->d_scal, sizeof(double) * 8)) != cudaSuccess ||
        (ce = cudaMemset(L->d_scal, 0, sizeof(double) * 8)) != cudaSuccess ||
        (ce = cudaStreamSynchronize(nullptr)) != cudaSuccess ||   // default-stream memsets vs the non-blocking stream
        (ce = cudaMallocHost(&L->h_scal, sizeof(double) * 8)) != cudaSuccess ||
        (ce = cudaEventCreate(&L->ev0)) != cudaSuccess || (ce = cudaEventCreate(&L->ev1)) != cudaSuccess) {
        set_error(std::string("level_create: ") + cudaGetErrorString(ce));
        delete L;
        return SOMAR_ERR_CUDA;
    }
    if (desc->nranks > 1 && borrowed) {
        L->comm = borrowed;
        L->comm_borrowed = true;
    } else if (desc->nranks > 1) {
        ncclUniqueId u;
        std::memcpy(&u, comm_id, 128);
        ncclResult_t ne = ncclCommInitRank(&L->comm, desc->nranks, u, desc->rank);
        if (ne != ncclSuccess) { set_error(std::string("ncclCommInitRank: ") + ncclGetErrorString(ne)); delete L; return SOMAR_ERR_COMM; }
    }
    if ((rc = level_build(L)) != SOMAR_OK) { delete L; return rc; }
    *out = L;
    return SOMAR_OK;
}

extern "C" {

int somar_level_destroy(SomarLevel* L) {
    if (!L) return SOMAR_OK;
    cudaSetDevice(L->device);
    cudaStreamSynchronize(L->cx.stream);
    for (auto& dp : L->depths) {
        for (int d = 0; d < 3; ++d)
            for (int c = 0; c < 3; ++c) cudaFree(dp.jg[d][c]);
        cudaFree(dp.jinv); cudaFree(dp.lapdiag); cudaFree(dp.tmp); cudaFree(dp.line_fact); cudaFree(dp.line_dmod);
        line_compact_free(dp);
        for (int s = 0; s < 2; ++s) { cudaFree(dp.sendbuf[s]); cudaFree(dp.recvbuf[s]); }
        if (dp.res) { cudaFree(dp.res->p); delete dp.res; }
        if (dp.cor) { cudaFree(dp.cor->p); delete dp.cor; }
    }
    for (auto* f : L->bwork) if (f) { cudaFree(f->p); delete f; }
    for (auto* f : L->swork) if (f) { cudaFree(f->p); delete f; }
    if (L->pwork) { cudaFree(L->pwork->p); delete L->pwork; }
    cudaFree(L->zeroPlane);
    for (auto* f : L->awork) if (f) { cudaFree(f->p); delete f; }
    halo_p2p_free(L);   // (collective: needs the communicator)
    if (L->comm && !L->comm_borrowed) ncclCommDestroy(L->comm);
    cudaFree(L->cx.scratch); cudaFree(L->cx.tickets); cudaFree(L->d_scal); cudaFreeHost(L->h_scal);
    if (L->d_bicg) cudaFree(L->d_bicg);
    if (L->h_bicg) cudaFreeHost(L->h_bicg);
    if (L->bicg_graph) cudaGraphExecDestroy(L->bicg_graph);
    gathered_bottom_free(L);
    cudaFree(L->stage);
    cudaEventDestroy(L->ev0); cudaEventDestroy(L->ev1);
    cudaStreamDestroy(L->cx.stream);
    delete L;
    return SOMAR_OK;
}

int somar_level_set_metric_fab(SomarLevel* L, const int lo[3], const int hi[3], int jgup_ncomp, const double* jgup0,
                               const double* jgup1, const double* jgup2, const double* jinv) {
    CHECK_LEVEL(L);
    cudaSetDevice(L->device);
    const int nd = L->desc.dim;
    Depth& dp = L->depths[0];
    const double* jg[3] = {jgup0, jgup1, jgup2};
    int rc;
    if (jgup_ncomp != 1 && jgup_ncomp != nd) { set_error("jgup_ncomp must be 1 or dim"); return SOMAR_ERR_ARG; }
    if (!L->desc.is_diagonal && jgup_ncomp != nd && jg[0]) { set_error("full metric needs all J g^{ij} components"); return SOMAR_ERR_ARG; }
    for (int d = 0; d < nd; ++d) {
        if (!jg[d]) continue;
        int fhi[3] = {hi[0], hi[1], hi[2]};
        fhi[d] += 1;
        const size_t comp = (size_t)(fhi[0] - lo[0] + 1) * (fhi[1] - lo[1] + 1) * (fhi[2] - lo[2] + 1);
        for (int c = 0; c < nd; ++c) {
            if (!dp.jg[d][c]) continue;          // diagonal metric stores J g^{dd} only
            const int srcComp = (jgup_ncomp == 1) ? 0 : c;
            if (jgup_ncomp == 1 && c != d) continue;
            if ((rc = copy3d(L, dp.g, dp.jg[d][c], const_cast<double*>(jg[d]) + comp * srcComp, lo, fhi, lo, fhi, true))) return rc;
        }
    }
    if (jinv && (rc = copy3d(L, dp.g, dp.jinv, const_cast<double*>(jinv), lo, hi, lo, hi, true))) return rc;
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    L->finalized = false;
    return SOMAR_OK;
}

int somar_level_set_metric_device(SomarLevel* L, int jgup_ncomp, const double* d_jgup0, const double* d_jgup1,
                                  const double* d_jgup2, const double* d_jinv) {
    CHECK_LEVEL(L);
    cudaSetDevice(L->device);
    const int nd = L->desc.dim;
    Depth& dp = L->depths[0];
    const Geom& g = dp.g;
    const double* jg[3] = {d_jgup0, d_jgup1, d_jgup2};
    if (jgup_ncomp != 1 && jgup_ncomp != nd) { set_error("jgup_ncomp must be 1 or dim"); return SOMAR_ERR_ARG; }
    for (int d = 0; d < nd; ++d) {
        if (!jg[d]) continue;
        int ext[3] = {d == 0, d == 1, d == 2};
        const size_t comp = (size_t)(g.n[0] + ext[0]) * (g.n[1] + ext[1]) * (g.n[2] + ext[2]);
        for (int c = 0; c < nd; ++c) {
            if (!dp.jg[d][c]) continue;
            if (jgup_ncomp == 1 && c != d) continue;
            const int srcComp = (jgup_ncomp == 1) ? 0 : c;
            launch_dense_copy(L->cx, g, dp.jg[d][c], const_cast<double*>(jg[d]) + comp * srcComp, ext, true);
        }
    }
    const int ext0[3] = {0, 0, 0};
    if (d_jinv) launch_dense_copy(L->cx, g, dp.jinv, const_cast<double*>(d_jinv), ext0, true);
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    L->finalized = false;
    return SOMAR_OK;
}

int somar_level_finalize(SomarLevel* L) {
    CHECK_LEVEL(L);
    cudaSetDevice(L->device);
    int rc = level_finalize(L);
    if (rc) return rc;
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    return SOMAR_OK;
}

int somar_level_set_alpha_beta(SomarLevel* L, double alpha, double beta) {
    CHECK_LEVEL(L);
    L->desc.alpha = alpha;
    L->desc.beta = beta;
    for (auto& dp : L->depths) { dp.g.alpha = alpha; dp.g.beta = beta; dp.line_valid = false; }
    return SOMAR_OK;   // lapDiag does not depend on alpha/beta (FILLMAPPEDLAPDIAG); the smoothers apply them
}

int somar_level_set_convergence_metric(SomarLevel* L, double metric) {
    CHECK_LEVEL(L);
    L->convergenceMetric = metric;
    return SOMAR_OK;
}

int somar_level_set_solver_params(SomarLevel* L, const SomarSolverParams* p) {
    CHECK_LEVEL(L);
    if (!p) { set_error("null params"); return SOMAR_ERR_ARG; }
    if (p->maxDepth != L->prm.maxDepth) {
        set_error("maxDepth fixes the multigrid schedule and cannot change after somar_level_create");
        return SOMAR_ERR_ARG;
    }
    L->prm = *p;
    for (auto& dp : L->depths) dp.line_valid = false;
    return SOMAR_OK;
}

int somar_level_num_depths(const SomarLevel* L, int* n) {
    CHECK_LEVEL(L);
    *n = L->ndepths();
    return SOMAR_OK;
}

int somar_level_depth_info(const SomarLevel* L, int depth, int n[3], int ratio[3], double dx[3], int* zero_avg,
                           int local_lo[3], int local_hi[3]) {
    CHECK_LEVEL(L);
    if (depth < 0 || depth >= L->ndepths()) { set_error("bad depth"); return SOMAR_ERR_ARG; }
    const Depth& dp = L->depths[depth];
    for (int d = 0; d < 3; ++d) {
        if (n) n[d] = dp.g.gn[d];
        if (ratio) ratio[d] = dp.ratio[d];
        if (dx) dx[d] = dp.g.dx[d];
        if (local_lo) local_lo[d] = dp.g.g0[d];
        if (local_hi) local_hi[d] = dp.g.g0[d] + dp.g.n[d] - 1;
    }
    if (zero_avg) *zero_avg = dp.zeroAvg;
    return SOMAR_OK;
}

int somar_level_sync(SomarLevel* L) {
    CHECK_LEVEL(L);
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    return SOMAR_OK;
}
int somar_level_stream(SomarLevel* L, void** s) {
    CHECK_LEVEL(L);
    *s = (void*)L->cx.stream;
    return SOMAR_OK;
}
int somar_level_launch_count(const SomarLevel* L, long long* n) {
    CHECK_LEVEL(L);
    *n = L->cx.launches;
    return SOMAR_OK;
}

int somar_level_relax_kernel(const SomarLevel* L, int depth, int mode, int* id) {
    CHECK_LEVEL(L);
    if (depth < 0 || depth >= L->ndepths() || !id) { set_error("bad depth"); return SOMAR_ERR_ARG; }
    const Depth& dp = L->depths[depth];
    if (mode == SOMAR_RELAX_LEVEL_GSRB)
        *id = (L->use_fused == 2 && gsrb_async_usable(dp.g, dp.m)) ? (gsrb_tma_usable(dp.g, dp.m) ? 7 : 5)
              : (L->use_fused && gsrb_fused_usable(dp.g, dp.m))     ? 1
                                                                    : (dp.m.diagonal ? 0 : 2);
    else if (mode == SOMAR_RELAX_LINE_GSRB) *id = line_compact_usable(L, dp) ? 8 : 3;
    else if (mode == SOMAR_RELAX_LOOSE_GSRB) *id = dp.m.diagonal ? 6 : -1;
    else if (mode == SOMAR_RELAX_JACOBI) *id = 4;
    else *id = -1;
    return SOMAR_OK;
}

int somar_level_halo_path(const SomarLevel* L, int* path) {
    CHECK_LEVEL(L);
    if (!path) { set_error("null out pointer"); return SOMAR_ERR_ARG; }
    *path = (L->desc.nranks <= 1) ? 0 : (L->halo.state == 1 ? 2 : 1);
    return SOMAR_OK;
}

int somar_level_get_metric(SomarLevel* L, int depth, double* jg0, double* jg1, double* jg2, double* jinv,
                           double* lapdiag) {
    CHECK_LEVEL(L);
    cudaSetDevice(L->device);
    if (depth < 0 || depth >= L->ndepths()) { set_error("bad depth"); return SOMAR_ERR_ARG; }
    Depth& dp = L->depths[depth];
    const Geom& g = dp.g;
    const int nd = L->desc.dim;
    double* jg[3] = {jg0, jg1, jg2};
    int lo[3], hi[3], rc;
    for (int d = 0; d < 3; ++d) { lo[d] = g.g0[d]; hi[d] = g.g0[d] + g.n[d] - 1; }
    for (int d = 0; d < nd; ++d) {
        if (!jg[d]) continue;
        int fhi[3] = {hi[0], hi[1], hi[2]};
        fhi[d] += 1;
        const size_t comp = (size_t)(fhi[0] - lo[0] + 1) * (fhi[1] - lo[1] + 1) * (fhi[2] - lo[2] + 1);
        for (int c = 0; c < nd; ++c) {
            if (!dp.jg[d][c]) { std::memset(jg[d] + comp * c, 0, comp * sizeof(double)); continue; }
            if ((rc = copy3d(L, g, dp.jg[d][c], jg[d] + comp * c, lo, fhi, lo, fhi, false))) return rc;
        }
    }
    if (jinv && (rc = copy3d(L, g, dp.jinv, jinv, lo, hi, lo, hi, false))) return rc;
    if (lapdiag && (rc = copy3d(L, g, dp.lapdiag, lapdiag, lo, hi, lo, hi, false))) return rc;
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    return SOMAR_OK;
}

// ---- fields ------------------------------------------------------------------------------------------------
int somar_field_create(SomarLevel* L, int depth, SomarField** out) {
    CHECK_LEVEL(L);
    cudaSetDevice(L->device);
    return field_alloc(L, depth, out);
}
int somar_field_destroy(SomarField* f) {
    if (!f) return SOMAR_OK;
    cudaSetDevice(f->lev->device);
    cudaStreamSynchronize(f->lev->cx.stream);
    cudaFree(f->p);
    delete f;
    return SOMAR_OK;
}
int somar_field_set_fab(SomarField* f, const double* fab, const int lo[3], const int hi[3], const int vlo[3],
                        const int vhi[3]) {
    if (!f || !fab) { set_error("null argument"); return SOMAR_ERR_ARG; }
    cudaSetDevice(f->lev->device);
    return copy3d(f->lev, f->lev->depths[f->depth].g, f->p, const_cast<double*>(fab), lo, hi, vlo, vhi, true);
}
int somar_field_get_fab(const SomarField* f, double* fab, const int lo[3], const int hi[3], const int vlo[3],
                        const int vhi[3]) {
    if (!f || !fab) { set_error("null argument"); return SOMAR_ERR_ARG; }
    cudaSetDevice(f->lev->device);
    int rc = copy3d(f->lev, f->lev->depths[f->depth].g, f->p, fab, lo, hi, vlo, vhi, false);
    if (rc) return rc;
    SOMAR_CUDA(cudaStreamSynchronize(f->lev->cx.stream));
    return SOMAR_OK;
}
static int field_dense(SomarField* f, double* data, int on_device, bool toDevice) {
    if (!f || !data) { set_error("null argument"); return SOMAR_ERR_ARG; }
    SomarLevel* L = f->lev;
    cudaSetDevice(L->device);
    const Geom& g = L->depths[f->depth].g;
    if (on_device) {
        const int ext[3] = {0, 0, 0};
        launch_dense_copy(L->cx, g, f->p, data, ext, toDevice);
        return SOMAR_OK;
    }
    int lo[3], hi[3];
    for (int d = 0; d < 3; ++d) { lo[d] = g.g0[d]; hi[d] = g.g0[d] + g.n[d] - 1; }
    int rc = copy3d(L, g, f->p, data, lo, hi, lo, hi, toDevice);
    if (rc) return rc;
    if (!toDevice) SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    return SOMAR_OK;
}
// Dense rank-local FACE array of direction dir (n[dir] + 1 points in that direction, x fastest).
static int field_dense_face(SomarField* f, double* data, int dir, int on_device, bool toDevice) {
    if (!f || !data || dir < 0 || dir > 2) { set_error("bad argument"); return SOMAR_ERR_ARG; }
    SomarLevel* L = f->lev;
    cudaSetDevice(L->device);
    const Geom& g = L->depths[f->depth].g;
    if (dir >= g.nd) { set_error("face direction is not active"); return SOMAR_ERR_ARG; }
    if (on_device) {
        int ext[3] = {0, 0, 0};
        ext[dir] = 1;
        launch_dense_copy(L->cx, g, f->p, data, ext, toDevice);
        return SOMAR_OK;
    }
    int lo[3], hi[3];
    for (int d = 0; d < 3; ++d) { lo[d] = g.g0[d]; hi[d] = g.g0[d] + g.n[d] - 1 + (d == dir); }
    int rc = copy3d(L, g, f->p, data, lo, hi, lo, hi, toDevice);
    if (rc) return rc;
    if (!toDevice) SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    return SOMAR_OK;
}
int somar_field_set_face(SomarField* f, const double* data, int dir, int on_device) {
    return field_dense_face(f, const_cast<double*>(data), dir, on_device, true);
}
int somar_field_get_face(const SomarField* f, double* data, int dir, int on_device) {
    return field_dense_face(const_cast<SomarField*>(f), data, dir, on_device, false);
}
int somar_field_set(SomarField* f, const double* data, int on_device) {
    return field_dense(f, const_cast<double*>(data), on_device, true);
}
int somar_field_get(const SomarField* f, double* data, int on_device) {
    return field_dense(const_cast<SomarField*>(f), data, on_device, false);
}

}  // extern "C"
