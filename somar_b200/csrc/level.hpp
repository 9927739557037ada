// level.hpp -- host-side objects behind the C ABI: one AMR level with all its multigrid depths resident in HBM.
// Mirrors, in role only, MappedAMRPoissonOp (one per depth) + MappedMultiGrid (depth stack, scratch) +
// MappedAMRMultiGrid (outer iteration) of the reference; see each member function for file:line.
#pragma once
#include "../../include/somar_b200.h"
#include "kernels.hpp"
#include <nccl.h>
#include <string>
#include <vector>

struct SomarField {
    SomarLevel* lev;
    int depth;
    double* p;       // brick of lev->depths[depth].g.total doubles
};

namespace somar {

struct Depth {
    Geom g;
    Metric m;
    int ratio[3] = {1, 1, 1};          // mgCrseRefRatio to the next coarser depth
    int coarsening[3] = {1, 1, 1};     // total coarsening from depth 0
    double* jg[3][3] = {{nullptr}};
    double* jinv = nullptr;
    double* lapdiag = nullptr;
    double* line_fact = nullptr;       // lazily built by line relaxation
    double* line_dmod = nullptr;
    bool line_valid = false;
    void* line_compact = nullptr;      // colour-split arrays of the line smoother (line_compact.cu), built on first use
    double* tmp = nullptr;             // ping-pong partner for out-of-place passes / Jacobi residual
    bool zeroAvg = false;
    double dxProduct = 1.0;
    SomarField* res = nullptr;         // MappedMultiGrid::m_residual[depth], m_correction[depth]
    SomarField* cor = nullptr;
    double* sendbuf[2] = {nullptr, nullptr};
    double* recvbuf[2] = {nullptr, nullptr};
};

// Peer-memory halo exchange (halo_p2p.cu): my arena, my y-neighbours' arenas mapped through CUDA IPC.
struct HaloArena {
    int state = 0;                     // 0 not set up yet, 1 in use, -1 unavailable (NCCL send/recv is used)
    char* base = nullptr;              // my arena: per depth 2 sides x 2 receive slots, then counters, then tickets
    char* peer[2] = {nullptr, nullptr};// the low / high neighbour's arena
    size_t bytes = 0, flagOff = 0, ticketOff = 0;
    std::vector<size_t> slotOff, slotBytes;          // per depth
    std::vector<unsigned long long> seq[2];          // per side, per depth: messages exchanged so far
};

// Full-domain copy of the bottom depth kept by every rank of a multi-rank level (bottom_gather.cu).
struct GatheredBottom {
    Geom g;                            // the bottom depth with all rows, no neighbours
    Metric m;
    double* jg[3] = {nullptr, nullptr, nullptr};
    double* jinv = nullptr;
    double* lapdiag = nullptr;
    double* phi = nullptr;             // correction / residual of the bottom visit on the full domain
    double* rhs = nullptr;
    double* work[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    double* sendbuf = nullptr;         // my slabs, dense
    double* allbuf = nullptr;          // every rank's slabs
    std::vector<int> row0, rows;       // first row / number of rows of each rank at the bottom depth
    int maxRows = 0;
};

}  // namespace somar

struct SomarLevel {
    SomarLevelDesc desc;
    std::vector<int> box_lo, box_hi, box_rank;
    SomarSolverParams prm;
    int device = 0;
    somar::Ctx cx;
    std::vector<somar::Depth> depths;
    bool finalized = false;
    bool metric_diag_only = true;
    int local_lo[3], local_hi[3];      // this rank's region at depth 0 (global indices)
    int box_size[3];                   // uniform tiling size at depth 0 (0 = not uniform)
    // communication
    ncclComm_t comm = nullptr;
    bool comm_borrowed = false;        // the communicator belongs to another level (leptic solver's derived levels)
    // scalars
    double* d_scal = nullptr;          // 8 doubles on device
    double* h_scal = nullptr;          // pinned mirror
    void* d_bicg = nullptr;            // BicgState of the bottom solver (bicgstab.cu) and its pinned mirror
    void* h_bicg = nullptr;
    int bottom_mode = 3;               // SOMAR_B200_BOTTOM: 0 host-sequenced BiCGStab, 1 scalars on the device,
                                       // 2 one persistent kernel per bottom visit where usable (else 1),
                                       // 3 (default) the on-chip-resident kernel where usable (else as 2)
    // launch-per-step bottom solver, one rank, full metric: a chunk of BiCGStab iterations as one CUDA graph (bicgstab.cu)
    cudaGraphExec_t bicg_graph = nullptr;
    unsigned char bicg_graph_key[160] = {0};
    int bicg_graph_launches = 0;       // kernel launches one replay stands for
    int bottom_resident_blocks = -1;   // co-resident blocks of k_bicg_resident (-1: not yet queried)
    int bottom_coop_blocks = -1;       // co-resident blocks of the persistent kernel (-1: not yet queried)
    bool bottom_iters_pending = false; // the device-side iteration counter holds uncollected counts
    somar::GatheredBottom* gb = nullptr; // multi-rank: replicated bottom depth (built at finalize when usable)
    somar::HaloArena halo;             // multi-rank: peer-memory halo exchange
    // staging for host FAB transfers
    double* stage = nullptr;
    size_t stage_bytes = 0;
    // bottom solver work fields (bottom depth)
    std::vector<SomarField*> bwork;
    // outer-solve work fields (depth 0)
    std::vector<SomarField*> swork;
    std::vector<SomarField*> awork;    // advection: WMinus/WPlus/WHalf scratch (depth 0, allocated on first use)
    double* zeroPlane = nullptr;       // off0 + sz zeros of depth 0: the input of a relaxation that starts from phi = 0
    SomarField* pwork = nullptr;       // projector: the divergence / right-hand side (depth 0)
    int projector_bc = 0;              // somar_level_set_projector_bc: 0 operator BCs (round 1), 1 the reference projectors' m_divBC / m_gradBC
    double bottomMetric = 0.0, bottomEps = 1e-6;
    double convergenceMetric = 0.0;    // MappedAMRMultiGrid::m_convergenceMetric (0: the initial residual norm)
    int bottom_iters_total = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // phase profiler (CUDA events on the level stream; off by default)
    int use_fused = 2;                 // SOMAR_B200_FUSED: 0 = two colour passes, 1 = k_gsrb_fused, 2 = k_gsrb_async
    bool prof_on = false;
    struct ProfRec { int phase; cudaEvent_t a, b; };
    std::vector<ProfRec> prof;          // recorded intervals
    std::vector<cudaEvent_t> prof_pool; // free events

    int ndepths() const { return (int)depths.size(); }
};

namespace somar {
enum { PH_SMOOTH0_KERNEL = 0, PH_SMOOTH_COARSE = 1, PH_RESTRICT = 2, PH_PROLONG = 3, PH_BOTTOM = 4, PH_OUTER = 5,
       PH_COUNT = 6 };
// RAII interval: records two events around a scope when profiling is on.
struct ProfScope {
    SomarLevel* L;
    int idx = -1;
    ProfScope(SomarLevel* L_, int phase) : L(L_) {
        if (!L->prof_on || phase < 0) return;
        cudaEvent_t e[2];
        for (int i = 0; i < 2; ++i) {
            if (!L->prof_pool.empty()) { e[i] = L->prof_pool.back(); L->prof_pool.pop_back(); }
            else cudaEventCreate(&e[i]);
        }
        cudaEventRecord(e[0], L->cx.stream);
        idx = (int)L->prof.size();
        L->prof.push_back({phase, e[0], e[1]});
    }
    ~ProfScope() {
        if (idx >= 0) cudaEventRecord(L->prof[idx].b, L->cx.stream);
    }
};
}  // namespace somar

namespace somar {

void set_error(const std::string& s);
#define SOMAR_CUDA(call)                                                                                   \
    do {                                                                                                    \
        cudaError_t e_ = (call);                                                                            \
        if (e_ != cudaSuccess) {                                                                            \
            somar::set_error(std::string(#call) + ": " + cudaGetErrorString(e_));                           \
            return SOMAR_ERR_CUDA;                                                                          \
        }                                                                                                   \
    } while (0)
#define SOMAR_NCCL(call)                                                                                   \
    do {                                                                                                    \
        ncclResult_t e_ = (call);                                                                           \
        if (e_ != ncclSuccess) {                                                                            \
            somar::set_error(std::string(#call) + ": " + ncclGetErrorString(e_));                           \
            return SOMAR_ERR_COMM;                                                                          \
        }                                                                                                   \
    } while (0)

// capi.cu
int level_create_impl(const SomarLevelDesc* desc, const SomarSolverParams* params, int device,
                      const unsigned char* comm_id, ncclComm_t borrowed, SomarLevel** out);
// level.cu
int level_build(SomarLevel* L);                      // schedule + allocation
int level_finalize(SomarLevel* L);                   // coarse metrics, lapDiag, null-space probe
int field_alloc(SomarLevel* L, int depth, SomarField** out);
int exchange(SomarLevel* L, int depth, double* x, int width);
int op_apply(SomarLevel* L, int depth, double* out, SomarField* phi, const double* rhs, int mode, bool homog);
int op_relax(SomarLevel* L, int depth, SomarField* phi, const SomarField* rhs, int iters, int modeOverride = -100,
             bool phiIsZero = false);   // phiIsZero: phi is logically zero and its memory has NOT been cleared
int op_precond(SomarLevel* L, int depth, SomarField* cor, const SomarField* res);
int op_restrict(SomarLevel* L, int depth, SomarField* crse, SomarField* phi, const SomarField* rhs);
int op_prolong(SomarLevel* L, int depth, SomarField* phi, const SomarField* crse);
int reduce_global(SomarLevel* L, int depth, const double* x, const double* y, int kind, double* out);
int residual_maxnorm(SomarLevel* L, SomarField* res, SomarField* phi, const SomarField* rhs, bool homog, double* out);
// solver.cu
int mg_cycle(SomarLevel* L, int depth, SomarField* e, const SomarField* r, bool eIsZero = false);
int bottom_solve(SomarLevel* L, SomarField* e, const SomarField* r, int* iters);
bool bottom_persistent_usable(SomarLevel* L);
int bottom_persistent(SomarLevel* L, SomarField* e, const SomarField* r);
int bottom_collect_iters(SomarLevel* L);
bool bottom_persistent_params_ok(SomarLevel* L, const Geom& g, bool full = false);
int launch_bottom_persistent(SomarLevel* L, const Geom& g, const Metric& m, double* phi, double* rhs,
                             double* const work[8]);
// bicg_resident.cu
bool bottom_resident_usable(SomarLevel* L, const Geom& g, const Metric& m);
int launch_bottom_resident(SomarLevel* L, const Geom& g, const Metric& m, double* phi, double* rhs, double* const work[8],
                           void* bicgBar, double* part, int* itersAccum);
// line_compact.cu
bool line_compact_usable(const SomarLevel* L, const Depth& dp);
int line_compact_relax(SomarLevel* L, Depth& dp, double* phi, const double* rhs, int iters, bool coefStale, int profPhase);
void line_compact_free(Depth& dp);
// projector.cu
int op_mac_gradient(SomarLevel* L, SomarField* const grad[3], SomarField* phi, bool homog, double scale, bool accumulate,
                    int gradBC = 0);
int op_wall_flux_bc(SomarLevel* L, SomarField* const flux[3]);
int op_cc_divergence(SomarLevel* L, SomarField* div, SomarField* const vel[3], SomarField* J, double divideBy);
int op_cc_gradient(SomarLevel* L, SomarField* const grad[3], SomarField* phi);
int project_cc(SomarLevel* L, SomarField* const vel[3], SomarField* J, SomarField* pressure, double dt, bool zeroPressure,
               bool forceHomog, SomarSolveStats* st);
int op_mac_divergence(SomarLevel* L, SomarField* div, const SomarField* const flux[3], double divideBy);
int project(SomarLevel* L, SomarField* const vel[3], SomarField* pressure, double dt, bool zeroPressure, bool forceHomog,
            SomarSolveStats* st);
// advect.cu
int advect_predict_scalar(SomarLevel* L, SomarField* const whalf[3], const SomarField* wold, const SomarField* source,
                          const SomarField* const oldVel[3], const SomarField* const advVel[3], const SomarField* ccJ,
                          double dt, const SomarAdvectParams* prm);
int advect_update(SomarLevel* L, SomarField* wnew, const SomarField* wold, const SomarField* const flux[3], double dt);
// leptic.cu
int leptic_vertical_solve(SomarLevel* L, SomarField* phi, SomarField* rhs, int loBC, int hiBC, double dzCrse,
                          const SomarField* bdryLo, const SomarField* bdryHi);
int leptic_vert_integral(SomarLevel* L, SomarField* flat, const SomarField* func, double dzScale);
int leptic_vertical_excess(SomarLevel* L, SomarField* excess, const SomarField* rhs, const SomarField* bdryLo,
                           const SomarField* bdryHi);
int leptic_add_extrusion(SomarLevel* L, SomarField* dest, const SomarField* flat);
int leptic_accum_div(SomarLevel* L, SomarField* flatRhs, const SomarField* flatFlux, int dir, double dxScale);
// halo_p2p.cu
int halo_p2p_exchange(SomarLevel* L, int depth, double* x, int width);   // 1 done, 0 use NCCL, < 0 error
void halo_p2p_free(SomarLevel* L);
// bottom_gather.cu
int gathered_bottom_build(SomarLevel* L);
void gathered_bottom_free(SomarLevel* L);
bool bottom_gathered_usable(SomarLevel* L);
int bottom_gathered(SomarLevel* L, SomarField* e, const SomarField* r);
int mg_solve(SomarLevel* L, SomarField* phi, const SomarField* rhs, bool zeroPhi, bool forceHomog,
             SomarSolveStats* st);

}  // namespace somar
