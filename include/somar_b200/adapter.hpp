// adapter.hpp -- C++ host side above the C ABI (include/somar_b200.h), shaped like the reference's classes so that
// a SOMAR maintainer can forward to it from thin subclasses (INTEGRATION.md shows the binding):
//
//   somar_b200::HostFab / LevelFabs   what the library needs of Chombo's FArrayBox / LevelData<FArrayBox>:
//                                     pointer + box incl. ghosts + valid box (the CHF_FRA convention)
//   somar_b200::PoissonOp             MappedAMRPoissonOp as seen by the MG drivers -- LinearOp<T> + MappedMGLevelOp<T>
//                                     (src/calculus/AMRElliptic/MappedAMRPoissonOp.H:187-292), one object per depth
//   somar_b200::PressureSolver        AMRPressureSolver (src/projection/AMRPressureSolver.H:43-104):
//                                     setAMRMGParameters / setBottomParameters / levelDefine / levelSolve / undefine
//
// Header-only, no CUDA headers: it only calls the extern "C" entry points.  Errors become somar_b200::Error
// (the reference calls MayDay::Error, which aborts; a subclass in SOMAR catches and forwards to MayDay).
// There is no CPU path behind any of these calls.
#pragma once
#include <array>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../somar_b200.h"

namespace somar_b200 {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};

inline void check(int rc, const char* where) {
    if (rc != SOMAR_OK) throw Error(rc, std::string(where) + ": " + somar_last_error());
}

using Int3 = std::array<int, 3>;

// A host FAB in the Fortran calling convention: data over [lo,hi] (ghosts included), x fastest, `ncomp` components
// outermost; `vlo,vhi` = the DisjointBoxLayout box this FAB belongs to (cells only).
struct HostFab {
    double* data = nullptr;
    Int3 lo{{0, 0, 0}}, hi{{0, 0, 0}};
    Int3 vlo{{0, 0, 0}}, vhi{{0, 0, 0}};
    int ncomp = 1;
};
using LevelFabs = std::vector<HostFab>;   // the FABs of the boxes THIS rank owns (a LevelData's DataIterator range)

// What levelDefine reads from LevelGeometry + BCMethodHolder + ProblemDomain (AMRPressureSolver.cpp:321-480).
struct LevelGeometryDesc {
    int dim = 3;
    Int3 n{{1, 1, 1}};
    Int3 periodic{{0, 0, 0}};
    std::vector<Int3> boxLo, boxHi;   // ALL boxes of the level
    std::vector<int> boxRank;         // owner ranks (empty = rank 0)
    std::array<double, 3> dx{{1, 1, 1}};
    bool isDiagonal = true;           // LevelGeometry::isDiagonal()
    std::array<int, 6> bcType{{SOMAR_BC_NEUM, SOMAR_BC_NEUM, SOMAR_BC_NEUM, SOMAR_BC_NEUM, SOMAR_BC_NEUM, SOMAR_BC_NEUM}};
    std::array<double, 6> bcValue{{0, 0, 0, 0, 0, 0}};
    int rank = 0, nranks = 1, device = 0;
    const unsigned char* commId = nullptr;   // 128-byte NCCL id from somar_comm_unique_id (nranks > 1)
};

class Level;

// Device field of one depth.  create()/clear() of LinearOp<T> (MappedAMRPoissonOp.H:214-216, 566-575).
class Field {
public:
    Field() = default;
    Field(Level& lev, int depth);
    ~Field() { clear(); }
    Field(Field&& o) noexcept : h_(o.h_), depth_(o.depth_) { o.h_ = nullptr; }
    Field& operator=(Field&& o) noexcept {
        if (this != &o) { clear(); h_ = o.h_; depth_ = o.depth_; o.h_ = nullptr; }
        return *this;
    }
    Field(const Field&) = delete;
    Field& operator=(const Field&) = delete;
    void clear() {
        if (h_) somar_field_destroy(h_);
        h_ = nullptr;
    }
    bool defined() const { return h_ != nullptr; }
    int depth() const { return depth_; }
    SomarField* handle() const { return h_; }
    // LevelData<FArrayBox> -> device and back, one FAB per owned box (ghost cells are never taken from the host:
    // the library fills them, as the reference's exchange/BC code overwrites them)
    void copyFrom(const LevelFabs& fabs) {
        for (const HostFab& f : fabs)
            check(somar_field_set_fab(h_, f.data, f.lo.data(), f.hi.data(), f.vlo.data(), f.vhi.data()), "Field::copyFrom");
    }
    void copyTo(LevelFabs& fabs) const {
        for (HostFab& f : fabs)
            check(somar_field_get_fab(h_, f.data, f.lo.data(), f.hi.data(), f.vlo.data(), f.vhi.data()), "Field::copyTo");
    }

private:
    SomarField* h_ = nullptr;
    int depth_ = 0;
};

// One AMR level with all its multigrid depths: what MappedAMRPoissonOpFactory::define + MappedMultiGrid::define build.
class Level {
public:
    Level(const LevelGeometryDesc& g, const SomarSolverParams& prm, double alpha = 0.0, double beta = 1.0) {
        SomarLevelDesc d{};
        d.dim = g.dim;
        std::vector<int> lo, hi;
        for (size_t b = 0; b < g.boxLo.size(); ++b)
            for (int k = 0; k < 3; ++k) { lo.push_back(g.boxLo[b][k]); hi.push_back(g.boxHi[b][k]); }
        for (int k = 0; k < 3; ++k) { d.n[k] = g.n[k]; d.periodic[k] = g.periodic[k]; d.dx[k] = g.dx[k]; }
        d.nboxes = (int)g.boxLo.size();
        d.box_lo = lo.data();
        d.box_hi = hi.data();
        d.box_rank = g.boxRank.empty() ? nullptr : g.boxRank.data();
        d.alpha = alpha;
        d.beta = beta;
        d.is_diagonal = g.isDiagonal ? 1 : 0;
        for (int k = 0; k < 6; ++k) { d.bc_type[k] = g.bcType[k]; d.bc_value[k] = g.bcValue[k]; }
        d.rank = g.rank;
        d.nranks = g.nranks;
        check(somar_level_create(&d, &prm, g.device, g.commId, &h_), "Level");
    }
    ~Level() {
        if (h_) somar_level_destroy(h_);
    }
    Level(const Level&) = delete;
    Level& operator=(const Level&) = delete;

    // LevelGeometry::getFCJgup() / getCCJinv() of one owned box (FluxBox: jgup[d] face-centred in d)
    void setMetric(const Int3& lo, const Int3& hi, int ncomp, const double* jgup0, const double* jgup1,
                   const double* jgup2, const double* jinv) {
        check(somar_level_set_metric_fab(h_, lo.data(), hi.data(), ncomp, jgup0, jgup1, jgup2, jinv), "Level::setMetric");
    }
    void finalize() { check(somar_level_finalize(h_), "Level::finalize"); }
    int numDepths() const {
        int n = 0;
        check(somar_level_num_depths(h_, &n), "Level::numDepths");
        return n;
    }
    SomarLevel* handle() const { return h_; }

private:
    SomarLevel* h_ = nullptr;
};

inline Field::Field(Level& lev, int depth) : depth_(depth) {
    check(somar_field_create(lev.handle(), depth, &h_), "Field");
}

// MappedAMRPoissonOp at one multigrid depth (MGnewOp(depth), MappedAMRPoissonOpFactory.cpp:362).
// Method names, argument order and meaning follow MappedAMRPoissonOp.H:187-292; T is a device Field.
class PoissonOp {
public:
    PoissonOp(Level& lev, int depth) : lev_(lev), depth_(depth) {}

    // --- LinearOp<T> ---
    void residual(Field& lhs, Field& phi, const Field& rhs, bool homogeneous = false) {
        check(somar_residual(L(), depth_, lhs.handle(), phi.handle(), rhs.handle(), homogeneous), "residual");
    }
    void preCond(Field& correction, const Field& residual) {
        check(somar_precond(L(), depth_, correction.handle(), residual.handle()), "preCond");
    }
    void applyOp(Field& lhs, Field& phi, bool homogeneous = false) {
        check(somar_apply_op(L(), depth_, lhs.handle(), phi.handle(), homogeneous), "applyOp");
    }
    void create(Field& lhs, const Field& /*rhs*/) { lhs = Field(lev_, depth_); }
    void clear(Field& lhs) { lhs.clear(); }
    void assign(Field& lhs, const Field& rhs) { check(somar_assign(L(), lhs.handle(), rhs.handle()), "assign"); }
    double dotProduct(const Field& a, const Field& b) {
        double v = 0;
        check(somar_dot(L(), a.handle(), b.handle(), &v), "dotProduct");
        return v;
    }
    void incr(Field& lhs, const Field& x, double scale) { check(somar_incr(L(), lhs.handle(), x.handle(), scale), "incr"); }
    void axby(Field& lhs, const Field& x, const Field& y, double a, double b) {
        check(somar_axby(L(), lhs.handle(), x.handle(), y.handle(), a, b), "axby");
    }
    void scale(Field& lhs, double s) { check(somar_scale(L(), lhs.handle(), s), "scale"); }
    double norm(const Field& x, int ord) {
        double v = 0;
        check(somar_norm(L(), x.handle(), ord, &v), "norm");
        return v;
    }
    double localMaxNorm(const Field& x) {
        double v = 0;
        check(somar_local_max_norm(L(), x.handle(), &v), "localMaxNorm");
        return v;
    }
    void setToZero(Field& x) { check(somar_set_to_zero(L(), x.handle()), "setToZero"); }

    // --- MappedMGLevelOp<T> ---
    void relax(Field& e, const Field& residual, int iterations) {
        check(somar_relax(L(), depth_, e.handle(), residual.handle(), iterations), "relax");
    }
    void createCoarser(Field& coarse, const Field& /*fine*/, bool /*ghosted*/) { coarse = Field(lev_, depth_ + 1); }
    void restrictResidual(Field& resCoarse, Field& phiFine, const Field& rhsFine) {
        check(somar_restrict_residual(L(), depth_, resCoarse.handle(), phiFine.handle(), rhsFine.handle()), "restrictResidual");
    }
    void prolongIncrement(Field& phiThisLevel, const Field& correctCoarse) {
        check(somar_prolong_increment(L(), depth_, phiThisLevel.handle(), correctCoarse.handle()), "prolongIncrement");
    }
    void setAlphaAndBeta(double alpha, double beta) {
        check(somar_level_set_alpha_beta(L(), alpha, beta), "setAlphaAndBeta");
    }
    int depth() const { return depth_; }

private:
    SomarLevel* L() const { return lev_.handle(); }
    Level& lev_;
    int depth_;
};

// AMRPressureSolver, level-solver flavour (levelDefine + levelSolve; AMRPressureSolver.H:43-112).
class PressureSolver {
public:
    PressureSolver() { somar_default_params(&prm_); }

    // AMRPressureSolver::setAMRMGParameters (same argument order, AMRPressureSolver.H:53-66)
    void setAMRMGParameters(int imin, int imax, double eps, int maxDepth, int num_precond_iters, int num_smooth_down,
                            int num_smooth_up, int num_smooth_bottom, int precondMode, int relaxMode, int numMG,
                            double hang, double norm_thresh, int verbosity) {
        requireUndefined("setAMRMGParameters");
        prm_.imin = imin; prm_.imax = imax; prm_.eps = eps; prm_.maxDepth = maxDepth;
        prm_.num_smooth_precond = num_precond_iters; prm_.num_smooth_down = num_smooth_down;
        prm_.num_smooth_up = num_smooth_up; prm_.num_smooth_bottom = num_smooth_bottom;
        prm_.precondMode = precondMode; prm_.relaxMode = relaxMode; prm_.numMG = numMG;
        prm_.hang = hang; prm_.normThresh = norm_thresh; prm_.verbosity = verbosity;
    }
    // AMRPressureSolver::setBottomParameters (AMRPressureSolver.H:70-77)
    void setBottomParameters(int imax, int numRestarts, double eps, double reps, double hang, double small_,
                             int normType, int /*verbosity*/) {
        requireUndefined("setBottomParameters");
        prm_.bottom_imax = imax; prm_.bottom_numRestarts = numRestarts; prm_.bottom_eps = eps;
        prm_.bottom_reps = reps; prm_.bottom_hang = hang; prm_.bottom_small = small_; prm_.bottom_normType = normType;
    }
    // AMRPressureSolver::levelDefine: the caller then hands over the metric of each owned box and calls ready().
    Level& levelDefine(const LevelGeometryDesc& geo) {
        lev_.reset(new Level(geo, prm_, 0.0, 1.0));   // pressure Poisson: alpha = 0, beta = 1 (AMRPressureSolver.cpp:404)
        return *lev_;
    }
    void ready() {
        lev_->finalize();
        phi_ = Field(*lev_, 0);
        rhs_ = Field(*lev_, 0);
    }
    void undefine() {
        phi_.clear();
        rhs_.clear();
        lev_.reset();
    }
    bool isDefined() const { return lev_ && phi_.defined(); }
    int getNumLevels() const { return 1; }

    // AMRPressureSolver::levelSolve (no coarser level): host LevelData in, host LevelData out.
    SomarSolveStats levelSolve(LevelFabs& phi, const LevelFabs& rhs, bool zeroPhi = true, bool forceHomogeneous = false) {
        if (!isDefined()) throw Error(SOMAR_ERR_ARG, "PressureSolver::levelSolve: not defined");
        rhs_.copyFrom(rhs);
        if (!zeroPhi) phi_.copyFrom(phi);
        SomarSolveStats st = solveDevice(phi_, rhs_, zeroPhi, forceHomogeneous);
        phi_.copyTo(phi);
        return st;
    }
    // Device-resident variant for callers that keep phi/rhs on the GPU between projections.
    SomarSolveStats solveDevice(Field& phi, const Field& rhs, bool zeroPhi = true, bool forceHomogeneous = false) {
        SomarSolveStats st{};
        const int rc = somar_solve(lev_->handle(), phi.handle(), rhs.handle(), zeroPhi, forceHomogeneous, &st);
        if (rc != SOMAR_OK && rc != SOMAR_ERR_DIVERGED) check(rc, "PressureSolver::solve");
        return st;   // non-convergence is not an error: st.exitStatus as MappedAMRMultiGrid::m_exitStatus
    }
    Level& level() { return *lev_; }
    const SomarSolverParams& params() const { return prm_; }

private:
    void requireUndefined(const char* who) const {
        if (lev_) throw Error(SOMAR_ERR_ARG, std::string(who) + ": can only be called before define");
    }
    SomarSolverParams prm_{};
    std::unique_ptr<Level> lev_;
    Field phi_, rhs_;
};

// LevelMACProjector / BaseProjector<FluxBox> on one level (src/projection/LevelMACProjector.cpp:156-240,
// BaseProjectorI.H:176-300) over device-resident fields.  The velocity is the flux J u^i on faces: one Field per
// direction (face (i,j,k) = low face of cell (i,j,k); upload a FluxBox FAB with its own surroundingNodes box
// through somar_field_set_fab).  multByJ / divByJ stay with the caller's LevelGeometry (a_velIsFlux = true).
class MACProjector {
public:
    explicit MACProjector(PressureSolver& solver) : solver_(solver) {}
    // computeDiv: div = Jinv d_i (J u^i)
    void computeDiv(Field& div, const Field& ux, const Field& uy, const Field* uz = nullptr) {
        check(somar_mac_divergence(L(), div.handle(), ux.handle(), uy.handle(), uz ? uz->handle() : nullptr, 0.0), "computeDiv");
    }
    // computeGrad: flux_d = J g^{dj} d_j phi with the pressure BCs
    void computeGrad(Field& gx, Field& gy, Field* gz, Field& phi) {
        check(somar_mac_gradient(L(), gx.handle(), gy.handle(), gz ? gz->handle() : nullptr, phi.handle(), 0), "computeGrad");
    }
    // applyCorrection: vel -= dt * Grad[phi] (dt == 0: vel -= Grad[phi]); the gradient is formed on the fly
    void applyCorrection(Field& ux, Field& uy, Field* uz, Field& phi, double dt) {
        check(somar_mac_apply_correction(L(), ux.handle(), uy.handle(), uz ? uz->handle() : nullptr, phi.handle(), dt), "applyCorrection");
    }
    // project: rhs = Div[vel]/dt; solve; vel -= dt Grad[pressure].  Non-convergence is reported in the stats.
    SomarSolveStats project(Field& ux, Field& uy, Field* uz, Field& pressure, double dt, bool zeroPressure = true,
                            bool forceHomogeneous = false) {
        SomarSolveStats st{};
        const int rc = somar_project(L(), ux.handle(), uy.handle(), uz ? uz->handle() : nullptr, pressure.handle(), dt,
                                     zeroPressure, forceHomogeneous, &st);
        if (rc != SOMAR_OK && rc != SOMAR_ERR_DIVERGED) check(rc, "MACProjector::project");
        return st;
    }

private:
    SomarLevel* L() const { return solver_.level().handle(); }
    PressureSolver& solver_;
};

// LevelCCProjector (src/projection/LevelCCProjector.H:38-100): the cell-centred ("approximate") projection of the time
// step (AMRNavierStokes::doCCProjection, AMRNavierStokesAdvancePPM.cpp:1513-1573).  The velocity components are CELL
// fields; J is the cell-centred Jacobian field the reference caches (LevelGeometry::getCCJ), or nullptr when the velocity
// is already a flux (a_velIsFlux = true).  multByJ / CellToEdge / the wall BC of m_divBC / EdgeToCell / divByJ are folded
// into the two sweeps either side of the solve.
class CCProjector {
public:
    explicit CCProjector(PressureSolver& solver) : solver_(solver) {}
    // computeDiv (LevelCCProjector.cpp:156-187) [of J u when J is given]
    void computeDiv(Field& div, Field& ux, Field& uy, Field* uz = nullptr, Field* J = nullptr) {
        check(somar_cc_divergence(L(), div.handle(), ux.handle(), uy.handle(), uz ? uz->handle() : nullptr,
                                  J ? J->handle() : nullptr, 0.0), "CCProjector::computeDiv");
    }
    // computeGrad (:199-231): cell-centred J g^{dj} d_j phi with the order-2 extrapolated ghosts of gradPiFuncBC
    void computeGrad(Field& gx, Field& gy, Field* gz, Field& phi) {
        check(somar_cc_gradient(L(), gx.handle(), gy.handle(), gz ? gz->handle() : nullptr, phi.handle()), "CCProjector::computeGrad");
    }
    // levelProject: [u *= J;] rhs = Div[u]/dt; solve; u -= dt Grad[pressure]; [u *= 1/J]
    SomarSolveStats project(Field& ux, Field& uy, Field* uz, Field& pressure, double dt, Field* J = nullptr,
                            bool zeroPressure = true, bool forceHomogeneous = false) {
        SomarSolveStats st{};
        const int rc = somar_project_cc(L(), ux.handle(), uy.handle(), uz ? uz->handle() : nullptr, J ? J->handle() : nullptr,
                                        pressure.handle(), dt, zeroPressure, forceHomogeneous, &st);
        if (rc != SOMAR_OK && rc != SOMAR_ERR_DIVERGED) check(rc, "CCProjector::project");
        return st;
    }

private:
    SomarLevel* L() const { return solver_.level().handle(); }
    PressureSolver& solver_;
};

// LevelLepticSolver (src/calculus/LepticSolver/LevelLepticSolver.H:40-347): the LinearSolver<LevelData<FArrayBox>> built
// on a LepticOperator.  define() takes the level that plays the operator's part (its metric, BCs and boxes are what
// LepticOperator::getFCJgup/getCCJinv/getBCs/smallestGridSize hand over, LepticOperator.H:33-44); the parameter setters
// mirror setParameters / setHorizMGParameters / setHorizBottomParameters / setFullMGParameters /
// setFullBottomParameters (LevelLepticSolver.cpp:524-640) and must be called before define, as in the reference.
class LevelLepticSolver {
public:
    LevelLepticSolver() { check(somar_leptic_default_params(&prm_), "LevelLepticSolver"); }
    ~LevelLepticSolver() { undefine(); }
    LevelLepticSolver(const LevelLepticSolver&) = delete;
    LevelLepticSolver& operator=(const LevelLepticSolver&) = delete;

    void setParameters(int maxOrder, double eps, double hang, int normType, double horizRhsTol) {
        prm_.max_order = maxOrder; prm_.eps = eps; prm_.hang = hang; prm_.norm_type = normType; prm_.horiz_rhs_tol = horizRhsTol;
    }
    SomarSolverParams& horizParameters() { return prm_.horiz; }   // multigrid + bottom solver of the flattened problem
    SomarSolverParams& fullParameters() { return prm_.full; }     // multigrid + bottom solver of the 3-D fall-back
    void setHomogeneous(bool homogeneous) { homogeneous_ = homogeneous; }

    void define(Level& op, bool homogeneous = false) {
        undefine();
        homogeneous_ = homogeneous;
        check(somar_leptic_create(op.handle(), &prm_, &h_), "LevelLepticSolver::define");
    }
    void undefine() {
        if (h_) somar_leptic_destroy(h_);
        h_ = nullptr;
    }
    bool isDefined() const { return h_ != nullptr; }

    // solve: phi += the leptic corrections for L[phi] = rhs.  Like the reference, running out of orders, hanging or
    // diverging is a status, not an error.
    SomarLepticStats solve(Field& phi, const Field& rhs) {
        if (!h_) throw Error(SOMAR_ERR_ARG, "LevelLepticSolver::solve before define");
        check(somar_leptic_solve(h_, phi.handle(), rhs.handle(), homogeneous_, &stats_), "LevelLepticSolver::solve");
        return stats_;
    }
    int exitStatusInt() const { return stats_.exit_status; }
    const char* exitStatusStr() const {
        static const char* names[] = {"NONE", "CONVERGE", "ITER", "HANG", "DIVERGE", "KABOOM"};
        const int s = stats_.exit_status;
        return (s >= -1 && s <= 4) ? names[s + 1] : "UNKNOWN!";
    }
    std::vector<double> resNorms() const { return std::vector<double>(stats_.res_norms, stats_.res_norms + stats_.num_res_norms); }

private:
    SomarLepticParams prm_{};
    SomarLepticStats stats_{0, -1, 0, 0, 0, 0, {0.0}, 0.0, 0};
    SomarLeptic* h_ = nullptr;
    bool homogeneous_ = false;
};

}  // namespace somar_b200
