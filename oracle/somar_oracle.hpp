// oracle/somar_oracle.hpp -- TEST INFRASTRUCTURE ONLY (CPU oracle; never the product path).
//
// CPU restatement of the C++ drivers on SOMAR's pressure-solve hot path (single AMR level):
//   MappedAMRPoissonOp (applyOpI/residualI/relax/restrictResidual/prolongIncrement/preCond),
//   RelaxationMethod + LevelGSRB/LooseGSRB/LineGSRB/Jacobi, FullWeightingPS, (ZeroAvg)ConstInterpPS,
//   MappedAMRPoissonOpFactory::MGnewOp (semicoarsening schedule + coarse metrics),
//   MappedMultiGrid::cycle, MappedAMRMultiGrid::solveNoInitResid, and a BiCGStab bottom solver.
// Each function cites the reference file:line it follows.  KERNELS PINNED to the reference's own Fortran: oracle/chf2py.py executes the .ChF sources from /root/reference and
// tests/test_reference_kernels.py demands identical bits (fixtures tests/golden/refkernels_{2,3}d.npz); the C++ DRIVERS above
// the kernels stay unpinned (the reference cannot be built here and ships no golden vectors).
#pragma once
#include "chombo_lite.hpp"
#include "kernels.hpp"
#include "line_transfer.hpp"
#include "smoothers.hpp"
#include <memory>
#include <string>

namespace orc {

struct Domain {
    Box box;
    bool periodic[3] = {false, false, false};
};

// One field on a layout: one FAB per box, `ghost` cells on every side (0 in inactive directions).
struct LevelData {
    std::vector<FAB> fabs;
    IntVect ghost;
    int nc = 1;
    void define(const std::vector<Box>& boxes, int ncomp, const IntVect& g) {
        ghost = g; nc = ncomp;
        fabs.resize(boxes.size());
        for (size_t b = 0; b < boxes.size(); ++b) fabs[b].define(grow(boxes[b], g), ncomp);
    }
    FAB& operator[](size_t b) { return fabs[b]; }
    const FAB& operator[](size_t b) const { return fabs[b]; }
    size_t size() const { return fabs.size(); }
};

// FluxBox: one face-centred FAB per direction.
struct FluxLevelData {
    std::vector<std::array<FAB, 3>> fabs;
    void define(const std::vector<Box>& boxes, int ncomp, int ndirs) {
        fabs.resize(boxes.size());
        for (size_t b = 0; b < boxes.size(); ++b)
            for (int d = 0; d < ndirs; ++d) fabs[b][d].define(surroundingNodes(boxes[b], d), ncomp);
    }
};

// BCMethodHolder stand-in (src/calculus/BCInterface): constant Neumann / Dirichlet per side.
// PhysBCUtil::basicPressureFuncBC (src/BCutil/PhysBCUtil.cpp:1404-1430) = homogeneous Neumann on
// every non-periodic side, for both the ghost and the flux method.
struct BCSpec {
    int type[3][2] = {{BC_Neum, BC_Neum}, {BC_Neum, BC_Neum}, {BC_Neum, BC_Neum}};
    double value[3][2] = {{0, 0}, {0, 0}, {0, 0}};
    int diriOrder = 1;  // EllipticConstDiriBCGhostClass uses order 1 (EllipticBCUtils.cpp:423)
    // BCDescriptor::stencil -- BCDescriptor.H:212-221
    int stencil(const Box& region, const Domain& dom, int dir, int side) const {
        if (!dom.periodic[dir] && region.sideEnd(side, dir) == dom.box.sideEnd(side, dir)) return type[dir][side];
        return BC_None;
    }
};

enum RelaxMode { RELAX_NONE = -1, RELAX_JACOBI = 0, RELAX_LEVEL_GSRB = 1, RELAX_LOOSE_GSRB = 2, RELAX_LINE_GSRB = 3 };

// RelaxationMethod::BoundaryBoxData -- RelaxationMethod.H
struct BoundaryBoxData {
    int index;
    Box valid, validBdry;
    int stencil[3][2];
};

struct SolverParams {
    // defaults: src/utils/ProblemContext.cpp:1147-1216
    double eps = 1e-6;
    int num_smooth_down = 2, num_smooth_up = 2, num_smooth_bottom = 2, num_smooth_precond = 2;
    int numMG = 1, imin = 5, imax = 20;
    double hang = 1e-15, normThresh = 1e-30;
    int maxDepth = -1;
    int relaxMode = RELAX_LEVEL_GSRB;
    int precondMode = 1;  // PrecondMode::DiagRelax
    double bottom_eps = 1e-6, bottom_reps = 1e-12, bottom_hang = 1e-7, bottom_small = 1e-30;
    int bottom_imax = 80, bottom_numRestarts = 5, bottom_normType = 2;
    int verbosity = 0;
};

struct SolveStats {
    int iters = 0;
    int exitStatus = 0;
    double initial_rnorm = 0, final_rnorm = 0;
    std::vector<double> history;  // residual max-norm after each V-cycle (history[0] = initial)
    int bottom_iters_total = 0;
};

// One MappedAMRPoissonOp at one multigrid depth.
struct Op {
    Domain domain;
    std::vector<Box> grids;
    double dx[3];
    double dxCrse[3];
    int nd = 3;        // number of active directions (SpaceDim, or SpaceDim-1 for a horizontal op)
    IntVect activeDirs;
    double alpha = 0.0, beta = 1.0;
    bool isDiagonal = true;
    BCSpec bc;
    FluxLevelData Jgup;
    LevelData Jinv, lapDiag;
    IntVect mgCrseRefRatio;
    Box validDomain;
    int relaxMode = RELAX_LEVEL_GSRB;
    int precondMode = 1, preCondSmoothIters = 2;
    bool zeroAvgProlong = false;
    double dxProduct = 1.0;
    // RelaxationMethod state
    LevelData extrap;
    std::vector<BoundaryBoxData> boundaryBoxData, simpleBoundaryBoxData;

    void defineCommon();       // MappedAMRPoissonOp::define4 + RelaxationMethod::define
    void collectBoundaryData();
    void computeLapDiag();

    void exchange(LevelData& phi, bool corners) const;
    void fillExtrapOp(FAB& ex, const FAB& phi, int order) const;                 // fillExtrap
    void setGhosts(FAB& phi, FAB* ex, const Box& valid, size_t bi, bool homog) const;
    void setFluxes(std::array<FAB, 3>& flux, const Box& valid, bool homog) const;
    void applyOp(LevelData& lhs, LevelData& phi, bool homog) const;
    void residual(LevelData& lhs, LevelData& phi, const LevelData& rhs, bool homog) const;
    void fillGhostsAndExtrapolate(LevelData& phi);
    // leptic solver pieces that need per-box ghosts (either metric): LevelLepticSolver::computeHorizRHS and
    // ::levelVertHorizGradient.  flat*: one value per (i, j), global x-fastest arrays.
    void lepticExtrapAllGhosts(LevelData& phi, int order) const;
    void lepticHorizRHS(double* flatRhs, LevelData& phi) const;
    void lepticVertHorizGradient(double* flatLo, double* flatHi, LevelData& phi, double scale) const;
    void relax(LevelData& phi, const LevelData& rhs, int iters);
    void relaxOnce(LevelData& phi, const LevelData& rhs);
    void fullStencilGSRB(FAB& phi, const FAB& rhs, const Box& region, size_t bi, int pass) const;
    void boundaryGSRB(LevelData& phi, const LevelData& rhs, int pass, bool all) const;
    void preCond(LevelData& phi, const LevelData& rhs);
    void restrictResidual(LevelData& resCoarse, LevelData& phiFine, const LevelData& rhsFine, const Op& crseOp) const;
    void prolongIncrement(LevelData& phiFine, const LevelData& corrCoarse) const;
    double norm(const LevelData& x, int ord) const;
    double dot(const LevelData& a, const LevelData& b) const;
    void setToZero(LevelData& x) const { for (auto& f : x.fabs) f.setVal(0.0); }
    void incr(LevelData& y, const LevelData& x, double s) const;
    void assign(LevelData& y, const LevelData& x) const;
    void scale(LevelData& y, double s) const;
    void create(LevelData& y, int ghosts) const;
    // projector calculus (Gradient::levelGradientMAC, Divergence::levelDivergenceMAC)
    // gradBC 0: the operator's BC holder (ghosts + boundary fluxes); 1: order-2 extrapolated ghosts and no flux
    // method -- PhysBCUtil::basicGradPressureFuncBC (PhysBCUtil.cpp), the m_gradBC of both reference projectors
    void macGradient(FluxLevelData& grad, LevelData& phi, bool homog, int gradBC = 0) const;
    void macDivergence(LevelData& div, const FluxLevelData& flux) const;
    // the projectors' m_divBC on solid walls (PhysBCUtil::uStarFuncBC -> BasicVelocityBCGhostClass, normal component,
    // face-centred branch of setSideDiriBC, EllipticBCUtils.cpp:41-118): boundary faces of direction d of flux[d] <- 0
    void wallFluxBC(FluxLevelData& flux) const;
    // Chombo AMRTools CellToEdge / EdgeToCell (component d of the cell data <-> faces of direction d)
    void cellToEdge(FluxLevelData& edge, std::array<LevelData, 3>& cell) const;
    void edgeToCell(std::array<LevelData, 3>& cell, const FluxLevelData& edge) const;
};

struct MG {
    std::vector<std::unique_ptr<Op>> ops;
    std::vector<LevelData> residual, correction;
    std::vector<IntVect> mgRefRatios;
    SolverParams p;
    long bottomCells = 0;
    // bottom solver convergence metrics (BiCGStabSolver::setConvergenceMetrics)
    double bottomMetric = 0.0, bottomEps = 1e-6;
    int bottom_iters_total = 0;
    double convergenceMetric = 0.0;   // MappedAMRMultiGrid::m_convergenceMetric (MappedAMRMultiGrid.H:1036-1038)
    bool blewUp = false;              // "solver blew up" abort of :1140-1144 (the reference calls MayDay::Error)
    int depth() const { return (int)ops.size(); }
    void cycle(int depth, LevelData& e, const LevelData& r);
    int bottomSolve(LevelData& e, const LevelData& r);
    void solve(LevelData& phi, const LevelData& rhs, bool zeroPhi, bool forceHomogeneous, SolveStats& st);
    void project(FluxLevelData& vel, LevelData& pressure, double dt, bool zeroPressure, bool forceHomogeneous,
                 SolveStats& st);   // BaseProjector::project, a_velIsFlux = true
    // BaseProjector<FArrayBox>::project through LevelCCProjector (LevelCCProjector.cpp:156-255); J == nullptr means
    // a_velIsFlux = true
    void projectCC(std::array<LevelData, 3>& vel, const LevelData* J, LevelData& pressure, double dt, bool zeroPressure,
                   bool forceHomogeneous, SolveStats& st);
};

// Factory: builds all depths (MGnewOp schedule + coarse metrics).
std::unique_ptr<MG> buildMG(const Domain& dom, const std::vector<Box>& grids, const double* dx, int nd, double alpha,
                            double beta, bool isDiagonal, const BCSpec& bc, const SolverParams& p,
                            FluxLevelData&& Jgup0, LevelData&& Jinv0);

}  // namespace orc
