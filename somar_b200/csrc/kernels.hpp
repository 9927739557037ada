// kernels.hpp -- host-callable launchers of the sm_100a kernels (one translation unit per kernel family).
#pragma once
#include "geom.cuh"
#include <cuda_runtime.h>

#define SOMAR_MAX_RED_BLOCKS 1184   // 148 SMs x 8

namespace somar {

struct Ctx {
    cudaStream_t stream = nullptr;
    long long launches = 0;   // kernels of this library launched so far
    int sms = 148;
    double* scratch = nullptr;        // device: 2 * SOMAR_MAX_RED_BLOCKS partial results
    unsigned int* tickets = nullptr;  // device: 8 zero-initialised counters
};

// ---- operator (stencil_ortho.cuh / stencil_full.cuh) ----
// mode 0: out = (alpha + beta L) phi;  mode 1: out = rhs - (alpha + beta L) phi
bool launch_residual_maxnorm(Ctx& cx, const Geom& g, const Metric& m, double* out, const double* phi, const double* rhs,
                             bool homog, double* maxOut);
void launch_apply(Ctx& cx, const Geom& g, const Metric& m, double* out, const double* phi, const double* rhs,
                  int mode, bool homog);
// one colour pass of point Gauss-Seidel (pass 0 = red: (i+j+k) even in global indices)
void launch_gsrb(Ctx& cx, const Geom& g, const Metric& m, double* phi, const double* rhs, int pass);
// LooseGSRB stages (diagonal metric): stage 0 = box interiors, stage 1 = box rims (boundary-kernel statements;
// pass 1 reads cross-box neighbours from `snap`)
void launch_gsrb_loose(Ctx& cx, const Geom& g, const Metric& m, double* phi, const double* snap, const double* rhs,
                       int pass, int stage);
// one whole iteration (red then black) in one sweep, out of place; see gsrb_fused.cu
bool gsrb_fused_usable(const Geom& g, const Metric& m);
void launch_gsrb_fused(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn,
                       const double* rhs);
// the same sweep as a warp-specialised pipeline: producer warps stream the planes into mbarrier-signalled
// shared-memory rings with asynchronous copies, consumers never touch global memory for loads; see gsrb_async.cu.
// Returns false (nothing launched) if the kernel cannot get its shared memory.
bool gsrb_async_usable(const Geom& g, const Metric& m);
bool launch_gsrb_async(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn,
                       const double* rhs, const double* zeroPlane = nullptr);
// k_gsrb_async with its rings fed by TMA tensor-map loads issued by one thread (gsrb_tma.cu); SOMAR_B200_TMA=0 turns
// it off.  Same contract and the same bits as launch_gsrb_async.
bool gsrb_tma_usable(const Geom& g, const Metric& m);
bool launch_gsrb_tma(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn, const double* rhs,
                     const double* zeroPlane = nullptr);
// full metric: out-of-place pass, phiOut gets the updated colour AND a copy of the other colour
void launch_gsrb_full(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn,
                      const double* rhs, int pass);
// phi += 0.5 * res / (alpha + beta lapDiag)     JACOBIITER, JacobiF.ChF:31-61
void launch_jacobi(Ctx& cx, const Geom& g, const Metric& m, double* phi, const double* res);
// phi = rhs / (alpha + beta lapDiag)            DIAGPRECOND, MappedAMRPoissonOpF.ChF:284-328
void launch_diag_precond(Ctx& cx, const Geom& g, const Metric& m, double* phi, const double* rhs);
// vertical line relaxation, one colour of columns (LineGSRBIter{2,3}D, GSRBF.ChF:1529-2042).
// fact/dmod hold the precomputed elimination of the (phi-independent) tridiagonal matrices.
void launch_line_factor(Ctx& cx, const Geom& g, const Metric& m, double* fact, double* dmod);
void launch_line_gsrb(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn,
                      const double* rhs, const double* fact, const double* dmod, int pass);

// ---- transfers ----
// coarse = J-weighted average of (rhs - L phi) over the ratio block   (restrictResidual + MAPPEDAVERAGE2)
void launch_restrict_residual(Ctx& cx, const Geom& gf, const Metric& mf, const Geom& gc, double* crse,
                              const double* phi, const double* rhs, const int ratio[3], double* scratch = nullptr);
// fine += coarse[i/m]; if sums != nullptr also sums[0] += sum(dV*fine), sums[1] += sum(dV)  (ConstInterp[WithAvg]PS)
void launch_prolong(Ctx& cx, const Geom& gf, const Metric& mf, const Geom& gc, double* fine, const double* crse,
                    const int ratio[3], double dxProduct, double* sums);
// x -= sums[0]/sums[1] over the whole brick (ZeroAvgConstInterpPS::prolongIncrement, ProlongationStrategy.cpp:163)
void launch_prolong_finish(Ctx& cx, const Geom& gf, const Geom& gc, double* fine, const double* crse, const int r[3],
                           const double* sums);

// ---- projector calculus (projector.cu) ----
// out[d](face) = J g^{dj} d_j phi with the operator's BCs; accumulate: out[d] += scale * that
void launch_mac_grad(Ctx& cx, const Geom& g, const Metric& m, double* const out[3], const double* phi, bool homog,
                     double scale, bool accumulate);
// div = Jinv * sum_d (F_d(+) - F_d(-)) / dx_d, divided by divideBy when that is non-zero
void launch_mac_div(Ctx& cx, const Geom& g, const Metric& m, double* div, const double* const flux[3], double divideBy);

// ---- set-up ----
void launch_lapdiag(Ctx& cx, const Geom& g, const Metric& m, double* lapdiag);           // FILLMAPPEDLAPDIAG*
void launch_avg_face(Ctx& cx, const Geom& gf, const Geom& gc, double* crse, const double* fine, int dir,
                     const int ratio[3]);                                                 // UNMAPPEDAVERAGEFACE
void launch_avg_harmonic(Ctx& cx, const Geom& gf, const Geom& gc, double* crse, const double* fine,
                         const int ratio[3]);                                             // UNMAPPEDAVERAGEHARMONIC

// ---- BLAS-1 over valid cells (LevelDataOps, MappedAMRPoissonOp.H:566-711) ----
void launch_fill(Ctx& cx, const Geom& g, double* x, double v);                      // whole brick
void launch_copy(Ctx& cx, const Geom& g, double* y, const double* x);               // whole brick
void launch_axpy(Ctx& cx, const Geom& g, double* y, const double* x, double a);     // y += a x
void launch_axby(Ctx& cx, const Geom& g, double* out, const double* x, const double* y, double a, double b);
void launch_scale(Ctx& cx, const Geom& g, double* x, double a);
// out[0] = reduction over valid cells: kind 0 max|x|, 1 sum|x|, 2 sum x^2, 3 sum x*y, 4 signed max x
void launch_reduce(Ctx& cx, const Geom& g, const double* x, const double* y, int kind, double* out);

// ---- halos ----
// copy periodic images inside the brick for direction d (width w ghost layers)
void launch_periodic_fill(Ctx& cx, const Geom& g, double* x, int dir, int w);
// pack / unpack w ghost rows in y (side 0 = low) to / from a contiguous buffer of w * nz_all * pitch doubles
void launch_pack_y(Ctx& cx, const Geom& g, const double* x, double* buf, int side, int w, bool unpack);
// brick <-> dense (rank-local, no ghosts, x fastest) array; ext = extra points per direction (faces)
void launch_dense_copy(Ctx& cx, const Geom& g, double* brick, double* dense, const int ext[3], bool toBrick);

}  // namespace somar
