// geom.cuh -- device-side data model of one multigrid depth of one AMR level (rank-local part).
//
// HBM layout (DESIGN.md "Data layout"): every cell-centred field AND every face-centred metric component of a
// depth uses the SAME padded brick, x fastest:
//     addr(i,j,k) = off0 + i + sy*j + sz*k            (i,j,k = rank-local cell indices, 0-based)
// with GY/GZ ghost planes on each side in y/z, XO (>= ghost, multiple of 4 doubles = one 32 B sector) pad cells in x
// and the pitch rounded to 128 B, so that cell i=0 of every row is sector aligned and a kernel can read any
// neighbour of any valid cell unconditionally.  Face (i,j,k) of direction d is the LOW face of cell (i,j,k); the
// high face of the last cell lives in the first ghost slot.  This replaces Chombo's per-box FArrayBox/FluxBox
// (x fastest, ghosts inside the box: SURVEY.md 8(b)); the boxes a rank owns are fused into this one brick.
#pragma once
#include <cstdint>

#define SOMAR_GHOST 2   // ghost planes in y and z (2 = enough for a fused red+black sweep)
#define SOMAR_XO 4      // pad cells in x on each side

namespace somar {

enum { BC_NONE = -1, BC_NEUM = 0, BC_DIRI = 1, BC_PERIODIC = 2,
       // order-2 extrapolated ghost, no flux method: EllipticExtrapBCGhostClass(2), the reference projectors' gradient BC
       // (PhysBCUtil::basicGradPressureFuncBC).  Only the gradient kernels of projector.cu run with it.
       BC_EXTRAP2 = 4 };

struct Geom {
    int n[3];        // rank-local valid cells
    int g0[3];       // global index of local cell 0
    int gn[3];       // domain cells at this depth
    int nd;          // active directions (2 or 3)
    int gh[3];       // ghost planes allocated per side (x: SOMAR_XO)
    long long sy, sz, off0, total;
    int bc[3][2];    // BC code where the local side is a non-periodic domain boundary, else BC_NONE
    int bct[3][2];   // BC code of the DOMAIN side (BC_PERIODIC for periodic directions)
    int bs[3];       // box size of the level's (uniform) DisjointBoxLayout tiling at this depth (full metric only)
    double bcval[3][2];
    double dx[3];
    double rdx[3];   // 1 / dx[d]            (exact IEEE quotients computed once on the host: kernels that
    double rdx2[3];  // 1 / (dx[d] * dx[d])   hoisted these divisions paid for them in registers)
    double alpha, beta;
    int diriOrder;   // EllipticConstDiriBCGhostClass order (1)
    int periodic[3];
    int nbr[2];      // neighbour rank on the low/high y side (-1: none; may be self for periodic single rank)
    __host__ __device__ inline long long idx(int i, int j, int k) const { return off0 + i + sy * j + sz * k; }
    __host__ __device__ inline long long cells() const { return (long long)n[0] * n[1] * n[2]; }
};

// Metric of one depth.  jg[d][c] = J g^{dc} on faces of direction d (c == d only when the metric is diagonal).
struct Metric {
    const double* jg[3][3];
    const double* jinv;
    const double* lapdiag;
    int diagonal;
};

}  // namespace somar
