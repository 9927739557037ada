// ops_ortho.cu -- diagonal-metric operator kernels: apply / residual, red-black point relaxation, Jacobi,
// diagonal preconditioner, fused residual+restriction.  HBM-bound fp64 streaming kernels (DESIGN.md "Kernels").
#include <algorithm>
#include <cstdlib>

#include "kernels.hpp"
#include "stencil_ortho.cuh"
#include "stencil_full.cuh"

namespace somar {

#define BX 64
#define BY 4
// k_march tuning, measured at 1024x1024x256 (unroll x CTAs/SM sweep): occupancy beats register comfort.
// residual (modes 0/1): unroll 4, 4 CTAs/SM (64 regs): 3.91 -> 2.86 ms; restrict (mode 2): no unroll, 5 CTAs/SM: 4.21 -> 3.36 ms
#define MARCH_UNROLL(MODE) ((MODE) == 2 ? 1 : 4)   // (MODE 3 = residual + max-norm: as MODE 1)
#define MARCH_MINB(MODE) ((MODE) == 2 ? 5 : 4)

// Full metric: is cell (i, j, k) strictly inside its box of the uniform tiling?  Then every `extrap` value its stencil
// reads is phi itself and the branch-free *_inner statement applies (same expressions, same bits as the general one).
template <int ND>
__device__ __forceinline__ bool full_inner_xy(const Geom& g, int i, int j) {
    return box_inner_dir<ND>(g, i, 0) && box_inner_dir<ND>(g, j, 1);
}

// ---------------------------------------------------------------------------------------------------------
// Full metric: the cells that touch a side of their box take the general statement (box-relative extrapolated views,
// ExView), which is an order of magnitude longer than the branch-free one.  In a thread-per-cell mapping the two x-edge
// cells of a row put it into that row's first and last warp with ONE active lane (ncu on k_gsrb_full,
// profiles/r2_ncu_k_gsrb_full.txt: 15 of 32 lanes active on average, 2.6 warps per scheduler), so the full-metric kernels
// run in two launches: the box-interior cells, and the rim cells mapped densely onto lanes -- three slabs, each rim cell
// in exactly one: the z-rim planes, all (i, j); then the y-rim rows of the other planes, all i; then the x-rim columns of
// what is left (lanes along j).  The m-th rim index of a direction is side m % 2 of local box m / 2 (a rank owns whole
// boxes); a one-cell-thick box has one.  Each rim cell is then dispatched by kind (cell_kind, stencil_full.cuh).
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ int rim_index(const Geom& g, int d, int m) {
    return (g.bs[d] > 1) ? (m >> 1) * g.bs[d] + ((m & 1) ? g.bs[d] - 1 : 0) : m;
}
__device__ __forceinline__ int rim_count(const Geom& g, int d) { return (g.bs[d] > 1) ? 2 * (g.n[d] / g.bs[d]) : g.n[d]; }

// One block row per rim line: `row` (blockIdx.y + gridDim.y * blockIdx.z) counts the nzr * n1 rows of the z-rim planes, then
// nyr rows for each plane, then nxr columns for each plane; blockIdx.x tiles the line.
#ifndef RIM_MINB
#define RIM_MINB 4
#endif
template <int ND>
__device__ __forceinline__ bool rim_cell(const Geom& g, int nzr, int nyr, int nxr, int& i, int& j, int& k, int pass = -1) {
    // pass >= 0: only the cells of that colour, two cells of the line per lane step
    const int lane0 = blockIdx.x * blockDim.x + threadIdx.x;
    const int gsum = g.g0[0] + g.g0[1] + ((ND > 2) ? g.g0[2] : 0) + pass;
#define RIM_LANE(a, b) ((pass < 0) ? lane0 : 2 * lane0 + ((gsum + (a) + (b)) & 1))
    const int planes = (ND > 2) ? g.n[2] : 1;
    int row = blockIdx.y + gridDim.y * blockIdx.z;
    if (row < nzr * g.n[1]) {                                    // slab Z
        k = rim_index(g, 2, row / g.n[1]);
        j = row % g.n[1];
        i = RIM_LANE(j, k);
        return i < g.n[0];
    }
    row -= nzr * g.n[1];
    if (row < planes * nyr) {                                    // slab Y
        k = row / nyr;
        j = rim_index(g, 1, row % nyr);
        i = RIM_LANE(j, k);
        if (ND > 2 && !box_inner_dir<ND>(g, k, 2)) return false;
        return i < g.n[0];
    }
    row -= planes * nyr;
    if (row >= planes * nxr) return false;
    k = row / nxr;                                               // slab X: lanes along j
    i = rim_index(g, 0, row % nxr);
    j = RIM_LANE(i, k);
    if (ND > 2 && !box_inner_dir<ND>(g, k, 2)) return false;
    return j < g.n[1] && box_inner_dir<ND>(g, j, 1);
#undef RIM_LANE
}

// host side: can this rank's part of the depth be split (whole boxes in every direction), and the rim kernel's grid
static bool rim_split_usable(const Geom& g) {
    static const bool on = []() { const char* e = getenv("SOMAR_B200_FULL_SPLIT"); return !(e && e[0] == '0'); }();
    if (!on) return false;
    for (int d = 0; d < g.nd; ++d)
        if (g.bs[d] <= 0 || g.n[d] % g.bs[d] != 0 || g.g0[d] % g.bs[d] != 0) return false;
    return true;
}
struct RimGrid {
    dim3 grid;
    int nzr, nyr, nxr;
};
static RimGrid rim_grid(const Geom& g, bool oneColour = false) {
    auto cnt = [&](int d) { return (g.bs[d] > 1) ? 2 * (g.n[d] / g.bs[d]) : g.n[d]; };
    RimGrid r;
    r.nzr = (g.nd == 3) ? cnt(2) : 0;
    r.nyr = cnt(1);
    r.nxr = cnt(0);
    const int planes = (g.nd == 3) ? g.n[2] : 1;
    const long long rows = (long long)r.nzr * g.n[1] + (long long)planes * (r.nyr + r.nxr);
    const int gy = (int)std::min<long long>(rows, 32768);
    const int line = std::max(g.n[0], g.n[1]);
    r.grid = dim3(((oneColour ? (line + 1) / 2 : line) + 127) / 128, gy, (unsigned)((rows + gy - 1) / gy));
    return r;
}

// ---------------------------------------------------------------------------------------------------------
// apply / residual: one thread per cell of an (BX x BY) tile, marching KZ planes in z.
// ---------------------------------------------------------------------------------------------------------
template <int ND, bool FULL, bool SPLIT = false>
__global__ void __launch_bounds__(BX* BY) k_apply(const Geom g, const Metric m, double* __restrict__ out,
                                                   const double* __restrict__ phi, const double* __restrict__ rhs,
                                                   int mode, bool homog, int kz) {
    const int i = blockIdx.x * BX + threadIdx.x;
    const int j = blockIdx.y * BY + threadIdx.y;
    if (i >= g.n[0] || j >= g.n[1]) return;
    const int k0 = blockIdx.z * kz;
    const int k1 = min(k0 + kz, g.n[2]);
    const bool innerXY = FULL && full_inner_xy<ND>(g, i, j);
    // SPLIT: every cell whose accessor is phi itself (CELL_FAST: box-interior cells, and cells on box-box interfaces away from
    // the domain boundary); the others are left to k_apply_full_rim
    const unsigned cxy = (FULL && SPLIT) ? (class_bit(cell_dir_class(g, i, 0)) | class_bit(cell_dir_class(g, j, 1))) : 0u;
    for (int k = k0; k < k1; ++k) {
        const long long c = g.idx(i, j, k);
        double v;
        if (FULL) {
            const bool fast = SPLIT ? classes_fast(cxy | ((ND > 2) ? class_bit(cell_dir_class(g, k, 2)) : 0u))
                                    : (innerXY && (ND < 3 || box_inner_dir<ND>(g, k, 2)));
            if (fast) v = apply_cell_full_inner<ND>(g, m, phi, i, j, k, c, homog);
            else if (SPLIT) continue;
            else v = apply_cell_full<ND>(g, m, phi, i, j, k, c, homog);
        } else {
            v = apply_cell_ortho<ND>(g, m, phi, i, j, k, c, homog);
        }
        out[c] = mode ? (rhs[c] - v) : v;
    }
}

template <int ND>
__global__ void __launch_bounds__(128, RIM_MINB) k_apply_full_rim(const Geom g, const Metric m, double* __restrict__ out,
                                                         const double* __restrict__ phi, const double* __restrict__ rhs,
                                                         int mode, bool homog, int nzr, int nyr, int nxr) {
    int i, j, k;
    if (!rim_cell<ND>(g, nzr, nyr, nxr, i, j, k)) return;
    const long long c = g.idx(i, j, k);
    const int kind = cell_kind<ND>(g, i, j, k);
    if (kind == CELL_FAST) return;                               // k_apply<ND, true, true> did it
    const double v = apply_cell_full_kind<ND>(g, m, phi, i, j, k, c, homog, kind);
    out[c] = mode ? (rhs[c] - v) : v;
}

// ---------------------------------------------------------------------------------------------------------
// Diagonal metric, 3-D: the same operator marching through z with the column's phi (planes k-1, k, k+1) and
// the low z-face coefficient in registers, so every array is read once per cell from HBM/L2 (x/y neighbours
// are L1 hits: the plane was fetched one step earlier).  MODE 0: out = L phi, 1: out = rhs - L phi,
// 2: restrictResidual for mgRefRatio (1,1,2): crse = J-weighted average of the residual of two planes
// (restrictResidual + MAPPEDAVERAGE2, MappedAMRPoissonOp.cpp:1281-1304, MappedCoarseAverageF.ChF:132-168);
// the accumulation order is the reference's (kk = 0, 1).
// ---------------------------------------------------------------------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(BX* BY, MARCH_MINB(MODE)) k_march(const Geom g, const Metric m, const Geom gc, double* __restrict__ out,
                                                   const double* __restrict__ phi, const double* __restrict__ rhs,
                                                   bool homog, int kz, unsigned long long* __restrict__ maxBits = nullptr) {
    // MODE 3 = MODE 1 + max |residual| folded into *maxBits (the bit pattern of a non-negative double orders like the
    // integer; a maximum does not depend on the order it is taken in, so this equals k_reduce's result bit for bit)
    const int i = blockIdx.x * BX + threadIdx.x;
    const int j = blockIdx.y * BY + threadIdx.y;
    const bool act = i < g.n[0] && j < g.n[1];
    if (MODE != 3 && !act) return;
    double rmax = 0.0;
    if (act) {
    const int k0 = blockIdx.z * kz;
    const int k1 = min(k0 + kz, g.n[2]);
    const double* __restrict__ jg0 = m.jg[0][0];
    const double* __restrict__ jg1 = m.jg[1][1];
    const double* __restrict__ jg2 = m.jg[2][2];
    const double* __restrict__ jinv = m.jinv;
    long long c = g.idx(i, j, k0);
    double pm = phi[c - g.sz], pc = phi[c], zlo = jg2[c];
    double s = 0.0, sj = 0.0;
    // side bits that do not depend on k
    unsigned smxy = 0;
    if (i == 0 && g.bc[0][0] != BC_NONE) smxy |= 1u;
    if (i == g.n[0] - 1 && g.bc[0][1] != BC_NONE) smxy |= 2u;
    if (j == 0 && g.bc[1][0] != BC_NONE) smxy |= 4u;
    if (j == g.n[1] - 1 && g.bc[1][1] != BC_NONE) smxy |= 8u;
    constexpr int UNR = MARCH_UNROLL(MODE);
#pragma unroll UNR
    for (int k = k0; k < k1; ++k, c += g.sz) {
        const double pn = phi[c + g.sz], zhi = jg2[c + g.sz];
        double nb[3][2], glo[3], ghi[3];
        nb[0][0] = phi[c - 1];
        nb[0][1] = phi[c + 1];
        nb[1][0] = phi[c - g.sy];
        nb[1][1] = phi[c + g.sy];
        nb[2][0] = pm;
        nb[2][1] = pn;
        glo[0] = jg0[c];
        ghi[0] = jg0[c + 1];
        glo[1] = jg1[c];
        ghi[1] = jg1[c + g.sy];
        glo[2] = zlo;
        ghi[2] = zhi;
        const double ji = jinv[c];
        unsigned sm = smxy;
        if (k == 0 && g.bc[2][0] != BC_NONE) sm |= 16u;
        if (k == g.n[2] - 1 && g.bc[2][1] != BC_NONE) sm |= 32u;
        const double v = apply_vals_ortho3(g, sm, homog, pc, nb, glo, ghi, ji);
        if (MODE == 0) {
            out[c] = v;
        } else if (MODE == 1) {
            out[c] = rhs[c] - v;
        } else if (MODE == 3) {
            const double res = rhs[c] - v;
            out[c] = res;
            rmax = fmax(rmax, fabs(res));
        } else {
            const double res = rhs[c] - v;
            s = s + res / ji;
            sj = sj + 1.0 / ji;
            if ((k - k0) & 1) {
                out[gc.idx(i, j, k >> 1)] = s / sj;
                s = 0.0;
                sj = 0.0;
            }
        }
        pm = pc;
        pc = pn;
        zlo = zhi;
    }
    }
    if (MODE == 3) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) rmax = fmax(rmax, __shfl_xor_sync(0xffffffffu, rmax, o));
        __shared__ double wmax[BX * BY / 32];
        const int t = threadIdx.y * BX + threadIdx.x;
        if ((t & 31) == 0) wmax[t >> 5] = rmax;
        __syncthreads();
        if (t == 0) {
#pragma unroll
            for (int w = 1; w < BX * BY / 32; ++w) rmax = fmax(rmax, wmax[w]);
            if (rmax > 0.0) atomicMax(maxBits, (unsigned long long)__double_as_longlong(rmax));   // one per block
        }
    }
}

static int march_kz(const Geom& g) { return (g.n[2] >= 32) ? 16 : ((g.n[2] + 1) / 2) * 2; }

// residual and its max-norm in one pass (the outer loop's "residual, then norm"); false when this geometry does not
// run k_march (the caller then takes the two-pass route).  *maxBits must be zero on the stream before the launch.
bool launch_residual_maxnorm(Ctx& cx, const Geom& g, const Metric& m, double* out, const double* phi, const double* rhs,
                             bool homog, double* maxOut) {
    if (!(m.diagonal && g.nd == 3)) return false;
    const int kzm = march_kz(g);
    dim3 gridm((g.n[0] + BX - 1) / BX, (g.n[1] + BY - 1) / BY, (g.n[2] + kzm - 1) / kzm), blockm(BX, BY);
    k_march<3><<<gridm, blockm, 0, cx.stream>>>(g, m, g, out, phi, rhs, homog, kzm, reinterpret_cast<unsigned long long*>(maxOut));
    cx.launches++;
    return true;
}

void launch_apply(Ctx& cx, const Geom& g, const Metric& m, double* out, const double* phi, const double* rhs,
                  int mode, bool homog) {
    if (m.diagonal && g.nd == 3) {
        const int kzm = march_kz(g);
        dim3 gridm((g.n[0] + BX - 1) / BX, (g.n[1] + BY - 1) / BY, (g.n[2] + kzm - 1) / kzm), blockm(BX, BY);
        if (mode) k_march<1><<<gridm, blockm, 0, cx.stream>>>(g, m, g, out, phi, rhs, homog, kzm);
        else k_march<0><<<gridm, blockm, 0, cx.stream>>>(g, m, g, out, phi, rhs, homog, kzm);
        cx.launches++;
        return;
    }
    const int kz = (g.n[2] >= 64) ? 8 : 1;
    dim3 grid((g.n[0] + BX - 1) / BX, (g.n[1] + BY - 1) / BY, (g.n[2] + kz - 1) / kz), block(BX, BY);
    if (m.diagonal) {
        if (g.nd == 3) k_apply<3, false><<<grid, block, 0, cx.stream>>>(g, m, out, phi, rhs, mode, homog, kz);
        else k_apply<2, false><<<grid, block, 0, cx.stream>>>(g, m, out, phi, rhs, mode, homog, kz);
    } else {
        const bool split = rim_split_usable(g);
        if (!split) {
            if (g.nd == 3) k_apply<3, true><<<grid, block, 0, cx.stream>>>(g, m, out, phi, rhs, mode, homog, kz);
            else k_apply<2, true><<<grid, block, 0, cx.stream>>>(g, m, out, phi, rhs, mode, homog, kz);
        } else {
            if (g.nd == 3) k_apply<3, true, true><<<grid, block, 0, cx.stream>>>(g, m, out, phi, rhs, mode, homog, kz);
            else k_apply<2, true, true><<<grid, block, 0, cx.stream>>>(g, m, out, phi, rhs, mode, homog, kz);
            const RimGrid r = rim_grid(g);
            if (g.nd == 3) k_apply_full_rim<3><<<r.grid, 128, 0, cx.stream>>>(g, m, out, phi, rhs, mode, homog, r.nzr, r.nyr, r.nxr);
            else k_apply_full_rim<2><<<r.grid, 128, 0, cx.stream>>>(g, m, out, phi, rhs, mode, homog, r.nzr, r.nyr, r.nxr);
            cx.launches++;
        }
    }
    cx.launches++;
}

// ---------------------------------------------------------------------------------------------------------
// GSRB colour pass, in place (diagonal metric: a cell only reads the other colour).
// Thread tx of a row handles cell i = 2*tx + parity.
// ---------------------------------------------------------------------------------------------------------
template <int ND>
__global__ void __launch_bounds__(BX* BY) k_gsrb_ortho(const Geom g, const Metric m, double* __restrict__ phi,
                                                        const double* __restrict__ rhs, int pass, int kz) {
    const int t = blockIdx.x * BX + threadIdx.x;
    const int j = blockIdx.y * BY + threadIdx.y;
    if (j >= g.n[1]) return;
    const int k0 = blockIdx.z * kz;
    const int k1 = min(k0 + kz, g.n[2]);
    for (int k = k0; k < k1; ++k) {
        const int par = (g.g0[0] + g.g0[1] + j + ((ND > 2) ? (g.g0[2] + k) : 0) + pass) & 1;
        const int i = 2 * t + par;
        if (i < g.n[0]) {
            const long long c = g.idx(i, j, k);
            phi[c] = gsrb_cell_ortho<ND>(g, m, phi, rhs, i, j, k, c);
        }
    }
}

void launch_gsrb(Ctx& cx, const Geom& g, const Metric& m, double* phi, const double* rhs, int pass) {
    const int kz = (g.n[2] >= 64) ? 8 : 1;
    const int half = (g.n[0] + 1) / 2;
    dim3 grid((half + BX - 1) / BX, (g.n[1] + BY - 1) / BY, (g.n[2] + kz - 1) / kz), block(BX, BY);
    if (g.nd == 3) k_gsrb_ortho<3><<<grid, block, 0, cx.stream>>>(g, m, phi, rhs, pass, kz);
    else k_gsrb_ortho<2><<<grid, block, 0, cx.stream>>>(g, m, phi, rhs, pass, kz);
    cx.launches++;
}

// ---------------------------------------------------------------------------------------------------------
// LooseGSRB::relax (GSRB.cpp:104-141), diagonal metric: one exchange per iteration instead of two.  Per box the
// reference relaxes the interior grow(box,-1) red then black with the interior kernel, finishes the exchange it began
// before, and then relaxes the 1-cell box rim red then black with the BOUNDARY kernel (boundaryGSRB(..., true),
// GSRBBOUNDARYITER*ORTHO).  The ghosts are not refreshed between the two rim passes, so a rim black cell reads the
// PRE-ITERATION value of red cells that belong to another box.  In the brick there are no per-box ghosts: stage 1
// reads those cross-box neighbours from `snap`, a copy of phi taken after the interior stages (the rim is still
// untouched then), everything else from phi in place.  stage 0 = interior, stage 1 = rim.
// ---------------------------------------------------------------------------------------------------------
template <int ND>
__global__ void __launch_bounds__(BX* BY) k_gsrb_loose(const Geom g, const Metric m, double* __restrict__ phi,
                                                        const double* __restrict__ snap,
                                                        const double* __restrict__ rhs, int pass, int stage, int kz) {
    const int t = blockIdx.x * BX + threadIdx.x;
    const int j = blockIdx.y * BY + threadIdx.y;
    if (j >= g.n[1]) return;
    const int k0 = blockIdx.z * kz;
    const int k1 = min(k0 + kz, g.n[2]);
    for (int k = k0; k < k1; ++k) {
        const int par = (g.g0[0] + g.g0[1] + j + ((ND > 2) ? (g.g0[2] + k) : 0) + pass) & 1;
        const int i = 2 * t + par;
        if (i >= g.n[0]) continue;
        const long long c = g.idx(i, j, k);
        const BoxOf b = box_of<ND>(g, i, j, k);
        if (stage == 0) {
            if (!b.rim) phi[c] = gsrb_cell_ortho<ND>(g, m, phi, rhs, i, j, k, c);
            continue;
        }
        if (!b.rim) continue;
        const int q[3] = {i, j, k};
        const long long st[3] = {1, g.sy, g.sz};
        double nb[3][2] = {{0.0, 0.0}, {0.0, 0.0}, {0.0, 0.0}};
#pragma unroll
        for (int d = 0; d < ND; ++d) {
#pragma unroll
            for (int s = 0; s < 2; ++s) {
                const bool atSide = q[d] == (s ? b.hi[d] : b.lo[d]);
                const bool cross = atSide && (b.st[d][s] == SIDE_INTERNAL || b.st[d][s] == BC_PERIODIC);
                const double* __restrict__ src = (pass == 1 && cross) ? snap : phi;
                nb[d][s] = src[s ? c + st[d] : c - st[d]];
            }
        }
        const GsrbCoef qc = gsrb_load_coef<ND>(g, m, rhs, c);
        phi[c] = gsrb_update_vals<ND>(g, qc, i, j, k, phi[c], nb, false, true);
    }
}

void launch_gsrb_loose(Ctx& cx, const Geom& g, const Metric& m, double* phi, const double* snap, const double* rhs,
                       int pass, int stage) {
    const int kz = (g.n[2] >= 64) ? 8 : 1;
    const int half = (g.n[0] + 1) / 2;
    dim3 grid((half + BX - 1) / BX, (g.n[1] + BY - 1) / BY, (g.n[2] + kz - 1) / kz), block(BX, BY);
    if (g.nd == 3) k_gsrb_loose<3><<<grid, block, 0, cx.stream>>>(g, m, phi, snap, rhs, pass, stage, kz);
    else k_gsrb_loose<2><<<grid, block, 0, cx.stream>>>(g, m, phi, snap, rhs, pass, stage, kz);
    cx.launches++;
}

// full metric: out of place (cross terms read same-colour cells of the pre-pass state = the `extrap` snapshot,
// RelaxationMethod.cpp:411).  Every cell is written: updated colour gets the new value, the other a copy.
// A thread owns the x-adjacent PAIR (2t, 2t+1) of a row: exactly one of the two cells is of the pass's colour in any
// plane, so every lane computes (a thread per cell idles the lanes of the other colour) and the pair leaves as one
// 16-byte store.  Box-interior cells take the branch-free statement.
// SPLIT: cells that touch a side of their box are left to k_gsrb_full_rim (their old value is copied).  The general
// statement (box-relative extrapolated views, ExView) is an order of magnitude longer than the branch-free one, and in
// this kernel's mapping the two x-edge cells of a row put it into two of the row's eight warps with ONE active lane:
// ncu (profiles/r2_ncu_k_gsrb_full.txt) shows 15 of 32 lanes active on average and 2.6 warps per scheduler.
template <int ND, bool SPLIT>
__global__ void __launch_bounds__(BX* BY) k_gsrb_full(const Geom g, const Metric m, double* __restrict__ phiOut,
                                                       const double* __restrict__ phiIn,
                                                       const double* __restrict__ rhs, int pass, int kz) {
    const int t = blockIdx.x * BX + threadIdx.x;
    const int j = blockIdx.y * BY + threadIdx.y;
    const int ia = 2 * t;
    if (ia >= g.n[0] || j >= g.n[1]) return;
    const bool two = ia + 1 < g.n[0];
    const int k0 = blockIdx.z * kz;
    const int k1 = min(k0 + kz, g.n[2]);
    const bool innerY = box_inner_dir<ND>(g, j, 1);
    const bool inner0 = innerY && box_inner_dir<ND>(g, ia, 0), inner1 = innerY && two && box_inner_dir<ND>(g, ia + 1, 0);
    for (int k = k0; k < k1; ++k) {
        const long long c = g.idx(ia, j, k);
        // cell ia is of this pass's colour iff par == 0
        const int par = (g.g0[0] + ia + g.g0[1] + j + ((ND > 2) ? (g.g0[2] + k) : 0) + pass) & 1;
        const int i = ia + par;
        double va = phiIn[c], vb = two ? phiIn[c + 1] : 0.0;
        if (i < g.n[0]) {
            const bool inner = (par ? inner1 : inner0) && (ND < 3 || box_inner_dir<ND>(g, k, 2));
            if (inner || !SPLIT) {
                const double v = inner ? gsrb_cell_full_inner<ND>(g, m, phiIn, rhs, i, j, k, c + par)
                                       : gsrb_cell_full<ND>(g, m, phiIn, rhs, i, j, k, c + par);
                if (par) vb = v; else va = v;
            }
        }
        if (two && ((c & 1) == 0)) {
            *reinterpret_cast<double2*>(phiOut + c) = make_double2(va, vb);
        } else {
            phiOut[c] = va;
            if (two) phiOut[c + 1] = vb;
        }
    }
}

template <int ND>
__global__ void __launch_bounds__(128, RIM_MINB) k_gsrb_full_rim(const Geom g, const Metric m, double* __restrict__ phiOut,
                                                        const double* __restrict__ phiIn, const double* __restrict__ rhs,
                                                        int pass, int nzr, int nyr, int nxr) {
    int i, j, k;
    if (!rim_cell<ND>(g, nzr, nyr, nxr, i, j, k, pass)) return;
    const long long c = g.idx(i, j, k);
    // (the box-interior test costs k_gsrb_full<ND, true> 80 registers, the CELL_FAST test 100 and 6 % of its rate: the
    // CELL_FAST cells on box-box interfaces stay here)
    phiOut[c] = gsrb_cell_full_kind<ND>(g, m, phiIn, rhs, i, j, k, c, cell_kind<ND>(g, i, j, k));
}

void launch_gsrb_full(Ctx& cx, const Geom& g, const Metric& m, double* phiOut, const double* phiIn,
                      const double* rhs, int pass) {
    const int kz = (g.n[2] >= 64) ? 8 : 1;
    const int half = (g.n[0] + 1) / 2;
    dim3 grid((half + BX - 1) / BX, (g.n[1] + BY - 1) / BY, (g.n[2] + kz - 1) / kz), block(BX, BY);
    if (!rim_split_usable(g)) {
        if (g.nd == 3) k_gsrb_full<3, false><<<grid, block, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, pass, kz);
        else k_gsrb_full<2, false><<<grid, block, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, pass, kz);
        cx.launches++;
        return;
    }
    if (g.nd == 3) k_gsrb_full<3, true><<<grid, block, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, pass, kz);
    else k_gsrb_full<2, true><<<grid, block, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, pass, kz);
    const RimGrid r = rim_grid(g, true);
    if (g.nd == 3) k_gsrb_full_rim<3><<<r.grid, 128, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, pass, r.nzr, r.nyr, r.nxr);
    else k_gsrb_full_rim<2><<<r.grid, 128, 0, cx.stream>>>(g, m, phiOut, phiIn, rhs, pass, r.nzr, r.nyr, r.nxr);
    cx.launches += 2;
}

// ---------------------------------------------------------------------------------------------------------
// Jacobi / diagonal preconditioner / lapDiag
// ---------------------------------------------------------------------------------------------------------
__global__ void k_pointwise_diag(const Geom g, const Metric m, double* __restrict__ phi,
                                 const double* __restrict__ src, int kind) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= g.n[0] || j >= g.n[1]) return;
    const long long c = g.idx(i, j, k);
    const double den = g.alpha + g.beta * m.lapdiag[c];
    if (kind == 0) phi[c] = phi[c] + 0.5 * src[c] / den;   // JACOBIITER
    else phi[c] = src[c] / den;                            // DIAGPRECOND
}

void launch_jacobi(Ctx& cx, const Geom& g, const Metric& m, double* phi, const double* res) {
    dim3 grid((g.n[0] + BX - 1) / BX, (g.n[1] + BY - 1) / BY, g.n[2]), block(BX, BY);
    k_pointwise_diag<<<grid, block, 0, cx.stream>>>(g, m, phi, res, 0);
    cx.launches++;
}
void launch_diag_precond(Ctx& cx, const Geom& g, const Metric& m, double* phi, const double* rhs) {
    dim3 grid((g.n[0] + BX - 1) / BX, (g.n[1] + BY - 1) / BY, g.n[2]), block(BX, BY);
    k_pointwise_diag<<<grid, block, 0, cx.stream>>>(g, m, phi, rhs, 1);
    cx.launches++;
}

// FILLMAPPEDLAPDIAG{2,3}D -- MappedAMRPoissonOpF.ChF:149-274
__global__ void k_lapdiag(const Geom g, const Metric m, double* __restrict__ lap) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = blockIdx.y * blockDim.y + threadIdx.y;
    const int k = blockIdx.z;
    if (i >= g.n[0] || j >= g.n[1]) return;
    const long long c = g.idx(i, j, k);
    const double s0 = 1.0 / (g.dx[0] * g.dx[0]), s1 = 1.0 / (g.dx[1] * g.dx[1]);
    double v;
    if (g.nd == 2) {
        v = -m.jinv[c] * ((m.jg[0][0][c + 1] + m.jg[0][0][c]) * s0 + (m.jg[1][1][c + g.sy] + m.jg[1][1][c]) * s1);
    } else {
        const double s2 = 1.0 / (g.dx[2] * g.dx[2]);
        v = -m.jinv[c] * ((m.jg[0][0][c + 1] + m.jg[0][0][c]) * s0 + (m.jg[1][1][c + g.sy] + m.jg[1][1][c]) * s1 +
                          (m.jg[2][2][c + g.sz] + m.jg[2][2][c]) * s2);
    }
    lap[c] = v;
}

void launch_lapdiag(Ctx& cx, const Geom& g, const Metric& m, double* lapdiag) {
    dim3 grid((g.n[0] + BX - 1) / BX, (g.n[1] + BY - 1) / BY, g.n[2]), block(BX, BY);
    k_lapdiag<<<grid, block, 0, cx.stream>>>(g, m, lapdiag);
    cx.launches++;
}

// ---------------------------------------------------------------------------------------------------------
// fused restrictResidual: crse = sum_block((rhs - L phi)/Jinv) / sum_block(1/Jinv), block swept x fastest
// (MappedAMRPoissonOp.cpp:1281-1304; FullWeightingPS::restrict, RestrictionStrategy.cpp:36-99; MAPPEDAVERAGE2,
// MappedCoarseAverageF.ChF:132-170).  One thread per coarse cell; the fine residual never touches HBM.
// ---------------------------------------------------------------------------------------------------------
template <int ND, bool FULL>
__global__ void __launch_bounds__(BX* BY) k_restrict(const Geom gf, const Metric mf, const Geom gc,
                                                      double* __restrict__ crse, const double* __restrict__ phi,
                                                      const double* __restrict__ rhs, int r0, int r1, int r2) {
    const int ic = blockIdx.x * BX + threadIdx.x;
    const int jc = blockIdx.y * BY + threadIdx.y;
    const int kc = blockIdx.z;
    if (ic >= gc.n[0] || jc >= gc.n[1]) return;
    double s = 0.0, sj = 0.0;
    for (int kk = 0; kk < r2; ++kk)
        for (int jj = 0; jj < r1; ++jj)
            for (int ii = 0; ii < r0; ++ii) {
                const int i = ic * r0 + ii, j = jc * r1 + jj, k = kc * r2 + kk;
                const long long c = gf.idx(i, j, k);
                double v;
                if (FULL) {
                    const bool inner = box_inner_dir<ND>(gf, i, 0) && box_inner_dir<ND>(gf, j, 1) &&
                                       (ND < 3 || box_inner_dir<ND>(gf, k, 2));
                    v = inner ? apply_cell_full_inner<ND>(gf, mf, phi, i, j, k, c, true)
                              : apply_cell_full<ND>(gf, mf, phi, i, j, k, c, true);
                } else {
                    v = apply_cell_ortho<ND>(gf, mf, phi, i, j, k, c, true);
                }
                const double res = rhs[c] - v;
                const double ji = mf.jinv[c];
                s = s + res / ji;
                sj = sj + 1.0 / ji;
            }
    crse[gc.idx(ic, jc, kc)] = s / sj;
}

// the averaging half alone, of a residual already in HBM (same sums, same order)
__global__ void __launch_bounds__(BX* BY) k_restrict_avg(const Geom gf, const Metric mf, const Geom gc,
                                                          double* __restrict__ crse, const double* __restrict__ res,
                                                          int r0, int r1, int r2) {
    const int ic = blockIdx.x * BX + threadIdx.x;
    const int jc = blockIdx.y * BY + threadIdx.y;
    const int kc = blockIdx.z;
    if (ic >= gc.n[0] || jc >= gc.n[1]) return;
    double s = 0.0, sj = 0.0;
    for (int kk = 0; kk < r2; ++kk)
        for (int jj = 0; jj < r1; ++jj)
            for (int ii = 0; ii < r0; ++ii) {
                const long long c = gf.idx(ic * r0 + ii, jc * r1 + jj, kc * r2 + kk);
                const double ji = mf.jinv[c];
                s = s + res[c] / ji;
                sj = sj + 1.0 / ji;
            }
    crse[gc.idx(ic, jc, kc)] = s / sj;
}

void launch_restrict_residual(Ctx& cx, const Geom& gf, const Metric& mf, const Geom& gc, double* crse,
                              const double* phi, const double* rhs, const int r[3], double* scratch) {
    if (!mf.diagonal && scratch && rim_split_usable(gf)) {
        // full metric: the residual through the inner / rim pair of launch_apply (the general cells densely packed), then
        // the average; 16 B per fine cell more traffic, several times less time than one thread per coarse cell running
        // the general statement for the rim cells of its block
        launch_apply(cx, gf, mf, scratch, phi, rhs, 1, true);
        dim3 grid((gc.n[0] + BX - 1) / BX, (gc.n[1] + BY - 1) / BY, gc.n[2]), block(BX, BY);
        k_restrict_avg<<<grid, block, 0, cx.stream>>>(gf, mf, gc, crse, scratch, r[0], r[1], r[2]);
        cx.launches++;
        return;
    }
    if (mf.diagonal && gf.nd == 3 && r[0] == 1 && r[1] == 1 && r[2] == 2 && (gf.n[2] & 1) == 0) {
        // vertical semicoarsening (the anisotropic case): z-marching kernel
        const int kzm = march_kz(gf);
        dim3 gridm((gf.n[0] + BX - 1) / BX, (gf.n[1] + BY - 1) / BY, (gf.n[2] + kzm - 1) / kzm), blockm(BX, BY);
        k_march<2><<<gridm, blockm, 0, cx.stream>>>(gf, mf, gc, crse, phi, rhs, true, kzm);
        cx.launches++;
        return;
    }
    dim3 grid((gc.n[0] + BX - 1) / BX, (gc.n[1] + BY - 1) / BY, gc.n[2]), block(BX, BY);
    if (mf.diagonal) {
        if (gf.nd == 3) k_restrict<3, false><<<grid, block, 0, cx.stream>>>(gf, mf, gc, crse, phi, rhs, r[0], r[1], r[2]);
        else k_restrict<2, false><<<grid, block, 0, cx.stream>>>(gf, mf, gc, crse, phi, rhs, r[0], r[1], r[2]);
    } else {
        if (gf.nd == 3) k_restrict<3, true><<<grid, block, 0, cx.stream>>>(gf, mf, gc, crse, phi, rhs, r[0], r[1], r[2]);
        else k_restrict<2, true><<<grid, block, 0, cx.stream>>>(gf, mf, gc, crse, phi, rhs, r[0], r[1], r[2]);
    }
    cx.launches++;
}

}  // namespace somar
