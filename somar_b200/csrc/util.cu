// util.cu -- BLAS-1 over valid cells, reductions (warp-shuffle + two-stage, deterministic), prolongation,
// metric coarsening, halo packing.  All streaming kernels; see DESIGN.md for the byte model of each.
#include "kernels.hpp"
#include "cells.cuh"
#include <cstdio>

namespace somar {

// whole-brick fill / copy: 16-byte accesses, two per thread and step (the brick size is a multiple of 16 doubles)
__global__ void k_fill_all(double* __restrict__ x, long long n, double v) {
    double2* __restrict__ x2 = reinterpret_cast<double2*>(x);
    const long long n2 = n >> 1, stride = (long long)gridDim.x * blockDim.x;
    const double2 vv = make_double2(v, v);
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n2; t += 2 * stride) {
        x2[t] = vv;
        if (t + stride < n2) x2[t + stride] = vv;
    }
}
__global__ void k_copy_all(double* __restrict__ y, const double* __restrict__ x, long long n) {
    double2* __restrict__ y2 = reinterpret_cast<double2*>(y);
    const double2* __restrict__ x2 = reinterpret_cast<const double2*>(x);
    const long long n2 = n >> 1, stride = (long long)gridDim.x * blockDim.x;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n2; t += 2 * stride) {
        const double2 a = x2[t];
        const bool more = t + stride < n2;
        double2 b = a;
        if (more) b = x2[t + stride];
        y2[t] = a;
        if (more) y2[t + stride] = b;
    }
}
static inline int flat_blocks(long long n, int sms) {
    long long nb = (n + TPB - 1) / TPB;
    const long long cap = (long long)sms * 16;
    return (int)(nb < 1 ? 1 : (nb > cap ? cap : nb));
}
void launch_fill(Ctx& cx, const Geom& g, double* x, double v) {
    k_fill_all<<<flat_blocks(g.total, cx.sms), TPB, 0, cx.stream>>>(x, g.total, v);
    cx.launches++;
}
void launch_copy(Ctx& cx, const Geom& g, double* y, const double* x) {
    k_copy_all<<<flat_blocks(g.total, cx.sms), TPB, 0, cx.stream>>>(y, x, g.total);
    cx.launches++;
}

__global__ void k_axpy(const Geom g, double* __restrict__ y, const double* __restrict__ x, double a) {
    double xv[4], yv[4];
    FOR_VALID_CELLS4(g, c, u, { xv[u] = x[c]; yv[u] = y[c]; }, { y[c] = yv[u] + a * xv[u]; })
}
__global__ void k_axby(const Geom g, double* __restrict__ out, const double* __restrict__ x,
                       const double* __restrict__ y, double a, double b) {
    double xv[4], yv[4];
    FOR_VALID_CELLS4(g, c, u, { xv[u] = x[c]; yv[u] = y[c]; }, { out[c] = a * xv[u] + b * yv[u]; })
}
__global__ void k_scale(const Geom g, double* __restrict__ x, double a) {
    double xv[4];
    FOR_VALID_CELLS4(g, c, u, { xv[u] = x[c]; }, { x[c] = xv[u] * a; })
}
void launch_axpy(Ctx& cx, const Geom& g, double* y, const double* x, double a) {
    k_axpy<<<row_blocks(g, cx.sms), TPB, 0, cx.stream>>>(g, y, x, a);
    cx.launches++;
}
void launch_axby(Ctx& cx, const Geom& g, double* out, const double* x, const double* y, double a, double b) {
    k_axby<<<row_blocks(g, cx.sms), TPB, 0, cx.stream>>>(g, out, x, y, a, b);
    cx.launches++;
}
void launch_scale(Ctx& cx, const Geom& g, double* x, double a) {
    k_scale<<<row_blocks(g, cx.sms), TPB, 0, cx.stream>>>(g, x, a);
    cx.launches++;
}


__global__ void k_reduce(const Geom g, const double* __restrict__ x, const double* __restrict__ y, int kind,
                         double* __restrict__ scratch, unsigned int* __restrict__ tickets,
                         double* __restrict__ out) {
    const bool isMax = (kind == 0 || kind == 4);
    const double ident = (kind == 4) ? -1e300 : 0.0;
    double acc = ident;
    double xv[4], yv[4] = {0.0, 0.0, 0.0, 0.0};
    FOR_VALID_CELLS4(g, c, u, { xv[u] = x[c]; if (kind == 3) yv[u] = y[c]; }, {
        const double v = xv[u];
        if (kind == 0) acc = fmax(acc, fabs(v));
        else if (kind == 1) acc += fabs(v);
        else if (kind == 2) acc += v * v;
        else if (kind == 3) acc += v * yv[u];
        else if (kind == 5) acc += v;                    // plain sum (LevelLepticSolver::setZeroAvg)
        else acc = fmax(acc, v);
    })
    acc = block_red(acc, isMax, ident);
    finish_reduce(acc, isMax, ident, scratch, out, tickets, 0, false);
}

void launch_reduce(Ctx& cx, const Geom& g, const double* x, const double* y, int kind, double* out) {
    k_reduce<<<row_blocks(g, cx.sms), TPB, 0, cx.stream>>>(g, x, y, kind, cx.scratch, cx.tickets, out);
    cx.launches++;
}

// ---- prolongation -----------------------------------------------------------------------------------------
// ConstInterpPS / ConstInterpWithAvgPS -- ProlongationStrategyF.ChF:37-160:  fine += coarse[i/m];
// optionally sums[0] += sum(dV*fine), sums[1] += sum(dV), dV = dxProduct / Jinv.
// SH: every refinement ratio is 1 or 2 (the semicoarsening schedule produces nothing else), so i / r is a shift
template <bool SH>
__global__ void k_prolong(const Geom gf, const double* __restrict__ jinv, const Geom gc, double* __restrict__ fine,
                          const double* __restrict__ crse, int r0, int r1, int r2, double dxProduct, bool withAvg,
                          double* __restrict__ scratch, unsigned int* __restrict__ tickets,
                          double* __restrict__ sums) {
    // Without the average: fine += coarse.  With it (zero-average strategy): this pass only FORMS the sums of
    // v = fine + coarse (nothing is written); k_prolong_finish then stores v - sums[0]/sums[1].  The values are the
    // ones the reference gets from "add, reduce, subtract"; the fine field is read twice but written once
    // (40 B/cell instead of 52).
    double aS = 0.0, aV = 0.0;
    double fv[4], cv[4], jv[4] = {1.0, 1.0, 1.0, 1.0};
    FOR_VALID_CELLS4(gf, c, u,
                     { fv[u] = fine[c]; cv[u] = crse[SH ? gc.idx(i >> (r0 - 1), j >> (r1 - 1), k >> (r2 - 1)) : gc.idx(i / r0, j / r1, k / r2)]; if (withAvg) jv[u] = jinv[c]; },
                     {
                         const double v = fv[u] + cv[u];
                         if (withAvg) {
                             const double dvol = dxProduct / jv[u];
                             aS = aS + dvol * v;
                             aV = aV + dvol;
                         } else {
                             fine[c] = v;
                         }
                     })
    if (withAvg) {
        aS = block_red(aS, false, 0.0);
        aV = block_red(aV, false, 0.0);
        // two independent tickets / scratch halves
        finish_reduce(aS, false, 0.0, scratch, sums, tickets, 1, false);
        finish_reduce(aV, false, 0.0, scratch + MAX_RED_BLOCKS, sums + 1, tickets, 2, false);
    }
}

void launch_prolong(Ctx& cx, const Geom& gf, const Metric& mf, const Geom& gc, double* fine, const double* crse,
                    const int r[3], double dxProduct, double* sums) {
    const bool sh = r[0] <= 2 && r[1] <= 2 && r[2] <= 2;
    if (sh) k_prolong<true><<<row_blocks(gf, cx.sms), TPB, 0, cx.stream>>>(gf, mf.jinv, gc, fine, crse, r[0], r[1], r[2], dxProduct,
                                                                           sums != nullptr, cx.scratch, cx.tickets, sums);
    else k_prolong<false><<<row_blocks(gf, cx.sms), TPB, 0, cx.stream>>>(gf, mf.jinv, gc, fine, crse, r[0], r[1], r[2], dxProduct,
                                                                         sums != nullptr, cx.scratch, cx.tickets, sums);
    cx.launches++;
}

// fine = (fine + coarse[i/m]) - sums[0]/sums[1]     (ZeroAvgConstInterpPS::prolongIncrement, ProlongationStrategy.cpp:163)
template <bool SH>
__global__ void k_prolong_finish(const Geom gf, const Geom gc, double* __restrict__ fine, const double* __restrict__ crse,
                                 int r0, int r1, int r2, const double* __restrict__ sums) {
    const double avg = sums[0] / sums[1];
    double fv[4], cv[4];
    FOR_VALID_CELLS4(gf, c, u, { fv[u] = fine[c]; cv[u] = crse[SH ? gc.idx(i >> (r0 - 1), j >> (r1 - 1), k >> (r2 - 1)) : gc.idx(i / r0, j / r1, k / r2)]; },
                     { fine[c] = (fv[u] + cv[u]) - avg; })
}
void launch_prolong_finish(Ctx& cx, const Geom& gf, const Geom& gc, double* fine, const double* crse, const int r[3],
                           const double* sums) {
    if (r[0] <= 2 && r[1] <= 2 && r[2] <= 2) k_prolong_finish<true><<<row_blocks(gf, cx.sms), TPB, 0, cx.stream>>>(gf, gc, fine, crse, r[0], r[1], r[2], sums);
    else k_prolong_finish<false><<<row_blocks(gf, cx.sms), TPB, 0, cx.stream>>>(gf, gc, fine, crse, r[0], r[1], r[2], sums);
    cx.launches++;
}

// ---- metric coarsening (fill_MGfields, MappedAMRPoissonOpFactory.cpp:1164-1180) ----------------------------
// UNMAPPEDAVERAGEFACE -- MappedCoarseAverageF.ChF:182-225 (faces of direction dir; no offset along dir)
__global__ void k_avg_face(const Geom gf, const Geom gc, double* __restrict__ crse, const double* __restrict__ fine,
                           int dir, int r0, int r1, int r2) {
    const int ic = blockIdx.x * blockDim.x + threadIdx.x;
    const int jc = blockIdx.y * blockDim.y + threadIdx.y;
    const int kc = blockIdx.z;
    const int e0 = (dir == 0), e1 = (dir == 1), e2 = (dir == 2);
    if (ic >= gc.n[0] + e0 || jc >= gc.n[1] + e1 || kc >= gc.n[2] + e2) return;
    const int rr0 = e0 ? 1 : r0, rr1 = e1 ? 1 : r1, rr2 = e2 ? 1 : r2;
    const int rd = e0 ? r0 : (e1 ? r1 : r2);
    const double refScale = double(rd) / double(r0 * r1 * r2);
    double s = 0.0;
    for (int kk = 0; kk < rr2; ++kk)
        for (int jj = 0; jj < rr1; ++jj)
            for (int ii = 0; ii < rr0; ++ii) s = s + fine[gf.idx(ic * r0 + ii, jc * r1 + jj, kc * r2 + kk)];
    crse[gc.idx(ic, jc, kc)] = refScale * s;
}
void launch_avg_face(Ctx& cx, const Geom& gf, const Geom& gc, double* crse, const double* fine, int dir,
                     const int r[3]) {
    dim3 block(64, 4), grid((gc.n[0] + 1 + 63) / 64, (gc.n[1] + 1 + 3) / 4, gc.n[2] + 1);
    k_avg_face<<<grid, block, 0, cx.stream>>>(gf, gc, crse, fine, dir, r[0], r[1], r[2]);
    cx.launches++;
}
// UNMAPPEDAVERAGEHARMONIC -- MappedCoarseAverageF.ChF:48-86
__global__ void k_avg_harm(const Geom gf, const Geom gc, double* __restrict__ crse, const double* __restrict__ fine,
                           int r0, int r1, int r2) {
    const int ic = blockIdx.x * blockDim.x + threadIdx.x;
    const int jc = blockIdx.y * blockDim.y + threadIdx.y;
    const int kc = blockIdx.z;
    if (ic >= gc.n[0] || jc >= gc.n[1]) return;
    const double refScale = 1.0 / double(r0 * r1 * r2);
    double s = 0.0;
    for (int kk = 0; kk < r2; ++kk)
        for (int jj = 0; jj < r1; ++jj)
            for (int ii = 0; ii < r0; ++ii) s = s + 1.0 / fine[gf.idx(ic * r0 + ii, jc * r1 + jj, kc * r2 + kk)];
    crse[gc.idx(ic, jc, kc)] = 1.0 / (s * refScale);
}
void launch_avg_harmonic(Ctx& cx, const Geom& gf, const Geom& gc, double* crse, const double* fine, const int r[3]) {
    dim3 block(64, 4), grid((gc.n[0] + 63) / 64, (gc.n[1] + 3) / 4, gc.n[2]);
    k_avg_harm<<<grid, block, 0, cx.stream>>>(gf, gc, crse, fine, r[0], r[1], r[2]);
    cx.launches++;
}

// ---- halos --------------------------------------------------------------------------------------------------
// Periodic images inside the brick for direction dir: ghost layer l (1..w) on the low side <- cell n-l, on the
// high side <- cell l-1.  Sweeps the full allocated extent of the other directions so that successive calls
// (y, then x, then z) also fill edges and corners, like LevelData::exchange with a corner copier.
__global__ void k_periodic_fill(const Geom g, double* __restrict__ x, int dir, int w) {
    // plane coordinates (u, v) = the two directions other than dir, over the allocated extent
    const int d1 = (dir == 0) ? 1 : 0, d2 = (dir == 2) ? 1 : 2;
    const int lo1 = -((d1 == 0) ? min(g.gh[0], SOMAR_GHOST) : g.gh[d1]);
    const int lo2 = -((d2 == 0) ? min(g.gh[0], SOMAR_GHOST) : g.gh[d2]);
    const int n1 = g.n[d1] - 2 * lo1, n2 = g.n[d2] - 2 * lo2;
    const long long tot = (long long)n1 * n2 * w * 2;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < tot; t += (long long)gridDim.x * blockDim.x) {
        const int u = (int)(t % n1) + lo1;
        long long r = t / n1;
        const int v = (int)(r % n2) + lo2;
        r /= n2;
        const int l = (int)(r % w) + 1;
        const int side = (int)(r / w);
        int p[3], q[3];
        p[d1] = q[d1] = u;
        p[d2] = q[d2] = v;
        if (side == 0) { p[dir] = -l; q[dir] = g.n[dir] - l; }
        else { p[dir] = g.n[dir] + l - 1; q[dir] = l - 1; }
        x[g.idx(p[0], p[1], p[2])] = x[g.idx(q[0], q[1], q[2])];
    }
}
void launch_periodic_fill(Ctx& cx, const Geom& g, double* x, int dir, int w) {
    const long long tot = (long long)(g.n[0] + 8) * (g.n[1] + 4) * (g.n[2] + 4) / (g.n[dir] + 4) * w * 2;
    k_periodic_fill<<<flat_blocks(tot, cx.sms), TPB, 0, cx.stream>>>(g, x, dir, w);
    cx.launches++;
}

// Pack the w outermost valid rows next to the low/high y side (or unpack into the ghost rows of that side).
// Buffer layout: [l][k over allocated z][i over -SOMAR_GHOST .. n0+SOMAR_GHOST).
__global__ void k_pack_y(const Geom g, double* __restrict__ x, double* __restrict__ buf, int side, int w,
                         bool unpack) {
    const int nxx = g.n[0] + 2 * SOMAR_GHOST;
    const int nzz = g.n[2] + 2 * g.gh[2];
    const long long tot = (long long)nxx * nzz * w;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < tot; t += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(t % nxx) - SOMAR_GHOST;
        long long r = t / nxx;
        const int k = (int)(r % nzz) - g.gh[2];
        const int l = (int)(r / nzz);
        int j;
        if (!unpack) j = (side == 0) ? l : g.n[1] - w + l;          // valid rows, ascending y
        else j = (side == 0) ? -w + l : g.n[1] + l;                  // ghost rows, ascending y
        if (unpack) x[g.idx(i, j, k)] = buf[t];
        else buf[t] = x[g.idx(i, j, k)];
    }
}
void launch_pack_y(Ctx& cx, const Geom& g, const double* x, double* buf, int side, int w, bool unpack) {
    const long long tot = (long long)(g.n[0] + 2 * SOMAR_GHOST) * (g.n[2] + 2 * g.gh[2]) * w;
    k_pack_y<<<flat_blocks(tot, cx.sms), TPB, 0, cx.stream>>>(g, const_cast<double*>(x), buf, side, w, unpack);
    cx.launches++;
}

// brick <-> dense rank-local array of (n0+ext0, n1+ext1, n2+ext2) points, x fastest.
__global__ void k_dense_copy(const Geom g, double* __restrict__ brick, double* __restrict__ dense, int e0, int e1,
                             int e2, bool toBrick) {
    const int m0 = g.n[0] + e0, m1 = g.n[1] + e1, m2 = g.n[2] + e2;
    const long long tot = (long long)m0 * m1 * m2;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < tot; t += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(t % m0);
        const long long r = t / m0;
        const int j = (int)(r % m1);
        const int k = (int)(r / m1);
        if (toBrick) brick[g.idx(i, j, k)] = dense[t];
        else dense[t] = brick[g.idx(i, j, k)];
    }
}
void launch_dense_copy(Ctx& cx, const Geom& g, double* brick, double* dense, const int ext[3], bool toBrick) {
    const long long tot = (long long)(g.n[0] + ext[0]) * (g.n[1] + ext[1]) * (g.n[2] + ext[2]);
    k_dense_copy<<<flat_blocks(tot, cx.sms), TPB, 0, cx.stream>>>(g, brick, dense, ext[0], ext[1], ext[2], toBrick);
    cx.launches++;
}

}  // namespace somar
