// stencil_full.cuh -- per-cell device functions for the FULL-metric operator (all J g^{ij} non-zero:
// terrain-following grids).  Normal derivatives read phi; cross derivatives read `extrap`, which in the
// reference is a copy of phi whose ghost cells were extrapolated (RelaxationMethod::fillGhostsAndExtrapolate,
// RelaxationMethod.cpp:376-435; MappedAMRPoissonOp::fillExtrap, MappedAMRPoissonOp.cpp:2244-2270) and then
// re-extrapolated at Neumann sides by the BC ghost method (setSideNeumBC -> ExtrapolateFaceAndCopy order 2,
// EllipticBCUtils.cpp:130-215; ExtrapolationUtils.cpp:109-163).  Here `extrap` is never materialised: ex_at()
// evaluates the value the reference would hold, RELATIVE TO THE BOX of the cell being updated:
//   - inside the box ........................ phi
//   - outside through sides S of the box:
//       * if a side in S is a Neumann domain side, or the point sits in the first layer next to a Neumann
//         domain side of the box in a direction not in S (the "nearBox" sweep of ExtrapolateFaceAndCopy),
//         every direction of S is extrapolated with order 2 from the box's own data -- even through
//         box-box interfaces and periodic sides (that is what the reference does);
//       * otherwise box-box interface -> neighbour's phi (halo); Dirichlet -> order 2 (operator) / 1 (smoother);
//         periodic -> wrapped phi (operator; validDomain is grown, MappedAMRPoissonOp.cpp:278-283) or order-1
//         extrapolation (smoother; RelaxationMethod.cpp:413 does not grow the domain).
// Extrapolations are composed lowest direction innermost.  The reference's last write at a few edge cells
// composes them in another order: mathematically identical, rounding differs (<= 1e-15 relative), so
// full-metric parity is tested to 1e-13 (BASELINE.json) while the diagonal-metric kernels are bit-exact.
// Formulas: EXTRAPOLATEFACENOEV ExtrapolationUtilsF.ChF:35-140; MAPPEDGETFLUX MappedAMRPoissonOpF.ChF:335-427;
// GSRBITER{2,3}D GSRBF.ChF:155-440; GSRBBOUNDARYITER{2,3}D GSRBF.ChF:858-1232.
#pragma once
#include "geom.cuh"
#include "stencil_ortho.cuh"

namespace somar {

// every function of this header also compiles for the host (SOMAR_HD, stencil_ortho.cuh): tests/cuda/full_accessors_host.cu
// runs the accessors and the statements on the CPU (DirView / TabView / FastView against ExView, cell by cell, bit for bit)

enum { EXMODE_SMOOTHER = 0, EXMODE_OP = 1 };
enum { SIDE_INTERNAL = -3 };

// The box (of the level's uniform tiling) that holds local cell q, in local indices.
struct BoxOf {
    int lo[3], hi[3];
    int st[3][2];   // side type: BC_NEUM / BC_DIRI / BC_PERIODIC / SIDE_INTERNAL
    bool rim;       // q touches a side of its box
};

template <int ND>
SOMAR_HD __forceinline__ BoxOf box_of(const Geom& g, int i, int j, int k) {
    BoxOf b;
    const int q[3] = {i, j, k};
    b.rim = false;
SOMAR_UNROLL
    for (int d = 0; d < 3; ++d) {
        if (d < ND) {
            const int gq = q[d] + g.g0[d];
            const int glo = (gq / g.bs[d]) * g.bs[d];
            const int ghi = glo + g.bs[d] - 1;
            b.lo[d] = glo - g.g0[d];
            b.hi[d] = ghi - g.g0[d];
            b.st[d][0] = (glo == 0) ? g.bct[d][0] : SIDE_INTERNAL;
            b.st[d][1] = (ghi == g.gn[d] - 1) ? g.bct[d][1] : SIDE_INTERNAL;
            b.rim = b.rim | (q[d] == b.lo[d]) | (q[d] == b.hi[d]);
        } else {
            b.lo[d] = b.hi[d] = q[d];
            b.st[d][0] = b.st[d][1] = SIDE_INTERNAL;
        }
    }
    return b;
}

template <int D>
struct ExRec {
    SOMAR_HD static double eval(const Geom& g, const double* __restrict__ phi, int p0, int p1, int p2,
                                  const int* o, const int* ord) {
        if (ord[D] == 0) return ExRec<D - 1>::eval(g, phi, p0, p1, p2, o, ord);
        int p[3] = {p0, p1, p2};
        const int s = o[D];
        p[D] -= s;
        const double s1 = ExRec<D - 1>::eval(g, phi, p[0], p[1], p[2], o, ord);
        p[D] -= s;
        const double s2 = ExRec<D - 1>::eval(g, phi, p[0], p[1], p[2], o, ord);
        if (ord[D] == 1) return 2.0 * s1 - s2;
        p[D] -= s;
        const double s3 = ExRec<D - 1>::eval(g, phi, p[0], p[1], p[2], o, ord);
        return 3.0 * (s1 - s2) + s3;
    }
};
template <>
struct ExRec<-1> {
    SOMAR_HD static double eval(const Geom& g, const double* __restrict__ phi, int p0, int p1, int p2, const int*,
                                  const int*) {
        return phi[g.idx(p0, p1, p2)];
    }
};

// extrap(p) as seen from box b.
template <int ND>
SOMAR_HD __noinline__ double ex_slow(const Geom& g, const double* __restrict__ phi, const BoxOf& b, int p0, int p1,
                                       int p2, int mode) {
    const int p[3] = {p0, p1, p2};
    int o[3] = {0, 0, 0}, ord[3] = {0, 0, 0}, t[3] = {0, 0, 0};
    bool force = false, any = false;
SOMAR_UNROLL
    for (int d = 0; d < ND; ++d) {
        if (p[d] < b.lo[d]) { o[d] = -1; t[d] = b.st[d][0]; }
        else if (p[d] > b.hi[d]) { o[d] = 1; t[d] = b.st[d][1]; }
        if (o[d] != 0) {
            any = true;
            if (t[d] == BC_NEUM) force = true;
        } else {
            if ((p[d] == b.lo[d] && b.st[d][0] == BC_NEUM) || (p[d] == b.hi[d] && b.st[d][1] == BC_NEUM)) force = true;
        }
    }
    if (!any) return phi[g.idx(p0, p1, p2)];
SOMAR_UNROLL
    for (int d = 0; d < ND; ++d) {
        if (o[d] == 0) continue;
        if (force) ord[d] = 2;
        else if (t[d] == SIDE_INTERNAL) ord[d] = 0;
        else if (t[d] == BC_PERIODIC) ord[d] = (mode == EXMODE_OP) ? 0 : 1;
        else ord[d] = (mode == EXMODE_OP) ? 2 : 1;
    }
    return ExRec<ND - 1>::eval(g, phi, p0, p1, p2, o, ord);
}

// Accessor: phi for normal terms, ex for cross terms, relative to the box of the cell being processed.
template <int ND>
struct ExView {
    const Geom& g;
    const double* __restrict__ phi;
    BoxOf b;
    int mode;
    SOMAR_HD ExView(const Geom& g_, const double* __restrict__ phi_, int i, int j, int k, int mode_)
        : g(g_), phi(phi_), b(box_of<ND>(g_, i, j, k)), mode(mode_) {}
    SOMAR_HD __forceinline__ double ex(int i, int j, int k) const {
        if (!b.rim) return phi[g.idx(i, j, k)];
        return ex_slow<ND>(g, phi, b, i, j, k, mode);
    }
};

// The same accessor for a cell that is INTERIOR TO ITS BOX (no side of the box, hence no side of the domain, within
// one cell): every ex() is phi itself, addressed by a compile-time offset from the cell -- no box search, no branches.
// The expressions instantiated with it are the ones ExView takes when b.rim is false, so the bits are the same.
template <int ND>
struct FastView {
    const double* __restrict__ p;   // phi + index of the cell (i0, j0, k0)
    int i0, j0, k0;
    long long sy, sz;
    SOMAR_HD __forceinline__ double ex(int i, int j, int k) const {
        return p[(i - i0) + sy * (j - j0) + sz * (k - k0)];
    }
};

// Is local cell (q0, q1, q2) strictly inside its box of the level's uniform tiling (not touching any box side)?
// x/y part and z part separately, so that a z-marching thread tests x/y once.
template <int ND>
SOMAR_HD __forceinline__ bool box_inner_dir(const Geom& g, int q, int d) {
    const int r = (q + g.g0[d]) % g.bs[d];
    return r != 0 && r != g.bs[d] - 1;
}

// ---------------------------------------------------------------------------------------------------------
// Which accessor a cell needs.  ex_slow's rules, read per direction d of the cell q being updated (every access is within one
// cell of q):
//   class 0  q[d] touches no side of its box and is at least two cells from a Neumann DOMAIN side
//   class 1  touches no side, but sits next to the first layer of (one cell from) a Neumann domain side
//   class 2  touches only box-box interfaces, and is not within one cell of a Neumann domain side
//   class 3  touches a domain side (any type), or touches an interface within one cell of a Neumann domain side
// An access leaves the box only in directions of class 2 / 3; `force` (order-2 extrapolation of every direction left) needs
// a Neumann side crossed (class 3) or a point in the first layer next to one while another direction is left (class 1 or 3
// together with a class 2 / 3).  Hence:
//   CELL_FAST     no class 3, and not (a class 2 together with a class 1): every ex() is phi (interfaces read the halo);
//   CELL_DIR0..2  exactly one direction of class 3, the others class 0: only that direction can be left, and `force` is
//                 "the side crossed is Neumann" -- DirView below;
//   CELL_GENERAL  the rest (edges and corners of the boxes next to the domain boundary): ExView.
// ---------------------------------------------------------------------------------------------------------
enum { CELL_FAST = 0, CELL_DIR0 = 1, CELL_DIR1 = 2, CELL_DIR2 = 3, CELL_GENERAL = 4 };

SOMAR_HD __forceinline__ int cell_dir_class(const Geom& g, int q, int d) {
    const int gq = q + g.g0[d];
    const int r = gq % g.bs[d];
    const bool touch = (r == 0) || (r == g.bs[d] - 1);
    const bool nearN = (gq <= 1 && g.bct[d][0] == BC_NEUM) || (gq >= g.gn[d] - 2 && g.bct[d][1] == BC_NEUM);
    if (!touch) return nearN ? 1 : 0;
    const bool dom = (gq == 0) || (gq == g.gn[d] - 1);
    return (dom || nearN) ? 3 : 2;
}

// CELL_FAST from the classes as a bit mask (bit c-1 for a direction of class c): no class 3, and not a 2 with a 1.
// A z-marching thread keeps the mask of x and y.
SOMAR_HD __forceinline__ unsigned class_bit(int c) { return c ? 1u << (c - 1) : 0u; }
SOMAR_HD __forceinline__ bool classes_fast(unsigned mask) { return mask < 3u; }

template <int ND>
SOMAR_HD __forceinline__ int cell_kind(const Geom& g, int i, int j, int k) {
    const int c0 = cell_dir_class(g, i, 0), c1 = cell_dir_class(g, j, 1), c2 = (ND > 2) ? cell_dir_class(g, k, 2) : 0;
    if (classes_fast(class_bit(c0) | class_bit(c1) | class_bit(c2))) return CELL_FAST;
    if (c0 == 3 && c1 == 0 && c2 == 0) return CELL_DIR0;
    if (c1 == 3 && c0 == 0 && c2 == 0) return CELL_DIR1;
    if (ND > 2 && c2 == 3 && c0 == 0 && c1 == 0) return CELL_DIR2;
    return CELL_GENERAL;
}

// value of an order-`ord` extrapolation from the three values one, two and three steps inside (ExRec's expressions)
SOMAR_HD __forceinline__ double ex_comb(int ord, double s1, double s2, double s3) {
    const double r1 = 2.0 * s1 - s2, r2 = 3.0 * (s1 - s2) + s3;
    return (ord == 0) ? s1 : (ord == 1 ? r1 : r2);
}
// how many steps inside the K-th value (K = 0, 1, 2) of an order-`ord` extrapolation sits; a value the order does not use
// repeats the last one it does, so that no address is read that ex_slow would not read
SOMAR_HD __forceinline__ int ex_steps(int ord, int K) { return (ord == 0) ? 0 : ((K < ord) ? K + 1 : ord + 1); }

// The accessors below are written for memory-level parallelism.  ex_slow / ExView reach every value through branches and
// calls, so the ~50 reads of a cell's statement are issued one behind the other: a thread needs tens of microseconds for its
// cell, and on a coarse depth (few warps) that latency IS the launch.  Here every read is unconditional -- a direction that is
// not extrapolated reads offset 0 three times -- and whether a direction CAN be left is decided at compile time from the
// point's offset from the cell (a constant at every call site of the statements).

// Accessor of a CELL_DIR<D> cell: ex_slow with every direction but D known to stay inside the box.
template <int ND, int D>
struct DirView {
    const double* __restrict__ phi;
    long long base;      // index of the cell
    long long sy, sz;
    int i0, j0, k0;
    int ordLo, ordHi;    // order of the extrapolation one cell below / above the cell in direction D; 0: phi itself (the
                         // point is inside the box, or behind a box-box interface: the neighbour's phi)
    SOMAR_HD __forceinline__ DirView(const Geom& g, const double* __restrict__ phi_, int i, int j, int k, long long c, int mode)
        : phi(phi_), base(c), sy(g.sy), sz(g.sz), i0(i), j0(j), k0(k) {
        const int q = (D == 0) ? i : (D == 1 ? j : k);
        const int gq = q + g.g0[D];
        const int glo = (gq / g.bs[D]) * g.bs[D], ghi = glo + g.bs[D] - 1;
        const int tLo = (glo == 0) ? g.bct[D][0] : SIDE_INTERNAL, tHi = (ghi == g.gn[D] - 1) ? g.bct[D][1] : SIDE_INTERNAL;
        ordLo = (gq == glo) ? side_order(tLo, mode) : 0;
        ordHi = (gq == ghi) ? side_order(tHi, mode) : 0;
    }
    SOMAR_HD static __forceinline__ int side_order(int t, int mode) {
        if (t == BC_NEUM) return 2;                                  // force
        if (t == SIDE_INTERNAL) return 0;
        if (t == BC_PERIODIC) return (mode == EXMODE_OP) ? 0 : 1;
        return (mode == EXMODE_OP) ? 2 : 1;
    }
    SOMAR_HD __forceinline__ double ex(int i, int j, int k) const {
        const int di = i - i0, dj = j - j0, dk = k - k0;            // constants at the call sites
        const long long a = base + di + sy * dj + sz * dk;
        const int dq = (D == 0) ? di : (D == 1 ? dj : dk);
        if (dq == 0) return phi[a];
        const int ord = (dq < 0) ? ordLo : ordHi;
        const long long st = (D == 0) ? 1 : (D == 1 ? sy : sz);
        const long long in = (dq < 0) ? st : -st;                    // one step back into the box
        const double s1 = phi[a + ex_steps(ord, 0) * in], s2 = phi[a + ex_steps(ord, 1) * in], s3 = phi[a + ex_steps(ord, 2) * in];
        return ex_comb(ord, s1, s2, s3);
    }
};

// Accessor of a CELL_GENERAL cell.  The 48 ex() reads of a cell's statement visit 18 points only (the 6 face and 12 edge
// neighbours; 8 in 2-D): each is evaluated once, ex_slow's rules with the nested extrapolation laid out flat -- 3 values for a
// face neighbour, 9 for an edge neighbour, loaded unconditionally and combined in ExRec's order (lowest direction innermost):
// the same expressions, the same bits.
template <int ND>
struct TabView {
    // the values as scalars (an array member keeps the whole view, its i0/j0/k0 included, in local memory: the compiler then
    // cannot fold ex()'s index); slot = (di+1) + 3 (dj+1) + 9 (dk+1)
    double e1, e3, e4, e5, e7;                                         // dk = -1
    double e9, e10, e11, e12, e14, e15, e16, e17;                      // dk = 0
    double e19, e21, e22, e23, e25;                                    // dk = +1
    int i0, j0, k0;
    template <int DI, int DJ, int DK>
    SOMAR_HD static __forceinline__ double eval(const Geom& g, const double* __restrict__ phi, const BoxOf& b, int i, int j,
                                                  int k, long long c, int mode) {
        constexpr int dd[3] = {DI, DJ, DK};
        const int p[3] = {i + DI, j + DJ, k + DK};
        int o[3] = {0, 0, 0}, t[3] = {0, 0, 0}, ord[3] = {0, 0, 0};
        bool force = false;
SOMAR_UNROLL
        for (int d = 0; d < ND; ++d) {
            if (dd[d] < 0 && p[d] < b.lo[d]) { o[d] = -1; t[d] = b.st[d][0]; }
            if (dd[d] > 0 && p[d] > b.hi[d]) { o[d] = 1; t[d] = b.st[d][1]; }
            if (o[d] != 0) force = force | (t[d] == BC_NEUM);
            else force = force | (p[d] == b.lo[d] && b.st[d][0] == BC_NEUM) | (p[d] == b.hi[d] && b.st[d][1] == BC_NEUM);
        }
SOMAR_UNROLL
        for (int d = 0; d < ND; ++d) {
            if (o[d] == 0) ord[d] = 0;
            else if (force) ord[d] = 2;
            else if (t[d] == SIDE_INTERNAL) ord[d] = 0;
            else if (t[d] == BC_PERIODIC) ord[d] = (mode == EXMODE_OP) ? 0 : 1;
            else ord[d] = (mode == EXMODE_OP) ? 2 : 1;
        }
        const long long str[3] = {1, g.sy, g.sz};
        const double* __restrict__ q = phi + (c + DI + g.sy * DJ + g.sz * DK);
        constexpr int n0 = DI ? 3 : 1, n1 = DJ ? 3 : 1, n2 = (ND > 2 && DK) ? 3 : 1;
        long long off[3][3];
SOMAR_UNROLL
        for (int d = 0; d < 3; ++d)
SOMAR_UNROLL
            for (int K = 0; K < 3; ++K) off[d][K] = (d < ND && dd[d] != 0) ? -(long long)o[d] * ex_steps(ord[d], K) * str[d] : 0;
        double vz[3] = {0.0, 0.0, 0.0};
SOMAR_UNROLL
        for (int cc = 0; cc < n2; ++cc) {
            double vy[3] = {0.0, 0.0, 0.0};
SOMAR_UNROLL
            for (int bb = 0; bb < n1; ++bb) {
                const long long a = off[1][bb] + off[2][cc];
                vy[bb] = (n0 == 1) ? q[a] : ex_comb(ord[0], q[a + off[0][0]], q[a + off[0][1]], q[a + off[0][2]]);
            }
            vz[cc] = (n1 == 1) ? vy[0] : ex_comb(ord[1], vy[0], vy[1], vy[2]);
        }
        return (n2 == 1) ? vz[0] : ex_comb(ord[2], vz[0], vz[1], vz[2]);
    }
    SOMAR_HD __forceinline__ TabView(const Geom& g, const double* __restrict__ phi, int i, int j, int k, long long c, int mode)
        : i0(i), j0(j), k0(k) {
        const BoxOf b = box_of<ND>(g, i, j, k);
#define TAB_SET(slot, di, dj, dk) e##slot = eval<di, dj, dk>(g, phi, b, i, j, k, c, mode)
        TAB_SET(12, -1, 0, 0); TAB_SET(14, 1, 0, 0); TAB_SET(10, 0, -1, 0); TAB_SET(16, 0, 1, 0);
        TAB_SET(9, -1, -1, 0); TAB_SET(11, 1, -1, 0); TAB_SET(15, -1, 1, 0); TAB_SET(17, 1, 1, 0);
        e1 = e3 = e4 = e5 = e7 = e19 = e21 = e22 = e23 = e25 = 0.0;
        if (ND > 2) {
            TAB_SET(4, 0, 0, -1); TAB_SET(22, 0, 0, 1);
            TAB_SET(3, -1, 0, -1); TAB_SET(5, 1, 0, -1); TAB_SET(1, 0, -1, -1); TAB_SET(7, 0, 1, -1);
            TAB_SET(21, -1, 0, 1); TAB_SET(23, 1, 0, 1); TAB_SET(19, 0, -1, 1); TAB_SET(25, 0, 1, 1);
        }
#undef TAB_SET
    }
    SOMAR_HD __forceinline__ double ex(int i, int j, int k) const {
        switch ((i - i0 + 1) + 3 * (j - j0 + 1) + 9 * (k - k0 + 1)) {   // a constant wherever the statements call it
            case 1: return e1; case 3: return e3; case 4: return e4; case 5: return e5; case 7: return e7;
            case 9: return e9; case 10: return e10; case 11: return e11; case 12: return e12;
            case 14: return e14; case 15: return e15; case 16: return e16; case 17: return e17;
            case 19: return e19; case 21: return e21; case 22: return e22; case 23: return e23; case 25: return e25;
            default: return 0.0;
        }
    }
};

// 4-point transverse difference used by every cross term: around the face between cell A=(a0,a1,a2) and
// cell B = A - e_n (n = face normal), difference taken in direction t:
//    ex(A+t) - ex(A-t) + ex(B+t) - ex(B-t)
template <int ND, class V>
SOMAR_HD __forceinline__ double cross4(const V& v, int a0, int a1, int a2, int n, int t) {
    const int e0 = (n == 0), e1 = (n == 1), e2 = (n == 2);
    const int t0 = (t == 0), t1 = (t == 1), t2 = (t == 2);
    return v.ex(a0 + t0, a1 + t1, a2 + t2) - v.ex(a0 - t0, a1 - t1, a2 - t2) +
           v.ex(a0 - e0 + t0, a1 - e1 + t1, a2 - e2 + t2) - v.ex(a0 - e0 - t0, a1 - e1 - t1, a2 - e2 - t2);
}

// beta * flux through the low (S=0) / high (S=1) face of cell q=(i,j,k) in direction A (MAPPEDGETFLUX + BCs).
template <int ND, int A, int S, class V>
SOMAR_HD __forceinline__ double face_flux_full(const Geom& g, const Metric& m, const V& v,
                                                 const double* __restrict__ phi, int i, int j, int k, long long c,
                                                 bool atBdry, bool homog, double pc) {
    constexpr int B = (A + 1) % ND;
    constexpr int C = (A + 2) % ND;
    const long long st = (A == 0) ? 1 : (A == 1 ? g.sy : g.sz);
    // No load sits behind a branch (a thread's reads then go out together, see the accessors above): the neighbour is read
    // even where it is a ghost cell, and a Neumann face computes its interior expression before discarding it.
    const bool neum = atBdry && g.bc[A][S] == BC_NEUM;
    const bool extrap2 = atBdry && g.bc[A][S] == BC_EXTRAP2;   // ELLIPTICEXTRAPBCGHOST order 2 (projector gradients only)
    const double nearV = S ? phi[c + st] : phi[c - st];
    const double farV = S ? phi[c - st] : phi[c + st];
    const double far2 = phi[extrap2 ? (S ? c - 2 * st : c + 2 * st) : c];
    double nb = nearV;
    if (atBdry) nb = extrap2 ? 3.0 * (pc - farV) + far2 : diri_ghost_val(pc, farV, homog ? 0.0 : g.bcval[A][S], g.diriOrder);
    const long long f = S ? c + st : c;
    // "cell A" of the face = the cell on its high side
    const int a0 = i + (S && A == 0), a1 = j + (S && A == 1), a2 = k + (S && A == 2);
    const double aS = 1.0 / g.dx[A];
    const double bS = 0.25 * 1.0 / g.dx[B];
    const double diff = S ? (nb - pc) : (pc - nb);
    double fl = aS * m.jg[A][A][f] * diff + bS * m.jg[A][B][f] * cross4<ND>(v, a0, a1, a2, A, B);
    if (ND > 2) {
        const double cS = 0.25 * 1.0 / g.dx[C];
        fl += cS * m.jg[A][C][f] * cross4<ND>(v, a0, a1, a2, A, C);
    }
    return neum ? (homog ? 0.0 : g.bcval[A][S]) * g.beta : fl * g.beta;
}

// (alpha I + beta L) phi at one cell, full metric (applyOpI, MappedAMRPoissonOp.cpp:772-898).
template <int ND, class V>
SOMAR_HD __forceinline__ double apply_cell_full_v(const Geom& g, const Metric& m, const V& v,
                                                    const double* __restrict__ phi, int i, int j, int k, long long c,
                                                    bool homog, unsigned sm) {
    const double pc = phi[c];
    const double f0l = face_flux_full<ND, 0, 0>(g, m, v, phi, i, j, k, c, sm & 1u, homog, pc);
    const double f0h = face_flux_full<ND, 0, 1>(g, m, v, phi, i, j, k, c, sm & 2u, homog, pc);
    const double f1l = face_flux_full<ND, 1, 0>(g, m, v, phi, i, j, k, c, sm & 4u, homog, pc);
    const double f1h = face_flux_full<ND, 1, 1>(g, m, v, phi, i, j, k, c, sm & 8u, homog, pc);
    double s = (f0h - f0l) * (1.0 / g.dx[0]);
    s = s + (f1h - f1l) * (1.0 / g.dx[1]);
    if (ND > 2) {
        const double f2l = face_flux_full<ND, (ND > 2 ? 2 : 0), 0>(g, m, v, phi, i, j, k, c, sm & 16u, homog, pc);
        const double f2h = face_flux_full<ND, (ND > 2 ? 2 : 0), 1>(g, m, v, phi, i, j, k, c, sm & 32u, homog, pc);
        s = s + (f2h - f2l) * (1.0 / g.dx[2]);
    }
    double r = m.jinv[c] * s;
    if (g.alpha != 0.0) r = g.alpha * pc + 1.0 * r;
    return r;
}
template <int ND>
SOMAR_HD __forceinline__ double apply_cell_full(const Geom& g, const Metric& m, const double* __restrict__ phi,
                                                  int i, int j, int k, long long c, bool homog) {
    const ExView<ND> v(g, phi, i, j, k, EXMODE_OP);
    return apply_cell_full_v<ND>(g, m, v, phi, i, j, k, c, homog, side_mask(g, i, j, k));
}
// a cell of known kind (cell_kind): the accessor that kind allows
template <int ND>
SOMAR_HD __forceinline__ double apply_cell_full_kind(const Geom& g, const Metric& m, const double* __restrict__ phi,
                                                       int i, int j, int k, long long c, bool homog, int kind) {
    if (kind == CELL_FAST) {
        const FastView<ND> v{phi + c, i, j, k, g.sy, g.sz};
        return apply_cell_full_v<ND>(g, m, v, phi, i, j, k, c, homog, 0u);
    }
    const unsigned sm = side_mask(g, i, j, k);
    if (kind == CELL_DIR0) {
        const DirView<ND, 0> v(g, phi, i, j, k, c, EXMODE_OP);
        return apply_cell_full_v<ND>(g, m, v, phi, i, j, k, c, homog, sm);
    }
    if (kind == CELL_DIR1) {
        const DirView<ND, 1> v(g, phi, i, j, k, c, EXMODE_OP);
        return apply_cell_full_v<ND>(g, m, v, phi, i, j, k, c, homog, sm);
    }
    if (ND > 2 && kind == CELL_DIR2) {
        const DirView<ND, (ND > 2 ? 2 : 0)> v(g, phi, i, j, k, c, EXMODE_OP);
        return apply_cell_full_v<ND>(g, m, v, phi, i, j, k, c, homog, sm);
    }
    const TabView<ND> v(g, phi, i, j, k, c, EXMODE_OP);
    return apply_cell_full_v<ND>(g, m, v, phi, i, j, k, c, homog, sm);
}
// box-interior cell: no boundary side, no extrapolated value
template <int ND>
SOMAR_HD __forceinline__ double apply_cell_full_inner(const Geom& g, const Metric& m, const double* __restrict__ phi,
                                                        int i, int j, int k, long long c, bool homog) {
    const FastView<ND> v{phi + c, i, j, k, g.sy, g.sz};
    return apply_cell_full_v<ND>(g, m, v, phi, i, j, k, c, homog, 0u);
}

// d_a ( J g^{ab} d_b ex ) without the 1/(4 dx_a dx_b) factor: high-face term minus low-face term.
template <int ND, class V>
SOMAR_HD __forceinline__ double cross_ab(const Metric& m, const V& v, int i, int j, int k, long long c,
                                           long long sta, int a, int b) {
    const int e0 = (a == 0), e1 = (a == 1), e2 = (a == 2);
    const double up = m.jg[a][b][c + sta] * cross4<ND>(v, i + e0, j + e1, k + e2, a, b);
    const double dn = m.jg[a][b][c] * cross4<ND>(v, i, j, k, a, b);
    return up - dn;
}

// One red-black update of one cell, full metric: phi = pre-pass state (normal terms read the other colour,
// cross terms the frozen snapshot).  GSRBITER{2,3}D / GSRBBOUNDARYITER{2,3}D.
// interior statement (GSRBITER{2,3}D) on any accessor
template <int ND, class V>
SOMAR_HD __forceinline__ double gsrb_interior_full(const Geom& g, const Metric& m, const V& v,
                                                     const double* __restrict__ phi, const double* __restrict__ rhs,
                                                     int i, int j, int k, long long c) {
    const double ji = m.jinv[c];
    {
        const double xx = 1.0 / (g.dx[0] * g.dx[0]), yy = 1.0 / (g.dx[1] * g.dx[1]);
        const double xy = 0.25 / (g.dx[0] * g.dx[1]);
        const double JDxx = m.jg[0][0][c + 1] * phi[c + 1] + m.jg[0][0][c] * phi[c - 1];
        const double JDyy = m.jg[1][1][c + g.sy] * phi[c + g.sy] + m.jg[1][1][c] * phi[c - g.sy];
        const double JDxy = cross_ab<ND>(m, v, i, j, k, c, 1, 0, 1);
        const double JDyx = cross_ab<ND>(m, v, i, j, k, c, g.sy, 1, 0);
        double lphi;
        if (ND == 2) {
            lphi = g.beta * ji * (JDxx * xx + JDyy * yy + (JDxy + JDyx) * xy);
        } else {
            const double zz = 1.0 / (g.dx[2] * g.dx[2]);
            const double yz = 0.25 / (g.dx[1] * g.dx[2]);
            const double zx = 0.25 / (g.dx[2] * g.dx[0]);
            const double JDzz = m.jg[2][2][c + g.sz] * phi[c + g.sz] + m.jg[2][2][c] * phi[c - g.sz];
            const double JDxz = cross_ab<ND>(m, v, i, j, k, c, 1, 0, 2);
            const double JDyz = cross_ab<ND>(m, v, i, j, k, c, g.sy, 1, 2);
            const double JDzx = cross_ab<ND>(m, v, i, j, k, c, g.sz, 2, 0);
            const double JDzy = cross_ab<ND>(m, v, i, j, k, c, g.sz, 2, 1);
            lphi = g.beta * ji *
                   (JDxx * xx + JDyy * yy + JDzz * zz + (JDxy + JDyx) * xy + (JDyz + JDzy) * yz + (JDzx + JDxz) * zx);
        }
        return (rhs[c] - lphi) / (g.alpha + g.beta * m.lapdiag[c]);
    }
}
// box-interior cell: every ex() is phi
template <int ND>
SOMAR_HD __forceinline__ double gsrb_cell_full_inner(const Geom& g, const Metric& m, const double* __restrict__ phi,
                                                       const double* __restrict__ rhs, int i, int j, int k, long long c) {
    const FastView<ND> v{phi + c, i, j, k, g.sy, g.sz};
    return gsrb_interior_full<ND>(g, m, v, phi, rhs, i, j, k, c);
}

template <int ND, class V>
SOMAR_HD __forceinline__ double gsrb_cell_full_v(const Geom& g, const Metric& m, const V& v,
                                                   const double* __restrict__ phi, const double* __restrict__ rhs, int i,
                                                   int j, int k, long long c) {
    if (!on_domain_rim<ND>(g, i, j, k)) return gsrb_interior_full<ND>(g, m, v, phi, rhs, i, j, k, c);
    const long long st[3] = {1, g.sy, g.sz};
    const double ji = m.jinv[c];
    // rim cell (GSRBBOUNDARYITER*): every non-Neumann face contributes its whole flux, Neumann faces nothing
    const unsigned sm = side_mask(g, i, j, k);
    const double pc = phi[c];
    double JD[3][2] = {{0.0, 0.0}, {0.0, 0.0}, {0.0, 0.0}};
    double lapDiag = 0.0;
SOMAR_UNROLL
    for (int d = 0; d < ND; ++d) {
SOMAR_UNROLL
        for (int s = 0; s < 2; ++s) {
            const bool atB = (sm >> (2 * d + s)) & 1u;
            const bool skip = atB && g.bc[d][s] == BC_NEUM;      // the face contributes nothing (no load behind a branch:
            const long long f = s ? c + st[d] : c;               // its expression is computed and discarded)
            const int a0 = i + (s && d == 0), a1 = j + (s && d == 1), a2 = k + (s && d == 2);
            double nb = s ? phi[c + st[d]] : phi[c - st[d]];
            const double farV = s ? phi[c - st[d]] : phi[c + st[d]];
            if (atB && g.bc[d][s] == BC_DIRI) nb = diri_ghost_val(pc, farV, 0.0, g.diriOrder);
            const double sgn = s ? 1.0 : -1.0;
            double val = 0.0;
SOMAR_UNROLL
            for (int t = 0; t < ND; ++t) {
                if (t == d) {
                    const double aa = 1.0 / (g.dx[d] * g.dx[d]);
                    val += aa * m.jg[d][d][f] * nb;
                } else {
                    const double ab = 0.25 / (g.dx[d] * g.dx[t]);
                    val += sgn * ab * m.jg[d][t][f] * cross4<ND>(v, a0, a1, a2, d, t);
                }
            }
            JD[d][s] = skip ? 0.0 : val;
            const double ld = lapDiag - (1.0 / (g.dx[d] * g.dx[d])) * m.jg[d][d][f];
            lapDiag = skip ? lapDiag : ld;
        }
    }
    lapDiag = lapDiag * ji;
    double s = JD[0][0] + JD[0][1] + JD[1][0] + JD[1][1];
    if (ND > 2) s = s + JD[2][0] + JD[2][1];
    const double lphi = g.beta * ji * s;
    return (rhs[c] - lphi) / (g.alpha + g.beta * lapDiag);
}
template <int ND>
SOMAR_HD __forceinline__ double gsrb_cell_full(const Geom& g, const Metric& m, const double* __restrict__ phi,
                                                 const double* __restrict__ rhs, int i, int j, int k, long long c) {
    const ExView<ND> v(g, phi, i, j, k, EXMODE_SMOOTHER);
    return gsrb_cell_full_v<ND>(g, m, v, phi, rhs, i, j, k, c);
}
// a cell of known kind (cell_kind)
template <int ND>
SOMAR_HD __forceinline__ double gsrb_cell_full_kind(const Geom& g, const Metric& m, const double* __restrict__ phi,
                                                      const double* __restrict__ rhs, int i, int j, int k, long long c,
                                                      int kind) {
    if (kind == CELL_FAST) return gsrb_cell_full_inner<ND>(g, m, phi, rhs, i, j, k, c);
    if (kind == CELL_DIR0) {
        const DirView<ND, 0> v(g, phi, i, j, k, c, EXMODE_SMOOTHER);
        return gsrb_cell_full_v<ND>(g, m, v, phi, rhs, i, j, k, c);
    }
    if (kind == CELL_DIR1) {
        const DirView<ND, 1> v(g, phi, i, j, k, c, EXMODE_SMOOTHER);
        return gsrb_cell_full_v<ND>(g, m, v, phi, rhs, i, j, k, c);
    }
    if (ND > 2 && kind == CELL_DIR2) {
        const DirView<ND, (ND > 2 ? 2 : 0)> v(g, phi, i, j, k, c, EXMODE_SMOOTHER);
        return gsrb_cell_full_v<ND>(g, m, v, phi, rhs, i, j, k, c);
    }
    const TabView<ND> v(g, phi, i, j, k, c, EXMODE_SMOOTHER);
    return gsrb_cell_full_v<ND>(g, m, v, phi, rhs, i, j, k, c);
}

}  // namespace somar
