// oracle/chombo_lite.hpp -- TEST INFRASTRUCTURE ONLY (CPU oracle). Never linked into the product.
//
// A minimal stand-in for the Chombo 3.1 containers the reference's hot path is written against
// (Chombo is NOT vendored in /root/reference; README.md:33).  Only the Box calculus and the
// Fortran-ordered FArrayBox that the restated kernels need are provided:
//   IntVect, Box (cell- or node-centred per direction), FAB (x fastest, then y, z, then component;
//   ghosts live inside the allocation -- the CHF_FRA calling convention).
// The Box member functions follow the Chombo semantics the reference relies on
// (adjCellBox, bdryBox, shiftHalf, surroundingNodes, grow, &): see the call sites cited next to
// each use in oracle/somar_oracle.cpp.
//
// 2-D problems are embedded as 3-D boxes one cell thick in direction 2 (that direction "inactive").
#pragma once
#include <algorithm>
#include <array>
#include <cassert>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace orc {

constexpr int D3 = 3;

struct IntVect {
    int v[3];
    IntVect() : v{0, 0, 0} {}
    IntVect(int a, int b, int c) : v{a, b, c} {}
    int& operator[](int i) { return v[i]; }
    int operator[](int i) const { return v[i]; }
    static IntVect Zero() { return IntVect(0, 0, 0); }
    static IntVect Unit() { return IntVect(1, 1, 1); }
    static IntVect Basis(int d) { IntVect r; r.v[d] = 1; return r; }
    IntVect operator+(const IntVect& o) const { return IntVect(v[0] + o.v[0], v[1] + o.v[1], v[2] + o.v[2]); }
    IntVect operator-(const IntVect& o) const { return IntVect(v[0] - o.v[0], v[1] - o.v[1], v[2] - o.v[2]); }
    IntVect operator*(const IntVect& o) const { return IntVect(v[0] * o.v[0], v[1] * o.v[1], v[2] * o.v[2]); }
    IntVect operator*(int s) const { return IntVect(v[0] * s, v[1] * s, v[2] * s); }
    bool operator==(const IntVect& o) const { return v[0] == o.v[0] && v[1] == o.v[1] && v[2] == o.v[2]; }
    int product() const { return v[0] * v[1] * v[2]; }
};

// floor division (Chombo coarsen semantics for negative indices)
inline int floordiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

enum Side { Lo = 0, Hi = 1 };
inline int sign(int side) { return side == Lo ? -1 : 1; }
inline int flip(int side) { return 1 - side; }

// Box: inclusive index range; type[d] = 0 cell-centred, 1 node-centred in direction d.
struct Box {
    IntVect lo, hi;
    IntVect type;
    Box() : lo(0, 0, 0), hi(-1, -1, -1), type(0, 0, 0) {}
    Box(const IntVect& l, const IntVect& h) : lo(l), hi(h), type(0, 0, 0) {}
    Box(const IntVect& l, const IntVect& h, const IntVect& t) : lo(l), hi(h), type(t) {}

    bool isEmpty() const { return hi[0] < lo[0] || hi[1] < lo[1] || hi[2] < lo[2]; }
    int size(int d) const { return hi[d] - lo[d] + 1; }
    IntVect size() const { return IntVect(size(0), size(1), size(2)); }
    long numPts() const { return isEmpty() ? 0 : (long)size(0) * size(1) * size(2); }
    int sideEnd(int side, int d) const { return side == Lo ? lo[d] : hi[d]; }
    bool contains(const Box& b) const {
        if (b.isEmpty()) return true;
        for (int d = 0; d < 3; ++d)
            if (b.lo[d] < lo[d] || b.hi[d] > hi[d]) return false;
        return true;
    }
    bool contains(int i, int j, int k) const {
        return i >= lo[0] && i <= hi[0] && j >= lo[1] && j <= hi[1] && k >= lo[2] && k <= hi[2];
    }
    bool intersects(const Box& b) const { return !((*this) & b).isEmpty(); }
    bool operator==(const Box& b) const { return lo == b.lo && hi == b.hi && type == b.type; }

    Box operator&(const Box& b) const {
        Box r(*this);
        for (int d = 0; d < 3; ++d) {
            r.lo[d] = std::max(lo[d], b.lo[d]);
            r.hi[d] = std::min(hi[d], b.hi[d]);
        }
        return r;
    }
    Box& operator&=(const Box& b) { *this = (*this) & b; return *this; }

    Box& grow(int d, int n) { lo[d] -= n; hi[d] += n; return *this; }
    Box& grow(const IntVect& n) { for (int d = 0; d < 3; ++d) grow(d, n[d]); return *this; }
    Box& growDir(int d, int side, int n) { if (side == Lo) lo[d] -= n; else hi[d] += n; return *this; }
    Box& shift(int d, int n) { lo[d] += n; hi[d] += n; return *this; }
    Box& shift(const IntVect& s) { for (int d = 0; d < 3; ++d) shift(d, s[d]); return *this; }

    // cell -> node in direction d: the faces surrounding the cells [lo, hi+1].
    Box& surroundingNodes(int d) { if (type[d] == 0) { hi[d] += 1; type[d] = 1; } return *this; }
    // node -> cell in direction d: the cells enclosed by the nodes [lo, hi-1].
    Box& enclosedCells(int d) { if (type[d] == 1) { hi[d] -= 1; type[d] = 0; } return *this; }

    // Shift by n half cells; an odd n toggles the centring.  Node i is the low face of cell i.
    Box& shiftHalf(int d, int n) {
        if (n == 0) return *this;
        const int nfull = floordiv(n, 2);          // whole cells (towards -inf)
        const bool odd = (n - 2 * nfull) != 0;    // remaining +1 half step
        shift(d, nfull);
        if (odd) {
            if (type[d] == 0) { shift(d, 1); type[d] = 1; }   // cell i + half -> node i+1
            else { type[d] = 0; }                              // node i + half -> cell i
        }
        return *this;
    }

    // Chombo coarsen / refine by per-direction ratios (cell-centred boxes).
    Box coarsened(const IntVect& r) const {
        Box b(*this);
        for (int d = 0; d < 3; ++d) {
            b.lo[d] = floordiv(lo[d], r[d]);
            b.hi[d] = (type[d] == 0) ? floordiv(hi[d], r[d]) : floordiv(hi[d] + r[d] - 1, r[d]);
        }
        return b;
    }
    Box refined(const IntVect& r) const {
        Box b(*this);
        for (int d = 0; d < 3; ++d) {
            b.lo[d] = lo[d] * r[d];
            b.hi[d] = (type[d] == 0) ? (hi[d] + 1) * r[d] - 1 : hi[d] * r[d];
        }
        return b;
    }
};

inline Box grow(const Box& b, const IntVect& n) { Box r(b); r.grow(n); return r; }
inline Box grow(const Box& b, int n) { Box r(b); r.grow(IntVect(n, n, n)); return r; }
inline Box surroundingNodes(const Box& b, int d) { Box r(b); r.surroundingNodes(d); return r; }

// Cell-centred box of |len| cells adjacent to (len>0: outside, len<0: inside) the given side of b.
inline Box adjCellBox(const Box& b, int d, int side, int len) {
    assert(b.type[d] == 0);
    Box r(b);
    if (len > 0) {
        if (side == Lo) { r.lo[d] = b.lo[d] - len; r.hi[d] = b.lo[d] - 1; }
        else { r.lo[d] = b.hi[d] + 1; r.hi[d] = b.hi[d] + len; }
    } else {
        if (side == Lo) { r.lo[d] = b.lo[d]; r.hi[d] = b.lo[d] - len - 1; }
        else { r.lo[d] = b.hi[d] + len + 1; r.hi[d] = b.hi[d]; }
    }
    return r;
}

// Node-centred (in d) box one face thick on the boundary of the cell-centred box b.
inline Box bdryBox(const Box& b, int d, int side, int /*len*/ = 1) {
    Box r(b);
    if (b.type[d] == 0) {
        const int node = (side == Lo) ? b.lo[d] : b.hi[d] + 1;
        r.lo[d] = r.hi[d] = node;
        r.type[d] = 1;
    } else {
        const int node = (side == Lo) ? b.lo[d] : b.hi[d];
        r.lo[d] = r.hi[d] = node;
    }
    return r;
}

// Fortran-ordered array over a box (ghosts included in the box), ncomp components outermost.
struct FAB {
    Box b;
    int nc = 0;
    long sx = 0, sy = 0, sz = 0, sn = 0;
    std::vector<double> d;
    FAB() {}
    FAB(const Box& box, int ncomp) { define(box, ncomp); }
    void define(const Box& box, int ncomp) {
        b = box; nc = ncomp;
        sx = 1; sy = box.size(0); sz = sy * box.size(1); sn = sz * box.size(2);
        d.assign((size_t)std::max<long>(0, sn * nc), 0.0);
    }
    const Box& box() const { return b; }
    int nComp() const { return nc; }
    inline long idx(int i, int j, int k, int n = 0) const {
        return (i - b.lo[0]) + sy * (j - b.lo[1]) + sz * (k - b.lo[2]) + sn * n;
    }
    inline double& operator()(int i, int j, int k, int n = 0) {
#ifdef ORC_BOUNDS
        if (!b.contains(i, j, k) || n < 0 || n >= nc) { fprintf(stderr, "FAB oob (%d,%d,%d,%d)\n", i, j, k, n); abort(); }
#endif
        return d[idx(i, j, k, n)];
    }
    inline double operator()(int i, int j, int k, int n = 0) const {
#ifdef ORC_BOUNDS
        if (!b.contains(i, j, k) || n < 0 || n >= nc) { fprintf(stderr, "FAB oob (%d,%d,%d,%d)\n", i, j, k, n); abort(); }
#endif
        return d[idx(i, j, k, n)];
    }
    void setVal(double x) { std::fill(d.begin(), d.end(), x); }
    void setVal(double x, const Box& r, int c0, int ncmp) {
        Box rr = r & b;
        for (int n = c0; n < c0 + ncmp; ++n)
            for (int k = rr.lo[2]; k <= rr.hi[2]; ++k)
                for (int j = rr.lo[1]; j <= rr.hi[1]; ++j)
                    for (int i = rr.lo[0]; i <= rr.hi[0]; ++i) (*this)(i, j, k, n) = x;
    }
    // copy src -> this over region (intersected with both boxes), all comps
    void copy(const FAB& src, const Box& region) {
        Box rr = region & b & src.b;
        if (rr.isEmpty()) return;
        for (int n = 0; n < nc; ++n)
            for (int k = rr.lo[2]; k <= rr.hi[2]; ++k)
                for (int j = rr.lo[1]; j <= rr.hi[1]; ++j)
                    for (int i = rr.lo[0]; i <= rr.hi[0]; ++i) (*this)(i, j, k, n) = src(i, j, k, n);
    }
    void copy(const FAB& src) { copy(src, src.b); }
    // copy with an index shift: this(p) = src(p - shift) for p in region
    void copyShifted(const FAB& src, const Box& region, const IntVect& sh) {
        for (int n = 0; n < nc; ++n)
            for (int k = region.lo[2]; k <= region.hi[2]; ++k)
                for (int j = region.lo[1]; j <= region.hi[1]; ++j)
                    for (int i = region.lo[0]; i <= region.hi[0]; ++i)
                        (*this)(i, j, k, n) = src(i - sh[0], j - sh[1], k - sh[2], n);
    }
};

}  // namespace orc
