// tests/cuda/full_accessors_host.cu -- runs ON THE HOST (no GPU): the full-metric accessors and cell statements of
// somar_b200/csrc/stencil_full.cuh, which are __host__ __device__, on random fields over a set of box layouts, boundary-
// condition mixes and rank offsets.  ExView / ex_slow is the accessor whose results were established against the oracle
// (1e-13, tests/test_gpu_parity.py); the kind-dispatched ones must give the same bits:
//   every cell, every one of the 18 (8 in 2-D) points a statement reads:   TabView::ex == ExView::ex
//   CELL_FAST cells:      FastView::ex == ExView::ex  (i.e. phi itself)
//   CELL_DIR<D> cells:    DirView<D>::ex == ExView::ex
//   every cell:           gsrb_cell_full_kind == gsrb_cell_full,  apply_cell_full_kind == apply_cell_full
// Built and run by tests/test_full_accessors_host.py:  nvcc -O1 -std=c++17 -fmad=false -I somar_b200/csrc ... && ./a.out
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "stencil_full.cuh"

using namespace somar;

struct Case {
    const char* name;
    int nd, gn[3], bs[3], bct[3][2];
    int g0y, ny;   // this "rank" owns rows g0y .. g0y + ny - 1 (ny = 0: the whole domain)
};

static unsigned long long rng_state = 0x9E3779B97F4A7C15ull;
static double rnd() {
    rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
    return (double)(rng_state >> 11) / 9007199254740992.0 * 2.0 - 1.0;
}

static const int GH = 4;

static Geom make_geom(const Case& c) {
    Geom g;
    std::memset(&g, 0, sizeof(g));
    g.nd = c.nd;
    for (int d = 0; d < 3; ++d) {
        g.gn[d] = (d < c.nd) ? c.gn[d] : 1;
        g.n[d] = g.gn[d];
        g.g0[d] = 0;
        g.bs[d] = (d < c.nd) ? c.bs[d] : 1;
        g.gh[d] = GH;
        g.dx[d] = 0.5 + 0.25 * d;
        g.rdx[d] = 1.0 / g.dx[d];
        g.rdx2[d] = 1.0 / (g.dx[d] * g.dx[d]);
        for (int s = 0; s < 2; ++s) {
            g.bct[d][s] = (d < c.nd) ? c.bct[d][s] : BC_NONE;
            g.bcval[d][s] = 0.0;
        }
        g.periodic[d] = (d < c.nd) && c.bct[d][0] == BC_PERIODIC;
    }
    if (c.ny > 0) { g.g0[1] = c.g0y; g.n[1] = c.ny; }
    for (int d = 0; d < 3; ++d)
        for (int s = 0; s < 2; ++s) {
            const bool atDomain = s ? (g.g0[d] + g.n[d] == g.gn[d]) : (g.g0[d] == 0);
            g.bc[d][s] = (d < c.nd && atDomain && g.bct[d][s] != BC_PERIODIC) ? g.bct[d][s] : BC_NONE;
        }
    g.sy = g.n[0] + 2 * GH;
    g.sz = g.sy * (g.n[1] + 2 * GH);
    g.off0 = GH + g.sy * GH + g.sz * GH;
    g.total = g.sz * (g.n[2] + 2 * GH);
    g.alpha = 0.0;
    g.beta = 1.0;
    g.diriOrder = 1;
    g.nbr[0] = g.nbr[1] = -1;
    return g;
}

static long long bad = 0, checked = 0;
static void expect(bool ok, const char* what, const Case& c, int mode, int i, int j, int k, double a, double b) {
    ++checked;
    if (ok) return;
    if (bad < 20) fprintf(stderr, "MISMATCH %s  case %s mode %d cell (%d,%d,%d): %.17g vs %.17g\n", what, c.name, mode, i, j, k, a, b);
    ++bad;
}
static bool same(double a, double b) { return std::memcmp(&a, &b, sizeof(double)) == 0; }

template <int ND>
static void run_case(const Case& c) {
    const Geom g = make_geom(c);
    std::vector<double> phi(g.total), rhs(g.total), jinv(g.total), lap(g.total);
    std::vector<double> jg[3][3];
    for (auto& v : phi) v = rnd();
    for (auto& v : rhs) v = rnd();
    for (auto& v : jinv) v = 1.0 + 0.25 * rnd();
    for (auto& v : lap) v = -4.0 + rnd();
    Metric m;
    std::memset(&m, 0, sizeof(m));
    for (int d = 0; d < ND; ++d)
        for (int e = 0; e < ND; ++e) {
            jg[d][e].resize(g.total);
            for (auto& v : jg[d][e]) v = (d == e) ? 1.0 + 0.25 * rnd() : 0.3 * rnd();
            m.jg[d][e] = jg[d][e].data();
        }
    m.jinv = jinv.data();
    m.lapdiag = lap.data();
    m.diagonal = false;
    long long kinds[5] = {0, 0, 0, 0, 0};
    for (int mode = 0; mode < 2; ++mode)
        for (int k = 0; k < g.n[2]; ++k)
            for (int j = 0; j < g.n[1]; ++j)
                for (int i = 0; i < g.n[0]; ++i) {
                    const long long cc = g.idx(i, j, k);
                    const int kind = cell_kind<ND>(g, i, j, k);
                    if (mode == 0) kinds[kind]++;
                    const ExView<ND> ref(g, phi.data(), i, j, k, mode);
                    const TabView<ND> tab(g, phi.data(), i, j, k, cc, mode);
                    const FastView<ND> fast{phi.data() + cc, i, j, k, g.sy, g.sz};
                    const DirView<ND, 0> d0(g, phi.data(), i, j, k, cc, mode);
                    const DirView<ND, 1> d1(g, phi.data(), i, j, k, cc, mode);
                    const DirView<ND, (ND > 2 ? 2 : 0)> d2(g, phi.data(), i, j, k, cc, mode);
                    for (int dk = (ND > 2 ? -1 : 0); dk <= (ND > 2 ? 1 : 0); ++dk)
                        for (int dj = -1; dj <= 1; ++dj)
                            for (int di = -1; di <= 1; ++di) {
                                const int nz = (di != 0) + (dj != 0) + (dk != 0);
                                if (nz != 1 && nz != 2) continue;
                                const double r = ref.ex(i + di, j + dj, k + dk);
                                const double t = tab.ex(i + di, j + dj, k + dk);
                                expect(same(r, t), "TabView", c, mode, i, j, k, t, r);
                                if (kind == CELL_FAST) {
                                    const double f = fast.ex(i + di, j + dj, k + dk);
                                    expect(same(r, f), "FastView", c, mode, i, j, k, f, r);
                                } else if (kind == CELL_DIR0) {
                                    const double f = d0.ex(i + di, j + dj, k + dk);
                                    expect(same(r, f), "DirView<0>", c, mode, i, j, k, f, r);
                                } else if (kind == CELL_DIR1) {
                                    const double f = d1.ex(i + di, j + dj, k + dk);
                                    expect(same(r, f), "DirView<1>", c, mode, i, j, k, f, r);
                                } else if (kind == CELL_DIR2) {
                                    const double f = d2.ex(i + di, j + dj, k + dk);
                                    expect(same(r, f), "DirView<2>", c, mode, i, j, k, f, r);
                                }
                            }
                    if (mode == EXMODE_SMOOTHER) {
                        const double a = gsrb_cell_full<ND>(g, m, phi.data(), rhs.data(), i, j, k, cc);
                        const double b = gsrb_cell_full_kind<ND>(g, m, phi.data(), rhs.data(), i, j, k, cc, kind);
                        expect(same(a, b), "gsrb_cell_full_kind", c, mode, i, j, k, b, a);
                    } else {
                        for (int homog = 0; homog < 2; ++homog) {
                            const double a = apply_cell_full<ND>(g, m, phi.data(), i, j, k, cc, homog != 0);
                            const double b = apply_cell_full_kind<ND>(g, m, phi.data(), i, j, k, cc, homog != 0, kind);
                            expect(same(a, b), "apply_cell_full_kind", c, mode, i, j, k, b, a);
                        }
                    }
                }
    printf("%-28s cells %6lld  fast %6lld  dir %5lld/%5lld/%5lld  general %5lld\n", c.name, (long long)g.cells(), kinds[0],
           kinds[1], kinds[2], kinds[3], kinds[4]);
}

int main() {
    const int N = BC_NEUM, D = BC_DIRI, P = BC_PERIODIC;
    const Case cases[] = {
        {"3d one box neumann", 3, {12, 10, 8}, {12, 10, 8}, {{N, N}, {N, N}, {N, N}}, 0, 0},
        {"3d boxes xyz neumann", 3, {12, 12, 8}, {6, 4, 4}, {{N, N}, {N, N}, {N, N}}, 0, 0},
        {"3d boxes y, dirichlet top", 3, {10, 12, 8}, {10, 3, 8}, {{N, N}, {N, N}, {N, D}}, 0, 0},
        {"3d periodic y boxes", 3, {8, 12, 6}, {8, 4, 6}, {{N, N}, {P, P}, {N, N}}, 0, 0},
        {"3d periodic xy one box", 3, {8, 8, 6}, {8, 8, 6}, {{P, P}, {P, P}, {N, D}}, 0, 0},
        {"3d mixed diri/neum boxes", 3, {12, 8, 8}, {4, 4, 4}, {{D, N}, {N, D}, {D, D}}, 0, 0},
        {"3d thin z (bottom depth)", 3, {16, 16, 4}, {16, 4, 4}, {{N, N}, {N, N}, {N, N}}, 0, 0},
        {"3d boxes of 2 cells", 3, {8, 8, 4}, {2, 2, 2}, {{N, N}, {N, D}, {N, N}}, 0, 0},
        {"3d boxes of 1 cell in z", 3, {6, 6, 3}, {3, 3, 1}, {{N, N}, {N, N}, {N, N}}, 0, 0},
        {"3d middle rank of y slabs", 3, {8, 16, 6}, {8, 4, 6}, {{N, N}, {N, N}, {N, N}}, 4, 8},
        {"3d last rank of y slabs", 3, {8, 16, 6}, {4, 4, 3}, {{N, N}, {N, D}, {N, N}}, 12, 4},
        {"3d first rank, periodic y", 3, {8, 16, 6}, {8, 4, 6}, {{N, N}, {P, P}, {N, N}}, 0, 8},
        {"2d one box neumann", 2, {16, 10, 1}, {16, 10, 1}, {{N, N}, {N, N}, {0, 0}}, 0, 0},
        {"2d boxes mixed", 2, {16, 12, 1}, {4, 3, 1}, {{N, D}, {D, N}, {0, 0}}, 0, 0},
        {"2d periodic xy", 2, {12, 12, 1}, {6, 6, 1}, {{P, P}, {P, P}, {0, 0}}, 0, 0},
    };
    for (const Case& c : cases) {
        if (c.nd == 3) run_case<3>(c);
        else run_case<2>(c);
    }
    printf("%lld comparisons, %lld mismatches\n", checked, bad);
    return bad ? 1 : 0;
}
