// tests/cuda/statements_host.cu -- TEST INFRASTRUCTURE.  The product's full-metric cell statements
// (somar_b200/csrc/stencil_full.cuh, __host__ __device__) compiled for the HOST behind a small C ABI, so that
// tests/test_statements_host.py can compare them with the CPU oracle without a GPU: the operator and the red-black
// smoother on a one-rank level, dense arrays in, dense arrays out (also the diagonal-metric statements of
// stencil_ortho.cuh: fh_apply_ortho / fh_relax_ortho).  `accessor` 0 runs the statements through ExView
// (ex_slow), 1 through the accessor cell_kind selects (what the rim kernels and the persistent bottom kernel run).
// Built by the test:  nvcc -shared -Xcompiler -fPIC -O1 -std=c++17 -fmad=false -I somar_b200/csrc ...
#include <cstring>
#include <vector>

#include "stencil_full.cuh"

using namespace somar;

namespace {
const int GH = 4;

struct HostLevel {
    Geom g;
    Metric m;
    std::vector<double> jg[3][3], jinv, lap;
};

// dense (nz, ny, nx) -> ghosted brick
void scatter(const Geom& g, const double* dense, std::vector<double>& out) {
    out.assign(g.total, 0.0);
    for (int k = 0; k < g.n[2]; ++k)
        for (int j = 0; j < g.n[1]; ++j)
            for (int i = 0; i < g.n[0]; ++i) out[g.idx(i, j, k)] = dense[((long long)k * g.n[1] + j) * g.n[0] + i];
}
// face array of direction d: dense (nz + [d==2], ny + [d==1], nx + [d==0]); the face at index q is the LOW face of cell q
void scatter_face(const Geom& g, const double* dense, int d, std::vector<double>& out) {
    out.assign(g.total, 0.0);
    const int e[3] = {d == 0, d == 1, d == 2};
    const int fx = g.n[0] + e[0], fy = g.n[1] + e[1], fz = g.n[2] + ((g.nd > 2) ? e[2] : 0);
    for (int k = 0; k < fz; ++k)
        for (int j = 0; j < fy; ++j)
            for (int i = 0; i < fx; ++i) out[g.idx(i, j, k)] = dense[((long long)k * fy + j) * fx + i];
}
// one ghost layer of every periodic direction, x then y then z over the already extended range (the corners follow)
void periodic_fill(const Geom& g, std::vector<double>& a) {
    for (int d = 0; d < g.nd; ++d) {
        if (!g.periodic[d]) continue;
        int lo[3], hi[3];
        for (int e = 0; e < 3; ++e) {
            const bool ext = e < g.nd && e != d;
            lo[e] = ext ? -1 : 0;
            hi[e] = (e == d) ? 0 : (ext ? g.n[e] : g.n[e] - 1);
        }
        for (int k = lo[2]; k <= hi[2]; ++k)
            for (int j = lo[1]; j <= hi[1]; ++j)
                for (int i = lo[0]; i <= hi[0]; ++i) {
                    int q[3] = {i, j, k};
                    int src[3] = {i, j, k};
                    q[d] = -1; src[d] = g.n[d] - 1;
                    a[g.idx(q[0], q[1], q[2])] = a[g.idx(src[0], src[1], src[2])];
                    q[d] = g.n[d]; src[d] = 0;
                    a[g.idx(q[0], q[1], q[2])] = a[g.idx(src[0], src[1], src[2])];
                }
    }
}

void build(HostLevel& L, int nd, const int* n, const int* bs, const int* bc_type, const double* bc_value, const int* periodic,
           const double* dx, double alpha, double beta, const double* const* jg /* [d * nd + e] */, const double* jinv,
           const double* lap) {
    Geom& g = L.g;
    std::memset(&g, 0, sizeof(g));
    g.nd = nd;
    for (int d = 0; d < 3; ++d) {
        const bool act = d < nd;
        g.n[d] = g.gn[d] = act ? n[d] : 1;
        g.bs[d] = act ? bs[d] : 1;
        g.gh[d] = GH;
        g.dx[d] = act ? dx[d] : 1.0;
        g.rdx[d] = 1.0 / g.dx[d];
        g.rdx2[d] = 1.0 / (g.dx[d] * g.dx[d]);
        g.periodic[d] = act && periodic[d];
        for (int s = 0; s < 2; ++s) {
            g.bct[d][s] = !act ? BC_NONE : (g.periodic[d] ? BC_PERIODIC : bc_type[2 * d + s]);
            g.bc[d][s] = (act && !g.periodic[d]) ? bc_type[2 * d + s] : BC_NONE;
            g.bcval[d][s] = bc_value[2 * d + s];
        }
    }
    g.sy = g.n[0] + 2 * GH;
    g.sz = g.sy * (g.n[1] + 2 * GH);
    g.off0 = GH + g.sy * GH + g.sz * GH;
    g.total = g.sz * (g.n[2] + 2 * GH);
    g.alpha = alpha;
    g.beta = beta;
    g.diriOrder = 1;
    g.nbr[0] = g.nbr[1] = -1;
    std::memset(&L.m, 0, sizeof(L.m));
    for (int d = 0; d < nd; ++d)
        for (int e = 0; e < nd; ++e) {
            scatter_face(g, jg[d * nd + e], d, L.jg[d][e]);
            L.m.jg[d][e] = L.jg[d][e].data();
        }
    scatter(g, jinv, L.jinv);
    scatter(g, lap, L.lap);
    L.m.jinv = L.jinv.data();
    L.m.lapdiag = L.lap.data();
    L.m.diagonal = 0;
}

template <int ND>
void apply_all(const HostLevel& L, const std::vector<double>& phi, double* outDense, bool homog, int accessor) {
    const Geom& g = L.g;
    for (int k = 0; k < g.n[2]; ++k)
        for (int j = 0; j < g.n[1]; ++j)
            for (int i = 0; i < g.n[0]; ++i) {
                const long long c = g.idx(i, j, k);
                const double v = accessor ? apply_cell_full_kind<ND>(g, L.m, phi.data(), i, j, k, c, homog, cell_kind<ND>(g, i, j, k))
                                          : apply_cell_full<ND>(g, L.m, phi.data(), i, j, k, c, homog);
                outDense[((long long)k * g.n[1] + j) * g.n[0] + i] = v;
            }
}

template <int ND>
void gsrb_pass(const HostLevel& L, std::vector<double>& phi, const std::vector<double>& rhs, int pass, int accessor) {
    const Geom& g = L.g;
    std::vector<double> out(phi);
    for (int k = 0; k < g.n[2]; ++k)
        for (int j = 0; j < g.n[1]; ++j)
            for (int i = 0; i < g.n[0]; ++i) {
                if (((i + j + ((ND > 2) ? k : 0) + pass) & 1) != 0) continue;
                const long long c = g.idx(i, j, k);
                out[c] = accessor ? gsrb_cell_full_kind<ND>(g, L.m, phi.data(), rhs.data(), i, j, k, c, cell_kind<ND>(g, i, j, k))
                                  : gsrb_cell_full<ND>(g, L.m, phi.data(), rhs.data(), i, j, k, c);
            }
    phi.swap(out);
}

// diagonal metric: apply_cell_ortho, and gsrb_cell_ortho IN PLACE (a cell only reads the other colour)
template <int ND>
void apply_all_ortho(const HostLevel& L, const std::vector<double>& phi, double* outDense, bool homog) {
    const Geom& g = L.g;
    for (int k = 0; k < g.n[2]; ++k)
        for (int j = 0; j < g.n[1]; ++j)
            for (int i = 0; i < g.n[0]; ++i)
                outDense[((long long)k * g.n[1] + j) * g.n[0] + i] = apply_cell_ortho<ND>(g, L.m, phi.data(), i, j, k, g.idx(i, j, k), homog);
}
template <int ND>
void gsrb_pass_ortho(const HostLevel& L, std::vector<double>& phi, const std::vector<double>& rhs, int pass) {
    const Geom& g = L.g;
    for (int k = 0; k < g.n[2]; ++k)
        for (int j = 0; j < g.n[1]; ++j)
            for (int i = 0; i < g.n[0]; ++i) {
                if (((i + j + ((ND > 2) ? k : 0) + pass) & 1) != 0) continue;
                const long long c = g.idx(i, j, k);
                phi[c] = gsrb_cell_ortho<ND>(g, L.m, phi.data(), rhs.data(), i, j, k, c);
            }
}
}  // namespace

extern "C" {

// the same two for a diagonal metric: jg holds nd arrays (J g^{dd} on the faces of direction d)
int fh_apply_ortho(int nd, const int* n, const int* bs, const int* bc_type, const double* bc_value, const int* periodic,
                   const double* dx, double alpha, double beta, const double* const* jg, const double* jinv,
                   const double* lap, const double* phi, double* out, int homog) {
    HostLevel L;
    std::vector<const double*> full(nd * nd, nullptr);
    for (int d = 0; d < nd; ++d)
        for (int e = 0; e < nd; ++e) full[d * nd + e] = jg[d];   // (off-diagonal slots are not read)
    build(L, nd, n, bs, bc_type, bc_value, periodic, dx, alpha, beta, full.data(), jinv, lap);
    L.m.diagonal = 1;
    std::vector<double> p;
    scatter(L.g, phi, p);
    periodic_fill(L.g, p);
    if (nd == 3) apply_all_ortho<3>(L, p, out, homog != 0);
    else apply_all_ortho<2>(L, p, out, homog != 0);
    return 0;
}
int fh_relax_ortho(int nd, const int* n, const int* bs, const int* bc_type, const double* bc_value, const int* periodic,
                   const double* dx, double alpha, double beta, const double* const* jg, const double* jinv,
                   const double* lap, double* phi, const double* rhs, int iters) {
    HostLevel L;
    std::vector<const double*> full(nd * nd, nullptr);
    for (int d = 0; d < nd; ++d)
        for (int e = 0; e < nd; ++e) full[d * nd + e] = jg[d];
    build(L, nd, n, bs, bc_type, bc_value, periodic, dx, alpha, beta, full.data(), jinv, lap);
    L.m.diagonal = 1;
    std::vector<double> p, r;
    scatter(L.g, phi, p);
    scatter(L.g, rhs, r);
    for (int it = 0; it < iters; ++it)
        for (int pass = 0; pass < 2; ++pass) {
            periodic_fill(L.g, p);
            if (nd == 3) gsrb_pass_ortho<3>(L, p, r, pass);
            else gsrb_pass_ortho<2>(L, p, r, pass);
        }
    const Geom& g = L.g;
    for (int k = 0; k < g.n[2]; ++k)
        for (int j = 0; j < g.n[1]; ++j)
            for (int i = 0; i < g.n[0]; ++i) phi[((long long)k * g.n[1] + j) * g.n[0] + i] = p[g.idx(i, j, k)];
    return 0;
}

// out = (alpha + beta L) phi, applyOpI with its exchange (here: the periodic ghost fill; boxes share one array)
int fh_apply(int nd, const int* n, const int* bs, const int* bc_type, const double* bc_value, const int* periodic,
             const double* dx, double alpha, double beta, const double* const* jg, const double* jinv, const double* lap,
             const double* phi, double* out, int homog, int accessor) {
    HostLevel L;
    build(L, nd, n, bs, bc_type, bc_value, periodic, dx, alpha, beta, jg, jinv, lap);
    std::vector<double> p;
    scatter(L.g, phi, p);
    periodic_fill(L.g, p);
    if (nd == 3) apply_all<3>(L, p, out, homog != 0, accessor);
    else apply_all<2>(L, p, out, homog != 0, accessor);
    return 0;
}

// `iters` LevelGSRB iterations in place (exchange, red pass, exchange, black pass -- GSRB.cpp:58-98)
int fh_relax(int nd, const int* n, const int* bs, const int* bc_type, const double* bc_value, const int* periodic,
             const double* dx, double alpha, double beta, const double* const* jg, const double* jinv, const double* lap,
             double* phi, const double* rhs, int iters, int accessor) {
    HostLevel L;
    build(L, nd, n, bs, bc_type, bc_value, periodic, dx, alpha, beta, jg, jinv, lap);
    std::vector<double> p, r;
    scatter(L.g, phi, p);
    scatter(L.g, rhs, r);
    for (int it = 0; it < iters; ++it)
        for (int pass = 0; pass < 2; ++pass) {
            periodic_fill(L.g, p);
            if (nd == 3) gsrb_pass<3>(L, p, r, pass, accessor);
            else gsrb_pass<2>(L, p, r, pass, accessor);
        }
    const Geom& g = L.g;
    for (int k = 0; k < g.n[2]; ++k)
        for (int j = 0; j < g.n[1]; ++j)
            for (int i = 0; i < g.n[0]; ++i) phi[((long long)k * g.n[1] + j) * g.n[0] + i] = p[g.idx(i, j, k)];
    return 0;
}

}  // extern "C"
