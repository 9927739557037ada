// advect.cu -- the PPM / PLM / corner-transport-upwind scalar predictor of SOMAR's advection step (SURVEY.md 8(a)
// "MappedAdvectionUtil::predictScalar" row, 8(f)-3) on a FULLY PERIODIC level (BASELINE.json's TaylorGreen config), one
// rank or y-slab ranks.  FOUR launches in 3-D (three in 2-D; the reference makes 18 passes and round 1 made 18 launches):
//   k_adv_states   the normal predictor of ALL directions in one pass over the cells, nothing on faces stored:
//                  PPM  MappedAdvectionUtil::PPMNormalPred / PPMFaceValues   src/advection/AdvectUtil.cpp:1158-1418 (both
//                       branches: m_useHighOrderLimiter -> FourthInterpFaces, MappedGodunovUtilitiesF.ChF:517-609; else
//                       vanLeerSlopes (MappedGodunovUtilities.cpp:49-102; SecondSlopeDiffsF, VanLeerLimiterF,
//                       MappedGodunovUtilitiesF.ChF:35-95,151-186) + PPMFaceValuesF (:353-414)), for J and J W, W = JW / J
//                       -- a cell forms its own two faces (the neighbour forms the shared one again: same statements, same
//                       bits); PPMLimiterF (:315-351) or GETSECONDDIFF + CheckCubicLimiterF (:470-512, 611-745);
//                       PPMNormalPredF (:416-468)
//                  PLM  ::PLMNormalPred (:1066-1152), CTU ::CTUNormalPred (:1043-1057)
//                  + the half-step source term (AdvectUtil.cpp:699-705)
//   k_adv_riemann  RiemannSolver / AvgStates (AdvectUtilF.ChF:28-132) with the transverse corrections of
//                  quasilinearUpdate (QuasilinearUpdateF, AdvectUtilF.ChF:134-159) applied on the fly to both states;
//                  one launch per STAGE of predictScalar, a stage being a list of independent solves:
//                  the first solves of all directions (AdvectUtil.cpp:707-713), the six intermediate 3-D solves
//                  (:748-775), the final solves (:818-866) with "flux = state * u_adv" (:911)
// Every stage is a pure stencil, so on a periodic level the box decomposition does not matter and the kernels index the
// rank's brick with wrapped indices in x and z (and in y on one rank); the final states are also written to the periodic
// image in the first ghost slot, where somar_mac_divergence expects the high face of the last cell.
// Several ranks (y-slabs): the y stencils read the brick's ghost rows -- W and J are exchanged two rows deep (the
// second differences of the extremum-preserving limiter, which would need a third row, travel as a field of their own:
// k_adv_d2 + one exchange -- GETSECONDDIFF forms that field in the reference too), the stage results one row deep.
// The arithmetic is the Fortran's, statement for statement (-fmad=false): bit-identical to oracle/advect_ref.py on
// any number of ranks.
// Not built: non-periodic sides (they need the physics BC holders scalarRiemannBC / scalarSlopeBC; the reference's own
// branch carries TODO() markers at every BC call) -- the entry point rejects them.
#include "level.hpp"

namespace somar {

#define ATPB 128

struct AGeom {
    int n[3];
    int nd;
    int wrapY;              // one rank: y is wrapped like x and z; several ranks: the ghost rows hold the neighbours' cells
    long long sy, sz, off0;
};

__device__ __forceinline__ int awrap(int i, int n) { return i < 0 ? i + n : (i >= n ? i - n : i); }
__device__ __forceinline__ long long aidx(const AGeom& g, int i, int j, int k) {
    return g.off0 + awrap(i, g.n[0]) + g.sy * (g.wrapY ? awrap(j, g.n[1]) : j) + g.sz * awrap(k, g.n[2]);
}
// cell (i,j,k) shifted by o along direction d
__device__ __forceinline__ long long aoff(const AGeom& g, int i, int j, int k, int d, int o) {
    return aidx(g, i + (d == 0 ? o : 0), j + (d == 1 ? o : 0), k + (d == 2 ? o : 0));
}
// thread -> cell: x from the block and thread index, y and z from the grid (no per-cell integer division: the
// 64-bit div/mod of a flattened index cost more issue slots than the stencil arithmetic, cf. k_prolong)
__device__ __forceinline__ bool acell(const AGeom& g, int& i, int& j, int& k) {
    i = blockIdx.x * blockDim.x + threadIdx.x;
    j = blockIdx.y;
    k = blockIdx.z;
    return i < g.n[0];
}

// vanLeerSlopes at one cell from its neighbours along the direction
template <bool LIM>
__device__ __forceinline__ double vl_slope(double xm, double x0, double xp) {
    const double dWR = xp - x0;
    const double dWL = x0 - xm;
    double dWC = 0.5 * (dWR + dWL);
    if (LIM) {
        double lim = fmin(2.0 * fabs(dWL), 2.0 * fabs(dWR));
        lim = fmin(lim, fabs(dWC));
        dWC = (dWL * dWR < 0.0) ? 0.0 : lim * copysign(1.0, dWL);
    }
    return dWC;
}
// vanLeerSlopes with the utilities' m_highOrderLimiter set (MappedAdvectionUtil::define hands it useHighOrderLimiter,
// AdvectUtil.cpp:104): slopeLimiterExtPreserving -> ExtPreservingVanLeerLimiterF, MappedGodunovUtilitiesF.ChF:190-262
__device__ __forceinline__ double vl_slope_ext(double xm2, double xm1, double x0, double xp1, double xp2) {
    const double dWL = x0 - xm1, dWLL = xm1 - xm2, dWR = xp1 - x0, dWRR = xp2 - xp1;
    const double dWC = 0.5 * (dWR + dWL);
    const double dp1 = dWL * dWR, dp2 = dWLL * dWRR;
    double dWlim;
    if (fmin(dp1, dp2) < 0.0) {
        const double dW2L = dWL - dWLL;
        const double dW2C = (dWR - dWL) * 0.5;
        const double dW2R = dWRR - dWR;
        const double sign2 = copysign(1.0, dW2C);
        const double dW2lim = fmin(fmin(fabs(dW2C), fmax(sign2 * dW2L, 0.0)), fmax(sign2 * dW2R, 0.0));
        const double cvl = 5.0 * 0.25;
        const double dWvl = cvl * 3.0 * 0.5 * dW2lim;
        dWlim = (sign2 * dWC < 0.0) ? fmin(dWvl, 2.0 * fabs(dWL)) : fmin(dWvl, 2.0 * fabs(dWR));
    } else {
        dWlim = 2.0 * fmin(fabs(dWL), fabs(dWR));
    }
    return copysign(1.0, dWC) * fmin(fabs(dWC), dWlim);
}
// PPMFaceValuesF (centre box): the face between cells -1 and 0
template <bool LIM>
__device__ __forceinline__ double ppm_face(double xm2, double xm1, double x0, double xp1) {
    const double d0 = vl_slope<LIM>(xm1, x0, xp1);
    const double dm1 = vl_slope<LIM>(xm2, xm1, x0);
    const double wr = x0 - d0 / 3.0;
    const double wl = xm1 + dm1 / 3.0;
    return (wl + wr) * 0.5;
}

// FourthInterpFaces (centre box, MappedGodunovUtilitiesF.ChF:559-566): the face between cells -1 and 0
__device__ __forceinline__ double fourth_face(double xm2, double xm1, double x0, double xp1) {
    return (7.0 * (xm1 + x0) - (xm2 + xp1)) / 12.0;
}

// PPMFaceValues (AdvectUtil.cpp:1246-1418) at one face from the four cells around it: (J W)_face / J_face
// HO = m_useHighOrderLimiter: FourthInterpFaces replaces van Leer slopes + PPMFaceValuesF (:1269-1292,1345-1368)
template <bool LIM, bool HO>
__device__ __forceinline__ double w_face(const double* jj, const double* jw) {
    const double edgeJ = HO ? fourth_face(jj[0], jj[1], jj[2], jj[3]) : ppm_face<LIM>(jj[0], jj[1], jj[2], jj[3]);
    const double fw = HO ? fourth_face(jw[0], jw[1], jw[2], jw[3]) : ppm_face<LIM>(jw[0], jw[1], jw[2], jw[3]);
    return fw / edgeJ;
}

// CheckCubicLimiterF, inner centre box (MappedGodunovUtilitiesF.ChF:611-745) with limitC = 1.25, C3 = 0.1, eps = 1e-12
// (MappedGodunovUtilities::PPMLimiter, MappedGodunovUtilities.cpp:381-436).  WLL, WC, WRR = W(i-2), W(i), W(i+2);
// d2[q] = d2W(i - 2 + q), d2W(i) = (W(i+1) - W(i)) - (W(i) - W(i-1)) is GETSECONDDIFF (:470-512); d2Wfcf = (dWM + dWP) * 6.
__device__ __forceinline__ void cubic_limiter(double WLL, double WC, double WRR, const double d2[5], double& dm, double& dp) {
    const double limitC = 1.25, C3 = 0.1, eps = 1.0e-12;
    const double dWM = dm, dWP = dp;
    const bool bigM = fabs(dWM) > 2.0 * fabs(dWP);
    const bool bigP = fabs(dWP) > 2.0 * fabs(dWM);
    const double dWavgM = WC - WLL;
    const double dWavgP = WRR - WC;
    const double prodE1 = dWM * dWP;
    const double prodE2 = dWavgM * dWavgP;
    if (prodE1 >= 0.0 || prodE2 <= 0.0) {
        const double d2WL = d2[1], d2WC = d2[2], d2WR = d2[3];
        double atfcf = dWM;
        atfcf = atfcf + dWP;
        atfcf = atfcf * 6.0;
        double rho = 0.0;
        if (fabs(atfcf) >= eps) {
            const double sL = copysign(1.0, d2WL), sC = copysign(1.0, d2WC), sR = copysign(1.0, d2WR);
            const double sF = copysign(1.0, atfcf);
            if (sL == sC && sR == sC && sF == sC) {
                const double d2Wlim = sC * fmin(fmin(fabs(atfcf), limitC * fabs(d2WC)),
                                                fmin(limitC * fabs(d2WL), limitC * fabs(d2WR)));
                rho = d2Wlim / atfcf;
            }
        }
        if (rho < (1.0 - eps)) {
            const double d2WLL = d2[0], d2WRR = d2[4];
            const double d3WLL = d2WL - d2WLL;
            const double d3WL = d2WC - d2WL;
            const double d3WR = d2WR - d2WC;
            const double d3WRR = d2WRR - d2WR;
            const double d3Wmin = fmin(fmin(d3WLL, d3WL), fmin(d3WR, d3WRR));
            const double d3Wmax = fmax(fmax(d3WLL, d3WL), fmax(d3WR, d3WRR));
            const double prodD3 = C3 * fmax(fabs(d3Wmax), fabs(d3Wmin)) - fabs(d3Wmax - d3Wmin);
            if (prodD3 <= 0.0) {
                if (prodE1 > 0.0) {
                    dm = dWM * rho;
                    dp = dWP * rho;
                } else if (bigM) {
                    dm = dWM * rho - 2.0 * dWP * (1.0 - rho);
                } else if (bigP) {
                    dp = dWP * rho - 2.0 * dWM * (1.0 - rho);
                }
            }
        }
    } else {
        if (bigM) dm = -2.0 * dWP;
        if (bigP) dp = -2.0 * dWM;
    }
}

// GETSECONDDIFF along d as a field (several ranks: it is exchanged, so that the limiter sees d2W two rows beyond the slab)
__global__ void __launch_bounds__(ATPB) k_adv_d2(const AGeom g, int d, const double* __restrict__ W, double* __restrict__ out) {
    int i, j, k;
    if (!acell(g, i, j, k)) return;
    const long long c = aidx(g, i, j, k);
    const double w = W[c];
    out[c] = (W[aoff(g, i, j, k, d, 1)] - w) - (w - W[aoff(g, i, j, k, d, -1)]);
}

struct AStates {
    const double* vel[3];   // cell-centred velocity component of each direction
    double* wminus[3];
    double* wplus[3];
    double dtbydx[3];
    int plmExt;             // PLM: the van Leer slopes are limited by the extremum-preserving limiter (useLimiting && useHighOrderLimiter)
};

// The normal predictor of every direction at one cell.  ORDER 2: PPM (X = m_useHighOrderLimiter); ORDER 1: PLM
// (X = m_useFourthOrderSlopes; the second limiter pass of AdvectUtil.cpp:1126-1128 quirk included: without X it is handed
// zeros as one-sided differences, so use_limiting alone flattens every slope); ORDER 0: CTU (both states are W).
// d2y != nullptr: second differences along y come from that (exchanged) field instead of W(j-3 .. j+3).
template <int ORDER, bool LIM, bool X>
__global__ void __launch_bounds__(ATPB) k_adv_states(const AGeom g, const double* __restrict__ W,
                                                     const double* __restrict__ J, const double* __restrict__ src,
                                                     double srcScale, const double* __restrict__ d2y, AStates a) {
    int i, j, k;
    if (!acell(g, i, j, k)) return;
    const long long c = aidx(g, i, j, k);
    const double w = W[c];
    const double s = src ? src[c] * srcScale : 0.0;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        if (d >= g.nd) break;
        double nm = w, np = w;
        if (ORDER == 2) {
            double ww[5], jj[5], jw[5];
#pragma unroll
            for (int o = 0; o < 5; ++o) {
                const long long q = (o == 2) ? c : aoff(g, i, j, k, d, o - 2);
                ww[o] = (o == 2) ? w : W[q];
                jj[o] = J[q];
                jw[o] = ww[o] * jj[o];
            }
            double dm = (0.0 - w) + w_face<LIM, X>(jj, jw);             // low face: cells -2 .. +1
            double dp = (0.0 - w) + w_face<LIM, X>(jj + 1, jw + 1);     // high face: cells -1 .. +2
            if (LIM && X) {
                double d2[5];
                if (d == 1 && d2y) {
#pragma unroll
                    for (int q = 0; q < 5; ++q) d2[q] = d2y[aoff(g, i, j, k, 1, q - 2)];
                } else {
                    const double wm3 = W[aoff(g, i, j, k, d, -3)], wp3 = W[aoff(g, i, j, k, d, 3)];
                    const double w7[7] = {wm3, ww[0], ww[1], ww[2], ww[3], ww[4], wp3};
#pragma unroll
                    for (int q = 0; q < 5; ++q) d2[q] = (w7[q + 2] - w7[q + 1]) - (w7[q + 1] - w7[q]);
                }
                cubic_limiter(ww[0], w, ww[4], d2, dm, dp);
            } else if (LIM) {   // PPMLimiterF
                if (dm * dp < 0.0) {
                    const double dwc = (dp + dm) * 0.5;
                    const double d2w = dp - dm;
                    const double sg = copysign(1.0, dwc);
                    if (dwc * d2w > 0.0) dp = sg * fmin(-2.0 * sg * dm, sg * dp);
                    else dm = sg * fmin(sg * dm, -2.0 * sg * dp);
                } else {
                    dp = 0.0;
                    dm = 0.0;
                }
            }
            // PPMNormalPredF, one slope
            const double lam = a.vel[d][c];
            const double dtbydx = a.dtbydx[d];
            double sigmin = lam * dtbydx;
            sigmin = fmin(sigmin, 0.0);
            sigmin = -sigmin;
            double sigmax = lam * dtbydx;
            sigmax = fmax(sigmax, 0.0);
            double sigM, sigP;
            if (lam > 0.0) { sigM = sigmin; sigP = lam * dtbydx; }
            else { sigM = -lam * dtbydx; sigP = sigmax; }
            nm = dm + sigM * ((dp - dm) - (dp + dm) * (3.0 - 2.0 * sigM)) * 0.5;
            np = dp + sigP * ((dm - dp) - (dp + dm) * (3.0 - 2.0 * sigP)) * 0.5;
            nm = nm + w;
            np = np + w;
        } else if (ORDER == 1) {
            const double wm1 = W[aoff(g, i, j, k, d, -1)], wp1 = W[aoff(g, i, j, k, d, 1)];
            double dW, dWm = 0.0, dWp = 0.0;
            const bool ext = LIM && a.plmExt;
            if (X) {   // FourthSlopeDiffsF (MappedGodunovUtilitiesF.ChF:96-146) + one-sided differences for the limiter
                const double wm2 = W[aoff(g, i, j, k, d, -2)], wp2 = W[aoff(g, i, j, k, d, 2)];
                double vlP, vlM;
                if (ext) {
                    const double wm3 = W[aoff(g, i, j, k, d, -3)], wp3 = W[aoff(g, i, j, k, d, 3)];
                    vlP = vl_slope_ext(wm1, w, wp1, wp2, wp3);
                    vlM = vl_slope_ext(wm3, wm2, wm1, w, wp1);
                } else {
                    vlP = vl_slope<LIM>(w, wp1, wp2);
                    vlM = vl_slope<LIM>(wm2, wm1, w);
                }
                const double dWR = wp1 - vlP * 0.25;
                const double dWL = wm1 + vlM * 0.25;
                dW = (2.0 / 3.0) * (dWR - dWL);
                dWm = w - wm1;
                dWp = wp1 - w;
            } else if (ext) {
                dW = vl_slope_ext(W[aoff(g, i, j, k, d, -2)], wm1, w, wp1, W[aoff(g, i, j, k, d, 2)]);
            } else {
                dW = vl_slope<LIM>(wm1, w, wp1);
            }
            if (LIM || X) {   // VanLeerLimiterF
                double lim = fmin(2.0 * fabs(dWm), 2.0 * fabs(dWp));
                lim = fmin(lim, fabs(dW));
                dW = (dWm * dWp < 0.0) ? 0.0 : lim * copysign(1.0, dWm);
            }
            // PLMNormalPredF (:266-311), one slope
            const double lam = a.vel[d][c];
            const double dtbydx = a.dtbydx[d];
            const double lmin = fmin(lam, 0.0), lmax = fmax(lam, 0.0);
            const double loFac = (lam < 0.0) ? lam : lmin;
            const double hiFac = (lam > 0.0) ? lam : lmax;
            const double dm = dW * (-0.5) * (1.0 + dtbydx * loFac);
            const double dp = dW * 0.5 * (1.0 - dtbydx * hiFac);
            nm = dm + w;
            np = dp + w;
        }
        if (src) {
            nm = nm + s;
            np = np + s;
        }
        a.wminus[d][c] = nm;
        a.wplus[d][c] = np;
    }
}

// up to two transverse corrections  A = scale * u_d2 * (WHalf(i + e_d2) - WHalf(i))  added to both states, in order
struct ACorr {
    const double* wh[2];
    const double* vel[2];
    double scale[2];
    int dir[2];
    int n;
};
// one Riemann solve of a stage: faces of direction d from the states of the cells either side
struct ATask {
    const double* wplus;
    const double* wminus;
    const double* advVel;
    double* out;
    ACorr cr;
    int d, timesVel, image;
};
struct AStage {
    ATask t[6];
    int n, upwind;
};

__global__ void __launch_bounds__(ATPB) k_adv_riemann(const AGeom g, const AStage st) {
    int i, j, k;
    if (!acell(g, i, j, k)) return;
    const long long cR = aidx(g, i, j, k);
    for (int q = 0; q < st.n; ++q) {
        const ATask& T = st.t[q];
        const int d = T.d;
        const int iL = i - (d == 0), jL = j - (d == 1), kL = k - (d == 2);
        const long long cL = aidx(g, iL, jL, kL);
        double wl = T.wplus[cL];
        double wr = T.wminus[cR];
        for (int t = 0; t < T.cr.n; ++t) {
            const int d2 = T.cr.dir[t];
            const double aL = T.cr.scale[t] * T.cr.vel[t][cL] * (T.cr.wh[t][aoff(g, iL, jL, kL, d2, 1)] - T.cr.wh[t][cL]);
            const double aR = T.cr.scale[t] * T.cr.vel[t][cR] * (T.cr.wh[t][aoff(g, i, j, k, d2, 1)] - T.cr.wh[t][cR]);
            wl = wl + aL;
            wr = wr + aR;
        }
        const double u = T.advVel[cR];
        double v;
        if (!st.upwind) v = 0.5 * (wl + wr);
        else if (u > 1.0e-9) v = wl;
        else if (u < -1.0e-9) v = wr;
        else v = 0.5 * (wl + wr);
        if (T.timesVel) v = v * u;
        T.out[cR] = v;
        if (T.image) {   // the periodic image of face 0 = the high face of the last cell (first ghost slot)
            const int p = (d == 0) ? i : (d == 1 ? j : k);
            if (p == 0) T.out[cR + (d == 0 ? (long long)g.n[0] : (d == 1 ? g.sy * g.n[1] : g.sz * g.n[2]))] = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
int advect_predict_scalar(SomarLevel* L, SomarField* const whalf[3], const SomarField* wold, const SomarField* source,
                          const SomarField* const oldVel[3], const SomarField* const advVel[3], const SomarField* ccJ,
                          double dt, const SomarAdvectParams* prm) {
    const Geom& G = L->depths[0].g;
    const int nd = G.nd;
    const bool multi = L->desc.nranks > 1;
    for (int d = 0; d < nd; ++d)
        if (!G.periodic[d]) {
            set_error("predictScalar: only fully periodic levels are built (non-periodic sides need the physics BC holders "
                      "scalarRiemannBC / scalarSlopeBC, AdvectUtil.cpp:715-740)");
            return SOMAR_ERR_ARG;
        }
    const int order = prm->normal_pred_order;
    if (order < 0 || order > 2) { set_error("predictScalar: the normal predictor order must be 0 (CTU), 1 (PLM) or 2 (PPM)"); return SOMAR_ERR_ARG; }
    if (order == 2 && prm->use_high_order_limiter)
        for (int d = 0; d < nd; ++d)
            if (G.n[d] < 4) { set_error("predictScalar: the high-order limiter needs at least 4 cells per direction"); return SOMAR_ERR_ARG; }
    if (multi && G.n[1] < 2) { set_error("predictScalar: a rank needs at least two rows"); return SOMAR_ERR_ARG; }
    AGeom g;
    for (int d = 0; d < 3; ++d) g.n[d] = G.n[d];
    g.nd = nd;
    g.wrapY = multi ? 0 : 1;
    g.sy = G.sy; g.sz = G.sz; g.off0 = G.off0;
    const size_t need = ((nd == 3) ? 15 : 6) + 1;   // WMinus, WPlus, WHalf1 per direction (+ six WHalf2 in 3-D) + d2W
    int rc;
    while (L->awork.size() < need) {
        SomarField* f = nullptr;
        if ((rc = field_alloc(L, 0, &f)) != SOMAR_OK) return rc;
        L->awork.push_back(f);
    }
    double *wm[3], *wp[3], *wh1[3], *wh2[3][3] = {};
    for (int d = 0; d < nd; ++d) { wm[d] = L->awork[d]->p; wp[d] = L->awork[nd + d]->p; wh1[d] = L->awork[2 * nd + d]->p; }
    if (nd == 3) {
        int q = 9;
        for (int d1 = 0; d1 < 3; ++d1)
            for (int d2 = 0; d2 < 3; ++d2)
                if (d2 != d1) wh2[d1][d2] = L->awork[q++]->p;
    }
    double* const d2w = L->awork[need - 1]->p;
    const dim3 nb((g.n[0] + ATPB - 1) / ATPB, g.n[1], g.n[2]);
    cudaStream_t st = L->cx.stream;
    const bool lim = prm->use_limiting != 0;
    const bool ho = prm->use_high_order_limiter != 0;
    const bool fourth = prm->use_fourth_order_slopes != 0;
    double dtondx[3] = {0.0, 0.0, 0.0};
    for (int d = 0; d < nd; ++d) dtondx[d] = dt / G.dx[d];
    // several ranks: the neighbours' rows of what a stage reads (the inputs' ghost rows are refreshed in place)
    auto halo = [&](const double* p, int width) { return multi ? exchange(L, 0, const_cast<double*>(p), width) : SOMAR_OK; };

    // ---- normal predictor, all directions
    const double* d2y = nullptr;
    if ((rc = halo(wold->p, 2)) != SOMAR_OK) return rc;
    if (order == 2 && (rc = halo(ccJ->p, 2)) != SOMAR_OK) return rc;
    if (multi && order == 2 && lim && ho) {
        k_adv_d2<<<nb, ATPB, 0, st>>>(g, 1, wold->p, d2w);
        L->cx.launches++;
        if ((rc = halo(d2w, 2)) != SOMAR_OK) return rc;
        d2y = d2w;
    }
    AStates a;
    for (int d = 0; d < 3; ++d) {
        a.vel[d] = d < nd ? oldVel[d]->p : nullptr;
        a.wminus[d] = d < nd ? wm[d] : nullptr;
        a.wplus[d] = d < nd ? wp[d] : nullptr;
        a.dtbydx[d] = dtondx[d];
    }
    a.plmExt = (order == 1 && lim && ho) ? 1 : 0;
    if (multi && a.plmExt && fourth) {
        set_error("predictScalar: PLM with fourth-order slopes and the extremum-preserving limiter reaches three rows: "
                  "one rank only");
        return SOMAR_ERR_ARG;
    }
    const double* src = source ? source->p : nullptr;
#define ADV_STATES(ORDER_, LIM_, X_) \
    k_adv_states<ORDER_, LIM_, X_><<<nb, ATPB, 0, st>>>(g, wold->p, ccJ->p, src, 0.5 * dt, d2y, a)
    if (order == 0) ADV_STATES(0, false, false);
    else if (order == 1 && lim && fourth) ADV_STATES(1, true, true);
    else if (order == 1 && lim) ADV_STATES(1, true, false);
    else if (order == 1 && fourth) ADV_STATES(1, false, true);
    else if (order == 1) ADV_STATES(1, false, false);
    else if (lim && ho) ADV_STATES(2, true, true);
    else if (lim) ADV_STATES(2, true, false);
    else if (ho) ADV_STATES(2, false, true);
    else ADV_STATES(2, false, false);
#undef ADV_STATES
    L->cx.launches++;
    if (multi) {
        // a y face reads the states of the row below; the transverse corrections read the velocities there
        if ((rc = halo(wp[1], 1)) != SOMAR_OK || (rc = halo(wm[1], 1)) != SOMAR_OK) return rc;
        for (int d = 0; d < nd; ++d)
            if ((rc = halo(oldVel[d]->p, 1)) != SOMAR_OK) return rc;
    }

    const ACorr none = {{nullptr, nullptr}, {nullptr, nullptr}, {0.0, 0.0}, {0, 0}, 0};
    auto task = [&](int d, const ACorr& c, double* out, int timesVel, int image) {
        ATask T;
        T.wplus = wp[d]; T.wminus = wm[d]; T.advVel = advVel[d]->p; T.out = out; T.cr = c;
        T.d = d; T.timesVel = timesVel; T.image = image;
        return T;
    };
    // ---- first Riemann solves (:707-713)
    AStage S;
    S.upwind = prm->use_upwinding != 0;
    S.n = 0;
    for (int d = 0; d < nd; ++d) S.t[S.n++] = task(d, none, wh1[d], 0, 0);
    k_adv_riemann<<<nb, ATPB, 0, st>>>(g, S);
    L->cx.launches++;
    for (int d = 0; d < nd && multi; ++d)
        if ((rc = halo(wh1[d], 1)) != SOMAR_OK) return rc;
    // ---- 3-D: the six intermediate solves, each corrected by the first solve of one other direction (:748-775)
    if (nd == 3) {
        S.n = 0;
        for (int d1 = 0; d1 < 3; ++d1)
            for (int d2 = 0; d2 < 3; ++d2) {
                if (d2 == d1) continue;
                ACorr c = none;
                c.n = 1;
                c.wh[0] = wh1[d2]; c.vel[0] = oldVel[d2]->p; c.scale[0] = -(1.0 / 3.0) * dtondx[d2]; c.dir[0] = d2;
                S.t[S.n++] = task(d1, c, wh2[d1][d2], 0, 0);
            }
        k_adv_riemann<<<nb, ATPB, 0, st>>>(g, S);
        L->cx.launches++;
        for (int d1 = 0; d1 < 3 && multi; ++d1)
            for (int d2 = 0; d2 < 3; ++d2)
                if (d2 != d1 && (rc = halo(wh2[d1][d2], 1)) != SOMAR_OK) return rc;
    }
    // ---- final solves (:818-866), flux = state * u_adv (:911)
    S.n = 0;
    for (int d1 = 0; d1 < nd; ++d1) {
        ACorr c = none;
        for (int d2 = 0; d2 < nd; ++d2) {
            if (d2 == d1) continue;
            const int t = c.n++;
            c.wh[t] = (nd == 2) ? wh1[d2] : wh2[d2][3 - d1 - d2];
            c.vel[t] = oldVel[d2]->p;
            c.scale[t] = -(1.0 / 2.0) * dtondx[d2];
            c.dir[t] = d2;
        }
        S.t[S.n++] = task(d1, c, whalf[d1]->p, prm->return_flux != 0, (multi && d1 == 1) ? 0 : 1);
    }
    k_adv_riemann<<<nb, ATPB, 0, st>>>(g, S);
    L->cx.launches++;
    if (multi && (rc = halo(whalf[1]->p, 1)) != SOMAR_OK) return rc;      // the high face of the slab's last row
    SOMAR_CUDA(cudaGetLastError());
    return SOMAR_OK;
}

// The finite-volume update of getNewScalar (src/NavierStokes/AMRNavierStokesAdvancePPM.cpp:561-584, no background,
// sponge or diffusive terms): adv = Div[flux] (MAPPEDFLUXDIVERGENCE) * -1.0;  new = old + adv * dt.
template <int ND>
__global__ void __launch_bounds__(ATPB) k_adv_update(const Geom g, const double* __restrict__ jinv,
                                                     double* __restrict__ wnew, const double* __restrict__ wold,
                                                     const double* __restrict__ f0, const double* __restrict__ f1,
                                                     const double* __restrict__ f2, double dt) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= g.cells()) return;
    const int i = (int)(t % g.n[0]);
    const long long r = t / g.n[0];
    const int j = (int)(r % g.n[1]), k = (int)(r / g.n[1]);
    const long long c = g.idx(i, j, k);
    double s = (f0[c + 1] - f0[c]) * g.rdx[0];
    s = s + (f1[c + g.sy] - f1[c]) * g.rdx[1];
    if (ND > 2) s = s + (f2[c + g.sz] - f2[c]) * g.rdx[2];
    double adv = jinv[c] * s;
    adv = adv * -1.0;
    wnew[c] = wold[c] + adv * dt;
}

int advect_update(SomarLevel* L, SomarField* wnew, const SomarField* wold, const SomarField* const flux[3], double dt) {
    const Depth& dp = L->depths[0];
    const Geom& g = dp.g;
    const unsigned nb = (unsigned)((g.cells() + ATPB - 1) / ATPB);
    if (g.nd == 3) k_adv_update<3><<<nb, ATPB, 0, L->cx.stream>>>(g, dp.m.jinv, wnew->p, wold->p, flux[0]->p, flux[1]->p, flux[2]->p, dt);
    else k_adv_update<2><<<nb, ATPB, 0, L->cx.stream>>>(g, dp.m.jinv, wnew->p, wold->p, flux[0]->p, flux[1]->p, nullptr, dt);
    L->cx.launches++;
    return SOMAR_OK;
}

}  // namespace somar
