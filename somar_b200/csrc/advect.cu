// advect.cu -- the PPM / corner-transport-upwind scalar predictor of SOMAR's advection step (SURVEY.md 8(a)
// "MappedAdvectionUtil::predictScalar" row, 8(f)-3) on a FULLY PERIODIC level (BASELINE.json's TaylorGreen config):
//   k_adv_face     MappedAdvectionUtil::PPMFaceValues   src/advection/AdvectUtil.cpp:1246-1418 (both branches:
//                  m_useHighOrderLimiter -> FourthInterpFaces, MappedGodunovUtilitiesF.ChF:517-609; else)
//                  -> vanLeerSlopes (MappedGodunovUtilities.cpp:49-102; SecondSlopeDiffsF, VanLeerLimiterF,
//                  MappedGodunovUtilitiesF.ChF:35-95,151-186) + PPMFaceValuesF (:353-414), for J and J W, then W = JW / J
//   k_adv_normal   MappedAdvectionUtil::PPMNormalPred   AdvectUtil.cpp:1158-1240 -> PPMLimiterF (:315-351) or, with the
//                  high-order limiter, GETSECONDDIFF + CheckCubicLimiterF (:470-512, 611-745),
//                  PPMNormalPredF (:416-468), + the half-step source term (AdvectUtil.cpp:699-705)
//   k_adv_riemann  RiemannSolver / AvgStates (AdvectUtilF.ChF:28-132) with the transverse corrections of
//                  quasilinearUpdate (QuasilinearUpdateF, AdvectUtilF.ChF:134-159) applied on the fly to both states:
//                  the three uses in predictScalar (AdvectUtil.cpp:707-713, 748-775, 818-866) and "flux = state * u_adv" (:911)
// Every stage is a pure stencil, so on a periodic level the box decomposition does not matter and the kernels
// index the rank's brick with wrapped indices (no ghost cells are read); the final states are also written to the
// periodic image in the first ghost slot, where somar_mac_divergence expects the high face of the last cell.
// The arithmetic is the Fortran's, statement for statement (-fmad=false): bit-identical to oracle/advect_ref.py.
//   k_adv_normal_low   the other two normal predictors: CTUNormalPred (:1043-1057) and PLMNormalPred (:1066-1152)
// Not built: non-periodic sides (they need the physics BC holders scalarRiemannBC / scalarSlopeBC), more than one rank
// -- the entry point rejects them.
// HBM-bound streaming passes (18 launches in 3-D); intermediate states live in scratch fields, not yet fused.
#include "level.hpp"

namespace somar {

#define ATPB 128

struct AGeom {
    int n[3];
    int nd;
    long long sy, sz, off0;
};

__device__ __forceinline__ int awrap(int i, int n) { return i < 0 ? i + n : (i >= n ? i - n : i); }
__device__ __forceinline__ long long aidx(const AGeom& g, int i, int j, int k) {
    return g.off0 + awrap(i, g.n[0]) + g.sy * awrap(j, g.n[1]) + g.sz * awrap(k, g.n[2]);
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

// HO = m_useHighOrderLimiter: FourthInterpFaces replaces van Leer slopes + PPMFaceValuesF (AdvectUtil.cpp:1269-1292,1345-1368)
template <bool LIM, bool HO>
__global__ void __launch_bounds__(ATPB) k_adv_face(const AGeom g, int d, const double* __restrict__ W,
                                                   const double* __restrict__ J, double* __restrict__ wface) {
    int i, j, k;
    if (!acell(g, i, j, k)) return;
    double jj[4], jw[4];
#pragma unroll
    for (int o = 0; o < 4; ++o) {
        const long long c = aoff(g, i, j, k, d, o - 2);
        jj[o] = J[c];
        jw[o] = W[c] * jj[o];
    }
    const double edgeJ = HO ? fourth_face(jj[0], jj[1], jj[2], jj[3]) : ppm_face<LIM>(jj[0], jj[1], jj[2], jj[3]);
    const double fw = HO ? fourth_face(jw[0], jw[1], jw[2], jw[3]) : ppm_face<LIM>(jw[0], jw[1], jw[2], jw[3]);
    wface[aidx(g, i, j, k)] = fw / edgeJ;
}

// CheckCubicLimiterF, inner centre box (MappedGodunovUtilitiesF.ChF:611-745) with limitC = 1.25, C3 = 0.1, eps = 1e-12
// (MappedGodunovUtilities::PPMLimiter, MappedGodunovUtilities.cpp:381-436).  w[q] = W(i - 3 + q), q = 0..6;
// d2W(i) = (W(i+1) - W(i)) - (W(i) - W(i-1)) is GETSECONDDIFF (:470-512); d2Wfcf = (dWM + dWP) * 6.
__device__ __forceinline__ void cubic_limiter(const double w[7], double& dm, double& dp) {
    const double limitC = 1.25, C3 = 0.1, eps = 1.0e-12;
    const double dWM = dm, dWP = dp;
    const bool bigM = fabs(dWM) > 2.0 * fabs(dWP);
    const bool bigP = fabs(dWP) > 2.0 * fabs(dWM);
    const double WLL = w[1], WC = w[3], WRR = w[5];
    const double dWavgM = WC - WLL;
    const double dWavgP = WRR - WC;
    const double prodE1 = dWM * dWP;
    const double prodE2 = dWavgM * dWavgP;
    if (prodE1 >= 0.0 || prodE2 <= 0.0) {
        double d2[5];                                  // d2W at i-2 .. i+2
#pragma unroll
        for (int q = 0; q < 5; ++q) d2[q] = (w[q + 2] - w[q + 1]) - (w[q + 1] - w[q]);
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

template <bool LIM, bool HO>
__global__ void __launch_bounds__(ATPB) k_adv_normal(const AGeom g, int d, const double* __restrict__ W,
                                                     const double* __restrict__ wface, const double* __restrict__ vel,
                                                     const double* __restrict__ src, double srcScale, double dtbydx,
                                                     double* __restrict__ wminus, double* __restrict__ wplus) {
    int i, j, k;
    if (!acell(g, i, j, k)) return;
    const long long c = aidx(g, i, j, k);
    const double w = W[c];
    double dm = (0.0 - w) + wface[c];
    double dp = (0.0 - w) + wface[aoff(g, i, j, k, d, 1)];
    if (LIM && HO) {
        double w7[7];
#pragma unroll
        for (int q = 0; q < 7; ++q) w7[q] = (q == 3) ? w : W[aoff(g, i, j, k, d, q - 3)];
        cubic_limiter(w7, dm, dp);
    } else if (LIM) {   // PPMLimiterF
        if (dm * dp < 0.0) {
            const double dwc = (dp + dm) * 0.5;
            const double d2w = dp - dm;
            const double s = copysign(1.0, dwc);
            if (dwc * d2w > 0.0) dp = s * fmin(-2.0 * s * dm, s * dp);
            else dm = s * fmin(s * dm, -2.0 * s * dp);
        } else {
            dp = 0.0;
            dm = 0.0;
        }
    }
    // PPMNormalPredF, one slope
    const double lam = vel[c];
    double sigmin = lam * dtbydx;
    sigmin = fmin(sigmin, 0.0);
    sigmin = -sigmin;
    double sigmax = lam * dtbydx;
    sigmax = fmax(sigmax, 0.0);
    double sigM, sigP;
    if (lam > 0.0) { sigM = sigmin; sigP = lam * dtbydx; }
    else { sigM = -lam * dtbydx; sigP = sigmax; }
    double nm = dm + sigM * ((dp - dm) - (dp + dm) * (3.0 - 2.0 * sigM)) * 0.5;
    double np = dp + sigP * ((dm - dp) - (dp + dm) * (3.0 - 2.0 * sigP)) * 0.5;
    nm = nm + w;
    np = np + w;
    if (src) {
        const double s = src[c] * srcScale;
        nm = nm + s;
        np = np + s;
    }
    wminus[c] = nm;
    wplus[c] = np;
}

// MappedAdvectionUtil::PLMNormalPred (AdvectUtil.cpp:1066-1152) on a periodic level: van Leer slopes, or FOURTH:
// FourthSlopeDiffsF (MappedGodunovUtilitiesF.ChF:96-146) from the van Leer slopes of the two neighbours with the one-sided
// differences as limiter inputs; the second limiter pass of :1126-1128 (quirk included: without FOURTH it is handed
// zeros as one-sided differences, so use_limiting alone flattens every slope); PLMNormalPredF (:266-311), one slope;
// "+= W"; the half-step source.  ORDER 0 is CTUNormalPred (:1043-1057): both states are W.
template <bool LIM, bool FOURTH, int ORDER>
__global__ void __launch_bounds__(ATPB) k_adv_normal_low(const AGeom g, int d, const double* __restrict__ W,
                                                         const double* __restrict__ vel,
                                                         const double* __restrict__ src, double srcScale, double dtbydx,
                                                         double* __restrict__ wminus, double* __restrict__ wplus) {
    int i, j, k;
    if (!acell(g, i, j, k)) return;
    const long long c = aidx(g, i, j, k);
    const double w = W[c];
    double nm = w, np = w;
    if (ORDER == 1) {
        const double wm1 = W[aoff(g, i, j, k, d, -1)], wp1 = W[aoff(g, i, j, k, d, 1)];
        double dW, dWm = 0.0, dWp = 0.0;
        if (FOURTH) {
            const double wm2 = W[aoff(g, i, j, k, d, -2)], wp2 = W[aoff(g, i, j, k, d, 2)];
            const double vlP = vl_slope<LIM>(w, wp1, wp2);
            const double vlM = vl_slope<LIM>(wm2, wm1, w);
            const double dWR = wp1 - vlP * 0.25;
            const double dWL = wm1 + vlM * 0.25;
            dW = (2.0 / 3.0) * (dWR - dWL);
            dWm = w - wm1;
            dWp = wp1 - w;
        } else {
            dW = vl_slope<LIM>(wm1, w, wp1);
        }
        if (LIM || FOURTH) {   // VanLeerLimiterF
            double lim = fmin(2.0 * fabs(dWm), 2.0 * fabs(dWp));
            lim = fmin(lim, fabs(dW));
            dW = (dWm * dWp < 0.0) ? 0.0 : lim * copysign(1.0, dWm);
        }
        const double lam = vel[c];
        const double lmin = fmin(lam, 0.0), lmax = fmax(lam, 0.0);
        const double loFac = (lam < 0.0) ? lam : lmin;
        const double hiFac = (lam > 0.0) ? lam : lmax;
        const double dm = dW * (-0.5) * (1.0 + dtbydx * loFac);
        const double dp = dW * 0.5 * (1.0 - dtbydx * hiFac);
        nm = dm + w;
        np = dp + w;
    }
    if (src) {
        const double s = src[c] * srcScale;
        nm = nm + s;
        np = np + s;
    }
    wminus[c] = nm;
    wplus[c] = np;
}

// up to two transverse corrections  A = scale * u_d2 * (WHalf(i + e_d2) - WHalf(i))  added to both states, in order
struct ACorr {
    const double* wh[2];
    const double* vel[2];
    double scale[2];
    int dir[2];
    int n;
};

__global__ void __launch_bounds__(ATPB) k_adv_riemann(const AGeom g, int d, const double* __restrict__ wplus,
                                                      const double* __restrict__ wminus,
                                                      const double* __restrict__ advVel, ACorr cr, int upwind,
                                                      int timesVel, int image, double* __restrict__ out) {
    int i, j, k;
    if (!acell(g, i, j, k)) return;
    const long long cR = aidx(g, i, j, k);
    const int iL = i - (d == 0), jL = j - (d == 1), kL = k - (d == 2);
    const long long cL = aidx(g, iL, jL, kL);
    double wl = wplus[cL];
    double wr = wminus[cR];
    for (int t = 0; t < cr.n; ++t) {
        const int d2 = cr.dir[t];
        const double aL = cr.scale[t] * cr.vel[t][cL] * (cr.wh[t][aoff(g, iL, jL, kL, d2, 1)] - cr.wh[t][cL]);
        const double aR = cr.scale[t] * cr.vel[t][cR] * (cr.wh[t][aoff(g, i, j, k, d2, 1)] - cr.wh[t][cR]);
        wl = wl + aL;
        wr = wr + aR;
    }
    const double u = advVel[cR];
    double v;
    if (!upwind) v = 0.5 * (wl + wr);
    else if (u > 1.0e-9) v = wl;
    else if (u < -1.0e-9) v = wr;
    else v = 0.5 * (wl + wr);
    if (timesVel) v = v * u;
    out[cR] = v;
    if (image) {   // the periodic image of face 0 = the high face of the last cell (first ghost slot)
        const int q = (d == 0) ? i : (d == 1 ? j : k);
        if (q == 0) out[cR + (d == 0 ? (long long)g.n[0] : (d == 1 ? g.sy * g.n[1] : g.sz * g.n[2]))] = v;
    }
}

// ---------------------------------------------------------------------------------------------------------
int advect_predict_scalar(SomarLevel* L, SomarField* const whalf[3], const SomarField* wold, const SomarField* source,
                          const SomarField* const oldVel[3], const SomarField* const advVel[3], const SomarField* ccJ,
                          double dt, const SomarAdvectParams* prm) {
    const Geom& G = L->depths[0].g;
    const int nd = G.nd;
    if (L->desc.nranks > 1) { set_error("predictScalar: more than one rank is not built"); return SOMAR_ERR_ARG; }
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
    AGeom g;
    for (int d = 0; d < 3; ++d) g.n[d] = G.n[d];
    g.nd = nd;
    g.sy = G.sy; g.sz = G.sz; g.off0 = G.off0;
    const size_t need = (nd == 3) ? 15 : 6;     // WMinus, WPlus, WHalf1 per direction (+ six WHalf2 in 3-D)
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
    const dim3 nb((g.n[0] + ATPB - 1) / ATPB, g.n[1], g.n[2]);
    cudaStream_t st = L->cx.stream;
    const bool lim = prm->use_limiting != 0;
    const bool ho = prm->use_high_order_limiter != 0;
    const bool fourth = prm->use_fourth_order_slopes != 0;
    const int up = prm->use_upwinding != 0;
    double dtondx[3];
    for (int d = 0; d < nd; ++d) dtondx[d] = dt / G.dx[d];
    const ACorr none = {{nullptr, nullptr}, {nullptr, nullptr}, {0.0, 0.0}, {0, 0}, 0};
    for (int d = 0; d < nd; ++d) {
        double* wface = whalf[d]->p;            // the output's memory holds the face values until they are consumed
        const double* src = source ? source->p : nullptr;
#define ADV_STAGE(LIM_, HO_)                                                                                         \
        do {                                                                                                         \
            k_adv_face<LIM_, HO_><<<nb, ATPB, 0, st>>>(g, d, wold->p, ccJ->p, wface);                                \
            k_adv_normal<LIM_, HO_><<<nb, ATPB, 0, st>>>(g, d, wold->p, wface, oldVel[d]->p, src, 0.5 * dt, dtondx[d], \
                                                         wm[d], wp[d]);                                              \
        } while (0)
#define ADV_LOW(LIM_, FOURTH_, ORDER_)                                                                               \
        k_adv_normal_low<LIM_, FOURTH_, ORDER_><<<nb, ATPB, 0, st>>>(g, d, wold->p, oldVel[d]->p, src, 0.5 * dt,      \
                                                                     dtondx[d], wm[d], wp[d])
        if (order == 0) ADV_LOW(false, false, 0);
        else if (order == 1 && lim && fourth) ADV_LOW(true, true, 1);
        else if (order == 1 && lim) ADV_LOW(true, false, 1);
        else if (order == 1 && fourth) ADV_LOW(false, true, 1);
        else if (order == 1) ADV_LOW(false, false, 1);
        else if (lim && ho) ADV_STAGE(true, true);
        else if (lim) ADV_STAGE(true, false);
        else if (ho) ADV_STAGE(false, true);
        else ADV_STAGE(false, false);
#undef ADV_STAGE
#undef ADV_LOW
        k_adv_riemann<<<nb, ATPB, 0, st>>>(g, d, wp[d], wm[d], advVel[d]->p, none, up, 0, 0, wh1[d]);
        L->cx.launches += (order == 2) ? 3 : 2;
    }
    if (nd == 3) {
        for (int d1 = 0; d1 < 3; ++d1)
            for (int d2 = 0; d2 < 3; ++d2) {
                if (d2 == d1) continue;
                ACorr c = none;
                c.n = 1;
                c.wh[0] = wh1[d2]; c.vel[0] = oldVel[d2]->p; c.scale[0] = -(1.0 / 3.0) * dtondx[d2]; c.dir[0] = d2;
                k_adv_riemann<<<nb, ATPB, 0, st>>>(g, d1, wp[d1], wm[d1], advVel[d1]->p, c, up, 0, 0, wh2[d1][d2]);
                L->cx.launches++;
            }
    }
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
        k_adv_riemann<<<nb, ATPB, 0, st>>>(g, d1, wp[d1], wm[d1], advVel[d1]->p, c, up, prm->return_flux != 0, 1, whalf[d1]->p);
        L->cx.launches++;
    }
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
