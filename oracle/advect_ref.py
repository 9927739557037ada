"""oracle/advect_ref.py -- TEST INFRASTRUCTURE ONLY (CPU oracle; never the product path).

numpy restatement of SOMAR's PPM/CTU scalar predictor (SURVEY.md 8(a) "MappedAdvectionUtil::predictScalar" row,
8(f)-3) for a FULLY PERIODIC level -- the TaylorGreen configuration of BASELINE.json.  Statement order follows
  MappedAdvectionUtil::predictScalar      src/advection/AdvectUtil.cpp:520-933   (the non-USE_OLD_METHOD version)
  ::PPMNormalPred / ::PPMFaceValues       AdvectUtil.cpp:1158-1418               (both m_useHighOrderLimiter branches)
  ::RiemannSolver / ::quasilinearUpdate   AdvectUtil.cpp:944-1037
  MappedGodunovUtilities::vanLeerSlopes   src/advection/MappedGodunovUtilities.cpp:49-102
  ::PPMLimiter / ::PPMNormalPred          MappedGodunovUtilities.cpp:373-500
  SecondSlopeDiffsF, VanLeerLimiterF, PPMLimiterF, PPMFaceValuesF, PPMNormalPredF
                                          src/advection/MappedGodunovUtilitiesF.ChF:35-95,151-186,315-468
  GETSECONDDIFF, FourthInterpFaces, CheckCubicLimiterF (the extremum-preserving limiter, centre boxes)
                                          MappedGodunovUtilitiesF.ChF:470-512,517-566,611-745
  RiemannSolver, AvgStates, QuasilinearUpdateF   src/advection/AdvectUtilF.ChF:28-159
With every direction periodic the lo/hi one-sided boxes of PeriodicLoHiCenter are empty and no a_BCValues method runs,
so each stage is one whole-domain array statement; numpy evaluates them elementwise in IEEE double without
contraction, in the Fortran's operation order.  A level of several boxes gives the same numbers (every stage is a
pure stencil on exchanged ghosts), so the restatement works on global arrays with wrapped indices.

All three normal predictors (normalPredOrder 0 CTU, 1 PLM with or without fourth-order slopes, 2 PPM).
NOT covered (the device path rejects them too): non-periodic sides (the physics BC holders scalarRiemannBC /
scalarSlopeBC; the reference itself carries TODO() markers there, AdvectUtil.cpp:713,780,874,905).
PINNED to the reference's own Fortran kernels (SecondSlopeDiffsF, VanLeerLimiterF, ExtPreservingVanLeerLimiterF,
FourthSlopeDiffsF, PPMFaceValuesF, FourthInterpFaces, PPMLimiterF, GETSECONDDIFF, CheckCubicLimiterF, PPMNormalPredF,
PLMNormalPredF, RiemannSolver, AvgStates, QuasilinearUpdateF), executed through oracle/chf2py.py in the call order of the
reference's C++: tests/test_reference_kernels.py (case "advect") demands identical bits for every direction, limiter family
and predictor in 2-D and 3-D.  The sequencing of predictScalar above the kernels is pinned by tests/test_advect.py to
exactness properties of the scheme (constants, uniform translation, conservation).

Arrays are (nz, ny, nx) C-ordered (nz = 1 in 2-D); direction d lives on numpy axis 2 - d.  Face f of direction d is
the LOW face of cell f (periodic: n faces per direction)."""
import numpy as np

SMALLNUMBER = 1.0e-9    # AdvectUtilF.ChF:26


def _sh(a, d, o):
    """a at index + o in direction d (periodic)"""
    return np.roll(a, -o, axis=2 - d)


def van_leer_slopes(W, d, use_limiting, ext_preserving=False):
    """vanLeerSlopes (MappedGodunovUtilities.cpp:49-102) -> slopes (SecondSlopeDiffsF, centre box) [+ slopeLimiter ->
    VanLeerLimiterF, or -- when the utilities' m_highOrderLimiter is set, which MappedAdvectionUtil::define does with
    useHighOrderLimiter (AdvectUtil.cpp:104) -- slopeLimiterExtPreserving -> ExtPreservingVanLeerLimiterF,
    MappedGodunovUtilitiesF.ChF:190-262]"""
    dWR = _sh(W, d, 1) - W
    dWL = W - _sh(W, d, -1)
    dWC = 0.5 * (dWR + dWL)
    if use_limiting and ext_preserving:
        dWLL = _sh(dWL, d, -1)
        dWRR = _sh(dWR, d, 1)
        dp1 = dWL * dWR
        dp2 = dWLL * dWRR
        dpmin = np.minimum(dp1, dp2)
        dW2L = dWL - dWLL
        dW2C = (dWR - dWL) * 0.5
        dW2R = dWRR - dWR
        sign2 = np.copysign(1.0, dW2C)
        dW2lim = np.minimum(np.minimum(np.abs(dW2C), np.maximum(sign2 * dW2L, 0.0)), np.maximum(sign2 * dW2R, 0.0))
        cvl = 5.0 * 0.25
        dWvl = cvl * 3.0 * 0.5 * dW2lim
        ext = np.where(sign2 * dWC < 0.0, np.minimum(dWvl, 2.0 * np.abs(dWL)), np.minimum(dWvl, 2.0 * np.abs(dWR)))
        dWlim = np.where(dpmin < 0.0, ext, 2.0 * np.minimum(np.abs(dWL), np.abs(dWR)))
        return np.copysign(1.0, dWC) * np.minimum(np.abs(dWC), dWlim)
    if use_limiting:
        lim = np.minimum(2.0 * np.abs(dWL), 2.0 * np.abs(dWR))
        lim = np.minimum(lim, np.abs(dWC))
        dWC = np.where(dWL * dWR < 0.0, 0.0, lim * np.copysign(1.0, dWL))
    return dWC


def _ppm_face(X, dX, d):
    """PPMFaceValuesF, centre box: face f between cells f-1 and f"""
    w_right = X - dX / 3.0
    w_left = _sh(X, d, -1) + _sh(dX, d, -1) / 3.0
    return (w_left + w_right) * 0.5


def _fourth_face(X, d):
    """FourthInterpFaces, centre box (MappedGodunovUtilitiesF.ChF:559-566): face f between cells f-1 and f"""
    return (7.0 * (_sh(X, d, -1) + X) - (_sh(X, d, -2) + _sh(X, d, 1))) / 12.0


def ppm_face_values(W, J, d, use_limiting, high_order=False):
    """MappedAdvectionUtil::PPMFaceValues: 4th-order face values of J W divided by those of J
    (m_useHighOrderLimiter: FourthInterpFaces, AdvectUtil.cpp:1269-1292,1345-1368; else van Leer slopes + PPMFaceValuesF)"""
    JW = W * J
    if high_order:
        return _fourth_face(JW, d) / _fourth_face(J, d)
    edge_j = _ppm_face(J, van_leer_slopes(J, d, use_limiting), d)
    w_face = _ppm_face(JW, van_leer_slopes(JW, d, use_limiting), d)
    return w_face / edge_j


def check_cubic_limiter(dm, dp, W, d, limit_c=1.25, c3=0.1, eps=1.0e-12):
    """MappedGodunovUtilities::PPMLimiter with m_highOrderLimiter (MappedGodunovUtilities.cpp:381-436):
    GETSECONDDIFF (MappedGodunovUtilitiesF.ChF:470-512) + CheckCubicLimiterF, inner centre box (:611-745)"""
    d2W = (_sh(W, d, 1) - W) - (W - _sh(W, d, -1))
    atfcf = (dm + dp) * 6.0
    big_m = np.abs(dm) > 2.0 * np.abs(dp)
    big_p = np.abs(dp) > 2.0 * np.abs(dm)
    dwavg_m = W - _sh(W, d, -2)
    dwavg_p = _sh(W, d, 2) - W
    prod_e1 = dm * dp
    prod_e2 = dwavg_m * dwavg_p
    extremum = (prod_e1 >= 0.0) | (prod_e2 <= 0.0)
    d2l, d2c, d2r = _sh(d2W, d, -1), d2W, _sh(d2W, d, 1)
    sl, sc, sr, sf = (np.copysign(1.0, x) for x in (d2l, d2c, d2r, atfcf))
    same = (sl == sc) & (sr == sc) & (sf == sc) & (np.abs(atfcf) >= eps)
    d2lim = sc * np.minimum(np.minimum(np.abs(atfcf), limit_c * np.abs(d2c)),
                            np.minimum(limit_c * np.abs(d2l), limit_c * np.abs(d2r)))
    with np.errstate(divide="ignore", invalid="ignore"):
        rho = np.where(same, d2lim / atfcf, 0.0)
    d2ll, d2rr = _sh(d2W, d, -2), _sh(d2W, d, 2)
    d3 = [d2l - d2ll, d2c - d2l, d2r - d2c, d2rr - d2r]
    d3min = np.minimum(np.minimum(d3[0], d3[1]), np.minimum(d3[2], d3[3]))
    d3max = np.maximum(np.maximum(d3[0], d3[1]), np.maximum(d3[2], d3[3]))
    prod_d3 = c3 * np.maximum(np.abs(d3max), np.abs(d3min)) - np.abs(d3max - d3min)
    apply = extremum & (rho < (1.0 - eps)) & (prod_d3 <= 0.0)
    case1 = apply & (prod_e1 > 0.0)
    case2 = apply & ~(prod_e1 > 0.0) & big_m
    case3 = apply & ~(prod_e1 > 0.0) & ~big_m & big_p
    new_m = np.where(case1, dm * rho, np.where(case2, dm * rho - 2.0 * dp * (1.0 - rho), dm))
    new_p = np.where(case1, dp * rho, np.where(case3, dp * rho - 2.0 * dm * (1.0 - rho), dp))
    # not an extremum: the PPM limiter
    new_m = np.where(~extremum & big_m, -2.0 * dp, new_m)
    new_p = np.where(~extremum & big_p, -2.0 * dm, new_p)
    return new_m, new_p


def ppm_normal_pred(W, J, vel_d, d, dtbydx, use_limiting, high_order=False):
    """MappedAdvectionUtil::PPMNormalPred: returns (WMinus, WPlus) on cells"""
    w_face = ppm_face_values(W, J, d, use_limiting, high_order)
    dm = (0.0 - W) + w_face                   # low face of the cell
    dp = (0.0 - W) + _sh(w_face, d, 1)        # high face
    if use_limiting and high_order:
        dm, dp = check_cubic_limiter(dm, dp, W, d)
    elif use_limiting:                        # PPMLimiterF
        opp = dm * dp < 0.0
        dwc = (dp + dm) * 0.5
        d2w = dp - dm
        s = np.copysign(1.0, dwc)
        up = dwc * d2w > 0.0
        new_p = s * np.minimum(-2.0 * s * dm, s * dp)
        new_m = s * np.minimum(s * dm, -2.0 * s * dp)
        dp2 = np.where(opp, np.where(up, new_p, dp), 0.0)
        dm2 = np.where(opp, np.where(up, dm, new_m), 0.0)
        dm, dp = dm2, dp2
    # PPMNormalPredF, nSlope = 1
    lam = vel_d
    sigmin = -np.minimum(lam * dtbydx, 0.0)
    sigmax = np.maximum(lam * dtbydx, 0.0)
    pos = lam > 0.0
    sig_m = np.where(pos, sigmin, -lam * dtbydx)
    sig_p = np.where(pos, lam * dtbydx, sigmax)
    new_m = dm + sig_m * ((dp - dm) - (dp + dm) * (3.0 - 2.0 * sig_m)) * 0.5
    new_p = dp + sig_p * ((dm - dp) - (dp + dm) * (3.0 - 2.0 * sig_p)) * 0.5
    return new_m + W, new_p + W


TWO3RD = 2.0 / 3.0     # Chombo CONSTANTS.H "two3rd" (BaseTools; third party, not under /root/reference)


def van_leer_limiter(dWc, dWl, dWr):
    """VanLeerLimiterF, MappedGodunovUtilitiesF.ChF:151-186"""
    lim = np.minimum(2.0 * np.abs(dWl), 2.0 * np.abs(dWr))
    lim = np.minimum(lim, np.abs(dWc))
    return np.where(dWl * dWr < 0.0, 0.0, lim * np.copysign(1.0, dWl))


def plm_normal_pred(W, vel_d, d, dtbydx, use_limiting, fourth_order_slopes=False, high_order=False):
    """MappedAdvectionUtil::PLMNormalPred, AdvectUtil.cpp:1066-1152 -> FourthSlopeDiffsF (MappedGodunovUtilitiesF.ChF:
    96-146), oneSidedDifferences / slopeLimiter (MappedGodunovUtilities.cpp:143-171,240-262), PLMNormalPredF (:266-311).
    The second limiter pass is the reference's, quirk included: WITHOUT fourth-order slopes it is handed the zeroed
    a_WMinus / a_WPlus as one-sided differences (AdvectUtil.cpp:1116-1128), so `use_limiting` alone flattens every slope
    and the predictor degenerates to CTU.  With useHighOrderLimiter the van Leer slopes are limited by the
    extremum-preserving limiter (the utilities' m_highOrderLimiter, AdvectUtil.cpp:104)."""
    if fourth_order_slopes:
        dWvL = van_leer_slopes(W, d, use_limiting, high_order)
        dWR = _sh(W, d, 1) - _sh(dWvL, d, 1) * 0.25
        dWL = _sh(W, d, -1) + _sh(dWvL, d, -1) * 0.25
        dW = TWO3RD * (dWR - dWL)
        dWm = W - _sh(W, d, -1)
        dWp = _sh(W, d, 1) - W
    else:
        dW = van_leer_slopes(W, d, use_limiting, high_order)
        dWm = np.zeros_like(W)
        dWp = np.zeros_like(W)
    if use_limiting or fourth_order_slopes:
        dW = van_leer_limiter(dW, dWm, dWp)
    lam = vel_d
    lmin = np.minimum(lam, 0.0)
    lmax = np.maximum(lam, 0.0)
    lo_fac = np.where(lam < 0.0, lam, lmin)     # lambdaK < 0: lambdaK, else lmin
    hi_fac = np.where(lam > 0.0, lam, lmax)     # lambdaK > 0: lambdaK, else lmax
    dm = dW * (-0.5) * (1.0 + dtbydx * lo_fac)
    dp = dW * 0.5 * (1.0 - dtbydx * hi_fac)
    return dm + W, dp + W


def normal_pred(W, J, vel_d, d, dtbydx, use_limiting, high_order, order, fourth_order_slopes):
    """the switch on m_normalPredOrder, AdvectUtil.cpp:660-692"""
    if order == 0:      # CTUNormalPred, AdvectUtil.cpp:1043-1057: straight copies
        return W.copy(), W.copy()
    if order == 1:
        return plm_normal_pred(W, vel_d, d, dtbydx, use_limiting, fourth_order_slopes, high_order)
    if order == 2:
        return ppm_normal_pred(W, J, vel_d, d, dtbydx, use_limiting, high_order)
    raise ValueError("normal predictor order must be 0 (CTU), 1 (PLM) or 2 (PPM)")


def riemann(w_plus, w_minus, adv_vel_d, d, upwind):
    """RiemannSolver / AvgStates at face f: left state = WPlus of cell f-1, right state = WMinus of cell f"""
    wl = _sh(w_plus, d, -1)
    wr = w_minus
    if not upwind:
        return 0.5 * (wl + wr)
    return np.where(adv_vel_d > SMALLNUMBER, wl, np.where(adv_vel_d < -SMALLNUMBER, wr, 0.5 * (wl + wr)))


def quasilinear_update(w_half_d, vel_d, scale, d):
    """QuasilinearUpdateF: AdWdx = scale * u_d * (WHalf(i + e_d) - WHalf(i))"""
    return scale * vel_d * (_sh(w_half_d, d, 1) - w_half_d)


def predict_scalar(W, J, old_vel, adv_vel, dt, dx, dim, source=None, use_limiting=False, use_upwinding=True,
                   return_flux=True, high_order=False, normal_pred_order=2, fourth_order_slopes=False):
    """old_vel[d]: cell-centred velocity components; adv_vel[d]: face-centred advecting velocity of direction d.
    Returns the list of face-centred, time-centred states (times adv_vel when return_flux)."""
    dtondx = [dt / dx[d] for d in range(dim)]
    wm, wp, wh1 = [], [], []
    for d in range(dim):
        m, p = normal_pred(W, J, old_vel[d], d, dt / dx[d], use_limiting, high_order, normal_pred_order,
                           fourth_order_slopes)
        if source is not None:
            m = m + source * (0.5 * dt)
            p = p + source * (0.5 * dt)
        wm.append(m)
        wp.append(p)
        wh1.append(riemann(p, m, adv_vel[d], d, use_upwinding))
    wh2 = {}
    if dim == 3:
        for d1 in range(3):
            for d2 in range(3):
                if d2 == d1:
                    continue
                a = quasilinear_update(wh1[d2], old_vel[d2], -(1.0 / 3.0) * dtondx[d2], d2)
                wh2[(d1, d2)] = riemann(wp[d1] + a, wm[d1] + a, adv_vel[d1], d1, use_upwinding)
    out = []
    for d1 in range(dim):
        m, p = wm[d1], wp[d1]
        for d2 in range(dim):
            if d2 == d1:
                continue
            if dim == 2:
                a = quasilinear_update(wh1[d2], old_vel[d2], -(1.0 / 2.0) * dtondx[d2], d2)
            else:
                d3 = 3 - d1 - d2
                a = quasilinear_update(wh2[(d2, d3)], old_vel[d2], -(1.0 / 2.0) * dtondx[d2], d2)
            m = m + a
            p = p + a
        wh = riemann(p, m, adv_vel[d1], d1, use_upwinding)
        if return_flux:
            wh = wh * adv_vel[d1]
        out.append(wh)
    return out
