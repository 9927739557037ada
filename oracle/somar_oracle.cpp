// oracle/somar_oracle.cpp -- TEST INFRASTRUCTURE ONLY (CPU oracle; never the product path).
// See somar_oracle.hpp.  Paths cited are relative to /root/reference.
#include "somar_oracle.hpp"
#include <cstring>

namespace orc {

// ----------------------------------------------------------------------------------------------
// ExtrapolateFaceNoEV / ExtrapolateFaceAndCopy -- src/calculus/extrapolation/ExtrapolationUtils.cpp:34-163
// (cell-centred data only; that is all the elliptic path uses).
// ----------------------------------------------------------------------------------------------
static void ExtrapolateFaceNoEV(FAB& dest, const FAB& state, const Box& valid, int dir, int side, int order) {
    const int isign = sign(side);
    Box toRegion = adjCellBox(valid, dir, side, 1);   // :52
    toRegion &= state.box();                          // :56
    if (toRegion.isEmpty()) return;
    extrapolate_face_noev(dest, state, toRegion, dir, isign, order);  // :60
}

static void ExtrapolateFaceAndCopy(FAB& dest, const FAB& state, const Box& valid, int dir, int side, int order,
                                   int nd, int numLayers = 1) {
    ExtrapolateFaceNoEV(dest, state, valid, dir, side, order);  // :124
    Box ghostBox = adjCellBox(valid, dir, side, 1);              // :126
    ghostBox &= state.box();
    Box nearBox = ghostBox;                                      // :129
    nearBox.shift(dir, -sign(side));
    nearBox.growDir(dir, flip(side), numLayers - 1);
    if (&dest != &state) dest.copy(state, nearBox);              // :134
    // :139 -- loops over all SpaceDim directions; an embedded 2-D box has no third direction.
    for (int edir = 0; edir < nd; ++edir) {
        if (edir == dir) continue;
        for (int eside = 0; eside < 2; ++eside) {
            ExtrapolateFaceNoEV(dest, dest, ghostBox, edir, eside, order);
            ExtrapolateFaceNoEV(dest, dest, nearBox, edir, eside, order);
        }
        ghostBox.grow(IntVect::Basis(edir));
        ghostBox &= state.box();
        nearBox.grow(IntVect::Basis(edir));
        nearBox &= state.box();
    }
}

// ----------------------------------------------------------------------------------------------
// Op set-up
// ----------------------------------------------------------------------------------------------
void Op::defineCommon() {
    activeDirs = IntVect(1, nd > 1, nd > 2);
    // MappedAMRPoissonOp::define4 -- MappedAMRPoissonOp.cpp:278-283
    validDomain = domain.box;
    for (int d = 0; d < nd; ++d)
        if (domain.periodic[d]) validDomain.grow(d, 1);
    // RelaxationMethod::define -- RelaxationMethod.cpp:70
    extrap.define(grids, 1, activeDirs);
    collectBoundaryData();
}

void Op::computeLapDiag() {
    lapDiag.define(grids, 1, IntVect::Zero());
    for (size_t b = 0; b < grids.size(); ++b)
        fill_lapdiag(lapDiag[b], Jgup.fabs[b].data(), Jinv[b], lapDiag[b].box(), dx, nd);
}

// RelaxationMethod::collectBoundaryData -- RelaxationMethod.cpp:81-330 (restated literally: faces, then
// edges with edir > fdir, then vertices with vdir > edir; `all` = the "simple" variant that visits every
// box boundary instead of only physical boundaries).
static void collectOne(const Op& op, bool all, std::vector<BoundaryBoxData>& out) {
    const Box& domBox = op.domain.box;
    Box domBdry[3][2];
    for (int dir = 0; dir < op.nd; ++dir)
        for (int s = 0; s < 2; ++s) {
            domBdry[dir][s] = bdryBox(domBox, dir, s, 1);
            domBdry[dir][s].shiftHalf(dir, -sign(s));
        }
    const Box domInterior = grow(domBox, op.activeDirs * -1);
    out.clear();
    for (size_t bi = 0; bi < op.grids.size(); ++bi) {
        const Box& valid = op.grids[bi];
        const Box interior = all ? grow(valid, op.activeDirs * -1) : (valid & domInterior);
        BoundaryBoxData data;
        data.index = (int)bi;
        data.valid = valid;
        auto push = [&](const Box& bx) {
            for (int d = 0; d < 3; ++d)
                for (int s = 0; s < 2; ++s)
                    data.stencil[d][s] = (d < op.nd) ? op.bc.stencil(bx, op.domain, d, s) : BC_Neum;
            data.validBdry = bx;
            out.push_back(data);
        };
        for (int fdir = 0; fdir < op.nd; ++fdir) {
            for (int fs = 0; fs < 2; ++fs) {
                Box faceBox = adjCellBox(interior, fdir, fs, 1) & valid;
                if (!all && !domBdry[fdir][fs].intersects(faceBox)) continue;
                if (all && faceBox.isEmpty()) continue;
                push(faceBox);
                for (int edir = fdir + 1; edir < op.nd; ++edir) {
                    for (int es = 0; es < 2; ++es) {
                        Box edgeBox = adjCellBox(faceBox, edir, es, 1) & valid;
                        if (!all && !domBdry[edir][es].intersects(edgeBox)) continue;
                        if (all && edgeBox.isEmpty()) continue;
                        push(edgeBox);
                        if (op.nd < 3) continue;
                        const int vdir = 3 - fdir - edir;
                        if (vdir <= edir) continue;  // process each vertex once
                        for (int vs = 0; vs < 2; ++vs) {
                            Box vertexBox = adjCellBox(edgeBox, vdir, vs, 1) & valid;
                            if (!all && !domBdry[vdir][vs].intersects(vertexBox)) continue;
                            if (all && vertexBox.isEmpty()) continue;
                            push(vertexBox);
                        }
                    }
                }
            }
        }
    }
}

void Op::collectBoundaryData() {
    collectOne(*this, false, boundaryBoxData);
    collectOne(*this, true, simpleBoundaryBoxData);
}

// ----------------------------------------------------------------------------------------------
// LevelData::exchange stand-in (Chombo, not in tree).  Fills ghost cells that lie inside another
// box of the layout or inside a periodic image of one.  corners == false reproduces
// Copier::trimEdges (faces only), used for diagonal metrics (MappedAMRPoissonOp.cpp:425-427).
// ----------------------------------------------------------------------------------------------
void Op::exchange(LevelData& phi, bool corners) const {
    const IntVect g = phi.ghost;
    if (g == IntVect::Zero()) return;
    std::vector<IntVect> shifts;
    for (int sz = -1; sz <= 1; ++sz)
        for (int sy = -1; sy <= 1; ++sy)
            for (int sx = -1; sx <= 1; ++sx) {
                if ((sx && !domain.periodic[0]) || (sy && !domain.periodic[1]) || (sz && !domain.periodic[2])) continue;
                shifts.push_back(IntVect(sx * domain.box.size(0), sy * domain.box.size(1), sz * domain.box.size(2)));
            }
#pragma omp parallel for schedule(dynamic)
    for (size_t di = 0; di < grids.size(); ++di) {
        FAB& dst = phi[di];
        const Box& dvalid = grids[di];
        const Box gbox = grow(dvalid, g);
        for (size_t si = 0; si < grids.size(); ++si) {
            for (const IntVect& sh : shifts) {
                if (si == di && sh == IntVect::Zero()) continue;
                Box simg = grids[si];
                simg.shift(sh);
                Box ov = gbox & simg;
                if (ov.isEmpty()) continue;
                if (!corners) {
                    // keep only cells displaced from dvalid along exactly one direction
                    for (int d = 0; d < 3; ++d) {
                        if (g[d] == 0) continue;
                        Box face = ov;
                        for (int o = 0; o < 3; ++o)
                            if (o != d) { face.lo[o] = std::max(face.lo[o], dvalid.lo[o]); face.hi[o] = std::min(face.hi[o], dvalid.hi[o]); }
                        if (!face.isEmpty()) dst.copyShifted(phi[si], face, sh);
                    }
                } else {
                    dst.copyShifted(phi[si], ov, sh);
                }
            }
        }
    }
}

// MappedAMRPoissonOp::fillExtrap -- MappedAMRPoissonOp.cpp:2244-2270
void Op::fillExtrapOp(FAB& ex, const FAB& phi, int order) const {
    if (isDiagonal) return;                    // :2253
    const Box& stateBox = phi.box();
    Box validPhi = stateBox & validDomain;     // :2257
    ex.copy(phi, validPhi);                    // :2261
    for (int fdir = 0; fdir < nd; ++fdir) {    // :2264
        ExtrapolateFaceAndCopy(ex, ex, validPhi, fdir, Lo, order, nd);
        ExtrapolateFaceAndCopy(ex, ex, validPhi, fdir, Hi, order, nd);
        validPhi.grow(fdir, 1);
        validPhi &= stateBox;
    }
}

// BCMethodHolder::setGhosts with EllipticConstNeumBCGhostClass / EllipticConstDiriBCGhostClass --
// EllipticBCUtils.cpp:380-480 -> setSideNeumBC (:130-215) / setSideDiriBC (:41-118).
void Op::setGhosts(FAB& phi, FAB* ex, const Box& valid, size_t bi, bool homog) const {
    for (int idir = 0; idir < nd; ++idir) {
        if (domain.periodic[idir]) continue;
        for (int iside = 0; iside < 2; ++iside) {
            const int isign = sign(iside);
            if (bc.type[idir][iside] == BC_Neum) {
                // setSideNeumBC
                const Box faceBox = bdryBox(valid, idir, iside, 1) & bdryBox(domain.box, idir, iside, 1);
                if (faceBox.isEmpty()) continue;
                Box ghostBox = faceBox;
                ghostBox.shiftHalf(idir, isign);
                ghostBox &= phi.box();
                if (ghostBox.isEmpty()) continue;
                // NOTE: setSideNeumBC passes a_value, not the homogeneous-aware value, to Fortran
                // (EllipticBCUtils.cpp:172,205); pressure BC values are 0 so the two agree.
                const double bcval = bc.value[idir][iside];
                const FAB& Jgi = Jgup.fabs[bi][idir];
                if (isDiagonal) {
                    neum_ghost_ortho(phi, Jgi, idir, ghostBox, bcval, idir, isign, dx[idir]);
                } else {
                    FAB local;
                    FAB* exPtr = ex;
                    if (!exPtr) { local.define(phi.box(), phi.nComp()); exPtr = &local; }
                    ExtrapolateFaceAndCopy(*exPtr, phi, valid, idir, iside, 2, nd);   // :196
                    neum_ghost_full(phi, *exPtr, Jgi, ghostBox, bcval, idir, isign, dx, nd);
                }
            } else if (bc.type[idir][iside] == BC_Diri) {
                // setSideDiriBC, cell-centred branch
                const double bcval = homog ? 0.0 : bc.value[idir][iside];
                Box destBox = bdryBox(domain.box, idir, iside, 1) & surroundingNodes(valid, idir);
                if (destBox.isEmpty()) continue;
                destBox.shiftHalf(idir, isign);
                destBox &= phi.box();
                if (destBox.isEmpty()) continue;
                diri_ghost(phi, destBox, bcval, idir, isign, bc.diriOrder);
            }
        }
    }
}

// BCMethodHolder::setFluxes with EllipticConstNeumBCFluxClass -- EllipticBCUtils.cpp:491-546:
// boundary faces of direction a_dir are overwritten with the (homogeneous-aware) BC value.
// Dirichlet sides have no flux method (the ghost carries the BC).
void Op::setFluxes(std::array<FAB, 3>& flux, const Box& valid, bool homog) const {
    for (int idir = 0; idir < nd; ++idir) {
        if (domain.periodic[idir]) continue;
        for (int iside = 0; iside < 2; ++iside) {
            if (bc.type[idir][iside] != BC_Neum) continue;
            const double bcval = homog ? 0.0 : bc.value[idir][iside];
            Box thisDom = domain.box;
            thisDom.surroundingNodes(idir);
            Box destBox = bdryBox(thisDom, idir, iside, 1) & surroundingNodes(valid, idir);
            if (destBox.isEmpty()) continue;
            flux[idir].setVal(bcval, destBox, 0, flux[idir].nComp());
        }
    }
}

// MappedAMRPoissonOp::applyOpI -- MappedAMRPoissonOp.cpp:772-898 (USE_SYNC_METHOD branch).
void Op::applyOp(LevelData& lhs, LevelData& phi, bool homog) const {
    exchange(phi, true);  // exchangeComplete :2222-2238 (faces + corners)
#pragma omp parallel for schedule(dynamic)
    for (size_t b = 0; b < grids.size(); ++b) {
        const Box& valid = grids[b];
        FAB& phiFAB = phi[b];
        std::array<FAB, 3> flux;
        for (int d = 0; d < nd; ++d) flux[d].define(surroundingNodes(valid, d), phi.nc);
        FAB ex(phiFAB.box(), phi.nc);
        fillExtrapOp(ex, phiFAB, 2);                                     // :819
        setGhosts(phiFAB, &ex, valid, b, homog);                        // :823
        for (int d = 0; d < nd; ++d) {                                  // getFluxComplete :2048-2122, a_ref = 1
            if (isDiagonal) getflux_ortho(flux[d], phiFAB, Jgup.fabs[b][d], d, flux[d].box(), 1.0 / dx[d]);
            else getflux_full(flux[d], phiFAB, ex, Jgup.fabs[b][d], d, flux[d].box(), 1.0, dx, nd);
        }
        setFluxes(flux, valid, homog);                                   // :838
        for (int d = 0; d < nd; ++d)                                     // fluxFB *= m_beta :841
            for (double& v : flux[d].d) v *= beta;
        flux_divergence(lhs[b], flux.data(), Jinv[b], valid, dx, nd);    // :847-880
        if (alpha != 0.0) axbyip(lhs[b], phiFAB, alpha, 1.0, valid);     // :889-897
    }
}

// MappedAMRPoissonOp::residualI -- MappedAMRPoissonOp.cpp:646-677
void Op::residual(LevelData& lhs, LevelData& phi, const LevelData& rhs, bool homog) const {
    applyOp(lhs, phi, homog);
#pragma omp parallel for schedule(dynamic)
    for (size_t b = 0; b < grids.size(); ++b) subtractop(lhs[b], rhs[b], lhs[b], grids[b]);
}


// ----------------------------------------------------------------------------------------------
// Relaxation
// ----------------------------------------------------------------------------------------------

// RelaxationMethod::fillGhostsAndExtrapolate -- RelaxationMethod.cpp:376-435 (single level: the
// homogeneousCFInterp call at :399 is a no-op).
void Op::fillGhostsAndExtrapolate(LevelData& phi) {
    const int extrapOrder = 1;
#pragma omp parallel for schedule(dynamic)
    for (size_t b = 0; b < grids.size(); ++b) {
        FAB& phiFAB = phi[b];
        FAB& exFAB = extrap[b];
        if (!isDiagonal) {
            exFAB.copy(phiFAB);                         // :411
            Box domValid = domain.box;                  // :413 (periodic directions are NOT grown here)
            domValid &= exFAB.box();
            for (int fdir = 0; fdir < nd; ++fdir) {
                for (int fs = 0; fs < 2; ++fs) ExtrapolateFaceAndCopy(exFAB, exFAB, domValid, fdir, fs, extrapOrder, nd);
                domValid.grow(fdir, 1);                 // :422 (not re-intersected with the FAB box)
            }
        }
        setGhosts(phiFAB, &exFAB, grids[b], b, true);   // :432
    }
}

// BaseGSRB::fullStencilGSRB -- GSRB.cpp:341-478
void Op::fullStencilGSRB(FAB& phi, const FAB& rhs, const Box& region, size_t bi, int pass) const {
    if (region.isEmpty()) return;
    if (isDiagonal)
        gsrb_ortho(phi, rhs, Jgup.fabs[bi].data(), Jinv[bi], lapDiag[bi], region, dx, alpha, beta, pass, nd);
    else
        gsrb_full(phi, extrap[bi], rhs, Jgup.fabs[bi].data(), Jinv[bi], lapDiag[bi], region, dx, alpha, beta, pass, nd);
}

// BaseGSRB::boundaryGSRB -- GSRB.cpp:489-653
void Op::boundaryGSRB(LevelData& phi, const LevelData& rhs, int pass, bool all) const {
    const std::vector<BoundaryBoxData>& ref = all ? simpleBoundaryBoxData : boundaryBoxData;
    // (the slabs are disjoint and a colour pass reads only the other colour: any order gives the same bits)
#pragma omp parallel for schedule(dynamic)
    for (size_t idx = 0; idx < ref.size(); ++idx) {
        const BoundaryBoxData& data = ref[idx];
        const size_t bi = (size_t)data.index;
        if (isDiagonal)
            gsrb_boundary_ortho(phi[bi], rhs[bi], Jgup.fabs[bi].data(), Jinv[bi], data.validBdry, dx, alpha, beta,
                                data.stencil, pass, nd);
        else
            gsrb_boundary_full(phi[bi], extrap[bi], rhs[bi], Jgup.fabs[bi].data(), Jinv[bi], data.validBdry, dx, alpha,
                               beta, data.stencil, pass, nd);
    }
}

void Op::relaxOnce(LevelData& phi, const LevelData& rhs) {
    const bool corners = !isDiagonal;  // m_exchangeCopier is edge-trimmed for diagonal metrics
    if (relaxMode == RELAX_LEVEL_GSRB) {
        // LevelGSRB::relax -- GSRB.cpp:58-98
        const Box domInterior = grow(domain.box, activeDirs * -1);
        for (int pass = 0; pass < 2; ++pass) {
            exchange(phi, corners);
            fillGhostsAndExtrapolate(phi);
#pragma omp parallel for schedule(dynamic)
            for (size_t b = 0; b < grids.size(); ++b) {
                Box validInterior = grids[b] & domInterior;
                fullStencilGSRB(phi[b], rhs[b], validInterior, b, pass);
            }
            boundaryGSRB(phi, rhs, pass, false);
        }
    } else if (relaxMode == RELAX_LOOSE_GSRB) {
        // LooseGSRB::relax -- GSRB.cpp:104-141.  The exchange is begun before, and finished after, the
        // interior sweeps.  The interior sweeps touch grow(valid,-1) only, and an exchange reads only the
        // outermost valid layer of each neighbour, so the data delivered at exchangeEnd equals that of an
        // exchange performed after the sweeps.  fillGhostsAndExtrapolate runs BEFORE the new ghost data
        // lands: extrap carries the previous iteration's box-interface ghosts (that is the "loose" part).
        fillGhostsAndExtrapolate(phi);
#pragma omp parallel for schedule(dynamic)
        for (size_t b = 0; b < grids.size(); ++b) {
            const Box interior = grow(grids[b], activeDirs * -1);
            fullStencilGSRB(phi[b], rhs[b], interior, b, 0);
            fullStencilGSRB(phi[b], rhs[b], interior, b, 1);
        }
        exchange(phi, corners);  // exchangeEnd
        boundaryGSRB(phi, rhs, 0, true);
        boundaryGSRB(phi, rhs, 1, true);
    } else if (relaxMode == RELAX_LINE_GSRB) {
        // LineGSRB::relax -- GSRB.cpp:148-330
        for (int pass = 0; pass < 2; ++pass) {
            exchange(phi, corners);
            fillGhostsAndExtrapolate(phi);
#pragma omp parallel for schedule(dynamic)
            for (size_t b = 0; b < grids.size(); ++b) {
                int st[3][2];
                for (int d = 0; d < 3; ++d)
                    for (int s = 0; s < 2; ++s) st[d][s] = (d < nd) ? bc.stencil(grids[b], domain, d, s) : BC_Neum;
                line_gsrb(phi[b], extrap[b], rhs[b], Jgup.fabs[b].data(), Jinv[b], grids[b], dx, dxCrse[nd - 1], alpha,
                          beta, pass, st, nd, isDiagonal);
            }
        }
    } else if (relaxMode == RELAX_JACOBI) {
        // Jacobi::relax -- Jacobi.cpp:50-90: res = rhs - L[phi] (homogeneous); phi += 0.5 res / (alpha + beta diag)
        LevelData res;
        res.define(grids, phi.nc, IntVect::Zero());
        residual(res, phi, rhs, true);
#pragma omp parallel for schedule(dynamic)
        for (size_t b = 0; b < grids.size(); ++b) jacobi_iter(phi[b], res[b], lapDiag[b], grids[b], alpha, beta);
    } else if (relaxMode == -1) {
        // ProblemContext::RelaxMode::NORELAX: the factory installs no relaxation method and relax() returns at once
        // (MappedAMRPoissonOp.cpp:1733-1743, m_relaxPtr == NULL)
    } else {
        fprintf(stderr, "orc::Op::relaxOnce: bad relaxMode %d\n", relaxMode);
        abort();
    }
}

// MappedAMRPoissonOp::relax -- MappedAMRPoissonOp.cpp:1733-1743
void Op::relax(LevelData& phi, const LevelData& rhs, int iters) {
    for (int i = 0; i < iters; ++i) relaxOnce(phi, rhs);
}

// MappedAMRPoissonOp::preCond -- MappedAMRPoissonOp.cpp:684-734
void Op::preCond(LevelData& phi, const LevelData& rhs) {
    if (preCondSmoothIters == 0 || precondMode == 0) {
        assign(phi, rhs);
        return;
    }
#pragma omp parallel for schedule(dynamic)
    for (size_t b = 0; b < grids.size(); ++b) diagprecond(phi[b], rhs[b], lapDiag[b], grids[b], alpha, beta);
    const int saveMode = relaxMode;
    if (precondMode == 2) relaxMode = RELAX_LINE_GSRB;  // PrecondMode::DiagLineRelax
    for (int it = 0; it < preCondSmoothIters; ++it) relaxOnce(phi, rhs);
    relaxMode = saveMode;
}

// ----------------------------------------------------------------------------------------------
// Transfers
// ----------------------------------------------------------------------------------------------

// MappedAMRPoissonOp::restrictResidual -- MappedAMRPoissonOp.cpp:1281-1304 + FullWeightingPS::restrict
// -- RestrictionStrategy.cpp:36-99 (J-weighted branch).
void Op::restrictResidual(LevelData& resCoarse, LevelData& phiFine, const LevelData& rhsFine, const Op& crseOp) const {
    LevelData resFine;
    resFine.define(grids, phiFine.nc, IntVect::Zero());
    residual(resFine, phiFine, rhsFine, true);
    const IntVect r = mgCrseRefRatio;
#pragma omp parallel for schedule(dynamic)
    for (size_t b = 0; b < grids.size(); ++b)
        mappedaverage2(resCoarse[b], resFine[b], Jinv[b], crseOp.grids[b], r);
}

// ConstInterpPS::prolongIncrement / ZeroAvgConstInterpPS::prolongIncrement -- ProlongationStrategy.cpp:49-165
void Op::prolongIncrement(LevelData& phiFine, const LevelData& corrCoarse) const {
    const IntVect r = mgCrseRefRatio;
    if (!zeroAvgProlong) {
#pragma omp parallel for schedule(dynamic)
        for (size_t b = 0; b < grids.size(); ++b) const_interp(phiFine[b], corrCoarse[b], grids[b], r);
        return;
    }
    double localSum = 0.0, localVol = 0.0;
    for (size_t b = 0; b < grids.size(); ++b)
        const_interp_with_avg(phiFine[b], corrCoarse[b], grids[b], r, Jinv[b], dxProduct, localVol, localSum);
    const double avgPhi = localSum / localVol;
#pragma omp parallel for schedule(dynamic)
    for (size_t b = 0; b < grids.size(); ++b)
        for (double& v : phiFine[b].d) v -= avgPhi;  // a_phiThisLevel[dit] -= avgPhi (whole FAB, :163)
}

// ----------------------------------------------------------------------------------------------
// LevelDataOps-style BLAS-1 (MappedAMRPoissonOp.H:566-711; Chombo norm()/dotProduct over valid cells)
// ----------------------------------------------------------------------------------------------
double Op::norm(const LevelData& x, int ord) const {
    double r = 0.0;
    if (ord == 0) {   // a maximum does not depend on the order it is taken in: boxes in parallel
#pragma omp parallel for schedule(dynamic) reduction(max : r)
        for (size_t b = 0; b < grids.size(); ++b) {
            const FAB& f = x[b];
            ORC_LOOP(grids[b], i, j, k) r = std::max(r, std::fabs(f(i, j, k)));
        }
        return r;
    }
    for (size_t b = 0; b < grids.size(); ++b) {
        const FAB& f = x[b];
        ORC_LOOP(grids[b], i, j, k) {
            const double v = std::fabs(f(i, j, k));
            if (ord == 0) r = std::max(r, v);
            else if (ord == 1) r += v;
            else r += v * v;
        }
    }
    if (ord == 2) r = std::sqrt(r);
    return r;
}
double Op::dot(const LevelData& a, const LevelData& b2) const {
    double r = 0.0;
    for (size_t b = 0; b < grids.size(); ++b) {
        const FAB &fa = a[b], &fb = b2[b];
        ORC_LOOP(grids[b], i, j, k) r += fa(i, j, k) * fb(i, j, k);
    }
    return r;
}
void Op::incr(LevelData& y, const LevelData& x, double s) const {
#pragma omp parallel for schedule(dynamic)
    for (size_t b = 0; b < grids.size(); ++b) {
        FAB& fy = y[b];
        const FAB& fx = x[b];
        ORC_LOOP(grids[b], i, j, k) fy(i, j, k) += s * fx(i, j, k);
    }
}
void Op::assign(LevelData& y, const LevelData& x) const {
#pragma omp parallel for schedule(dynamic)
    for (size_t b = 0; b < grids.size(); ++b) y[b].copy(x[b]);  // whole FAB (MappedAMRPoissonOp.H:577-581)
}
void Op::scale(LevelData& y, double s) const {
#pragma omp parallel for schedule(dynamic)
    for (size_t b = 0; b < grids.size(); ++b) {
        FAB& fy = y[b];
        ORC_LOOP(grids[b], i, j, k) fy(i, j, k) *= s;
    }
}
void Op::create(LevelData& y, int ghosts) const { y.define(grids, 1, activeDirs * ghosts); }


// ----------------------------------------------------------------------------------------------
// Projector calculus: the MAC gradient and divergence around the pressure solve.
// ----------------------------------------------------------------------------------------------

// MAPPEDMACGRADORTHO, normal derivative -- DivCurlGradF.ChF:221-262.
void macgrad_ortho(FAB& grad, const FAB& phi, const FAB& Jga, const Box& edgeBox, double dxDir, int dir) {
    const int gi = dir == 0, gj = dir == 1, gk = dir == 2;
    const double dxinv = 1.0 / dxDir;
    ORC_LOOP(edgeBox, i, j, k) grad(i, j, k) = dxinv * Jga(i, j, k, dir) * (phi(i, j, k) - phi(i - gi, j - gj, k - gk));
}

// MAPPEDMACGRAD, normal derivative (gradDir == edgeDir) -- DivCurlGradF.ChF:87-147.
void macgrad_full(FAB& grad, const FAB& phi, const FAB& ex, const FAB& Jga, const Box& edgeBox, const double* dx,
                         int adir, int nd) {
    const int bdir = (adir + 1) % nd, cdir = (adir + 2) % nd;
    const int ai = adir == 0, aj = adir == 1, ak = adir == 2;
    const int bi = bdir == 0, bj = bdir == 1, bk = bdir == 2;
    const int ci = cdir == 0, cj = cdir == 1, ck = cdir == 2;
    const double dxinva = 1.0 / dx[adir];
    const double dxinvb = 0.25 / dx[bdir];
    const double dxinvc = (nd > 2) ? 0.25 / dx[cdir] : 0.0;
    ORC_LOOP(edgeBox, i, j, k) {
        double f = Jga(i, j, k, adir) * dxinva * (phi(i, j, k) - phi(i - ai, j - aj, k - ak)) +
                   Jga(i, j, k, bdir) * dxinvb *
                       (ex(i + bi, j + bj, k + bk) - ex(i - bi, j - bj, k - bk) + ex(i + bi - ai, j + bj - aj, k + bk - ak) -
                        ex(i - bi - ai, j - bj - aj, k - bk - ak));
        if (nd > 2)
            f += Jga(i, j, k, cdir) * dxinvc *
                 (ex(i + ci, j + cj, k + ck) - ex(i - ci, j - cj, k - ck) + ex(i + ci - ai, j + cj - aj, k + ck - ak) -
                  ex(i - ci - ai, j - cj - aj, k - ck - ak));
        grad(i, j, k) = f;
    }
}

// Gradient::levelGradientMAC, one component = normal gradients only -- Gradient.cpp:85-170, with
// singleBoxMacGrad -- :946-1100 (the non-USE_OLD_EXTRAP branch): exchange with corners, order-2
// extrapolation of the transverse-derivative copy past the DOMAIN (validPhi = FAB box & validDomain),
// BC ghosts (a_fluxBC->setGhosts, non-homogeneous unless asked), MAPPEDMACGRAD[ORTHO], BC fluxes.
// LevelMACProjector::computeGrad -- LevelMACProjector.cpp:186-218 calls it with the pressure BC.
// EllipticExtrapBCGhostClass(order 2) -- EllipticBCUtils.cpp (operator() -> setSideExtrapBC :224-310, cell-centred
// branch) -> ELLIPTICEXTRAPBCGHOST, EllipticBCUtilsF.ChF:115-173: ghost = 3 (s1 - s2) + s3 counted inward.
static void extrapGhostBC(FAB& phi, const Box& valid, const Domain& domain, int nd) {
    for (int idir = 0; idir < nd; ++idir) {
        if (domain.periodic[idir]) continue;
        for (int iside = 0; iside < 2; ++iside) {
            const int isign = sign(iside);
            Box destBox = bdryBox(domain.box, idir, iside, 1) & surroundingNodes(valid, idir);
            if (destBox.isEmpty()) continue;
            destBox.shiftHalf(idir, isign);
            destBox &= phi.box();
            if (destBox.isEmpty()) continue;
            const int e0 = (idir == 0) ? -isign : 0, e1 = (idir == 1) ? -isign : 0, e2 = (idir == 2) ? -isign : 0;
            ORC_LOOP(destBox, i, j, k)
                phi(i, j, k) = 3.0 * (phi(i + e0, j + e1, k + e2) - phi(i + 2 * e0, j + 2 * e1, k + 2 * e2)) +
                               phi(i + 3 * e0, j + 3 * e1, k + 3 * e2);
        }
    }
}

void Op::macGradient(FluxLevelData& grad, LevelData& phi, bool homog, int gradBC) const {
    exchange(phi, true);                                               // Gradient.cpp:120-121
    Box vdom = domain.box;                                             // :117,123-127
    for (int d = 0; d < nd; ++d)
        if (domain.periodic[d]) vdom.grow(d, 1);
    for (size_t b = 0; b < grids.size(); ++b) {
        FAB& phiFAB = phi[b];
        const Box validPhi = phiFAB.box() & vdom;                      // :137-141
        for (int dir = 0; dir < nd; ++dir) {
            FAB& g = grad.fabs[b][dir];
            const Box edgeBox = surroundingNodes(grids[b], dir);      // :148-149
            FAB ex(phiFAB.box(), 1);
            if (!isDiagonal) {                                         // :984-998
                ex.copy(phiFAB);
                Box valid = validPhi;
                for (int fdir = 0; fdir < nd; ++fdir) {
                    ExtrapolateFaceAndCopy(ex, ex, valid, fdir, Lo, 2, nd);
                    ExtrapolateFaceAndCopy(ex, ex, valid, fdir, Hi, 2, nd);
                    valid.grow(fdir, 1);
                }
            }
            if (gradBC == 1) extrapGhostBC(phiFAB, grids[b], domain, nd);
            else setGhosts(phiFAB, &ex, grids[b], b, homog);           // :1004-1020
            if (isDiagonal) macgrad_ortho(g, phiFAB, Jgup.fabs[b][dir], edgeBox, dx[dir], dir);        // :1070-1078
            else macgrad_full(g, phiFAB, ex, Jgup.fabs[b][dir], edgeBox, dx, dir, nd);                 // :1060-1067
        }
        if (gradBC != 1) setFluxes(grad.fabs[b], grids[b], homog);     // :1082-1096 (each direction touches its own FAB)
    }
}

// ----------------------------------------------------------------------------------------------
// Leptic solver pieces that depend on per-box ghost cells (full metrics; the diagonal case reduces to
// oracle/leptic_ref.py's global-array statements and is tested against them).
// ----------------------------------------------------------------------------------------------
// extrapAllGhosts -- src/calculus/extrapolation/ExtrapolationUtils.cpp:388-421: per box, before any exchange;
// direction by direction, the valid box growing so that edges and corners are reached.
void Op::lepticExtrapAllGhosts(LevelData& phi, int order) const {
    int numghosts = phi.ghost[0];
    for (int d = 1; d < nd; ++d) numghosts = std::min(numghosts, phi.ghost[d]);
    for (size_t b = 0; b < grids.size(); ++b) {
        Box valid = grids[b];
        for (int ghost = 0; ghost < numghosts; ++ghost)
            for (int dir = 0; dir < nd; ++dir) {
                ExtrapolateFaceNoEV(phi[b], phi[b], valid, dir, Lo, order);
                ExtrapolateFaceNoEV(phi[b], phi[b], valid, dir, Hi, order);
                valid.grow(dir, 1);
            }
    }
}

// LevelLepticSolver::computeHorizRHS -- LevelLepticSolver.cpp:981-1097: every ghost extrapolated (order 2), exchange
// (+ corners for a full metric), MAPPEDMACGRAD[ORTHO] of the horizontal directions on the faces interior to the DOMAIN
// (boundary faces take the BoundaryData values: zero in the horizontal directions), vertical average
// (UNMAPPEDVERTINTEGRAL, scale 1/Nz), -divergence (LEPTICACCUMDIV).  Boxes span the vertical (columns are whole).
void Op::lepticHorizRHS(double* flatRhs, LevelData& phi) const {
    const int vd = nd - 1;
    const Box& dom = domain.box;
    const int Nz = dom.size(vd);
    const long nx = dom.size(0);
    lepticExtrapAllGhosts(phi, 2);                                         // :1004
    exchange(phi, false);                                                  // :1006
    if (!isDiagonal) exchange(phi, true);                                  // :1008-1014 (no CF region on one level)
    for (size_t b = 0; b < grids.size(); ++b) {
        const Box& valid = grids[b];
        if (valid.lo[vd] != dom.lo[vd] || valid.hi[vd] != dom.hi[vd]) { fprintf(stderr, "lepticHorizRHS: box does not span the vertical\n"); abort(); }
        Box flat = valid;
        flat.hi[vd] = flat.lo[vd];
        FAB rhs(flat, 1);                                                  // setValLevel(a_rhs, 0) :1017
        for (int FCdir = 0; FCdir < vd; ++FCdir) {
            Box domInterior = surroundingNodes(dom, FCdir);
            domInterior.grow(FCdir, -1);                                   // :997-1001
            const Box faceBox = surroundingNodes(valid, FCdir);
            const Box interiorFaceBox = faceBox & domInterior;
            FAB grad(faceBox, 1);                                          // boundary faces: bdryData = 0 (:1063-1069)
            if (!interiorFaceBox.isEmpty()) {
                if (isDiagonal) macgrad_ortho(grad, phi[b], Jgup.fabs[b][FCdir], interiorFaceBox, dx[FCdir], FCdir);
                else macgrad_full(grad, phi[b], phi[b], Jgup.fabs[b][FCdir], interiorFaceBox, dx, FCdir, nd);
            }
            const Box flatFace = surroundingNodes(flat, FCdir);
            FAB avg(flatFace, 1);
            const double dzScale = 1.0 / double(Nz);
            const int k0 = flat.lo[vd];
            ORC_LOOP(faceBox, i, j, k) {                                   // the vertical index ascends for every column
                int a[3] = {i, j, k};
                a[vd] = k0;
                avg(a[0], a[1], a[2]) = avg(a[0], a[1], a[2]) + grad(i, j, k) * dzScale;
            }
            const double dxScale = -1.0 / dx[FCdir];
            const int ei = FCdir == 0, ej = FCdir == 1;
            ORC_LOOP(flat, i, j, k) rhs(i, j, k) = rhs(i, j, k) + (avg(i + ei, j + ej, k) - avg(i, j, k)) * dxScale;
        }
        ORC_LOOP(flat, i, j, k) {
            const long col = (nd > 2) ? (i - dom.lo[0]) + nx * (j - dom.lo[1]) : (i - dom.lo[0]);
            flatRhs[col] = rhs(i, j, k);
        }
    }
}

// LevelLepticSolver::levelVertHorizGradient -- LevelLepticSolver.cpp:1107-1176 + LEPTICVERTHORIZGRAD,
// LevelLepticSolverF.ChF:74-122: scale * sum_m J g^{z m} d_m phi on the lower / upper DOMAIN faces, from phi with
// its vertical domain ghosts extrapolated in place (order 2, ExtrapolateFaceAndCopy: faces, then the edges of the
// ghost layer).  Zero for a diagonal metric (:1125).  No exchange is made ("non-blocking, fast, and seems to work").
void Op::lepticVertHorizGradient(double* flatLo, double* flatHi, LevelData& phi, double scale) const {
    const int vd = nd - 1;
    const Box& dom = domain.box;
    const long nx = dom.size(0);
    const long ncol = (nd > 2) ? nx * dom.size(1) : nx;
    for (long c = 0; c < ncol; ++c) flatLo[c] = flatHi[c] = 0.0;           // a_vertBdryData.setVal(0.0) :1122
    if (isDiagonal) return;
    for (int side = 0; side < 2; ++side) {
        const int isign = side == Lo ? -1 : 1;
        double* out = side == Lo ? flatLo : flatHi;
        for (size_t b = 0; b < grids.size(); ++b) {
            const Box& valid = grids[b];
            if (valid.sideEnd(side, vd) != dom.sideEnd(side, vd)) continue;
            FAB& p = phi[b];
            const Box domValid = p.box() & dom;                             // :1147-1148
            ExtrapolateFaceAndCopy(p, p, domValid, vd, side, 2, nd);       // :1149
            Box dest = valid;                                               // bdryBox(valid, vd, side, 1): FC faces
            dest.surroundingNodes(vd);
            if (side == Lo) dest.hi[vd] = dest.lo[vd]; else dest.lo[vd] = dest.hi[vd];
            const FAB& Jgz = Jgup.fabs[b][vd];
            const double dxinv0 = scale * 0.25 / dx[0];
            const double dxinv1 = (nd > 2) ? scale * 0.25 / dx[1] : 0.0;
            // offsets from the face index to the ghost-side and valid-side cells (CHF_AUTOID[gi; ...; -(1-isign)/2])
            const int go = -(1 - isign) / 2, vo = -(1 + isign) / 2;
            ORC_LOOP(dest, i, j, k) {
                double v;
                if (nd > 2) {
                    v = Jgz(i, j, k, 0) * dxinv0 * (p(i + 1, j, k + go) - p(i - 1, j, k + go) + p(i + 1, j, k + vo) - p(i - 1, j, k + vo)) +
                        Jgz(i, j, k, 1) * dxinv1 * (p(i, j + 1, k + go) - p(i, j - 1, k + go) + p(i, j + 1, k + vo) - p(i, j - 1, k + vo));
                    out[(i - dom.lo[0]) + nx * (j - dom.lo[1])] = v;
                } else {
                    v = Jgz(i, j, k, 0) * dxinv0 * (p(i + 1, j + go, k) - p(i - 1, j + go, k) + p(i + 1, j + vo, k) - p(i - 1, j + vo, k));
                    out[i - dom.lo[0]] = v;
                }
            }
        }
    }
}

// Divergence::levelDivergenceMAC with a_fluxBC == NULL (the caller's fluxes carry their boundary values)
// -- Divergence.cpp:44-127 -> MAPPEDFLUXDIVERGENCE{2,3}D, DivCurlGradF.ChF:1034-1214.
void Op::macDivergence(LevelData& div, const FluxLevelData& flux) const {
    for (size_t b = 0; b < grids.size(); ++b) flux_divergence(div[b], flux.fabs[b].data(), Jinv[b], grids[b], dx, nd);
}

void Op::wallFluxBC(FluxLevelData& flux) const {
    for (size_t b = 0; b < grids.size(); ++b)
        for (int d = 0; d < nd; ++d) {
            if (domain.periodic[d]) continue;
            for (int iside = 0; iside < 2; ++iside) {
                Box thisDom = domain.box;
                thisDom.surroundingNodes(d);
                const Box destBox = bdryBox(thisDom, d, iside, 1) & surroundingNodes(grids[b], d);
                if (destBox.isEmpty()) continue;
                flux.fabs[b][d].setVal(0.0, destBox, 0, 1);
            }
        }
}

// CellToEdge(LevelData<FArrayBox>, LevelData<FluxBox>) with SpaceDim cell components per face component -- Chombo 3.2
// lib/src/AMRTools/CellToEdge.cpp + CellToEdgeF.ChF (third party, not under /root/reference; published algorithm):
//   edge_d(i) = half * (cell_d(i) + cell_d(i - e_d))   on the faces of the valid box; one ghost cell of cell_d is read.
void Op::cellToEdge(FluxLevelData& edge, std::array<LevelData, 3>& cell) const {
    for (size_t b = 0; b < grids.size(); ++b)
        for (int d = 0; d < nd; ++d) {
            const FAB& c = cell[d][b];
            FAB& e = edge.fabs[b][d];
            const int e0 = (d == 0), e1 = (d == 1), e2 = (d == 2);
            ORC_LOOP(surroundingNodes(grids[b], d), i, j, k) e(i, j, k) = 0.5 * (c(i, j, k) + c(i - e0, j - e1, k - e2));
        }
}

// EdgeToCell -- Chombo 3.2 lib/src/AMRTools/EdgeToCell.cpp + EdgeToCellF.ChF:
//   cell_d(i) = half * (edge_d(i) + edge_d(i + e_d))
void Op::edgeToCell(std::array<LevelData, 3>& cell, const FluxLevelData& edge) const {
    for (size_t b = 0; b < grids.size(); ++b)
        for (int d = 0; d < nd; ++d) {
            FAB& c = cell[d][b];
            const FAB& e = edge.fabs[b][d];
            const int e0 = (d == 0), e1 = (d == 1), e2 = (d == 2);
            ORC_LOOP(grids[b], i, j, k) c(i, j, k) = 0.5 * (e(i, j, k) + e(i + e0, j + e1, k + e2));
        }
}

// BaseProjector<FArrayBox>::project on one level (BaseProjectorI.H:176-300) with LevelCCProjector's virtuals
// (LevelCCProjector.cpp:156-255):
//   [vel *= J]                                           LevelGeometry::multByJ, LevelGeometryUtil.cpp:302-340 (cached CCJ)
//   rhs = levelDivergenceCC(vel) / dt                    Divergence.cpp:359-396 (USE_SIMPLE_STENCIL is off):
//                                                        CellToEdge, then levelDivergenceMAC with m_divBC = uStarFuncBC
//   solve L[p] = rhs                                     m_solverBC = LevelPressureFuncBC
//   corr = levelGradientCC(p)                            Gradient.cpp:469-494: levelGradientMAC with m_gradBC =
//                                                        gradPiFuncBC (order-2 extrapolated ghosts), then EdgeToCell
//   vel += corr * dtScale                                LevelCCProjector::applyCorrection :236-254
//   [vel *= Jinv]                                        LevelGeometry::divByJ :464-500
// The caller has filled the ghosts of vel (AMRNavierStokes::doCCProjection sets BCs and exchanges first,
// AMRNavierStokesAdvancePPM.cpp:1513-1530); here box-interface and periodic ghosts come from an exchange and the
// ghosts beyond solid walls are never used (the wall faces are overwritten by m_divBC).
void MG::projectCC(std::array<LevelData, 3>& vel, const LevelData* J, LevelData& pressure, double dt, bool zeroPressure,
                   bool forceHomogeneous, SolveStats& st) {
    Op& op = *ops[0];
    const int nd = op.nd;
    if (J)
        for (int d = 0; d < nd; ++d)
            for (size_t b = 0; b < op.grids.size(); ++b) ORC_LOOP(op.grids[b], i, j, k) vel[d][b](i, j, k) *= (*J)[b](i, j, k);
    for (int d = 0; d < nd; ++d) op.exchange(vel[d], false);
    FluxLevelData edge;
    edge.define(op.grids, 1, nd);
    op.cellToEdge(edge, vel);
    op.wallFluxBC(edge);
    LevelData rhs;
    op.create(rhs, 0);
    op.macDivergence(rhs, edge);
    if (dt != 0.0)
        for (size_t b = 0; b < op.grids.size(); ++b) ORC_LOOP(op.grids[b], i, j, k) rhs[b](i, j, k) /= dt;
    solve(pressure, rhs, zeroPressure, forceHomogeneous, st);
    FluxLevelData eg;
    eg.define(op.grids, 1, nd);
    op.macGradient(eg, pressure, false, 1);
    std::array<LevelData, 3> corr;
    for (int d = 0; d < nd; ++d) op.create(corr[d], 0);
    op.edgeToCell(corr, eg);
    const double dtScale = (dt == 0.0) ? -1.0 : -dt;
    for (int d = 0; d < nd; ++d)
        for (size_t b = 0; b < op.grids.size(); ++b) {
            ORC_LOOP(op.grids[b], i, j, k) vel[d][b](i, j, k) += corr[d][b](i, j, k) * dtScale;
            if (J) ORC_LOOP(op.grids[b], i, j, k) vel[d][b](i, j, k) *= op.Jinv[b](i, j, k);
        }
}

// BaseProjector::project with a_velIsFlux = true on one level -- BaseProjectorI.H:176-300:
// rhs = Div[J u] / dt; solve L[p] = rhs; J u -= dt * (J g^{ij} d_j p)  (applyCorrection,
// LevelMACProjector.cpp:221-240: dtScale = -dt, or -1 when dt == 0).
void MG::project(FluxLevelData& vel, LevelData& pressure, double dt, bool zeroPressure, bool forceHomogeneous,
                 SolveStats& st) {
    Op& op = *ops[0];
    LevelData rhs;
    op.create(rhs, 0);
    op.macDivergence(rhs, vel);
    if (dt != 0.0)
        for (size_t b = 0; b < op.grids.size(); ++b)
            ORC_LOOP(op.grids[b], i, j, k) rhs[b](i, j, k) /= dt;      // FArrayBox::operator/=(Real)
    solve(pressure, rhs, zeroPressure, forceHomogeneous, st);
    FluxLevelData corr;
    corr.define(op.grids, 1, op.nd);
    op.macGradient(corr, pressure, false);
    const double dtScale = (dt == 0.0) ? -1.0 : -dt;
    for (size_t b = 0; b < op.grids.size(); ++b)
        for (int d = 0; d < op.nd; ++d) {
            FAB& v = vel.fabs[b][d];
            const FAB& c = corr.fabs[b][d];
            ORC_LOOP(v.box(), i, j, k) v(i, j, k) += c(i, j, k) * dtScale;   // FArrayBox::plus(src, scale)
        }
}


// ----------------------------------------------------------------------------------------------
// Factory: MappedAMRPoissonOpFactory::MGnewOp -- MappedAMRPoissonOpFactory.cpp:362-702 (single AMR
// level, unforced ratios) and fill_MGfields -- :1069-1236.
// ----------------------------------------------------------------------------------------------
static bool coarsenable(const std::vector<Box>& boxes, const IntVect& r) {
    for (const Box& b : boxes)
        if (!(b.coarsened(r).refined(r) == b)) return false;
    return true;
}

// Returns false when the factory would return NULL for this depth.  On success `coarsening` and `dx`
// describe the new depth and the ratio (depth-1 -> depth) is appended to `ratios`.
static bool mgSchedule(int depth, const std::vector<Box>& boxes0, const double* dx0, int nd, int maxDepth,
                       std::vector<IntVect>& ratios, IntVect& coarsening, double* dx) {
    if (maxDepth >= 0 && depth > maxDepth) return false;                       // :373
    const IntVect maskedMaxCoarse(4, nd > 1 ? 4 : 1, nd > 2 ? 4 : 1);          // Factory.cpp:131-132, s_maxCoarse = 4
    for (int d = 0; d < 3; ++d) dx[d] = dx0[d];
    coarsening = IntVect::Unit();
    IntVect mgRefRatio = IntVect::Unit();
    for (int i = 0; i < depth; ++i) {                                          // :463
        if (i < depth - 1) {
            mgRefRatio = ratios[i];
        } else {
            double maxDx = 0.0;                                                // :476-489
            for (int dir = 0; dir < nd; ++dir) maxDx = std::max(maxDx, dx[dir]);
            mgRefRatio = IntVect::Unit();
            for (int dir = 0; dir < nd; ++dir)
                if (dx[dir] <= maxDx / 2.0) mgRefRatio[dir] = 2;
            if (mgRefRatio.product() == 1) mgRefRatio = IntVect(2, nd > 1 ? 2 : 1, nd > 2 ? 2 : 1);
        }
        for (int d = 0; d < 3; ++d) dx[d] *= mgRefRatio[d];
        coarsening = coarsening * mgRefRatio;
    }
    if (coarsening.product() > 1 && !coarsenable(boxes0, coarsening * maskedMaxCoarse)) {   // :504
        for (int d = 0; d < 3; ++d) { dx[d] /= mgRefRatio[d]; coarsening[d] /= mgRefRatio[d]; }  // 1. refine back
        mgRefRatio = IntVect::Unit();                                                     // 2. enough cells?
        for (int dir = 0; dir < nd; ++dir) {
            mgRefRatio[dir] = 2;
            if (!coarsenable(boxes0, coarsening * maskedMaxCoarse * mgRefRatio)) mgRefRatio[dir] = 1;
        }
        if (mgRefRatio.product() == 1) return false;
        int refDir;                                                                       // 3. worth it?
        for (refDir = 0; refDir < nd; ++refDir)
            if (mgRefRatio[refDir] == 1) break;
        // NOTE: if every direction can be coarsened refDir == nd and the reference reads dx[dims]
        // (out of range, Factory.cpp:531-535).  That cannot happen here: the first attempt would have
        // succeeded with all directions unless the anisotropic choice failed; we then use the last dir.
        if (refDir >= nd) refDir = nd - 1;
        for (int dir = 0; dir < nd; ++dir)
            if (mgRefRatio[dir] > 1 && dx[dir] / dx[refDir] > 0.5) mgRefRatio[dir] = 1;
        if (mgRefRatio.product() == 1) return false;
        for (int d = 0; d < 3; ++d) dx[d] *= mgRefRatio[d];                               // 4. coarsen
        coarsening = coarsening * mgRefRatio;
        if (coarsening.product() > 1 && !coarsenable(boxes0, coarsening * maskedMaxCoarse)) return false;
    }
    if (depth > 0) ratios.push_back(mgRefRatio);                                          // :553
    return true;
}

std::unique_ptr<MG> buildMG(const Domain& dom, const std::vector<Box>& grids, const double* dx0, int nd, double alpha,
                            double beta, bool isDiagonal, const BCSpec& bc, const SolverParams& p,
                            FluxLevelData&& Jgup0, LevelData&& Jinv0) {
    std::unique_ptr<MG> mg(new MG);
    mg->p = p;
    std::vector<IntVect> ratios;
    for (int depth = 0;; ++depth) {
        IntVect coarsening;
        double dx[3];
        if (!mgSchedule(depth, grids, dx0, nd, p.maxDepth, ratios, coarsening, dx)) break;
        std::unique_ptr<Op> op(new Op);
        op->nd = nd;
        op->domain = dom;
        op->domain.box = dom.box.coarsened(coarsening);
        op->grids.resize(grids.size());
        for (size_t b = 0; b < grids.size(); ++b) op->grids[b] = grids[b].coarsened(coarsening);
        for (int d = 0; d < 3; ++d) { op->dx[d] = dx[d]; op->dxCrse[d] = std::nan(""); }
        op->alpha = alpha; op->beta = beta; op->isDiagonal = isDiagonal; op->bc = bc;
        op->relaxMode = p.relaxMode; op->precondMode = p.precondMode; op->preCondSmoothIters = p.num_smooth_precond;
        op->mgCrseRefRatio = (depth > 0) ? ratios.back() : IntVect::Unit();
        if (depth == 0) {
            op->Jgup = std::move(Jgup0);
            op->Jinv = std::move(Jinv0);
        } else {
            // fill_MGfields: face average of Jgup, harmonic average of Jinv  (:1164-1180)
            const Op& fine = *mg->ops.back();
            const IntVect r = ratios.back();
            const int ncJ = fine.Jgup.fabs[0][0].nComp();
            op->Jgup.define(op->grids, ncJ, nd);
            op->Jinv.define(op->grids, 1, IntVect::Zero());
            for (size_t b = 0; b < grids.size(); ++b) {
                for (int d = 0; d < nd; ++d) average_face(op->Jgup.fabs[b][d], fine.Jgup.fabs[b][d], op->Jgup.fabs[b][d].box(), d, r);
                average_harmonic(op->Jinv[b], fine.Jinv[b], op->grids[b], r);
            }
        }
        op->computeLapDiag();
        op->defineCommon();
        // Prolongation choice: null-space probe L[1]-L[0] -- :659-693
        {
            LevelData phi, rhs, res;
            phi.define(op->grids, 1, op->activeDirs);
            rhs.define(op->grids, 1, IntVect::Zero());
            res.define(op->grids, 1, IntVect::Zero());
            op->applyOp(rhs, phi, true);
            for (auto& f : phi.fabs) f.setVal(1.0);
            op->residual(res, phi, rhs, true);
            double mx = -1e300;  // computeMax: signed maximum, then abs()
            for (size_t b = 0; b < op->grids.size(); ++b) {
                const FAB& f = res[b];
                ORC_LOOP(op->grids[b], i, j, k) mx = std::max(mx, f(i, j, k));
            }
            const double maxNorm = std::fabs(mx);
            const double thresh = 0.01 * p.eps;
            op->zeroAvgProlong = (maxNorm < thresh);
            op->dxProduct = dx[0] * (nd > 1 ? dx[1] : 1.0) * (nd > 2 ? dx[2] : 1.0);
        }
        // the finer op learns its ratio to this depth (MappedMultiGrid::define :383-386)
        if (depth > 0) mg->ops.back()->mgCrseRefRatio = ratios.back();
        mg->ops.push_back(std::move(op));
    }
    mg->mgRefRatios = ratios;
    const int nDepth = mg->depth();
    mg->residual.resize(nDepth);
    mg->correction.resize(nDepth);
    for (int d = 1; d < nDepth; ++d) {   // MappedMultiGrid::init -- MappedMultiGrid.H:421-434
        mg->residual[d].define(mg->ops[d]->grids, 1, IntVect::Zero());
        mg->correction[d].define(mg->ops[d]->grids, 1, mg->ops[d]->activeDirs);
    }
    mg->bottomCells = mg->ops.back()->domain.box.numPts();
    return mg;
}

// ----------------------------------------------------------------------------------------------
// MappedMultiGrid<T>::cycle -- MappedMultiGrid.H:555-653
// ----------------------------------------------------------------------------------------------
void MG::cycle(int depth, LevelData& e, const LevelData& r) {
    Op& op = *ops[depth];
    if (depth == this->depth() - 1) {
        if (bottomCells == 1) {
            op.relax(e, r, 1);
        } else {
            op.relax(e, r, p.num_smooth_bottom);
            bottom_iters_total += bottomSolve(e, r);
        }
        return;
    }
    int cycles = p.numMG;
    if (cycles < 0) {
        cycles = -cycles;  // F-cycle branch :578-619
        op.restrictResidual(residual[depth + 1], e, r, *ops[depth + 1]);
        op.setToZero(correction[depth + 1]);
        cycle(depth + 1, correction[depth + 1], residual[depth + 1]);
        op.prolongIncrement(e, correction[depth + 1]);
        op.relax(e, r, p.num_smooth_down);
        for (int img = 0; img < cycles; ++img) {
            op.restrictResidual(residual[depth + 1], e, r, *ops[depth + 1]);
            op.setToZero(correction[depth + 1]);
            const int save = p.numMG;
            p.numMG = 1;
            cycle(depth + 1, correction[depth + 1], residual[depth + 1]);
            p.numMG = save;
            op.prolongIncrement(e, correction[depth + 1]);
        }
        op.relax(e, r, p.num_smooth_up);
    } else {
        op.relax(e, r, p.num_smooth_down);                                         // :622
        op.restrictResidual(residual[depth + 1], e, r, *ops[depth + 1]);           // :628
        op.setToZero(correction[depth + 1]);                                       // :629
        for (int img = 0; img < cycles; ++img) cycle(depth + 1, correction[depth + 1], residual[depth + 1]);
        op.prolongIncrement(e, correction[depth + 1]);                             // :638
        op.relax(e, r, p.num_smooth_up);                                           // :641
    }
}

// ----------------------------------------------------------------------------------------------
// Bottom solver: Chombo 3.1 BiCGStabSolver<T>::solve.  Chombo is NOT under /root/reference
// (AMRPressureSolver.cpp:253-262 only configures it): restated from the published preconditioned
// BiCGStab (van der Vorst 1992) in the form Chombo uses -- convergence against
// eps * convergenceMetric (setConvergenceMetrics, MappedAMRMultiGrid.H:1047) or reps * initial residual,
// hang detection with restarts.  PARITY UNPINNED for this routine.
// Returns the number of iterations.
// ----------------------------------------------------------------------------------------------
int MG::bottomSolve(LevelData& a_phi, const LevelData& a_rhs) {
    Op& op = *ops.back();
    LevelData r, r_tilde, e, pv, p_tilde, s_tilde, t, v;
    op.create(r, 0); op.create(r_tilde, 0); op.create(e, 1); op.create(pv, 0);
    op.create(p_tilde, 1); op.create(s_tilde, 1); op.create(t, 0); op.create(v, 0);
    int recount = 0;
    op.residual(r, a_phi, a_rhs, true);
    op.assign(r_tilde, r);
    op.setToZero(e); op.setToZero(p_tilde); op.setToZero(s_tilde);
    int i = 0;
    double rho[4] = {0, 0, 0, 0};
    double nrm[2];
    nrm[0] = op.norm(r, p.bottom_normType);
    double initial_norm = nrm[0];
    const double initial_rnorm = nrm[0];
    nrm[1] = nrm[0];
    double alph[2] = {0, 0}, bet[2] = {0, 0}, omega[2] = {0, 0};
    bool init = true;
    int restarts = 0;
    if (bottomMetric > 0) initial_norm = bottomMetric;
    const double eps = bottomEps;
    while ((i < p.bottom_imax && nrm[0] > eps * nrm[1]) && (nrm[1] > 0)) {
        ++i;
        nrm[1] = nrm[0];
        alph[1] = alph[0]; bet[1] = bet[0]; omega[1] = omega[0];
        rho[3] = rho[2]; rho[2] = rho[1];
        rho[1] = op.dot(r_tilde, r);
        if (rho[1] == 0.0) break;  // will not converge any further
        if (init) {
            op.assign(pv, r);
            init = false;
        } else {
            bet[1] = (rho[1] / rho[2]) * (alph[1] / omega[1]);
            op.scale(pv, bet[1]);
            op.incr(pv, v, -bet[1] * omega[1]);
            op.incr(pv, r, 1.0);
        }
        op.preCond(p_tilde, pv);
        op.applyOp(v, p_tilde, true);
        const double m = op.dot(r_tilde, v);
        alph[0] = rho[1] / m;
        if (std::fabs(m) > p.bottom_small * std::fabs(rho[1])) {
            op.incr(r, v, -alph[0]);
            nrm[0] = op.norm(r, p.bottom_normType);
            op.incr(e, p_tilde, alph[0]);
        } else {
            op.setToZero(r);
            nrm[0] = 0.0;
        }
        if (nrm[0] > eps * initial_norm && nrm[0] > p.bottom_reps * initial_rnorm) {
            op.preCond(s_tilde, r);
            op.applyOp(t, s_tilde, true);
            omega[0] = op.dot(t, r) / op.dot(t, t);
            op.incr(e, s_tilde, omega[0]);
            op.incr(r, t, -omega[0]);
            nrm[0] = op.norm(r, p.bottom_normType);
        }
        if (nrm[0] <= eps * initial_norm || nrm[0] <= p.bottom_reps * initial_rnorm) break;
        if (omega[0] == 0.0 || nrm[0] > (1.0 - p.bottom_hang) * nrm[1]) {
            if (recount == 0) {
                recount = 1;
            } else {
                recount = 0;
                op.incr(a_phi, e, 1.0);
                if (restarts == p.bottom_numRestarts) { op.setToZero(e); break; }
                op.residual(r, a_phi, a_rhs, true);
                nrm[0] = op.norm(r, p.bottom_normType);
                rho[0] = rho[1] = rho[2] = rho[3] = 0.0;
                alph[0] = alph[1] = bet[0] = bet[1] = omega[0] = omega[1] = 0.0;
                op.setToZero(e);
                ++restarts;
                init = true;
                op.assign(r_tilde, r);
            }
        }
    }
    op.incr(a_phi, e, 1.0);
    return i;
}

// ----------------------------------------------------------------------------------------------
// MappedAMRMultiGrid<T>::solveNoInitResid -- MappedAMRMultiGrid.H:979-1186 with l_base == l_max == 0:
// AMRVCycle (:1510-1513) -> MappedMultiGrid::oneCycle (homogeneous) -> cycle(0, corr, resid);
// computeAMRResidual (:793-839) = residual + max-norm; postVCycleOps (:1189-1215).
// ----------------------------------------------------------------------------------------------
void MG::solve(LevelData& phi, const LevelData& rhs, bool zeroPhi, bool forceHomogeneous, SolveStats& st) {
    Op& op = *ops[0];
    LevelData uberCorrection, uberResidual, bestPhi;
    op.create(uberCorrection, 1);
    op.create(uberResidual, 0);
    op.create(bestPhi, 1);
    op.setToZero(uberCorrection);
    if (zeroPhi) op.setToZero(phi);
    op.assign(bestPhi, phi);
    auto computeResidual = [&]() {
        op.residual(uberResidual, phi, rhs, forceHomogeneous);
        return op.norm(uberResidual, 0);
    };
    double initial_rnorm = computeResidual();
    if (convergenceMetric != 0.0) initial_rnorm = convergenceMetric;   // :1036-1038
    bool somethingConverged = false;                                   // :1043
    double rnorm = initial_rnorm, norm_last = 2 * initial_rnorm, best_rnorm = rnorm;
    bool useBestPhi = false;
    bottomMetric = initial_rnorm;   // setConvergenceMetrics(initial_rnorm, cushion(=1) * eps) :1047
    bottomEps = 1.0 * p.eps;
    bottom_iters_total = 0;
    st.history.clear();
    st.history.push_back(rnorm);
    int iter = 0;
    bool goNorm = rnorm > p.normThresh;
    bool goRedu = rnorm > p.eps * initial_rnorm;
    bool goIter = iter < p.imax;
    bool goHang = iter < p.imin || rnorm < (1 - p.hang) * norm_last;
    while (goIter && goRedu && goHang && goNorm) {
        norm_last = rnorm;
        cycle(0, uberCorrection, uberResidual);          // AMRVCycle, single level
        op.incr(phi, uberCorrection, 1.0);               // postVCycleOps
        op.setToZero(uberCorrection);
        rnorm = computeResidual();
        ++iter;
        st.history.push_back(rnorm);
        if (rnorm <= best_rnorm) {
            best_rnorm = rnorm;
            op.assign(bestPhi, phi);
            useBestPhi = false;
            somethingConverged = true;
        } else {
            useBestPhi = true;
        }
        if (p.verbosity >= 3)
            fprintf(stderr, "   %s oracle MG iteration = %d, residual norm = %.6e, rate = %.3f\n", useBestPhi ? "[D]" : "   ",
                    iter, rnorm, rnorm > 0 ? norm_last / rnorm : 0.0);
        goNorm = rnorm > p.normThresh;
        goRedu = rnorm > p.eps * initial_rnorm;
        goIter = iter < p.imax;
        goHang = iter < p.imin || rnorm < (1 - p.hang) * norm_last;
    }
    if (useBestPhi) {
        rnorm = best_rnorm;
        op.assign(phi, bestPhi);
    }
    blewUp = (!somethingConverged && rnorm >= initial_rnorm && rnorm >= p.eps);   // :1140-1144
    st.iters = iter;
    st.initial_rnorm = initial_rnorm;
    st.final_rnorm = rnorm;
    st.exitStatus = int(!goRedu) + int(!goIter) * 2 + int(!goHang) * 4 + int(!goNorm) * 8;   // :1148
    st.bottom_iters_total = bottom_iters_total;
}

}  // namespace orc
