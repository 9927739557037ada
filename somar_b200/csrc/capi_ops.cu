// capi_ops.cu -- extern "C" operator / driver entry points (LinearOp, MappedMGLevelOp, solver).
#include "level.hpp"
#include <cmath>

using namespace somar;

#define CHECK_FINAL(L) do { if (!(L)) { set_error("null level"); return SOMAR_ERR_ARG; } \
                            if (!(L)->finalized) { set_error("level not finalized (call somar_level_finalize)"); return SOMAR_ERR_ARG; } \
                            cudaSetDevice((L)->device); } while (0)
#define CHECK_F(f, L, dpt) do { if (!(f) || (f)->lev != (L) || (f)->depth != (dpt)) { set_error("field does not belong to this level/depth"); return SOMAR_ERR_ARG; } } while (0)

extern "C" {

int somar_level_profile_enable(SomarLevel* L, int on) {
    if (!L) { set_error("null level"); return SOMAR_ERR_ARG; }
    L->prof_on = on != 0;
    return SOMAR_OK;
}
int somar_level_profile_read(SomarLevel* L, long long count[6], double ms[6]) {
    if (!L) { set_error("null level"); return SOMAR_ERR_ARG; }
    cudaSetDevice(L->device);
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    for (int i = 0; i < 6; ++i) { count[i] = 0; ms[i] = 0.0; }
    for (auto& r : L->prof) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess && r.phase >= 0 && r.phase < 6) {
            count[r.phase]++;
            ms[r.phase] += t;
        }
        L->prof_pool.push_back(r.a);
        L->prof_pool.push_back(r.b);
    }
    L->prof.clear();
    return SOMAR_OK;
}

int somar_apply_op(SomarLevel* L, int depth, SomarField* lhs, SomarField* phi, int homogeneous) {
    CHECK_FINAL(L); CHECK_F(lhs, L, depth); CHECK_F(phi, L, depth);
    return op_apply(L, depth, lhs->p, phi, nullptr, 0, homogeneous != 0);
}
int somar_residual(SomarLevel* L, int depth, SomarField* res, SomarField* phi, const SomarField* rhs, int homogeneous) {
    CHECK_FINAL(L); CHECK_F(res, L, depth); CHECK_F(phi, L, depth); CHECK_F(rhs, L, depth);
    if (res == rhs) { set_error("residual: lhs must not alias rhs"); return SOMAR_ERR_ARG; }
    return op_apply(L, depth, res->p, phi, rhs->p, 1, homogeneous != 0);
}
int somar_precond(SomarLevel* L, int depth, SomarField* cor, const SomarField* res) {
    CHECK_FINAL(L); CHECK_F(cor, L, depth); CHECK_F(res, L, depth);
    return op_precond(L, depth, cor, res);
}
int somar_set_to_zero(SomarLevel* L, SomarField* x) {
    CHECK_FINAL(L); CHECK_F(x, L, x ? x->depth : 0);
    launch_fill(L->cx, L->depths[x->depth].g, x->p, 0.0);
    return SOMAR_OK;
}
int somar_assign(SomarLevel* L, SomarField* y, const SomarField* x) {
    CHECK_FINAL(L); CHECK_F(x, L, x ? x->depth : 0); CHECK_F(y, L, x->depth);
    launch_copy(L->cx, L->depths[x->depth].g, y->p, x->p);
    return SOMAR_OK;
}
int somar_incr(SomarLevel* L, SomarField* y, const SomarField* x, double scale) {
    CHECK_FINAL(L); CHECK_F(x, L, x ? x->depth : 0); CHECK_F(y, L, x->depth);
    launch_axpy(L->cx, L->depths[x->depth].g, y->p, x->p, scale);
    return SOMAR_OK;
}
int somar_axby(SomarLevel* L, SomarField* lhs, const SomarField* x, const SomarField* y, double a, double b) {
    CHECK_FINAL(L); CHECK_F(x, L, x ? x->depth : 0); CHECK_F(y, L, x->depth); CHECK_F(lhs, L, x->depth);
    launch_axby(L->cx, L->depths[x->depth].g, lhs->p, x->p, y->p, a, b);
    return SOMAR_OK;
}
int somar_scale(SomarLevel* L, SomarField* x, double scale) {
    CHECK_FINAL(L); CHECK_F(x, L, x ? x->depth : 0);
    launch_scale(L->cx, L->depths[x->depth].g, x->p, scale);
    return SOMAR_OK;
}
// CH_XD::norm semantics (MappedAMRPoissonOp.H:698-702): ord 0 = max |x|, 1 = sum |x|, 2 = sqrt(sum x^2)
int somar_norm(SomarLevel* L, const SomarField* x, int ord, double* out) {
    CHECK_FINAL(L); CHECK_F(x, L, x ? x->depth : 0);
    if (!out || ord < 0 || ord > 2) { set_error("norm: bad argument"); return SOMAR_ERR_ARG; }
    int rc = reduce_global(L, x->depth, x->p, nullptr, ord, out);
    if (rc) return rc;
    if (ord == 2) *out = std::sqrt(*out);
    return SOMAR_OK;
}
int somar_local_max_norm(SomarLevel* L, const SomarField* x, double* out) {
    CHECK_FINAL(L); CHECK_F(x, L, x ? x->depth : 0);
    launch_reduce(L->cx, L->depths[x->depth].g, x->p, nullptr, 0, L->d_scal);
    SOMAR_CUDA(cudaMemcpyAsync(L->h_scal, L->d_scal, sizeof(double), cudaMemcpyDeviceToHost, L->cx.stream));
    SOMAR_CUDA(cudaStreamSynchronize(L->cx.stream));
    *out = L->h_scal[0];
    return SOMAR_OK;
}
int somar_dot(SomarLevel* L, const SomarField* a, const SomarField* b, double* out) {
    CHECK_FINAL(L); CHECK_F(a, L, a ? a->depth : 0); CHECK_F(b, L, a->depth);
    return reduce_global(L, a->depth, a->p, b->p, 3, out);
}

int somar_relax(SomarLevel* L, int depth, SomarField* phi, const SomarField* rhs, int iters) {
    CHECK_FINAL(L); CHECK_F(phi, L, depth); CHECK_F(rhs, L, depth);
    return op_relax(L, depth, phi, rhs, iters);
}
int somar_restrict_residual(SomarLevel* L, int depth, SomarField* res_coarse, SomarField* phi_fine,
                            const SomarField* rhs_fine) {
    CHECK_FINAL(L); CHECK_F(res_coarse, L, depth + 1); CHECK_F(phi_fine, L, depth); CHECK_F(rhs_fine, L, depth);
    return op_restrict(L, depth, res_coarse, phi_fine, rhs_fine);
}
int somar_prolong_increment(SomarLevel* L, int depth, SomarField* phi_fine, const SomarField* cor_coarse) {
    CHECK_FINAL(L); CHECK_F(phi_fine, L, depth); CHECK_F(cor_coarse, L, depth + 1);
    return op_prolong(L, depth, phi_fine, cor_coarse);
}

int somar_mg_cycle(SomarLevel* L, int depth, SomarField* e, const SomarField* r) {
    CHECK_FINAL(L); CHECK_F(e, L, depth); CHECK_F(r, L, depth);
    return mg_cycle(L, depth, e, r);
}
int somar_solve(SomarLevel* L, SomarField* phi, const SomarField* rhs, int zero_phi, int force_homogeneous,
                SomarSolveStats* stats) {
    CHECK_FINAL(L); CHECK_F(phi, L, 0); CHECK_F(rhs, L, 0);
    if (!stats) { set_error("null stats"); return SOMAR_ERR_ARG; }
    return mg_solve(L, phi, rhs, zero_phi != 0, force_homogeneous != 0, stats);
}

// ---- projector calculus ------------------------------------------------------------------
static int check_faces(SomarLevel* L, SomarField* const f[3]) {
    const int nd = L->desc.dim;
    for (int d = 0; d < nd; ++d)
        if (!f[d] || f[d]->lev != L || f[d]->depth != 0) { set_error("face field does not belong to depth 0 of this level"); return SOMAR_ERR_ARG; }
    for (int a = 0; a < nd; ++a)
        for (int b = a + 1; b < nd; ++b)
            if (f[a] == f[b]) { set_error("face fields must be distinct"); return SOMAR_ERR_ARG; }
    return SOMAR_OK;
}
int somar_mac_gradient(SomarLevel* L, SomarField* gx, SomarField* gy, SomarField* gz, SomarField* phi, int homogeneous) {
    CHECK_FINAL(L); CHECK_F(phi, L, 0);
    SomarField* g[3] = {gx, gy, gz};
    int rc = check_faces(L, g);
    if (rc) return rc;
    for (int d = 0; d < L->desc.dim; ++d) if (g[d] == phi) { set_error("macGradient: output aliases phi"); return SOMAR_ERR_ARG; }
    return op_mac_gradient(L, g, phi, homogeneous != 0, 1.0, false);
}
int somar_mac_divergence(SomarLevel* L, SomarField* div, const SomarField* fx, const SomarField* fy, const SomarField* fz,
                         double divide_by) {
    CHECK_FINAL(L); CHECK_F(div, L, 0);
    SomarField* f[3] = {const_cast<SomarField*>(fx), const_cast<SomarField*>(fy), const_cast<SomarField*>(fz)};
    int rc = check_faces(L, f);
    if (rc) return rc;
    for (int d = 0; d < L->desc.dim; ++d) if (f[d] == div) { set_error("macDivergence: output aliases a flux"); return SOMAR_ERR_ARG; }
    return op_mac_divergence(L, div, f, divide_by);
}
int somar_mac_apply_correction(SomarLevel* L, SomarField* vx, SomarField* vy, SomarField* vz, SomarField* phi, double dt) {
    CHECK_FINAL(L); CHECK_F(phi, L, 0);
    SomarField* v[3] = {vx, vy, vz};
    int rc = check_faces(L, v);
    if (rc) return rc;
    const double dtScale = (dt == 0.0) ? -1.0 : -dt;
    return op_mac_gradient(L, v, phi, false, dtScale, true);
}
int somar_project(SomarLevel* L, SomarField* vx, SomarField* vy, SomarField* vz, SomarField* pressure, double dt,
                  int zero_pressure, int force_homogeneous, SomarSolveStats* stats) {
    CHECK_FINAL(L); CHECK_F(pressure, L, 0);
    if (!stats) { set_error("null stats"); return SOMAR_ERR_ARG; }
    SomarField* v[3] = {vx, vy, vz};
    int rc = check_faces(L, v);
    if (rc) return rc;
    for (int d = 0; d < L->desc.dim; ++d) if (v[d] == pressure) { set_error("project: pressure aliases a velocity component"); return SOMAR_ERR_ARG; }
    return project(L, v, pressure, dt, zero_pressure != 0, force_homogeneous != 0, stats);
}

int somar_level_set_projector_bc(SomarLevel* L, int mode) {
    if (!L) { set_error("null level"); return SOMAR_ERR_ARG; }
    if (mode != 0 && mode != 1) { set_error("projector BC mode must be 0 or 1"); return SOMAR_ERR_ARG; }
    L->projector_bc = mode;
    return SOMAR_OK;
}
int somar_mac_gradient_bc(SomarLevel* L, SomarField* gx, SomarField* gy, SomarField* gz, SomarField* phi, int homogeneous,
                          int grad_bc) {
    CHECK_FINAL(L); CHECK_F(phi, L, 0);
    if (grad_bc != 0 && grad_bc != 1) { set_error("grad_bc must be 0 or 1"); return SOMAR_ERR_ARG; }
    SomarField* g[3] = {gx, gy, gz};
    int rc = check_faces(L, g);
    if (rc) return rc;
    for (int d = 0; d < L->desc.dim; ++d) if (g[d] == phi) { set_error("macGradient: output aliases phi"); return SOMAR_ERR_ARG; }
    return op_mac_gradient(L, g, phi, homogeneous != 0, 1.0, false, grad_bc);
}
int somar_mac_wall_flux_bc(SomarLevel* L, SomarField* fx, SomarField* fy, SomarField* fz) {
    CHECK_FINAL(L);
    SomarField* f[3] = {fx, fy, fz};
    int rc = check_faces(L, f);
    if (rc) return rc;
    return op_wall_flux_bc(L, f);
}
// ---- cell-centred projector (LevelCCProjector) -----------------------------------------------------------
int somar_cc_divergence(SomarLevel* L, SomarField* div, SomarField* ux, SomarField* uy, SomarField* uz, SomarField* J,
                        double divide_by) {
    CHECK_FINAL(L); CHECK_F(div, L, 0);
    if (J) CHECK_F(J, L, 0);
    SomarField* u[3] = {ux, uy, uz};
    int rc = check_faces(L, u);
    if (rc) return rc;
    for (int d = 0; d < L->desc.dim; ++d)
        if (u[d] == div || u[d] == J) { set_error("ccDivergence: a velocity component aliases div or J"); return SOMAR_ERR_ARG; }
    if (J == div) { set_error("ccDivergence: J aliases div"); return SOMAR_ERR_ARG; }
    return op_cc_divergence(L, div, u, J, divide_by);
}
int somar_cc_gradient(SomarLevel* L, SomarField* gx, SomarField* gy, SomarField* gz, SomarField* phi) {
    CHECK_FINAL(L); CHECK_F(phi, L, 0);
    SomarField* g[3] = {gx, gy, gz};
    int rc = check_faces(L, g);
    if (rc) return rc;
    for (int d = 0; d < L->desc.dim; ++d) if (g[d] == phi) { set_error("ccGradient: output aliases phi"); return SOMAR_ERR_ARG; }
    return op_cc_gradient(L, g, phi);
}
int somar_project_cc(SomarLevel* L, SomarField* ux, SomarField* uy, SomarField* uz, SomarField* J, SomarField* pressure,
                     double dt, int zero_pressure, int force_homogeneous, SomarSolveStats* stats) {
    CHECK_FINAL(L); CHECK_F(pressure, L, 0);
    if (J) CHECK_F(J, L, 0);
    if (!stats) { set_error("null stats"); return SOMAR_ERR_ARG; }
    SomarField* u[3] = {ux, uy, uz};
    int rc = check_faces(L, u);
    if (rc) return rc;
    for (int d = 0; d < L->desc.dim; ++d)
        if (u[d] == pressure || u[d] == J) { set_error("projectCC: a velocity component aliases the pressure or J"); return SOMAR_ERR_ARG; }
    if (J == pressure) { set_error("projectCC: J aliases the pressure"); return SOMAR_ERR_ARG; }
    return project_cc(L, u, J, pressure, dt, zero_pressure != 0, force_homogeneous != 0, stats);
}

// ---- leptic-solver sweeps (SURVEY 8(a) LevelLepticSolver row) ---------------------------------------------
#define CHECK_F0(f, L) do { if (!(f) || (f)->lev != (L) || (f)->depth != 0) { set_error("field does not belong to depth 0 of this level"); return SOMAR_ERR_ARG; } } while (0)
int somar_leptic_vertical_solve(SomarLevel* L, SomarField* phi, SomarField* rhs, int lo_bc, int hi_bc, double dz_crse,
                                const SomarField* bdry_lo, const SomarField* bdry_hi) {
    CHECK_FINAL(L); CHECK_F0(phi, L); CHECK_F0(rhs, L);
    if (phi == rhs) { set_error("leptic vertical solve: phi aliases rhs"); return SOMAR_ERR_ARG; }
    if (bdry_lo) CHECK_F0(bdry_lo, L);
    if (bdry_hi) CHECK_F0(bdry_hi, L);
    return leptic_vertical_solve(L, phi, rhs, lo_bc, hi_bc, dz_crse, bdry_lo, bdry_hi);
}
int somar_leptic_vert_integral(SomarLevel* L, SomarField* flat_sum, const SomarField* func, double dz_scale) {
    CHECK_FINAL(L); CHECK_F0(flat_sum, L); CHECK_F0(func, L);
    if (flat_sum == func) { set_error("vertical integral: output aliases input"); return SOMAR_ERR_ARG; }
    return leptic_vert_integral(L, flat_sum, func, dz_scale);
}
int somar_leptic_vertical_excess(SomarLevel* L, SomarField* excess, const SomarField* rhs, const SomarField* bdry_lo,
                                 const SomarField* bdry_hi) {
    CHECK_FINAL(L); CHECK_F0(excess, L); CHECK_F0(rhs, L); CHECK_F0(bdry_lo, L); CHECK_F0(bdry_hi, L);
    if (excess == rhs) { set_error("vertical excess: output aliases rhs"); return SOMAR_ERR_ARG; }
    return leptic_vertical_excess(L, excess, rhs, bdry_lo, bdry_hi);
}
int somar_leptic_add_extrusion(SomarLevel* L, SomarField* dest, const SomarField* flat_src) {
    CHECK_FINAL(L); CHECK_F0(dest, L); CHECK_F0(flat_src, L);
    if (dest == flat_src) { set_error("add extrusion: dest aliases src"); return SOMAR_ERR_ARG; }
    return leptic_add_extrusion(L, dest, flat_src);
}
int somar_leptic_accum_div(SomarLevel* L, SomarField* flat_rhs, const SomarField* flat_flux, int fc_dir, double dx_scale) {
    CHECK_FINAL(L); CHECK_F0(flat_rhs, L); CHECK_F0(flat_flux, L);
    if (flat_rhs == flat_flux) { set_error("accum div: rhs aliases flux"); return SOMAR_ERR_ARG; }
    return leptic_accum_div(L, flat_rhs, flat_flux, fc_dir, dx_scale);
}

// ---- advection predictor (SURVEY 8(a) predictScalar row) --------------------------------------------------
int somar_advect_predict_scalar(SomarLevel* L, SomarField* whalf_x, SomarField* whalf_y, SomarField* whalf_z,
                                const SomarField* w_old, const SomarField* source, const SomarField* old_vel_x,
                                const SomarField* old_vel_y, const SomarField* old_vel_z, const SomarField* adv_vel_x,
                                const SomarField* adv_vel_y, const SomarField* adv_vel_z, const SomarField* cc_j,
                                double dt, const SomarAdvectParams* prm) {
    CHECK_FINAL(L); CHECK_F0(w_old, L); CHECK_F0(cc_j, L);
    if (!prm) { set_error("null parameters"); return SOMAR_ERR_ARG; }
    if (source) CHECK_F0(source, L);
    SomarField* wh[3] = {whalf_x, whalf_y, whalf_z};
    const SomarField* ov[3] = {old_vel_x, old_vel_y, old_vel_z};
    const SomarField* av[3] = {adv_vel_x, adv_vel_y, adv_vel_z};
    const int nd = L->desc.dim;
    for (int d = 0; d < nd; ++d) {
        CHECK_F0(wh[d], L); CHECK_F0(ov[d], L); CHECK_F0(av[d], L);
        if (wh[d] == w_old || wh[d] == cc_j || wh[d] == source) { set_error("predictScalar: an output aliases an input"); return SOMAR_ERR_ARG; }
        for (int e = 0; e < nd; ++e)
            if (wh[d] == ov[e] || wh[d] == av[e] || (e != d && wh[d] == wh[e])) { set_error("predictScalar: an output aliases another field"); return SOMAR_ERR_ARG; }
    }
    return advect_predict_scalar(L, wh, w_old, source, ov, av, cc_j, dt, prm);
}

int somar_advect_update(SomarLevel* L, SomarField* w_new, const SomarField* w_old, const SomarField* flux_x,
                        const SomarField* flux_y, const SomarField* flux_z, double dt) {
    CHECK_FINAL(L); CHECK_F0(w_new, L); CHECK_F0(w_old, L);
    const SomarField* f[3] = {flux_x, flux_y, flux_z};
    for (int d = 0; d < L->desc.dim; ++d) {
        CHECK_F0(f[d], L);
        if (f[d] == w_new) { set_error("advect update: the output aliases a flux"); return SOMAR_ERR_ARG; }
    }
    return advect_update(L, w_new, w_old, f, dt);
}

}  // extern "C"
