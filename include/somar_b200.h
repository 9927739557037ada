/* somar_b200.h -- C ABI of libsomar_b200.so: SOMAR's mapped-Poisson multigrid hot path on B200 (sm_100a).
 *
 * The reference (UNC-CFD/somar) has no FFI for this path; its boundary is the C++ virtual-class contract
 *   LinearOp<T> -> MappedMGLevelOp<T> -> MappedAMRLevelOp<T> -> MappedAMRPoissonOp        (T = LevelData<FArrayBox>)
 * plus the Chombo-Fortran kernel calling convention (double* + lo/hi + ncomp, x fastest, ghosts inside the box;
 * SURVEY.md 8(b)).  Every entry point below names the reference interface it replaces (paths relative to the
 * reference tree).  A maintainer binds these from the adapter classes in include/somar_b200/ (C++ headers).
 *
 * Conventions
 *   - All functions return SOMAR_OK (0) or a negative error code; somar_last_error() gives the text.
 *     (The reference aborts through MayDay::Error; a C ABI cannot, so the adapter turns codes into MayDay calls.)
 *   - One caller thread per level; work is stream-ordered on the level's CUDA stream.  Functions that return a
 *     scalar to the host (norm, dot, solve) synchronise that stream.
 *   - Fields are opaque device-resident objects created per multigrid depth.  Host data crosses the boundary in the
 *     CHF_FRA convention: `const double* fab, lo[3], hi[3]` = a Fortran-ordered array over the box [lo,hi]
 *     (ghost cells included), absolute (global) cell indices.  2-D problems use n[2] == 1.
 *   - Lifetime of sources: somar_field_set / set_face / somar_level_set_metric_device with a DEVICE (or pinned host)
 *     source only enqueue the copy on the level stream and return.  The caller keeps the source alive and unmodified
 *     until somar_level_sync() (or an event recorded on somar_level_stream()) says the copy has run.
 *   - There is no CPU fallback: every call fails with SOMAR_ERR_CUDA when no sm_100 device is usable.
 */
#ifndef SOMAR_B200_H
#define SOMAR_B200_H

#ifdef __cplusplus
extern "C" {
#endif

#define SOMAR_OK 0
#define SOMAR_ERR_ARG (-1)
#define SOMAR_ERR_CUDA (-2)
#define SOMAR_ERR_LAYOUT (-3)
#define SOMAR_ERR_COMM (-4)
#define SOMAR_ERR_DIVERGED (-5) /* residual grew 10x ("kaboom", MappedAMRMultiGrid.H:1134-1138), or no V-cycle improved
                                   on the initial guess ("solver blew up", :1140-1144) */

/* BCType codes: src/calculus/BCInterface/BCDescriptor.H:33-38 */
#define SOMAR_BC_NONE (-1)
#define SOMAR_BC_NEUM 0
#define SOMAR_BC_DIRI 1
#define SOMAR_BC_PERIODIC 2

/* AMRMG.relax_mode: src/utils/ProblemContext.H:324-338 */
#define SOMAR_RELAX_NONE (-1)
#define SOMAR_RELAX_JACOBI 0
#define SOMAR_RELAX_LEVEL_GSRB 1
#define SOMAR_RELAX_LOOSE_GSRB 2
#define SOMAR_RELAX_LINE_GSRB 3

typedef struct SomarLevel SomarLevel; /* one AMR level: all multigrid depths, metrics, scratch, halo plans */
typedef struct SomarField SomarField; /* one cell-centred scalar field (1 component) at one depth           */

/* What MappedAMRPoissonOpFactory::define + MGnewOp receive
 * (src/calculus/AMRElliptic/MappedAMRPoissonOpFactory.cpp:60-200,362-702) flattened into plain C. */
typedef struct SomarLevelDesc {
    int dim;           /* 2 or 3 (CH_SPACEDIM) */
    int n[3];          /* domain cells; domain box is [0, n-1] */
    int periodic[3];   /* ProblemDomain::isPeriodic */
    int nboxes;        /* DisjointBoxLayout of this level: ALL boxes of the level (they fix the MG schedule) */
    const int* box_lo; /* nboxes x 3 */
    const int* box_hi; /* nboxes x 3, inclusive */
    const int* box_rank; /* nboxes owner ranks, or NULL = all on rank 0.  The boxes a rank owns must tile one
                            rectangular region that spans the whole domain in x and z (columns stay whole:
                            GSRB.H:90, ProblemContext.cpp:249-255). */
    double dx[3];      /* computational cell size (LevelGeometry::getDx) */
    double alpha, beta;/* (alpha I + beta L): MappedAMRPoissonOp::setAlphaAndBeta */
    int is_diagonal;   /* LevelGeometry::isDiagonal(): J g^{ij} = 0 for i != j */
    int bc_type[6];    /* [dir*2+side]: BCMethodHolder / BCDescriptor codes for non-periodic sides */
    double bc_value[6];/* constant Neumann flux / Dirichlet value (EllipticConstNeumBC*, EllipticConstDiriBC*) */
    int rank, nranks;  /* this process and the job size (1 = single GPU) */
} SomarLevelDesc;

/* AMRMG.* and bottom.* parameters: src/utils/ProblemContext.cpp:1147-1216, AMRPressureSolver.cpp:46-81 */
typedef struct SomarSolverParams {
    double eps;
    int num_smooth_down, num_smooth_up, num_smooth_bottom, num_smooth_precond;
    int numMG, imin, imax;
    double hang, normThresh;
    int maxDepth, relaxMode, precondMode;
    double bottom_eps, bottom_reps, bottom_hang, bottom_small;
    int bottom_imax, bottom_numRestarts, bottom_normType;
    int verbosity;
} SomarSolverParams;

/* MappedAMRMultiGrid<T>::m_exitStatus, m_iterations and the per-iteration residual trace
 * (MappedAMRMultiGrid.H:1050-1055,1105-1117,1148). */
typedef struct SomarSolveStats {
    int iters, exitStatus;
    double initial_rnorm, final_rnorm;
    int bottom_iters_total;
    int history_len;
    double history[128];
    double solve_ms;       /* device time of the solve (CUDA events on the level stream) */
    long long launches;    /* kernels of this library launched during the solve */
} SomarSolveStats;

/* ---- library ------------------------------------------------------------------------------------------- */
const char* somar_last_error(void);
const char* somar_version(void);
int somar_default_params(SomarSolverParams* p);       /* ProblemContext.cpp:1147-1216 defaults */
int somar_device_count(int* count);

/* ---- communicator (replaces Chombo_MPI::comm for LevelData::exchange and the scalar MPI_Allreduce calls) --
 * 128-byte NCCL unique id made on rank 0 and sent to the other ranks by the caller (MPI_Bcast / torch.distributed). */
int somar_comm_unique_id(unsigned char id[128]);

/* ---- level ---------------------------------------------------------------------------------------------
 * somar_level_create replaces MappedAMRPoissonOpFactory::define + MappedMultiGrid::define
 * (MappedAMRPoissonOpFactory.cpp:60-200; MappedMultiGrid.H:328-404): builds every multigrid depth
 * (semicoarsening schedule, Factory.cpp:466-553), allocates device metrics and scratch.  `device` is the CUDA
 * ordinal; `comm_id` is NULL when nranks == 1.  Metrics are supplied afterwards with somar_level_set_metric_fab
 * and frozen by somar_level_finalize, which builds the coarse metrics (fill_MGfields, Factory.cpp:1069-1236),
 * lapDiag (FILLMAPPEDLAPDIAG*, MappedAMRPoissonOpF.ChF:91-274) and runs the null-space probe (:659-693). */
int somar_level_create(const SomarLevelDesc* desc, const SomarSolverParams* params, int device,
                       const unsigned char* comm_id, SomarLevel** out);
int somar_level_destroy(SomarLevel* lev);
/* Host metric FABs of one box: jgup[d] is the FluxBox component d (face-centred in d, `jgup_ncomp` comps outermost,
 * comp c = J g^{dc}; ncomp==1 means only J g^{dd}), box [lo,hi] with hi[d]+1 faces in d; jinv is cell-centred.
 * Replaces LevelGeometry::getFCJgup()/getCCJinv() reads (LevelGeometry.H:130-286). Either pointer set may be NULL. */
int somar_level_set_metric_fab(SomarLevel* lev, const int lo[3], const int hi[3], int jgup_ncomp,
                               const double* jgup0, const double* jgup1, const double* jgup2, const double* jinv);
/* Same, from DEVICE arrays covering the rank-local region without ghosts (x fastest), for device-resident callers. */
int somar_level_set_metric_device(SomarLevel* lev, int jgup_ncomp, const double* d_jgup0, const double* d_jgup1,
                                  const double* d_jgup2, const double* d_jinv);
int somar_level_finalize(SomarLevel* lev);
int somar_level_set_alpha_beta(SomarLevel* lev, double alpha, double beta); /* MappedAMRPoissonOp::setAlphaAndBeta */
/* MappedAMRMultiGrid<T>::m_convergenceMetric (MappedAMRMultiGrid.H:1036-1038): when non-zero it replaces the initial
 * residual norm in the relative stopping test of somar_solve and as the bottom solver's metric; 0 restores the default. */
int somar_level_set_convergence_metric(SomarLevel* lev, double metric);
/* AMRPressureSolver::setAMRMGParameters / setBottomParameters (src/projection/AMRPressureSolver.cpp:46-81) on an
 * existing level: every AMRMG.* / bottom.* knob except maxDepth (it fixes the schedule built at create time). */
int somar_level_set_solver_params(SomarLevel* lev, const SomarSolverParams* params);
int somar_level_num_depths(const SomarLevel* lev, int* ndepths);
/* n = domain cells at that depth, ratio = mgCrseRefRatio to the next coarser depth, local_lo/hi = this rank's region */
int somar_level_depth_info(const SomarLevel* lev, int depth, int n[3], int ratio[3], double dx[3], int* zero_avg,
                           int local_lo[3], int local_hi[3]);
int somar_level_sync(SomarLevel* lev);                 /* cudaStreamSynchronize on the level stream */
int somar_level_stream(SomarLevel* lev, void** cuda_stream);
int somar_level_launch_count(const SomarLevel* lev, long long* launches);
/* Which kernel does relax(mode) run at `depth` (benchmark bookkeeping): 0 = k_gsrb_ortho, one launch per colour
 * pass; 1 = k_gsrb_fused, one launch per red+black iteration; 2 = k_gsrb_full (colour pass, full metric);
 * 3 = k_line_gsrb (colour pass); 4 = apply + k_jacobi; 5 = k_gsrb_async (as 1, pipelined); 6 = k_gsrb_loose (four
 * stage launches per iteration, diagonal metric); 7 = k_gsrb_tma (as 5, rings fed by TMA tensor-map loads);
 * 8 = k_line_pass_c (colour pass of the line smoother in the colour-split layout); -1 = none. */
int somar_level_relax_kernel(const SomarLevel* lev, int depth, int relax_mode, int* kernel_id);
/* Which path do y-halo exchanges take on this level: 0 = none needed (one rank), 1 = pack + ncclSend/Recv + unpack,
 * 2 = peer-memory push/pull kernels over CUDA IPC (halo_p2p.cu); valid after somar_level_finalize. */
int somar_level_halo_path(const SomarLevel* lev, int* path);
/* Reads back a depth's metric into rank-local arrays without ghosts (tests): comp-outermost for jg. */
int somar_level_get_metric(SomarLevel* lev, int depth, double* jg0, double* jg1, double* jg2, double* jinv,
                           double* lapdiag);
/* Phase timers on the level stream (CUDA events; replaces the CH_TIMERS of MappedAMRPoissonOp.cpp:777-781 and
 * MappedAMRMultiGrid.H:988).  Phases: 0 = depth-0 smoother colour-pass KERNEL (per launch), 1 = smoothing at the
 * intermediate depths, 2 = restrictResidual, 3 = prolongIncrement, 4 = bottom (relax + BiCGStab), 5 = outer loop
 * (phi += e, residual, norm, best-phi copy).  profile_read synchronises, fills count[6]/ms[6] and clears. */
int somar_level_profile_enable(SomarLevel* lev, int on);
int somar_level_profile_read(SomarLevel* lev, long long count[6], double ms[6]);

/* ---- fields: LinearOp<T>::create / clear (MappedAMRPoissonOp.H:566-575) ---------------------------------- */
int somar_field_create(SomarLevel* lev, int depth, SomarField** out);
int somar_field_destroy(SomarField* f);
/* host FAB [lo,hi] (ghosts included) -> device, only cells inside `valid` (the DisjointBoxLayout box) are taken */
int somar_field_set_fab(SomarField* f, const double* fab, const int lo[3], const int hi[3], const int vlo[3],
                        const int vhi[3]);
int somar_field_get_fab(const SomarField* f, double* fab, const int lo[3], const int hi[3], const int vlo[3],
                        const int vhi[3]);
/* rank-local region, no ghosts, x fastest; `on_device` != 0 means the pointer is a device pointer */
int somar_field_set(SomarField* f, const double* data, int on_device);
int somar_field_get(const SomarField* f, double* data, int on_device);

/* ---- LinearOp<T> (MappedAMRPoissonOp.H:187-266, .cpp:628-898) -------------------------------------------- */
int somar_apply_op(SomarLevel* lev, int depth, SomarField* lhs, SomarField* phi, int homogeneous);   /* applyOpI :772 */
int somar_residual(SomarLevel* lev, int depth, SomarField* res, SomarField* phi, const SomarField* rhs,
                   int homogeneous);                                                                /* residualI :646 */
int somar_precond(SomarLevel* lev, int depth, SomarField* cor, const SomarField* res);              /* preCond :684 */
int somar_set_to_zero(SomarLevel* lev, SomarField* x);
int somar_assign(SomarLevel* lev, SomarField* y, const SomarField* x);
int somar_incr(SomarLevel* lev, SomarField* y, const SomarField* x, double scale);
int somar_axby(SomarLevel* lev, SomarField* lhs, const SomarField* x, const SomarField* y, double a, double b);
int somar_scale(SomarLevel* lev, SomarField* x, double scale);
int somar_norm(SomarLevel* lev, const SomarField* x, int ord, double* out);       /* CH_XD::norm, all ranks */
int somar_local_max_norm(SomarLevel* lev, const SomarField* x, double* out);      /* localMaxNorm */
int somar_dot(SomarLevel* lev, const SomarField* a, const SomarField* b, double* out);

/* ---- allocation-free norm ---------------------------------------------------------------------------------
 * somar_norm_noalloc has NO counterpart in the reference: it was added for
 * callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid allocation).
 * It is the reduction behind somar_norm (ord 0: max |x|, 1: sum |x|, 2: sqrt(sum x^2)) over a DENSE device array of
 * n doubles, with the result written to the device address d_out, enqueued on `cuda_stream` (NULL = the default
 * stream) of the CURRENT device; it does not synchronise.  Its scratch (per-block partials + ticket) is a file-scope
 * static __device__ array of 128 KiB + 256 B that every GPU loading the library holds until exit, so
 * n <= SOMAR_NOALLOC_MAX_N; larger n returns SOMAR_ERR_ARG.  No cudaMalloc / cudaFree / allocating routine is called.
 * All calls share that array: calls that overlap on one GPU (several streams or host threads) race -- serialise them. */
#define SOMAR_NOALLOC_MAX_N 67108864LL /* 16384 blocks x 4096 doubles */
int somar_norm_noalloc(const double* d_x, long long n, int ord, double* d_out, void* cuda_stream);

/* ---- MappedMGLevelOp<T> (MappedMultiGrid.H:43-114) -------------------------------------------------------- */
int somar_relax(SomarLevel* lev, int depth, SomarField* phi, const SomarField* rhs, int iters);     /* relax :1733 */
int somar_restrict_residual(SomarLevel* lev, int depth, SomarField* res_coarse, SomarField* phi_fine,
                            const SomarField* rhs_fine);                                  /* restrictResidual :1281 */
int somar_prolong_increment(SomarLevel* lev, int depth, SomarField* phi_fine, const SomarField* cor_coarse);

/* ---- projector calculus: the callers either side of the solve ----------------------------------
 * Face fields are ordinary SomarFields of depth 0 holding face-centred data: face (i,j,k) of direction d is the low
 * face of cell (i,j,k), the high face of a row's last cell sits in the first ghost slot.  Upload / download them
 * with somar_field_set_fab/get_fab and the FluxBox FAB's own (surroundingNodes) box, or densely with
 * somar_field_set_face/get_face (rank-local array with n[dir]+1 points in direction dir).
 * somar_mac_gradient:   g_d = J g^{dj} d_j phi with the level's BCs      Gradient::levelGradientMAC, Gradient.cpp:85-170,
 *                       946-1100 -> MAPPEDMACGRAD[ORTHO], DivCurlGradF.ChF:87-262 (LevelMACProjector.cpp:186-218)
 * somar_mac_divergence: div = Jinv sum_d (F_d(+) - F_d(-)) / dx_d, then / divide_by unless that is 0
 *                       Divergence::levelDivergenceMAC, Divergence.cpp:44-127 -> MAPPEDFLUXDIVERGENCE*, DivCurlGradF.ChF:1034-1214
 * somar_mac_apply_correction: v_d -= dt * (J g^{dj} d_j phi)  (dt == 0: scale -1); gradient never stored
 *                       LevelMACProjector::applyCorrection, LevelMACProjector.cpp:221-240
 * somar_project:        BaseProjector::project with a_velIsFlux = true on one level (BaseProjectorI.H:176-300):
 *                       rhs = Div[v]/dt, solve L[pressure] = rhs, v -= dt Grad[pressure]. */
int somar_field_set_face(SomarField* f, const double* data, int dir, int on_device);
int somar_field_get_face(const SomarField* f, double* data, int dir, int on_device);
int somar_mac_gradient(SomarLevel* lev, SomarField* gx, SomarField* gy, SomarField* gz, SomarField* phi, int homogeneous);
int somar_mac_divergence(SomarLevel* lev, SomarField* div, const SomarField* fx, const SomarField* fy,
                         const SomarField* fz, double divide_by);
int somar_mac_apply_correction(SomarLevel* lev, SomarField* vx, SomarField* vy, SomarField* vz, SomarField* phi, double dt);
/* The reference projectors' own BC holders (LevelMACProjector.cpp:99-101, LevelCCProjector.cpp:103-105):
 *   m_divBC  = PhysBCUtil::uStarFuncBC: on solid walls the normal face value is set to zero IN PLACE
 *              (BasicVelocityBCGhostClass -> setSideDiriBC face-centred branch, EllipticBCUtils.cpp:41-118; inflow /
 *              outflow / tidal sides are not built)                                  -> somar_mac_wall_flux_bc
 *   m_gradBC = PhysBCUtil::basicGradPressureFuncBC: EllipticExtrapBCGhostClass(2), ghost = 3 (p1 - p2) + p3 and no
 *              flux method (ELLIPTICEXTRAPBCGHOST, EllipticBCUtilsF.ChF:165-173)    -> grad_bc = 1
 * somar_mac_gradient keeps the operator's BC holder (grad_bc = 0: with homogeneous Neumann pressure BCs a wall face gets
 * no correction, which is what the reference's velocity BCs restore right after the projection,
 * AMRNavierStokesAdvancePPM.cpp:287-294).  somar_level_set_projector_bc(lev, 1) makes somar_project apply both holders
 * itself, as LevelMACProjector::computeDiv / computeGrad do; 0 (default) is the round-1 behaviour. */
int somar_level_set_projector_bc(SomarLevel* lev, int mode);
int somar_mac_gradient_bc(SomarLevel* lev, SomarField* gx, SomarField* gy, SomarField* gz, SomarField* phi, int homogeneous,
                          int grad_bc);
int somar_mac_wall_flux_bc(SomarLevel* lev, SomarField* fx, SomarField* fy, SomarField* fz);

/* ---- cell-centred projector: LevelCCProjector (src/projection/LevelCCProjector.cpp:156-255), the second projection of
 * the reference's time step (AMRNavierStokes::doCCProjection, AMRNavierStokesAdvancePPM.cpp:1513-1573) ----------------
 * u* are CELL fields holding the velocity components; J is the cell-centred Jacobian field the reference caches
 * (LevelGeometry::getCCJ) or NULL when the velocity is already a flux (a_velIsFlux = true).  The library exchanges the
 * halos of u* and J itself; ghosts beyond solid walls are not read.
 * somar_cc_divergence: Divergence::levelDivergenceCC (Divergence.cpp:359-396): [multByJ,] CellToEdge, m_divBC,
 *                      MAPPEDFLUXDIVERGENCE, then / divide_by unless that is 0 -- one pass, no face arrays
 * somar_cc_gradient:   Gradient::levelGradientCC (Gradient.cpp:469-494): levelGradientMAC with m_gradBC, EdgeToCell
 * somar_project_cc:    BaseProjector<FArrayBox>::project on one level (BaseProjectorI.H:176-300):
 *                      [u *= J;] rhs = Div[u]/dt; solve; u -= dt Grad[p]; [u *= Jinv]  -- two sweeps around the solve */
int somar_cc_divergence(SomarLevel* lev, SomarField* div, SomarField* ux, SomarField* uy, SomarField* uz, SomarField* J,
                        double divide_by);
int somar_cc_gradient(SomarLevel* lev, SomarField* gx, SomarField* gy, SomarField* gz, SomarField* phi);
int somar_project_cc(SomarLevel* lev, SomarField* ux, SomarField* uy, SomarField* uz, SomarField* J, SomarField* pressure,
                     double dt, int zero_pressure, int force_homogeneous, SomarSolveStats* stats);

/* ---- leptic-solver sweeps (SURVEY 8(a) LevelLepticSolver row; the iteration that drives them is somar_leptic_solve below) --
 * "Flat" arguments are depth-0 fields whose vertical index 0 holds a horizontal array (the reference's VertShifter,
 * LevelLepticSolver.cpp:1213-1226).  BC codes are the reference's BCType values: 0 Neum, 1 Diri, 3 CF.
 * somar_leptic_vertical_solve:  per column, d_z (J g^zz d_z phi) = rhs with homogeneous vertical BCs; Neumann boundary
 *      VALUES (bdry_lo / bdry_hi, flat, may be NULL) are rolled into rhs as rhs -/+ bc/dz and unrolled afterwards.
 *      LevelLepticSolver::verticalLineSolver, LevelLepticSolver.cpp:1247-1425 -> TRIDIAGPOISSONNN1DFAB
 *      (src/utils/TridiagUtilsF.ChF:85-164) for Neum-Neum, else LEPTICLAPACKVERTICALSOLVER -> LAPACK dptsv
 *      (LevelLepticSolverF.ChF:161-289).  A column whose factorisation breaks down before the last pivot is an error.
 * somar_leptic_vert_integral:   flat += sum_k func * dz_scale               UNMAPPEDVERTINTEGRAL, SubspaceF.ChF:33-58
 * somar_leptic_vertical_excess: excess = bdry_hi - bdry_lo - Integral[rhs]  computeVerticalExcess, LevelLepticSolver.cpp:1183-1240
 * somar_leptic_add_extrusion:   dest(i,j,k) += flat(i,j)                    ADDEXTRUSION, SubspaceF.ChF:66-110
 * somar_leptic_accum_div:       flat_rhs += (flux(i+e) - flux(i)) * dx_scale LEPTICACCUMDIV, LevelLepticSolverF.ChF:291-331 */
int somar_leptic_vertical_solve(SomarLevel* lev, SomarField* phi, SomarField* rhs, int lo_bc, int hi_bc, double dz_crse,
                                const SomarField* bdry_lo, const SomarField* bdry_hi);
int somar_leptic_vert_integral(SomarLevel* lev, SomarField* flat_sum, const SomarField* func, double dz_scale);
int somar_leptic_vertical_excess(SomarLevel* lev, SomarField* excess, const SomarField* rhs, const SomarField* bdry_lo,
                                 const SomarField* bdry_hi);
int somar_leptic_add_extrusion(SomarLevel* lev, SomarField* dest, const SomarField* flat_src);
int somar_leptic_accum_div(SomarLevel* lev, SomarField* flat_rhs, const SomarField* flat_flux, int fc_dir, double dx_scale);

/* ---- the leptic iteration: LevelLepticSolver (src/calculus/LepticSolver/LevelLepticSolver.cpp) ------------------------
 * somar_leptic_create  = LevelLepticSolver::define (:147-437) on a finalized level (the LepticOperator handed to define):
 *      builds the "vertical" level (same J g^{ij}, Jinv = 1, :197-205; it carries m_opPtr and the 3-D multigrid of the
 *      fall-back, :219-300, with the `full` parameters) and, when both vertical sides are Neumann
 *      (gatherVerticalBCTypes, :1522-1644), the flattened "horizontal" level with vertically averaged J g^{ab}
 *      (LevelGeometry::createVertAvgFCJgupPtr, src/geometry/LevelGeometryBasics.cpp:499-571), zero-Neumann BCs (:363-368)
 *      and the `horiz` multigrid parameters.  Everything is built on the device from the level's own metric.
 * somar_leptic_solve   = LevelLepticSolver::solve (:646-956): J*residual to the vertical problem, then per order the
 *      excess bookkeeping, the batched vertical solves, the horizontal right-hand side (computeHorizRHS, :981-1097) and
 *      multigrid solve, the new residual, the multigrid fall-back at the last order when the iteration hangs, and
 *      phi += sum of corrections.  `homogeneous` is setHomogeneous (:444).
 * Scope: one level without a coarser one; a diagonal metric on one rank or y-slab ranks (the vertical level shares the
 * level's communicator, the small horizontal level is replicated and its rows all-gathered: destroy the solver BEFORE
 * the level), a full (terrain-following) metric on one rank (3-D: levelVertHorizGradient :1107-1176, MAPPEDMACGRAD in
 * computeHorizRHS, a 2-D full-metric horizontal level, a horizontal solve at every order); vertical sides Neumann or
 * Dirichlet; with a
 * horizontal solve the level must be 3-D with non-periodic horizontal directions (the reference leaves the boundary
 * faces of a periodic direction unset in computeHorizRHS).  Anything else is an error, never a CPU path. */
typedef struct SomarLepticParams {
    int max_order;             /* setParameters (:524-541); defaults setDefaultParameters (:463-514) */
    double eps, hang;
    int norm_type;
    double horiz_rhs_tol;
    SomarSolverParams horiz;   /* setHorizMGParameters + setHorizBottomParameters */
    SomarSolverParams full;    /* setFullMGParameters + setFullBottomParameters */
} SomarLepticParams;
typedef struct SomarLepticStats {
    int orders;                /* O(eps^n) steps taken */
    int exit_status;           /* LevelLepticSolver::ExitStatus (LevelLepticSolver.H:156-165): -1 NONE, 0 CONVERGE, 1 ITER,
                                  2 HANG, 3 DIVERGE, 4 KABOOM */
    int horiz_solves, krylov_used;
    int horiz_iters_total;     /* V-cycles of the horizontal solves */
    int num_res_norms;         /* m_resNorms: [0] initial, then one per order */
    double res_norms[16];
    double solve_ms;           /* device time (CUDA events on the vertical level's stream, host-synchronised) */
    long long launches;        /* kernels launched on the three levels during the solve */
} SomarLepticStats;
typedef struct SomarLeptic SomarLeptic;
int somar_leptic_default_params(SomarLepticParams* p);
int somar_leptic_create(SomarLevel* lev, const SomarLepticParams* params, SomarLeptic** out);
int somar_leptic_destroy(SomarLeptic* s);
int somar_leptic_solve(SomarLeptic* s, SomarField* phi, const SomarField* rhs, int homogeneous, SomarLepticStats* stats);
/* The derived levels (tests, diagnostics): which = 0 vertical, 1 horizontal (NULL when no horizontal solve is needed). */
int somar_leptic_level(SomarLeptic* s, int which, SomarLevel** out);

/* ---- advection: the PPM / corner-transport-upwind scalar predictor (SURVEY 8(a) predictScalar row, 8(f)-3) -----------
 * MappedAdvectionUtil::predictScalar, src/advection/AdvectUtil.cpp:520-933 (PPMNormalPred :1158, PPMFaceValues :1246,
 * RiemannSolver :944, quasilinearUpdate :1009; kernels MappedGodunovUtilitiesF.ChF:35-468, AdvectUtilF.ChF:28-159).
 * whalf_d (face fields, out): time-centred face states, times adv_vel_d when return_flux (the flux whose
 * somar_mac_divergence gives the advective update, AMRNavierStokesAdvancePPM.cpp:379-620).  w_old, source (may be NULL),
 * old_vel_d (cell-centred components) and cc_j (cell-centred J, LevelGeometry::fill_J) are cell fields, adv_vel_d face
 * fields.  Built for FULLY PERIODIC levels on one rank or y-slab ranks (the ghost rows of the INPUT fields are refreshed
 * in place by the halo exchanges), the three normal predictors (CTU, PLM, PPM) and both limiter families; anything else
 * returns SOMAR_ERR_ARG (non-periodic sides need the physics BC holders, which are outside this library). */
typedef struct SomarAdvectParams {
    int normal_pred_order;        /* ProblemContext normalPredOrder{Vel,Scal}: 0 = CTU, 1 = PLM, 2 = PPM */
    int use_limiting;             /* useLimiting{Vel,Scal}: van Leer + PPM limiters */
    int use_high_order_limiter;   /* useHighOrderLimiter{Vel,Scal}: FourthInterpFaces + CheckCubicLimiterF
                                     (MappedGodunovUtilitiesF.ChF:517-566,611-745) instead of van Leer/PPMLimiterF */
    int use_upwinding;            /* useUpwinding{Vel,Scal}: 1 = RiemannSolver, 0 = AvgStates */
    int return_flux;              /* a_returnFlux */
    int use_fourth_order_slopes;  /* useFourthOrderSlopes{Vel,Scal}: PLM only (FourthSlopeDiffsF + the van Leer limiter on
                                     one-sided differences, AdvectUtil.cpp:1094-1128) */
} SomarAdvectParams;
int somar_advect_predict_scalar(SomarLevel* lev, SomarField* whalf_x, SomarField* whalf_y, SomarField* whalf_z,
                                const SomarField* w_old, const SomarField* source, const SomarField* old_vel_x,
                                const SomarField* old_vel_y, const SomarField* old_vel_z, const SomarField* adv_vel_x,
                                const SomarField* adv_vel_y, const SomarField* adv_vel_z, const SomarField* cc_j,
                                double dt, const SomarAdvectParams* prm);
/* The finite-volume update of AMRNavierStokes::getNewScalar (AMRNavierStokesAdvancePPM.cpp:561-584, no background /
 * sponge / diffusive sources):  w_new = w_old + (-Div[flux]) * dt,  Div = MAPPEDFLUXDIVERGENCE with the level's Jinv.
 * flux_d are face fields (high face of the last cell in the first ghost slot, as predict_scalar leaves them). */
int somar_advect_update(SomarLevel* lev, SomarField* w_new, const SomarField* w_old, const SomarField* flux_x,
                        const SomarField* flux_y, const SomarField* flux_z, double dt);

/* ---- drivers --------------------------------------------------------------------------------------------
 * somar_mg_cycle: MappedMultiGrid<T>::cycle (MappedMultiGrid.H:555-653) from `depth` down.
 * somar_solve: AMRPressureSolver::solve -> MappedAMRMultiGrid<T>::solveNoInitResid, single level
 * (AMRPressureSolver.cpp:494-560; MappedAMRMultiGrid.H:979-1186). */
int somar_mg_cycle(SomarLevel* lev, int depth, SomarField* e, const SomarField* r);
int somar_solve(SomarLevel* lev, SomarField* phi, const SomarField* rhs, int zero_phi, int force_homogeneous,
                SomarSolveStats* stats);
int somar_project(SomarLevel* lev, SomarField* vx, SomarField* vy, SomarField* vz, SomarField* pressure, double dt,
                  int zero_pressure, int force_homogeneous, SomarSolveStats* stats);

#ifdef __cplusplus
}
#endif
#endif /* SOMAR_B200_H */
