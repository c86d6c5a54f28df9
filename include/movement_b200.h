/* movement_b200.h -- C ABI of the B200-native mesh-relocation hot path.
 *
 * The reference (mesh-adaptation/movement) has no FFI: every arithmetic step is a
 * Firedrake `assemble` / `interpolate` or a PETSc KSP/SNES solve reached from Python.
 * Each entry point below replaces one such call site; the citation after "replaces:" is
 * the reference file:line (relative to the reference checkout) whose arithmetic the
 * function performs.  The Python package `movement_b200` binds these with ctypes
 * (movement_b200/_lib.py); INTEGRATION.md shows the stub a reference maintainer adds.
 *
 * Conventions
 *  - All array arguments are plain pointers.  Unless stated otherwise they may be
 *    DEVICE pointers (torch-owned) or HOST pointers: the library inspects them with
 *    cudaPointerGetAttributes and stages host buffers itself (inside the call).
 *  - Layouts are Firedrake's `dat.data` layouts: scalar [V], vector [V,d], tensor
 *    [V,d,d] (row-major), P0 [C]; cell-node map [C,d+1] int32; all reals are fp64.
 *  - Every function returns an mv_status; mv_last_error() gives the message.
 *  - Handles are not thread-safe; one handle per GPU / process.  Work is enqueued on the
 *    stream set by mv_set_stream (default: the legacy default stream); a function syncs
 *    the stream only when it returns a scalar to the host.
 *  - There is no CPU fallback: without a CUDA device every compute call fails with
 *    MV_ERR_CUDA.
 */
#ifndef MOVEMENT_B200_H
#define MOVEMENT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  MV_OK = 0,
  MV_ERR_BADARG = 1,
  MV_ERR_CUDA = 2,
  MV_ERR_NOT_CONVERGED = 3, /* inner linear solve hit maxit */
  MV_ERR_VALENCE = 4,       /* a vertex has more neighbours than MV_MAX_ROW */
  MV_ERR_COMM = 5,
  MV_ERR_UNSUPPORTED = 6
} mv_status;

typedef struct mv_mesh mv_mesh; /* ingested topology + constant operators on the xi-mesh */
typedef struct mv_ma mv_ma;     /* Monge-Ampere mover state */

/* operators a linear-algebra call can name */
enum { MV_OP_MASS = 0, MV_OP_STIFFNESS = 1, MV_OP_MASS_PLUS_STIFFNESS = 2 };
/* preconditioner of the stiffness-type operators (the mass solves always use Jacobi):
 * replaces the `pc_type` choices of movement/solver_parameters.py:62-71 (bjacobi/ilu, gamg) */
enum { MV_PC_JACOBI = 0, MV_PC_AMG = 1 };
enum { MV_MASS_PCG = 0, MV_MASS_CHEBYSHEV = 1, MV_MASS_CHEBYSHEV_FP64 = 2 };

/* built-in (fused) monitor functions -- replaces: movement/monitor.py:138-143,
 * test/monitors.py:9-16, demos/monge_ampere_helmholtz.py:54-61,115-122,
 * demos/monge_ampere_3d.py:13-15,52-69 */
enum {
  MV_MON_CONST = 0,     /* m = 1 */
  MV_MON_BALL = 1,      /* params: c[3], radius, amplitude, width : 1 + a/cosh^2(w(|x-c|^2 - r^2)) */
  MV_MON_BALL_RAW = 2,  /* as BALL but (|x-c|^2 - r) unsquared radius (test/monitors.py) */
  MV_MON_GAUSS_HESS = 3,/* params: x0,y0,w,alpha : 1 + alpha |Hess exp(-|x-x0|^2/w^2)|_F / max */
  MV_MON_SINATAN3 = 4,  /* params: alpha : 1 + alpha (H:H)/max, H = Hess sinatan3 */
  MV_MON_USER = 100     /* caller writes the monitor field itself */
};

/* fields of an mv_ma that the caller owns and binds (device pointers, fp64) */
enum {
  MV_F_XI = 0,      /* [V,d]  computational coordinates            mover.py:97-99    */
  MV_F_X = 1,       /* [V,d]  physical coordinates                 mover.py:96       */
  MV_F_PHI = 2,     /* [V]    monge_ampere.py:412-417 */
  MV_F_PHI_OLD = 3, /* [V]    */
  MV_F_H = 4,       /* [V,d,d]*/
  MV_F_H_OLD = 5,   /* [V,d,d]*/
  MV_F_GRAD_PHI = 6,/* [V,d]  monge_ampere.py:178-179 */
  MV_F_MONITOR = 7, /* [V]    monge_ampere.py:177     */
  MV_F_VOLUME = 8,  /* [C]    mover.py:101-102        */
  MV_F_ORIG_VOLUME = 9, /* [C] monge_ampere.py:182    */
  MV_F_RESIDUAL = 10,   /* [V] assembled a - theta*m_l (monge_ampere.py:409) */
  MV_F_COUNT = 11
};

/* what mv_mesh_get can copy out (host or device destination) -- used by tests */
enum {
  MV_GET_ROW_NNZ = 0,   /* int32 [V]            */
  MV_GET_SLICE_PTR = 1, /* int64 [nslices+1]    */
  MV_GET_COLS = 2,      /* int32 [sell_len]     */
  MV_GET_MASS = 3,      /* f64   [sell_len]     */
  MV_GET_STIFF = 4,     /* f64   [sell_len]     */
  MV_GET_LUMPED = 5,    /* f64   [V]            */
  MV_GET_CELL_VOL = 6,  /* f64   [C]            */
  MV_GET_BF_NORMAL = 7, /* f64   [F,d]          */
  MV_GET_BF_AREA = 8,   /* f64   [F]            */
  MV_GET_BF_CELL = 9    /* int32 [F]            */
};

/* ---- library --------------------------------------------------------------------- */
int mv_version(void);
const char* mv_last_error(void);
int mv_device_count(void);
int mv_set_device(int device);
int mv_set_stream(void* cuda_stream);
int mv_sync(void);
/* Measurement hooks (bench.py): when enabled, CUDA events on the launching stream bracket the
 * SpMV launches inside PCG.  mv_prof_get: accumulated SpMV milliseconds / launch counts / PCG
 * iterations, indexed by the number of interleaved right-hand sides (arrays of 16), and the number
 * of kernels this library has launched since mv_prof_reset.  Any pointer may be NULL.           */
int mv_prof_enable(int on);
int mv_prof_reset(void);
int mv_prof_get(double* spmv_ms16, int64_t* spmv_count16, int64_t* pcg_iters16, int64_t* launches);

/* True-residual audit of the preconditioned CG solves (the KSPs of movement/monge_ampere.py:441-448 and
 * laplacian_smoothing.py:75 report the recurrence residual; PETSc's -ksp_true_residual_norm is the analogue).
 * When enabled, every mv_pcg_solve / Poisson / Laplacian solve recomputes, at exit and with one extra SpMV, the
 * Euclidean ||P(b - A x)|| / ||P b|| of the iterate it RETURNS (P: Dirichlet rows removed, mean removed for the
 * singular operator), maximum over its right-hand sides.  Arrays of 16, indexed by the number of right-hand
 * sides of the solve: running maximum, value of the last solve, number of audited solves.  Enabling resets.  */
int mv_true_residual_enable(int on);
int mv_true_residual_get(double* max_rel16, double* last_rel16, int64_t* count16);

/* ---- ingest (replaces: movement/mover.py:49-102 -- mesh copy, function spaces, cell
 * volumes; monge_ampere.py:176-184 -- total volume; and the operator assembly Firedrake
 * does inside every LinearVariationalSolver: monge_ampere.py:327,437,476,
 * laplacian_smoothing.py:68).
 *   n_owned_vertices / n_owned_cells: rows / cells this rank owns (== V / C on one GPU);
 *   owned entities come first in the local numbering, ghosts after.               */
int mv_mesh_create(int dim, int64_t V, int64_t C, const double* coords, const int32_t* cells,
                   int64_t F, const int32_t* bfacet_verts, const int32_t* bfacet_tags,
                   int64_t n_owned_vertices, int64_t n_owned_cells, mv_mesh** out);
/* The same for a mesh whose coordinate field is DG1 (coord_dg = 1): Firedrake's layout of PERIODIC meshes
 * (movement/mover.py:70-75, monge_ampere.py:236-239).  `cells` is the periodic P1 cell-node map over V vertices;
 * `coords` has C*(dim+1) rows, row c*(dim+1)+j = corner j of cell c.  No boundary facets, not partitioned.
 * Every coordinate-field argument of the other entry points then has C*(dim+1) rows.                            */
int mv_mesh_create_ex(int dim, int64_t V, int64_t C, const double* coords, const int32_t* cells, int64_t F,
                      const int32_t* bfacet_verts, const int32_t* bfacet_tags, int64_t n_owned_vertices,
                      int64_t n_owned_cells, int coord_dg, mv_mesh** out);
void mv_mesh_destroy(mv_mesh* m);
/* info[0..11] = V, C, F, nnz, sell_len, nslices, max_row, dim, n_owned_vertices, n_owned_cells,
 * slice schedule of the gather kernels (0 not built yet, 1 identity, 2 lattice tiles), its length */
int mv_mesh_info(const mv_mesh* m, int64_t* info);
/* scalars[0..2] = total_volume (owned cells), sum m_l^2 (owned), sum m_l (owned) */
int mv_mesh_scalars(const mv_mesh* m, double* scalars);
int mv_mesh_get(const mv_mesh* m, int what, void* dst);

/* ---- tangling (replaces: movement/tangling.py:26-29 and :39-53) --------------------
 * detJ = det[x1-x0 ... xd-x0] in cell-node-map order, evaluated without FMA contraction so
 * the sign agrees bit-for-bit with the CPU oracle; ratio = detJ/detJ0; count(ratio <= 0)
 * over owned cells.  `ratio` may be NULL.                                             */
int mv_tangle_init(mv_mesh* m, const double* coords, double* detJ0);
int mv_tangle_check(mv_mesh* m, const double* coords, const double* detJ0, double* ratio,
                    int64_t* count);

/* ---- sparse linear algebra (replaces: PETSc KSP cg + bjacobi/ilu,
 * movement/solver_parameters.py:62-71 as used at monge_ampere.py:340,441-448,483 and
 * laplacian_smoothing.py:75) -----------------------------------------------------
 * y = A x for ncomp interleaved components [V,ncomp].                                 */
int mv_spmv(mv_mesh* m, int op, int ncomp, const double* x, double* y);
/* Jacobi-preconditioned CG, ncomp independent right-hand sides sharing one matrix
 * stream.  x holds the initial guess on entry.  mask (uint16 per vertex, bit c set =
 * component c is a Dirichlet dof whose value in x is kept) may be NULL.  rhs may be NULL
 * (zero).  Stops when sqrt(r.D^-1 r) <= max(rtol*ref, atol) for every component, with
 * ref = max over {rhs, initial residual} of the same norm.  project_mean: treat the
 * operator as singular with constant nullspace (rhs and solution projected to zero mean
 * over all ranks).  iters / relres (per component, length ncomp) may be NULL.        */
int mv_pcg_solve(mv_mesh* m, int op, int ncomp, const double* rhs, double* x,
                 const uint16_t* mask, double rtol, double atol, int maxit, int project_mean,
                 int* iters, double* relres);
/* Preconditioner for MV_OP_STIFFNESS / MV_OP_MASS_PLUS_STIFFNESS solves: MV_PC_AMG (default) is a
 * smoothed-aggregation multigrid V-cycle whose hierarchy is built on the device at the first
 * solve and cached per (operator, constrained vertex set) -- replaces: PETSc `gamg`,
 * movement/solver_parameters.py:68-71 (monge_ampere.py:441-448), and the ILU of
 * laplacian_smoothing.py:55-77.  With MV_PC_AMG the stopping norm of mv_pcg_solve is
 * sqrt(r.M^-1 r), M^-1 = one V-cycle.                                                      */
int mv_set_preconditioner(mv_mesh* m, int kind);
/* Solver behind MV_OP_MASS in mv_pcg_solve and the Monge-Ampere mass solves (replaces: KSP cg + bjacobi/ilu,
 * movement/solver_parameters.py:62-66 at monge_ampere.py:340,483).  MV_MASS_CHEBYSHEV (default): one fused kernel
 * per iteration on Wathen's a-priori interval [1/2, 1 + d/2] of diag(M)^-1 M for P1 simplices, same measured
 * stopping rule as the PCG (sqrt(r.D^-1 r) <= max(rtol*ref, atol)), carried out as iterative refinement: residual
 * passes and the solution in double, sweeps of <= 16 Chebyshev iterations on the correction in single precision;
 * MV_MASS_CHEBYSHEV_FP64: every iteration in double; MV_MASS_PCG: Jacobi-preconditioned CG.  Process-wide.      */
int mv_set_mass_solver(int kind);
/* Hierarchy of the cached multigrid preconditioner of `op` (after a solve): info[0] = levels
 * (0 = none built), info[1] = rows of the coarsest level, info[2] = 1 if that level is solved
 * densely, then (rows, nnz) per level; rho[l] = bound of spec(D^-1 A_l); both may be NULL.   */
int mv_amg_info(mv_mesh* m, int op, int64_t* info, int max_levels, double* rho, double* setup_ms);
/* Copy out one SELL-32 matrix of the hierarchy (inspection / tests): which = 0 operator A_l,
 * 1 prolongator P_l, 2 restriction R_l = P_l^T.  size[0..3] = rows, slices, padded length, nnz;
 * the destination pointers (host or device) may be NULL (call once for the sizes).          */
int mv_amg_level_get(mv_mesh* m, int op, int level, int which, int64_t* size, int64_t* slice_ptr,
                     int32_t* cols, double* vals, int32_t* row_nnz);

/* ---- recovery operators of the solution-based monitors ---------------------------------------------------------
 * replaces: animate.recovery as called from movement/monitor.py:232-240 and :286-293 (third-party, un-pinned; restated
 * in oracle/movement_oracle.py).  Device pointers; the fields live on the mesh `m` was ingested with.
 * mv_recover_gradient_l2: g [V,d] = L2 projection of grad(u), u [V] in P1 (consistent mass solve to rtol).
 * mv_clement_gradient:    out [V,ncomp,d] = volume-weighted mean over the cells around each vertex of the cellwise
 *                         gradient of u [V,ncomp]; ncomp = 1 (gradient) or d (Hessian of a recovered gradient).
 * Ghost entries of u are refreshed first on a partitioned mesh (hence non-const).                                 */
int mv_recover_gradient_l2(mv_mesh* m, double* u, double* g, double rtol, int maxit, int* iters);
int mv_clement_gradient(mv_mesh* m, int ncomp, double* u, double* out);
/* replaces: movement/monitor.py:191-201, `Function(fs).project(self.solution)` -- the solution carried onto the mesh the
 * monitor is asked for (a mover's moved copy).  Firedrake projects through a supermesh (third-party, unpinned); restated
 * as the Galerkin L2 projection M_t u_t = int psi_i^t u_s with the (d+1)-point degree-2 rule on every target cell and
 * u_s evaluated by point location in the source mesh (walk through face neighbours from the same-numbered cell): the
 * identity when both meshes coincide, exact for globally linear fields, second order otherwise.  `src` / `tgt`: the two
 * meshes as ingested (their xi coordinates), same dimension and domain, P1 coordinates, one GPU.  u_src [V_src], u_tgt
 * [V_tgt] device pointers.  n_outside (optional): quadrature points the walk could not place inside a source cell
 * (domains that differ); they take the linear extension of the last cell visited.                                    */
int mv_project_p1(mv_mesh* src, const double* u_src, mv_mesh* tgt, double* u_tgt, double rtol, int maxit, int* iters,
                  int64_t* n_outside);

/* ---- Monge-Ampere pieces ---------------------------------------------------------- */
int mv_ma_create(mv_mesh* m, mv_ma** out);
void mv_ma_destroy(mv_ma* ma);
int mv_ma_bind(mv_ma* ma, int field, double* device_ptr);
/* replaces: monge_ampere.py:246-309, the boundary conditions of the L2 projector, decided per boundary tag like
 * _l2_projector_bcs does: tags in fixed_tags (and every tag in 1D) get a full zero Dirichlet condition; a free
 * axis-aligned segment (assemble(abs(n_i) ds) close to zero for all but one i) has that component fixed; any other free
 * segment takes the EquationBC branch (:274-309): before every projection the tangential gradient of phi_old is
 * L2-projected onto the segment with its boundary mass matrix, nodes shared with another segment pinned to zero, and
 * the result is Dirichlet data of the mass solve (csrc/mv_eqbc.cu).  The straight-line / plane check of :288-299 is the
 * host mirror's (movement_b200/monge_ampere.py).
 * _ex: all_tags = every boundary tag of the whole mesh (the same list on every rank): the abs(n_i) integrals are
 * summed over the ranks before the decision, so all ranks take the same branch; kinds (optional, one per sorted
 * unique tag) receives 0 = Dirichlet, 1 + i = component i fixed, -1 = EquationBC.  EquationBC segments on a
 * partitioned mesh: MV_ERR_UNSUPPORTED on every rank.                                                             */
int mv_ma_set_boundary(mv_ma* ma, int n_fixed, const int32_t* fixed_tags);
int mv_ma_set_boundary_ex(mv_ma* ma, int n_fixed, const int32_t* fixed_tags, int n_all, const int32_t* all_tags,
                          int32_t* kinds);
/* uint16 per owned vertex, bit c set = component c of grad_phi is a Dirichlet dof */
int mv_ma_get_bcmask(mv_ma* ma, uint16_t* dst);
int mv_ma_set_solver(mv_ma* ma, double rtol, double atol, int maxit_mass, int maxit_poisson);
int mv_ma_set_monitor(mv_ma* ma, int kind, int nparams, const double* params);
/* replaces: monge_ampere.py:311-342 + :501 -- M g = int psi grad(phi_old), d components */
int mv_ma_l2_project(mv_ma* ma, int* iters);
/* replaces: monge_ampere.py:226-241 (+ :505-506 when the monitor is built in):
 * x = xi + grad_phi, and m(x) evaluated and normalised in the same pass.             */
int mv_ma_update_coordinates(mv_ma* ma);
/* evaluate the built-in monitor at arbitrary coordinates (monge_ampere.py:181)        */
int mv_ma_eval_monitor(mv_ma* ma, const double* coords, double* m_out);
/* replaces: monge_ampere.py:507-517 + mover.py:199-217.
 * out[0..5] = theta, relative_l2_residual, vol_max, vol_min, vol_mean, vol_cov        */
int mv_ma_diagnostics(mv_ma* ma, int use_H_old, double* out);
/* replaces: monge_ampere.py:419-449 + :535 -- K dphi = dt * residual, phi = phi_old+dphi */
int mv_ma_pseudotimestep(mv_ma* ma, double pseudo_dt, int* iters);
/* Initial guess of the Poisson solve: 0 = zero, as a KSP without KSPSetInitialGuessNonzero starts
 * (monge_ampere.py:441-448); 1 (default) = Galerkin combination of the mover's previous solves at the default depth (3);
 * k >= 2 = depth k (at most 8).  Process-wide; the solve and its stopping rule are the same for every setting.    */
int mv_set_poisson_history(int on);
/* replaces: monge_ampere.py:451-485 + :536 -- M H = -int div(tau).grad(phi) + bdry      */
int mv_ma_equidistribute(mv_ma* ma, int* iters);
/* replaces: monge_ampere.py:537-538 */
int mv_ma_advance(mv_ma* ma);
/* K relaxation iterations without convergence tests (bench / profiling helper);
 * trace[K*6] receives mv_ma_diagnostics' output per iteration, cg_iters[K*3] the inner
 * iteration counts (l2 projection, poisson, hessian).                                */
int mv_ma_relax_iterate(mv_ma* ma, double pseudo_dt, int niter, double* trace, int* cg_iters);
/* replaces: the loop of monge_ampere.py:488-540 with its three exits (:518 converged, :525 diverged, :530 maxiter),
 * for a built-in monitor: iterations i0, i0+1, ... (at most nmax in this call), each = l2 projection, x update +
 * monitor, diagnostics, [exit tests], pseudo-timestep, Hessian recovery, advance (by pointer exchange).  res0: in/out
 * initial residual (written at i == 0).  trace[nmax*6], cg_iters[nmax*3] as in mv_ma_relax_iterate.  info[0] = index
 * i of the last iteration, info[1] = 0 converged / 1 diverged / 2 maxiter reached / 3 nmax done (call again with
 * i0 = info[0] + 1), info[2] = number of trace lines written.  The tangling test of :244 looks at the xi-mesh, which
 * is constant during the loop: the caller performs it once.                                                     */
int mv_ma_relax_move(mv_ma* ma, double pseudo_dt, int i0, int nmax, int maxiter, double rtol, double dtol, double* res0,
                     double* trace, int* cg_iters, int* info);

/* ---- quasi-Newton (replaces: PETSc SNES newtonls + KSP gmres + PC fieldsplit behind
 * movement/monge_ampere.py:640-645,670-675,686-700 and movement/solver_parameters.py:1-47).
 * Vectors F, Y, z are device arrays of n_owned_vertices*(1+d*d) doubles laid out [phi | H].
 * mv_qn_residual: F(phi,H) with the bound monitor: F_psi = -(a - theta m_l), F_tau = M H - S phi;
 *   out[0..2] = ||F||^2, theta, relative l2 residual.  The caller performs update_monitor
 *   (monge_ampere.py:648-662) first with mv_ma_l2_project / mv_ma_update_coordinates.
 * mv_qn_solve: J Y = F by flexible GMRES(restart) right-preconditioned with the multiplicative
 *   fieldsplit [(M+K)^-1 ; M^-1] (inexact inner PCG: pc_rtol, pc_maxit); info[0..2] = iterations,
 *   relative residual, restart length used (the requested one -- the reference's 200, solver_parameters.py:13 -- capped
 *   by ksp_maxit and by the device memory the 2*restart+1 basis vectors need).  Returns MV_ERR_NOT_CONVERGED when the
 *   residual test is not met after ksp_maxit iterations (PETSc: KSP_DIVERGED_ITS -> SNES_DIVERGED_LINEAR_SOLVE).  mv_qn_save / mv_qn_set_iterate: (phi,H) = saved - lambda*Y (line search). */
int mv_qn_residual(mv_ma* ma, double* F, double* out);
int mv_qn_jacobian_apply(mv_ma* ma, const double* z, double* y);
int mv_qn_solve(mv_ma* ma, const double* F, double* Y, double ksp_rtol, int ksp_maxit, int restart,
                double pc_rtol, int pc_maxit, double* info);
int mv_qn_save(mv_ma* ma);
int mv_qn_set_iterate(mv_ma* ma, double lambda, const double* Y);

/* raw pieces exposed for parity tests */
int mv_ma_l2_rhs(mv_mesh* m, const double* phi, double* b);            /* monge_ampere.py:328 */
int mv_ma_hess_rhs(mv_mesh* m, const double* phi, double* c);          /* monge_ampere.py:477-480 */
int mv_ma_assemble_a(mv_mesh* m, const double* mon, const double* H, double* a); /* :406-409 */

/* ---- cell volumes + diagnostics for Laplacian / spring movers
 * (replaces: laplacian_smoothing.py:119, spring.py:274, mover.py:199-217)
 * stats[0..3] = max, min, mean, coefficient of variation                              */
int mv_cell_volumes(mv_mesh* m, const double* coords, double* volume, double* stats);
/* The same diagnostic for callers that cannot allocate device memory.  (No counterpart in the reference: this entry
 * exists because of how the library is deployed, not because of a reference call site.)
 * Who that is: a rank whose neighbour's peer-memory kernel is spinning on a flag -- cudaMalloc / cudaFree synchronise
 * the device and would wait on that kernel; a process whose HBM is already owned by torch's caching allocator or by a
 * full-size mover, where a diagnostic must not be the call that runs out of memory; a monitoring thread that may not
 * block on the driver's allocator.  mv_cell_volumes needs an mv_mesh (whose creation allocates) and takes result slots
 * from cudaMallocAsync per call; this variant needs no handle and calls no cudaMalloc, cudaMallocAsync, cudaFree or
 * any routine that allocates.  Its scratch is one file-scope __device__ array of ~768 KB (csrc/mv_static.cu) that
 * every GPU loading the library holds until the process exits.
 *   dim 1..3; cells [C,dim+1] int32 and coords [*,dim] fp64 (P1 coordinates, not the DG layout of periodic meshes) are
 *   DEVICE pointers and so is the output volume [C]; host pointers are refused with MV_ERR_BADARG because staging
 *   them would allocate.  stats is a HOST array of 4: max, min, mean, coefficient of variation, over the C cells
 *   given (no reduction over ranks).  C > 134 217 728 (2^27: 32768 CTAs of 4096 cells, the size of the scratch) is
 *   refused with MV_ERR_UNSUPPORTED and a message naming the limit.
 * The first call on each device looks up the device address of the scratch (cudaGetSymbolAddress, with that device
 * current), which also makes the driver load the module there: make that first call before any stream capture
 * begins.  All calls on a GPU share the one scratch, which keeps the previous call's partial results; the single word
 * that must start at zero (the last-block ticket) is cleared at the start of every call.  Two calls that overlap on
 * the GPU (different streams, different threads) race on it: serialise them.  Work goes to the mv_set_stream stream;
 * the call returns after synchronising it.  Sums are folded in a fixed order that does not depend on the SM count.  */
int mv_cell_volumes_static(int dim, int64_t C, const double* coords, const int32_t* cells, double* volume,
                           double* stats);

/* ---- Laplacian smoothing (replaces: laplacian_smoothing.py:64-77 + :109-117)
 * Solves K v = 0 per component with v prescribed where mask bit is set (values already in
 * v), then x += v*dt.  displacement_norm = ||v*dt||_2.                                  */
int mv_laplace_move(mv_mesh* m, double* v, const uint16_t* mask, double dt, double* x,
                    double rtol, double atol, int maxit, int* iters, double* displacement_norm);

/* ---- lineal springs (replaces: spring.py:70-146 edge geometry, :148-181 stiffness,
 * :183-235 Dirichlet rows, :266 dense solve)                                          */
int mv_spring_edge_count(mv_mesh* m, int64_t* n_edges);
/* edges [E,2] int32 (i<j, sorted), length [E], tangent [E,2], angle [E]; any may be NULL */
int mv_spring_edge_geometry(mv_mesh* m, const double* coords, int32_t* edges, double* length,
                            double* tangent, double* angle);
/* u = K(coords)^-1 forcing with rows of masked vertices replaced by identity;
 * coords += u*dt; displacement [V,2] = u*dt.                                          */
int mv_spring_move(mv_mesh* m, double* coords, const double* forcing, const uint16_t* mask,
                   double dt, double* displacement, double rtol, double atol, int maxit,
                   int* iters, double* displacement_norm);
/* y = K(coords) x for the row-replaced spring matrix (tests: spring.py:184-235); mask bit c
 * set = dof (v,c) is a Dirichlet row (identity).  mask may be NULL.                    */
int mv_spring_apply(mv_mesh* m, const double* coords, const uint16_t* mask, const double* x,
                    double* y);
/* Solver of mv_spring_move on this mesh (replaces: the solver choice at spring.py:266, a dense numpy solve there).  Meshes
 * of at least MV_SPRING_AMG_MIN (4096) vertices assemble the xx / xy / yy blocks on the mesh pattern and precondition CG
 * with one multigrid V-cycle per diagonal block; smaller ones run the matrix-free Jacobi-PCG.
 * info[0], info[1] = levels of the xx / yy hierarchies (0: Jacobi path so far), info[2] = 1 if the last solve rebuilt
 * them, info[3] = PCG iterations right after the last set-up, info[4] = solves since, info[5] = iterations of the last solve */
int mv_spring_solver_info(const mv_mesh* m, int64_t* info);

/* ---- multi-GPU (replaces: the MPI halo exchange / allreduce implicit in PyOP2 and PETSc;
 * SURVEY 5.8).  nccl_lib_path: path of libnccl.so.2; unique_id: 128 bytes from
 * mv_comm_unique_id on rank 0, broadcast by the caller.                               */
int mv_comm_unique_id(const char* nccl_lib_path, void* unique_id_128);
int mv_comm_init(const char* nccl_lib_path, const void* unique_id_128, int rank, int nranks);
int mv_comm_finalize(void);
/* info[0] = 1 when the peer-memory collectives are active (halo exchange and scalar allreduce of the inner solvers as one
 * kernel each over CUDA-IPC windows, NVLink peer stores; csrc/mv_p2p.cu; MV_P2P=0 keeps everything on NCCL),
 * info[1], info[2] = halo exchanges / allreduces carried that way so far, info[3] = ranks.                          */
int mv_comm_info(int64_t* info);
/* halo plan: for each neighbour rank, the local owned vertices to send (send_idx grouped by
 * neighbour via send_ptr) and the contiguous ghost range received from it.             */
int mv_mesh_set_halo(mv_mesh* m, int n_neighbours, const int32_t* neighbour_ranks,
                     const int64_t* send_ptr, const int32_t* send_idx, const int64_t* recv_ptr);
/* vec [V,ncomp] interleaved: ghost rows receive their owners' values.  With the peer-memory path active the call
 * synchronises the stream and returns MV_ERR_COMM if a wait on a neighbour timed out (MV_P2P_TIMEOUT_S seconds, default
 * 600, 0 = never): the ghosts of THIS call are then void; the status is reported once and the next exchange pairs up
 * again as soon as the neighbour is back.                                                                       */
int mv_halo_exchange(mv_mesh* m, int ncomp, double* vec);
/* global numbers of the local cells [C] (host or device): the '+' cell of an interior edge (spring.py:113-127) is its
 * lowest-numbered incident cell in the GLOBAL numbering, which the local order (owned cells first) does not preserve */
int mv_mesh_set_cell_order(mv_mesh* m, const int64_t* global_cell);
int mv_allreduce_sum(double* device_scalars, int n);

#ifdef __cplusplus
}
#endif
#endif /* MOVEMENT_B200_H */
