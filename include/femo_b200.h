/* femo_b200.h -- C ABI of libfemo_b200.so: the B200 (sm_100a) replacement for the
 * dolfinx/PETSc/MUMPS calls behind femo_alpha's Reissner-Mindlin shell hot path.
 *
 * Every entry point returns 0 on success; on failure it returns non-zero and
 * femo_last_error() describes why.  Plain pointers and sizes only.  Pointers
 * flagged `on_device` may be CUDA device pointers (e.g. torch tensor
 * data_ptr()); otherwise they are host pointers borrowed for the call.  All
 * calls are synchronous with respect to the context's stream unless stated.
 *
 * The "replaces" notes cite /root/reference/femo_alpha/... (file:line).
 *
 * DOF numbering: 3x3 "block nodes".  Q2 displacement nodes first -- vertices
 * [0,nn), edge mid-points [nn,nn+ne), cell centres [nn+ne,nq2) -- then the Q1
 * rotation nodes nq2+v.  dof = 3*block + component.  (The reference's order is
 * dolfinx-internal and opaque to its callers.)
 */
#ifndef FEMO_B200_H
#define FEMO_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct femo_ctx femo_ctx;

/* inputs of the RM shell FEA: rm_shell/rm_shell_model.py:209-214 (fea.add_input) */
enum femo_input {
  FEMO_THICKNESS = 0, FEMO_F_SOLID = 1, FEMO_E = 2, FEMO_NU = 3, FEMO_DENSITY = 4, FEMO_UHAT = 5
};
/* scalar outputs: rm_shell/rm_shell_model.py:219-230 (fea.add_output) */
enum femo_output_id {
  FEMO_COMPLIANCE = 0, FEMO_MASS = 1, FEMO_ELASTIC_ENERGY = 2, FEMO_PNORM_STRESS = 3,
  FEMO_VOLUME = 4   /* int t dx, dynamic_rm_shell/volume_operation.py:57-70 */
};
/* pseudo-argument for femo_output_grad: derivative w.r.t. the state 'disp_solid' */
#define FEMO_ARG_STATE 100

typedef struct femo_shell_opts {
  int32_t element_wise_material;   /* VT = DG0 instead of CG1     rm_shell_pde.py:36-40 */
  int32_t elementwise_pressure;    /* VF = [DG0]^3 instead of CG1 rm_shell_pde.py:41-44 */
  int32_t quadrature_degree;       /* Gauss-Legendre tensor rule, m=(deg+2)/2 points/dir; 0 -> 6 (4x4) */
  int32_t reserved;
} femo_shell_opts;

typedef struct femo_sizes {
  int64_t n_nodes, n_edges, n_cells;   /* nn, ne, nc */
  int64_t n_blocks, n_dof;             /* nb = nq2+nn, n = 3 nb */
  int64_t nnz_blocks;                  /* BSR 3x3 blocks of dR/du */
  int64_t n_contrib;                   /* element-block contributions (169 per quad) */
  int64_t n_material, n_force;         /* sizes of thickness/E/nu/density and of F_solid */
  int32_t n_qp, all_planar;
} femo_sizes;

const char* femo_last_error(void);
int femo_version(void);

/* ---- context / mesh ------------------------------------------------------- */
int femo_ctx_create(int device, femo_ctx** out);
int femo_ctx_destroy(femo_ctx* ctx);
/* run all kernels of this context on `cuda_stream` (a cudaStream_t); NULL = default stream */
int femo_set_stream(femo_ctx* ctx, void* cuda_stream);

/* replaces reconstructFEAMesh / readFEAMesh (fea/utils_dolfinx.py:34-50,653-668):
 * xyz (n_nodes,3) row-major; conn (n_cells, nodes_per_cell) int32 in VTK/meshio cyclic
 * order; nodes_per_cell must be 4 (quadrilaterals). */
int femo_mesh_set(femo_ctx* ctx, int64_t n_nodes, const double* xyz,
                  int64_t n_cells, int nodes_per_cell, const int32_t* conn);

/* replaces RMShellPDE.__init__ + ShellElement (rm_shell_pde.py:25-47,
 * linear_shell_model.py:26-106): builds edges, the Q2xQ1 block dofmap, the BSR pattern of
 * dR/du and the cell->nnz contribution permutation used by the deterministic assembly. */
int femo_shell_setup(femo_ctx* ctx, const femo_shell_opts* opts);
int femo_get_sizes(femo_ctx* ctx, femo_sizes* out);
/* coordinates of all n_blocks block nodes (dof coordinates of W.sub(0) / W.sub(1)),
 * host buffer (n_blocks,3); used to evaluate shell_bc_func (rm_shell_model.py:168-180) */
int femo_get_block_coords(femo_ctx* ctx, double* xyz_out);
/* cell -> 13 block ids [9 Q2 (4 vertices, 4 edges, centre), 4 Q1], host (n_cells,13) */
int femo_get_cell_blocks(femo_ctx* ctx, int32_t* out);

/* ---- boundary conditions ---------------------------------------------------- */
/* strong Dirichlet (PENALTY_BC=False, rm_shell_model.py:168-180): list of clamped dofs */
int femo_set_strong_bc(femo_ctx* ctx, int64_t n, const int32_t* dofs);
/* penalty Dirichlet (PENALTY_BC=True, rm_shell_model.py:83-98, linear_shell_model.py:323-333):
 * vertex_on[v] != 0 where shell_bc_func holds; every mesh edge with both vertices on gets
 * beta/h_E * int (w - g).dw ds from each adjacent cell. */
int femo_set_penalty_bc(femo_ctx* ctx, int64_t n_nodes, const uint8_t* vertex_on);

/* ---- inputs / state --------------------------------------------------------- */
/* replaces utils_dolfinx.update (fea/utils_dolfinx.py:319-330): n == 1 broadcasts */
int femo_set_input(femo_ctx* ctx, int which, const double* data, int64_t n, int on_device);
int femo_get_input(femo_ctx* ctx, int which, double* out, int64_t n, int on_device);
int femo_set_state(femo_ctx* ctx, const double* w, int on_device);
int femo_get_state(femo_ctx* ctx, double* w, int on_device);

/* ---- assembly ------------------------------------------------------------------ */
/* replaces assembleMatrix(dR_du) / assembleSystem / NonlinearProblem.J
 * (fea/utils_dolfinx.py:200-221; csdl_alpha_opt/state_operation.py:289-294): recomputes the
 * BSR values of dR/du (incl. penalty terms) from the current inputs; BCs are applied on the
 * fly by the operator (rows+cols masked, unit diagonal). */
int femo_assemble_jacobian(femo_ctx* ctx);
/* replaces assembleVector(residual_form) (fea/utils_dolfinx.py:194-198):
 * R = dR/du * w - load(F_solid); w = NULL uses the stored state. */
int femo_assemble_residual(femo_ctx* ctx, const double* w, double* R, int on_device);
/* One residual evaluation through HOST buffers with the copies overlapped -- what StateOperation does per call (update
 * every input, then assembleVector(R): csdl_alpha_opt/state_operation.py:86-131, fea/utils_dolfinx.py:174-198) as ONE
 * call: the inputs dR/du depends on go up first and the Jacobian is assembled while F_solid / density / the state are
 * still on their way up on a second stream; R = K w - F comes back in row ranges while the SpMV still works on the next
 * one.  which / data / sizes: n_in input ids with their host arrays (size 1 = fill, like femo_set_input); w_host: the
 * state (n_dof) or NULL for the resident one; R_host: n_dof doubles.  Pinned host memory makes the overlap real.
 * On a rank of a multi-GPU context the arrays are the rank's local ones (as for femo_set_input / femo_assemble_residual:
 * owned rows of R are complete, no collective). */
int femo_evaluate_residual(femo_ctx* ctx, int n_in, const int* which, const double* const* data, const int64_t* sizes,
                           const double* w_host, double* R_host);
/* load vector int f . du J dx only */
int femo_assemble_load(femo_ctx* ctx, double* F, int on_device);
/* y = dR/du x (no BCs) / y = A x (with BCs); replaces computeMatVecProductFwd (utils_dolfinx.py:275-283) */
int femo_jacobian_apply(femo_ctx* ctx, const double* x, double* y, int with_bc, int on_device);
/* diagnostic/test: copy the BSR structure+values to host.  rowptr (nb+1) int32, colidx (nnzb) int32,
 * vals (nnzb,3,3) fp64; any pointer may be NULL.  with_bc applies the dolfinx BC convention. */
int femo_get_bsr(femo_ctx* ctx, int32_t* rowptr, int32_t* colidx, double* vals, int with_bc);
/* selected 3x3 blocks of dR/du (row-major, no BCs) by their BSR positions: out (n, 9).  Diagnostics / tests at sizes where
 * femo_get_bsr would move gigabytes (the reference's getValuesCSR on a sub-matrix, fea/utils_dolfinx.py:34-50 callers) */
int femo_get_blocks(femo_ctx* ctx, int64_t n, const int64_t* positions, double* out);

/* ---- solves ---------------------------------------------------------------------- */
/* replaces FEA.solve -> solveNonlinear -> NewtonSolver (fea/fea_dolfinx.py:159-170,
 * fea/utils_dolfinx.py:338-352,438-468): assembles dR/du and the load, then PCG (block-Jacobi,
 * 6x6 vertex / 3x3 edge+centre blocks) on A w = b, warm-started from the stored state. */
int femo_solve_state(femo_ctx* ctx, double rtol, int max_it, int* iters, double* relres);
/* replaces FEA.solveLinearBwd + BC zeroing (fea/fea_dolfinx.py:189-203,
 * csdl_alpha_opt/state_operation.py:206-218): A lam = rhs (A symmetric), lam[bc] = 0. */
int femo_solve_adjoint(femo_ctx* ctx, const double* rhs, double* lam, double rtol, int max_it,
                       int* iters, double* relres, int on_device);
/* replaces FEA.solveLinearFwd (fea/fea_dolfinx.py:173-187): A du = rhs */
int femo_solve_forward(femo_ctx* ctx, const double* rhs, double* du, double rtol, int max_it,
                       int* iters, double* relres, int on_device);

/* ---- partials of the residual (matrix-free) ---------------------------------------- */
/* out = (dR/d arg)^T lam ; replaces assembleMatrix(dR_df) + computeMatVecProductBwd
 * (state_operation.py:183-184,282-286; utils_dolfinx.py:294-306).  arg in {THICKNESS,F_SOLID,E,NU} */
int femo_dRdf_T_apply(femo_ctx* ctx, int arg, const double* lam, double* out, int on_device);
/* out = (dR/d arg) df ; replaces computeMatVecProductFwd on dRdf (state_operation.py:163-171) */
int femo_dRdf_apply(femo_ctx* ctx, int arg, const double* df, double* out, int on_device);

/* ---- outputs ------------------------------------------------------------------------- */
/* replaces OutputOperation.compute -> assemble_scalar (output_operation.py:51-56) */
int femo_output(femo_ctx* ctx, int which, double* value);
/* replaces OutputOperation.compute_derivatives (output_operation.py:58-69):
 * g = d output / d arg, arg = femo_input or FEMO_ARG_STATE */
int femo_output_grad(femo_ctx* ctx, int which, int arg, double* g, int on_device);

/* ---- stress functionals -------------------------------------------------- */
/* p-norm parameters m, rho of RMShellModel (rm_shell_model.py:63): pnorm = 1/Area int (m sigma_vm)^rho J dx */
int femo_set_stress_params(femo_ctx* ctx, double m, double rho);
/* per-cell sub-domain ids (rm_shell_model.py:101-133 mesh_tags); host int32[n_cells], -1 = untagged */
int femo_set_cell_tags(femo_ctx* ctx, int64_t n_cells, const int32_t* tags);
/* replaces assemble_scalar(pnorm_stress_form) (rm_shell_pde.py:112-128), degree-4 measure, von Mises
 * stress on the top surface xi2 = h/2 (linear_shell_model.py:393-467); tag < 0 = whole surface,
 * otherwise the sub-domain `tag` with its own area normalisation */
int femo_pnorm_stress(femo_ctx* ctx, int tag, double* value);
int femo_pnorm_stress_grad(femo_ctx* ctx, int tag, int arg, double* g, int on_device);
/* replaces FEA.projectFieldOutput -> project(von_Mises, DG1) (fea/utils_dolfinx.py:568-602):
 * cell-wise L2 projection, out (n_cells,4) in tensor vertex order; surface 0 Top, 1 Mid, 2 Bot */
int femo_stress_field(femo_ctx* ctx, int surface, double* out, int on_device);

/* ---- multi-GPU: one process per GPU (the reference is serial) ------------------- */
/* rank 0 creates the NCCL id (128 bytes); the host broadcasts it (torch.distributed / MPI) */
int femo_nccl_unique_id(void* id128);
/* partition of this rank (femo_alpha_b200/partition.py): neighbour ranks, per-neighbour offsets (n_nbr+1,
 * in block nodes) into the key-ordered send / receive block lists, owner flag per local block node, and the
 * number of owned cells (owned cells come first in the local cell order; the rest is the ghost layer).
 * After this call dot products are owner-masked, scalar outputs are summed over owned cells and over ranks,
 * and every operator application is preceded by a halo exchange. */
int femo_dist_setup(femo_ctx* ctx, int rank, int world, int n_nbr, const int32_t* nbr_ranks,
                    const int32_t* send_off, const int32_t* send_blocks, const int32_t* recv_off,
                    const int32_t* recv_blocks, const uint8_t* own_block, int64_t n_owned_cells);
/* collective: ncclCommInitRank with the broadcast id */
int femo_dist_init(femo_ctx* ctx, const void* id128);
/* use the communicator of `owner` (same rank / world) instead of creating one: the distributed levels of a multigrid
 * hierarchy share the communicator of the finest level.  `owner` must outlive the solves on ctx. */
int femo_dist_share_comm(femo_ctx* ctx, femo_ctx* owner);
/* ghost entries of x (n_dof) <- owners' values; collective */
int femo_halo_exchange(femo_ctx* ctx, double* x, int on_device);
/* reverse exchange: owned entries of x += the values their ghost copies hold on the other ranks (ghost entries are
 * left unchanged); collective.  The restriction step of a distributed V-cycle (no reference counterpart: the
 * reference is serial). */
int femo_halo_accumulate(femo_ctx* ctx, double* x, int on_device);
/* sum n <= 8 host scalars over ranks; collective */
int femo_dist_allreduce(femo_ctx* ctx, double* vals, int n);
int femo_dist_counts(femo_ctx* ctx, int64_t* halo_exchanges, int64_t* allreduces);
/* diagnostics (set FEMO_TRACE_OVERLAP=1 before femo_dist_setup): CUDA-event times of the last overlapped operator
 * application in ms after its start: [0] halo exchange complete (communication stream), [1] interior rows complete,
 * [2] boundary rows complete.  [0] <= [1]: the exchange was hidden behind the interior rows. */
int femo_dist_overlap_info(femo_ctx* ctx, double* ms3);

/* ---- geometric multigrid preconditioner for the state / adjoint solves --------------------------------
 * Replaces the sparse LU (MUMPS through PETSc `preonly + lu`, femo_alpha/fea/utils_dolfinx.py:438-468 setUpKSP /
 * NewtonSolver and :514-531 setUpKSP_MUMPS; fea_dolfinx.py:159-203 FEA.solve / solveLinearBwd) on meshes that come
 * with a hierarchy of uniformly refined quad meshes: level 0 = the mesh of the analysis, level l+1 = its parent.
 * Every level is an ordinary context (same inputs sampled on its own nodes, same BC predicate).
 * N GPUs: the finest k levels may be multi-GPU contexts (femo_dist_setup with NESTED partitions: a fine cell lives on the
 * rank of its parent), the coarser ones are replicated; smoothing exchanges halos per level, restriction between two
 * distributed levels ends in a reverse halo exchange, into a replicated level in one all-reduce. */
typedef struct femo_mg femo_mg;
int femo_mg_create(int n_levels, femo_ctx** level_ctxs, femo_mg** out);
int femo_mg_destroy(femo_mg* mg);
/* transfer between level l (fine) and l+1 (coarse) in block-node numbering: P (rows = fine blocks) and
 * R = P^T (rows = coarse blocks) as CSR with one scalar weight per (fine block, coarse block) pair */
int femo_mg_set_transfer(femo_mg* mg, int level, const int32_t* P_rowptr, const int32_t* P_col, const double* P_w,
                         const int32_t* R_rowptr, const int32_t* R_col, const double* R_w);
/* (re)build after the inputs changed: assembles every level, block-Jacobi inverses, spectral bounds of the
 * Chebyshev smoother (degree cheb_degree, default 3), direct factorisation of the coarsest operator (femo_direct_*).
 * reuse_coarse_inverse != 0 keeps the coarse factor of the previous set-up (lagged: still SPD, cheaper set-up
 * inside an optimisation loop with small steps, a few more iterations) */
int femo_mg_setup(femo_mg* mg, int cheb_degree, int reuse_coarse_inverse);
int femo_mg_get_lambda(femo_mg* mg, int level, double* lambda_max);
/* iterative refinement around femo_mg_solve's PCG: `steps` (0..8, default 0) corrections x += PCG(b - A x) with the residual
 * accumulated in double-double, each solved to inner_rtol relative to its own right-hand side; femo_mg_solve then reports
 * the TRUE relative residual of its result (bounded below by eps |A| |x| / |b|: x is stored in fp64) instead of CG's
 * recurrence residual.  Stands in for the direct solve behind the reference's KSP (femo_alpha/fea/utils_dolfinx.py:514-531,
 * MUMPS).  On a shell with kappa ~ 1e9 it removes the solver's share of the forward error; what remains against another
 * implementation (~1e-8) is the sensitivity of x to last-bit differences of the matrix entries. */
int femo_mg_set_refine(femo_mg* mg, int steps, double inner_rtol);
/* PCG with one V-cycle per iteration on A x = rhs (A = dR/du with BCs, x0 = 0); zero_bc as femo_solve_adjoint */
int femo_mg_solve(femo_mg* mg, const double* rhs, double* sol, double rtol, int max_it, int* iters, double* relres,
                  int on_device, int zero_bc);

/* ---- allocation-free variant of femo_dRdf_T_apply (thickness, E, nu) -----------------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid allocation).
 * out_dev (n_material doubles) = (dR/darg)^T lam_dev (n_dof doubles) at the context's current state and inputs; both
 * pointers are DEVICE pointers.  The per-cell scratch of the element kernel lives in a static device array of
 * FEMO_NOALLOC_MAX_CELLS x 4 doubles (2 MiB per GPU that loads the library, held until exit): meshes with more cells are
 * rejected with an error.  The call allocates and frees nothing and does not synchronise: its launches are queued on
 * the context's stream (femo_set_stream) and out_dev is valid once that stream has reached them.  Every call on a GPU
 * shares the one array (cleared at the start of each call): calls that overlap on the GPU race. */
#define FEMO_NOALLOC_MAX_CELLS 65536
int femo_dRdf_T_apply_noalloc(femo_ctx* ctx, int arg, const double* lam_dev, double* out_dev);

/* ---- sparse direct solver -----------------------------------------------------------------------------------
 * The reference solves EVERY system with a sparse LU (MUMPS: fea/utils_dolfinx.py:465-466 NewtonSolver "preonly + lu",
 * :514-531 setUpKSP_MUMPS).  Here: breadth-first level sets of the block-node graph make A block tridiagonal; block
 * LDL^T with explicit inverses of the Schur complements (hand-written blocked Cholesky + GEMM kernels, no library), a
 * solve = two sweeps of small matrix-vector kernels (or one product with the explicit inverse up to 8192 dofs).  Used
 * for the coarsest multigrid level, for the dense stepper of femo_dyn_*, and directly for meshes without a hierarchy
 * whose fronts fit in memory (the reference's wing box: 32 226 dofs, 71 fronts of <= 843 dofs, 0.4 GB).
 * Single-GPU contexts.  femo_direct_setup: symbolic analysis on first use + numeric factorisation of the CURRENT
 * operator (call again after inputs / BCs change).  femo_direct_solve: A x = rhs (clamped entries of rhs ignored) with
 * `refine` steps of iterative refinement on a residual accumulated in double-double; relres = |rhs - A x| / |rhs| of the returned x.
 * femo_direct_info: [level sets, largest front (dofs), bytes of the factor, 1 if the explicit inverse is used]. */
int femo_direct_setup(femo_ctx* ctx);
int femo_direct_solve(femo_ctx* ctx, const double* rhs, double* sol, int refine, double* relres, int on_device, int zero_bc);
int femo_direct_info(femo_ctx* ctx, double* info4);

/* ---- dynamic (time-stepped) RM shell: SURVEY 8 f1, BASELINE config 4 ------------------------------------------
 * Replaces femo_alpha/dynamic_rm_shell: PlateSim.solve_dynamic_problem (plate_sim.py:281-361), the residual
 * M wddot + K w_mid - F - W of PlateSim.SVK_residual (:190-215) under the implicit midpoint rule (:135-140), and the
 * forward / reverse sweeps of StateOperation.apply_inverse_jacobian and .compute_jacvec_product
 * (state_operation_dynamic.py:140-432,447-702).  The reference assembles and LU-factorises the (step-invariant)
 * Jacobian in every step and keeps N^2 work vectors in the reverse sweep; here A_eff = (2/dt^2) M + K/2 is assembled
 * once (direct factorisation up to 46 000 dofs, else block-Jacobi PCG or V-cycle PCG on the levels of an attached
 * multigrid), every step is one K product, one M product and one solve, the whole state history stays in HBM
 * ((n_steps+1) x n_dof doubles, level-major like utils.py:9-16) and the reverse sweep is the two-field (w, wdot)
 * recurrence with O(1) work vectors.  Mass form: linear_shell_model.py:335-348, 4x4 Gauss rule; density = the
 * FEMO_DENSITY input.  Homogeneous strong Dirichlet BCs (femo_set_strong_bc) only.
 * While a femo_dyn exists its context is in "dynamic mode": the operator of every solve on it is A_eff. */
typedef struct femo_dyn femo_dyn;
/* add_self_weight: body force rho t g e_z with g = g_factor * 9.81 (plate_sim.py:203-213; g_factor -1 = downwards) */
int femo_dyn_create(femo_ctx* ctx, double dt, int n_steps, int add_self_weight, double g_factor, femo_dyn** out);
int femo_dyn_destroy(femo_dyn* d);
/* optional: V-cycle PCG for the step solves; every level context of `mg` is switched to dynamic mode with the same dt */
int femo_dyn_set_mg(femo_dyn* d, femo_mg* mg);
/* (re)build after the inputs changed: K, M, A_eff, preconditioner / dense inverse.
 * solver: 0 auto (dense if n_dof <= 46000, else multigrid if attached, else PCG), 1 dense, 2 PCG, 3 multigrid.
 * On a multi-GPU context (femo_dist_setup) every rank steps its own rows with the distributed block-Jacobi PCG
 * (0 and 2 accepted, 1 and 3 rejected); histories and sweeps are then in the rank's local block order and every
 * femo_dyn_* call that solves or reduces is collective. */
int femo_dyn_setup(femo_dyn* d, int solver);
/* time loop from zero initial conditions; f_history (n_rows, n_force) row-major, rows beyond the last repeat it
 * (plate_sim.py:314-319).  Fills the resident history and the strain energy of every level (plate_sim.py:337).
 * The FEMO_F_SOLID input of the context is left at the last row used (as PlateSim.update_f does). */
int femo_dyn_solve(femo_dyn* d, const double* f_history, int64_t n_rows, double rtol, int max_it,
                   int64_t* total_iters, int on_device);
int femo_dyn_get_history(femo_dyn* d, double* w_hist, int on_device);      /* (n_steps+1, n_dof) */
int femo_dyn_set_history(femo_dyn* d, const double* w_hist, int on_device);
int femo_dyn_get_energies(femo_dyn* d, double* energies);                  /* host, n_steps+1 */
/* apply_inverse_jacobian 'rev' (state_operation_dynamic.py:562-690): (dR/dw)^T lam = d_out, clamped entries and
 * level 0 of lam are zero; 'fwd' (:516-560): (dR/dw) x = d_res */
int femo_dyn_solve_adjoint(femo_dyn* d, const double* d_out_hist, double* lam_hist, double rtol, int max_it,
                           int64_t* total_iters, int on_device);
int femo_dyn_solve_forward(femo_dyn* d, const double* d_res_hist, double* x_hist, double rtol, int max_it,
                           int64_t* total_iters, int on_device);
/* compute_jacvec_product on the state (state_operation_dynamic.py:224-275,316-372): y = (dR/dw) x or (dR/dw)^T x */
int femo_dyn_dRdw_apply(femo_dyn* d, const double* x_hist, double* y_hist, int transpose, int on_device);
/* compute_jacvec_product on the thickness (:277-309,374-404), at the resident history:
 * out = sum_i (dR_i/dh)^T lam_i  (n_material)  /  out_hist[i] = (dR_i/dh) dh */
int femo_dyn_dRdh_T_apply(femo_dyn* d, const double* lam_hist, double* out, int on_device);
int femo_dyn_dRdh_apply(femo_dyn* d, const double* dh, double* out_hist, int on_device);
/* TotalStrainEnergyOperation (total_strain_energy_operation.py:56-125) at the resident history;
 * arg = FEMO_THICKNESS (n_material) or FEMO_ARG_STATE ((n_steps+1) x n_dof) */
int femo_dyn_total_strain_energy(femo_dyn* d, double* value);
int femo_dyn_total_strain_energy_grad(femo_dyn* d, int arg, double* g, int on_device);
/* y = M x (no BCs; test / diagnostic) */
int femo_dyn_mass_apply(femo_dyn* d, const double* x, double* y, int on_device);
/* device time of the last femo_dyn_solve / femo_dyn_solve_adjoint (ms) and the solver in use (1 dense, 2 PCG, 3 MG) */
int femo_dyn_info(femo_dyn* d, double* ms_forward, double* ms_adjoint, int* solver);

/* ---- generic 3-D forms on trilinear hexahedra (BASELINE config 5) --------------------------------
 * Replaces, for the named forms below, what the reference's generic FEA does with an arbitrary UFL form
 * (femo_alpha/fea/fea_dolfinx.py:71-95 add_state / add_output, :159-203 solve / solveLinearBwd;
 * femo_alpha/fea/utils_dolfinx.py:194-221 assembleVector / assembleMatrix / assembleSystem):
 *   form 1  linear elasticity   R(u; rho, f) = K(rho) u - M f,  K_e = rho_e^p K0_e(E0, nu)   (SIMP)
 *   form 0  vector Poisson      the same with K0_e = kappa int grad N_a . grad N_b I3 (kappa = E0)
 * Q1 hexahedra (VTK_HEXAHEDRON node order), 2x2x2 Gauss; state = 3 dofs per node, dof = 3 node + component.
 * A context set up with femo_hex_setup is a "solid" context: femo_set_strong_bc, femo_set_state / femo_get_state,
 * femo_assemble_jacobian / _load / _residual, femo_jacobian_apply, femo_get_bsr, femo_get_sizes, femo_solve_state /
 * _adjoint / _forward, femo_pcg_fixed, femo_launch_count and femo_last_timings work on it unchanged; the RM-shell
 * entry points (inputs, partials, outputs, penalty BC, dist, mg, dyn) refuse it. */
int femo_hex_setup(femo_ctx* ctx, int64_t n_nodes, const double* xyz, int64_t n_cells, const int32_t* conn /* n_cells*8 */,
                   int form, double E0, double nu, double simp_p);
/* which: 0 cell density rho (n_cells values, or 1 host value broadcast), 1 nodal body force f (3 n_nodes) */
int femo_hex_set_input(femo_ctx* ctx, int which, const double* data, int64_t n, int on_device);
int femo_hex_get_input(femo_ctx* ctx, int which, double* out, int64_t n, int on_device);
/* matrix-free partials at the context's state (state_operation.py:163-171,183-184 with the matrices of :282-286):
 * arg 0: (dR/drho)^T lam = p rho_e^(p-1) lam_e^T K0_e u_e per cell,  (dR/drho) drho (3 n_nodes);  arg 1: -/+ M */
int femo_hex_dRdf_T_apply(femo_ctx* ctx, int arg, const double* lam, double* out, int on_device);
int femo_hex_dRdf_apply(femo_ctx* ctx, int arg, const double* df, double* out, int on_device);
/* outputs (output_operation.py:51-69): which 0 compliance int f.u dx, 1 volume sum rho_e |e|;
 * gradients w.r.t. arg 0 (density), 1 (body force) or FEMO_ARG_STATE */
int femo_hex_output(femo_ctx* ctx, int which, double* value);
int femo_hex_output_grad(femo_ctx* ctx, int which, int arg, double* grad, int on_device);
int femo_hex_get_mesh(femo_ctx* ctx, double* xyz /* 3 n_nodes or NULL */, int32_t* conn /* 8 n_cells or NULL */);

/* ---- instrumentation ---------------------------------------------------------------- */
/* number of kernels this context has launched so far */
int femo_launch_count(femo_ctx* ctx, int64_t* n);
/* device time (ms, CUDA events on the context's stream) of the last call of each phase:
 * [0] cell records, [1] BSR gather-assembly, [2] last PCG solve, [3] per-PCG-iteration mean */
int femo_last_timings(femo_ctx* ctx, double* ms4);
/* run `n_iter` PCG iterations on A x = b without convergence test (bench / roofline use) */
int femo_pcg_fixed(femo_ctx* ctx, int n_iter, double* ms_total);
/* micro-benchmarks for the roofline denominators: fp64 FMA throughput (TFLOP/s) and a device
 * copy (GB/s) measured on this GPU */
int femo_measure_peaks(femo_ctx* ctx, double* fp64_tflops, double* copy_gbs);
/* fp64 tensor-core (mma.sync m8n8k4 f64) rate measured by the last femo_measure_peaks call, TFLOP/s */
int femo_last_dmma_tflops(femo_ctx* ctx, double* tflops);

#ifdef __cplusplus
}
#endif
#endif
