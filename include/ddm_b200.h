/* ddm_b200.h -- C ABI of the B200-native solve phase of two-level overlapping Schwarz.
 *
 * This is the drop-in boundary for the hot path of nilsfriess/dune-ddm.  Each entry point names
 * the reference interface it stands in for (paths relative to the reference tree).  All functions
 * return 0 on success and a non-zero status otherwise; ddm_last_error() gives the message (the
 * C++ adaptors in dune-ddm_b200/dune/ddm_b200/adaptors.hh turn a non-zero status into
 * DUNE_THROW(Dune::Exception), the reference's error convention: dune/ddm/schwarz.hh:83,91,190).
 *
 * One ddm_ctx drives one GPU from one host thread (the reference is single-threaded per rank:
 * one MPI rank = one subdomain; here one context = one GPU = many subdomains).  With several
 * GPUs every process creates its own context and joins them with ddm_comm_init; halo exchange,
 * the coarse-residual allgather and the scalar allreduce then go over NCCL in place of the
 * reference's MPI communicators.
 *
 * Index spaces.  A context stores vectors over its *local* DOFs: first the n_own DOFs it owns
 * (contiguous block of the unique global numbering), then n_ghost copies of remote DOFs that its
 * subdomains or matrix rows touch.  Vector handles (ddm_vec) have n_own + n_ghost entries; only
 * the owned part carries meaning between calls.  Host-pointer entry points (*_host) take the
 * owned part only.
 *
 * There is no CPU fallback behind this interface: every compute entry runs CUDA kernels for
 * sm_100a and fails if no device is present.
 */
#ifndef DDM_B200_H
#define DDM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ddm_ctx ddm_ctx;
typedef struct ddm_vec ddm_vec;

/* SchwarzType, dune/ddm/schwarz.hh:35 (config key schwarz.type, :80-83). */
enum { DDM_SCHWARZ_STANDARD = 0, DDM_SCHWARZ_RESTRICTED = 1 };
/* ApplyMode, dune/ddm/combined_preconditioner.hh:21 (key combined_preconditioner.mode, :57-69). */
enum { DDM_COMBINE_ADDITIVE = 0, DDM_COMBINE_MULTIPLICATIVE = 1 };

/* Dune::InverseOperatorResult, extern/dune-istl/dune/istl/solver.hh:48-84. */
typedef struct {
    int32_t iterations;
    int32_t converged;
    double reduction;
    double conv_rate;
    double elapsed; /* seconds, device time of the solve loop */
    double def0;    /* initial defect norm */
} ddm_result;

const char *ddm_last_error(void);
const char *ddm_version(void);

/* ---- lifetime -------------------------------------------------------------------------- */
int ddm_ctx_create(int device, ddm_ctx **out);
int ddm_ctx_destroy(ddm_ctx *ctx);

/* Joins `nranks` contexts (one per process/GPU).  `nccl_unique_id` is the 128-byte ncclUniqueId
 * made on rank 0 (ddm_comm_unique_id) and distributed by the caller (the reference gets its
 * communicator from MPIHelper: examples/poisson.cc:165).  n_own_offset is the global id of this
 * rank's first owned DOF. */
int ddm_comm_unique_id(uint8_t out[128]);
int ddm_comm_init(ddm_ctx *ctx, int rank, int nranks, const uint8_t nccl_unique_id[128]);

/* ---- setup products, uploaded once ----------------------------------------------------- */

/* The operator of NonOverlappingOperator (dune/ddm/nonoverlapping_operator.hh:12-50): rows of
 * the assembled, Dirichlet-eliminated matrix for the owned DOFs, scalar CSR as in
 * Dune::BCRSMatrix<FieldMatrix<double,1,1>> (columns ascending, local numbering owned|ghost). */
int ddm_upload_operator(ddm_ctx *ctx, int64_t n_own, int64_t n_ghost, const int64_t *rowptr, const int32_t *col,
                        const double *val);

/* The same operator for Dune::BCRSMatrix<FieldMatrix<double,3,3>> (linear elasticity, examples/linearelasticity.hh): block
 * rows of the owned nodes, block-column indices in local node numbering, 9 doubles per block (row-major).  Vectors are
 * node-major (BlockVector<FieldVector<double,3>>): n_own = 3 nb_own, n_ghost = 3 nb_ghost scalar entries. */
int ddm_upload_operator_bcsr3(ddm_ctx *ctx, int64_t nb_own, int64_t nb_ghost, const int64_t *rowptr, const int32_t *col,
                              const double *val);

/* Halo plan replacing the OwnerOverlapCopyCommunication interfaces
 * (extern/dune-istl/dune/istl/owneroverlapcopy.hh:311-387).  Ghost entries are stored grouped by
 * owning peer: ghosts of peers[p] occupy local indices n_own + recv_ptr[p] .. n_own + recv_ptr[p+1].
 * send_idx[send_ptr[p] .. send_ptr[p+1]) are the owned local indices peer p holds as ghosts, in
 * the order of that peer's ghost block. */
int ddm_upload_halo(ddm_ctx *ctx, int32_t npeers, const int32_t *peers, const int64_t *send_ptr,
                    const int32_t *send_idx, const int64_t *recv_ptr);

/* Overlapping index lists R_i (dune/ddm/overlap_extension.hh:54-285) in local numbering and the
 * partition of unity D_i (dune/ddm/pou.hh:199, read at dune/ddm/schwarz.hh:141), for the nsub
 * subdomains this context hosts.  pou may be NULL (schwarz.hh:140). */
int ddm_upload_decomposition(ddm_ctx *ctx, int32_t nsub, const int64_t *sub_ptr, const int32_t *sub_dofs,
                             const double *pou);

/* Factor of A_i = R_i A R_i^T for local subdomain `sub`, in the layout of a supernodal LL^T
 * cholmod_factor -- what Dune::Cholmod::cholmodFactor() returns
 * (extern/dune-istl/dune/istl/cholmod.hh:495): super[nsuper+1], pi[nsuper+1], px[nsuper+1],
 * s[pi[nsuper]], x[px[nsuper]] (column-major panels, diagonal block included), perm[n]. */
int ddm_upload_factor(ddm_ctx *ctx, int32_t sub, int32_t n, int32_t nsuper, const int32_t *super, const int64_t *pi,
                      const int64_t *px, const int32_t *s, const double *x, const int32_t *perm);

/* Non-symmetric subdomain matrices: LU factor of A_i in the same supernodal layout (one symbolic structure for L and U^T).
 * Stands in for the numeric object of Dune::UMFPack, the solver the reference picks for non-symmetric problems
 * (dune/ddm/schwarz.hh:85-92; extern/dune-istl/dune/istl/umfpack.hh: umfpack_dl_numeric / umfpack_dl_solve, whose L, U, P, Q
 * umfpack_dl_get_numeric exports).  xl: panels of the unit-lower L (diagonal stored), xu: panels of U^T (top block = U_JJ^T,
 * lower triangular, pivots on its diagonal); perm is applied to rows and columns alike. */
int ddm_upload_factor_lu(ddm_ctx *ctx, int32_t sub, int32_t n, int32_t nsuper, const int32_t *super, const int64_t *pi,
                         const int64_t *px, const int32_t *s, const double *xl, const double *xu, const int32_t *perm);

/* Numeric factorisation of A_i on the device with a given symbolic analysis, then the same re-layout as
 * ddm_upload_factor: the cholmod_l_factorize half of Dune::Cholmod::setMatrix
 * (extern/dune-istl/dune/istl/cholmod.hh:469-472; the cholmod_l_analyze half -- perm, super, pi, px, s -- stays with the
 * caller), and the numeric re-factorisation of TwoLevelSchwarzSolver's update branch (dune/ddm/twolevel_schwarz.hh:99-104).
 * (a_rowptr, a_col, a_val): A_i = R_i A R_i^T in CSR, the subdomain's own numbering, symmetric positive definite, both
 * triangles stored, every entry inside the symbolic structure.  x_out (may be NULL) receives the LL^T panels in
 * cholmod_factor layout (px[nsuper] doubles), i.e. what ddm_upload_factor would have been given.  Setup only: multifrontal,
 * level by level; cuSOLVER/cuBLAS for the wide fronts. */
int ddm_factorize_subdomain(ddm_ctx *ctx, int32_t sub, int32_t n, int32_t nsuper, const int32_t *super, const int64_t *pi,
                            const int64_t *px, const int32_t *s, const int32_t *perm, const int64_t *a_rowptr,
                            const int32_t *a_col, const double *a_val, double *x_out);

/* Coarse space (dune/ddm/galerkin_preconditioner.hh:119-144,220-349): per hosted subdomain k[i]
 * basis vectors, Z = concatenation of column-major n_i x k[i] blocks (restr_vecs, :353), the
 * global coarse dimension N0, this context's first coarse index k_offset, and the dense
 * Galerkin matrix A0 (N0 x N0 row-major, replicated on every context; symmetric positive definite -> Cholesky, otherwise LU). */
int ddm_upload_coarse(ddm_ctx *ctx, const int32_t *k, const double *Z, int32_t N0, int32_t k_offset,
                      const double *A0);

/* Reproducible subdomain solves (optional, before ddm_finalize): the triangular sweeps then use no floating-point atomics --
 * every task result goes to a private slot and per-supernode "assemble" tasks add the slots in a fixed order -- so that two
 * runs give bit-identical results, as cholmod_l_solve does on one rank (extern/dune-istl/dune/istl/cholmod.hh:293).  Costs a
 * few percent of sweep time and 12 bytes per stored factor row; the default (off) accumulates with FP64 atomics. */
int ddm_set_deterministic(ddm_ctx *ctx, int on);

/* Builds the device schedules; call after all uploads. */
int ddm_finalize(ddm_ctx *ctx);

/* Galerkin product A0 = Z^T A Z on the device (GalerkinPreconditioner::build_solver, dune/ddm/galerkin_preconditioner.hh:246-328,
 * which the reference flags as a bottleneck for large runs, :217-218): call after ddm_finalize when ddm_upload_coarse was given
 * A0 = NULL.  Every column is restrict(A * prolong(e_j)) with the solve-phase kernels, all-reduced over the ranks; the matrix
 * is then factored like an uploaded one.  A0_out (N0 x N0 row-major, may be NULL) receives a copy. */
int ddm_galerkin_assemble(ddm_ctx *ctx, double *A0_out);

/* Optional, before ddm_galerkin_assemble: a colouring of ALL subdomains of the job such that no subdomain is coupled through A to
 * two subdomains of one colour (the reference's build_solver likewise touches neighbour blocks only,
 * galerkin_preconditioner.hh:279-328).  The product then takes ncolors x max k_i passes instead of N0: the columns (i, k) of all
 * subdomains i of a colour go through prolong -> A -> restrict together.  color[nsub_global], k_global[nsub_global] (basis sizes in
 * global subdomain order), owner_of[m * ncolors + c] = the subdomain of colour c coupled to m (m itself included) or -1.
 * nsub_global = 0 removes the colouring. */
int ddm_set_coarse_coloring(ddm_ctx *ctx, int32_t nsub_global, int32_t ncolors, const int32_t *color, const int32_t *k_global,
                            const int32_t *owner_of);

/* ---- value updates with unchanged structure (Newton loops) -------------------------------
 * TwoLevelSchwarzSolver::apply keeps the overlapping index sets and only refreshes matrix values on every call after
 * the first (dune/ddm/twolevel_schwarz.hh:99-104).  Here: new operator values in the uploaded sparsity pattern; a
 * ddm_upload_factor / ddm_upload_factor_lu call whose symbolic arrays equal the uploaded ones converts the new panels in
 * place and keeps the device schedules (no ddm_finalize needed); the coarse matrix is replaced (A0 != NULL) or
 * re-assembled on the device from the new operator (A0 == NULL). */
int ddm_update_operator_values(ddm_ctx *ctx, const double *val);
int ddm_update_coarse_matrix(ddm_ctx *ctx, const double *A0);

/* ---- batched GenEO eigensolve support (setup) -------------------------------
 * For a block-diagonal context (operator = blockdiag(C_i), subdomain i = the contiguous vector range
 * [sub_ptr[i], sub_ptr[i+1]), factors of K_i = A_neu,i + shift C_i) all subdomains run the shift-invert Lanczos
 * recurrence of their generalised eigenproblem A_neu v = lambda C v in lock step (the reference runs one
 * Spectra::SymGEigsShiftSolver per rank, dune/ddm/eigensolvers/spectra.hh:112-215, coarse_spaces.hh:319-331):
 * ddm_blockdiag_solve applies all K_i^{-1} with the triangular-sweep kernels, ddm_op_apply all C_i, and the two
 * segmented operations below do the per-subdomain inner products and updates. */
int ddm_blockdiag_solve(ddm_ctx *ctx, ddm_vec *d, ddm_vec *x);
int ddm_seg_dot(ddm_ctx *ctx, const ddm_vec *x, const ddm_vec *y, double *out_per_subdomain);
int ddm_seg_axpby(ddm_ctx *ctx, const double *alpha_per_subdomain, const ddm_vec *x, const double *beta_per_subdomain_or_null,
                  ddm_vec *y);

/* ---- vectors: the operations Dune's Krylov solvers need from X --------------------------
 * (extern/dune-istl/dune/istl/solvers.hh:292-335,933-1012: copy, =scalar, =X, axpy, *=, +=) */
int ddm_vec_create(ddm_ctx *ctx, ddm_vec **out);
int ddm_vec_destroy(ddm_vec *v);
int ddm_vec_upload(ddm_vec *v, const double *host_owned);
int ddm_vec_download(const ddm_vec *v, double *host_owned);
int ddm_vec_fill(ddm_vec *v, double value);
int ddm_vec_copy(ddm_vec *dst, const ddm_vec *src);
int ddm_vec_axpy(ddm_vec *y, double a, const ddm_vec *x); /* y += a x */
int ddm_vec_scale(ddm_vec *y, double a);                  /* y *= a   */
int ddm_vec_add(ddm_vec *y, const ddm_vec *x);            /* y += x   */
int ddm_vec_xpay(ddm_vec *y, double a, const ddm_vec *x); /* y = x + a y (p *= beta; p += q fused) */
double *ddm_vec_device_ptr(ddm_vec *v);

/* ---- operator and scalar product -------------------------------------------------------- */
/* NonOverlappingOperator::apply / applyscaleadd, dune/ddm/nonoverlapping_operator.hh:34-50. */
int ddm_op_apply(ddm_ctx *ctx, ddm_vec *x, ddm_vec *y);
int ddm_op_applyscaleadd(ddm_ctx *ctx, double alpha, ddm_vec *x, ddm_vec *y);
/* NonOverlappingScalarProduct::dot / norm, dune/ddm/nonoverlapping_operator.hh:76-83. */
int ddm_dot(ddm_ctx *ctx, const ddm_vec *x, const ddm_vec *y, double *out);
int ddm_norm(ddm_ctx *ctx, const ddm_vec *x, double *out);

/* Dot product without any allocation.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same sum as
 * ddm_dot over the first n entries of two device arrays, result stored in *out ON THE DEVICE (no host synchronisation); runs on
 * the context's stream.  x, y, out are device pointers on the context's GPU (e.g. ddm_vec_device_ptr).  Rank-local: with several
 * ranks the caller adds the per-rank results.  The call allocates and frees nothing: its per-block partial sums live in a static
 * device array of 16 384 doubles (128 KB, held by every GPU that loads the library until the process exits), one partial per
 * chunk of 8 192 entries, so n is limited to 134 217 728 entries; a longer vector is refused with an error (use ddm_dot).
 * All calls on a GPU share that array and it keeps the last call's contents (a call only reads what it has just written, so
 * nothing needs clearing); calls that overlap on the GPU -- other streams, other contexts -- race on it. */
int ddm_dot_noalloc(ddm_ctx *ctx, int64_t n, const double *x, const double *y, double *out);

/* ---- preconditioners ------------------------------------------------------------------- */
/* SchwarzPreconditioner::apply, dune/ddm/schwarz.hh:115-149. */
int ddm_schwarz_apply(ddm_ctx *ctx, int schwarz_type, ddm_vec *d, ddm_vec *x);
/* GalerkinPreconditioner::apply, dune/ddm/galerkin_preconditioner.hh:151-194. */
int ddm_coarse_apply(ddm_ctx *ctx, ddm_vec *d, ddm_vec *x);
/* CombinedPreconditioner::apply over {Schwarz, Galerkin}, dune/ddm/combined_preconditioner.hh:127-163.
 * with_coarse = 0 gives the one-level method. */
int ddm_twolevel_apply(ddm_ctx *ctx, int schwarz_type, int combine_mode, int with_coarse, ddm_vec *d, ddm_vec *x);

/* ---- device-resident Krylov loops -------------------------------------------------------
 * Same recurrences and stopping rule as Dune::CGSolver::apply (extern/dune-istl/dune/istl/solvers.hh:279-340,
 * solver.hh:457-480); b is overwritten by the defect, x is the initial guess and the result.
 * history (may be NULL) receives the defect norms, maxit+1 entries. */
int ddm_pcg_solve(ddm_ctx *ctx, ddm_vec *b, ddm_vec *x, double reduction, int32_t maxit, int schwarz_type,
                  int combine_mode, int with_coarse, ddm_result *res, double *history);

/* Dune::RestartedGMResSolver::apply (extern/dune-istl/dune/istl/solvers.hh:923-1035): left-preconditioned GMRES(restart),
 * modified Gram-Schmidt, Givens rotations (:1072-1111); the defect is the preconditioned residual norm |s[i+1]|.
 * A breakdown (|w| < 1e-80, :983) returns a non-zero status (the reference throws SolverAbort). */
int ddm_gmres_solve(ddm_ctx *ctx, ddm_vec *b, ddm_vec *x, double reduction, int32_t restart, int32_t maxit, int schwarz_type,
                    int combine_mode, int with_coarse, ddm_result *res, double *history);

/* ---- host-pointer forms (stage through pinned memory; owned entries only) ---------------
 * These are what an adaptor with X = Dune::BlockVector<FieldVector<double,1>> calls. */
/* Optional: page-lock a caller's vector in place (cudaHostRegister) so that the host-pointer entries copy from / to it
 * directly instead of through the pinned staging area; a driver does this once for its BlockVector storage. */
int ddm_host_register(ddm_ctx *ctx, double *host, int64_t count);
int ddm_host_unregister(ddm_ctx *ctx, double *host);
int ddm_op_apply_host(ddm_ctx *ctx, const double *x, double *y);
int ddm_op_applyscaleadd_host(ddm_ctx *ctx, double alpha, const double *x, double *y);
int ddm_dot_host(ddm_ctx *ctx, const double *x, const double *y, double *out);
int ddm_twolevel_apply_host(ddm_ctx *ctx, int schwarz_type, int combine_mode, int with_coarse, const double *d,
                            double *x);
int ddm_pcg_solve_host(ddm_ctx *ctx, double *b, double *x, double reduction, int32_t maxit, int schwarz_type,
                       int combine_mode, int with_coarse, ddm_result *res, double *history);

int ddm_gmres_solve_host(ddm_ctx *ctx, double *b, double *x, double reduction, int32_t restart, int32_t maxit,
                         int schwarz_type, int combine_mode, int with_coarse, ddm_result *res, double *history);

/* ---- introspection ---------------------------------------------------------------------- */
/* Algorithmic bytes per call of each kernel family and launch counters, for roofline reports.
 * keys: "spmv", "restrict", "trisolve_fwd", "trisolve_bwd", "prolong", "coarse", "dot", "axpy". */
int ddm_get_stat(ddm_ctx *ctx, const char *key, double *value);
/* Event timers named after the reference's Logger events (dune/ddm/schwarz.hh:178-182,
 * galerkin_preconditioner.hh:203-204): enable, run, then read accumulated milliseconds and counts. */
int ddm_timers_enable(ddm_ctx *ctx, int on);
int ddm_timers_reset(ddm_ctx *ctx);
int ddm_get_timer(ddm_ctx *ctx, const char *event, double *total_ms, int64_t *calls);
/* CUDA-event stopwatch on the context's stream (the stream every kernel of this library is launched on). */
int ddm_stopwatch_start(ddm_ctx *ctx);
int ddm_stopwatch_stop(ddm_ctx *ctx, double *elapsed_ms);
int64_t ddm_kernel_launches(ddm_ctx *ctx);
/* cudaProfilerStart / cudaProfilerStop around a region of interest (ncu --profile-from-start off). */
int ddm_profiler_range(ddm_ctx *ctx, int on);
int ddm_sync(ddm_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* DDM_B200_H */
