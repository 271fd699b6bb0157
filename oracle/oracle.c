/* CPU ORACLE -- TEST INFRASTRUCTURE, NOT THE PRODUCT.
 *
 * Plain-C restatement of the solve-phase hot path of dune-ddm, used only as the checker in
 * tests/, in __graft_entry__.smoke() and as bench.py's cpu_baseline / --impl reference leg.
 * Nothing under dune-ddm_b200/ may call into this file.
 *
 * Parity status: the reference ships no test of the apply path, so the
 * apply-path results are pinned by this restatement plus (a) the reference's own golden
 * literals for the setup (tests/test_galerkin_coarse_matrix.cc, checked in tests/test_golden.py),
 * (b) the residual identity ||A_i x_i - d_i|| <= 1e-12 ||d_i|| at the direct-solver boundary
 * (SuiteSparse is an un-vendored, un-pinned system dependency of the reference:
 * CMakeLists.txt:43), and (c) the reference's real Dune::CGSolver compiled from
 * /root/reference (oracle/_ref, see oracle/Makefile).
 *
 * Vectors are indexed by unique global DOF; a reference rank's consistent local vector is the
 * restriction of that vector to the rank's index list.
 */
#include "oracle.h"
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* y = A x, row by row, columns ascending: BCRSMatrix::mv,
 * /root/reference/extern/dune-istl/dune/istl/bcrsmatrix.hh:1641-1661 */
void oracle_spmv(int64_t n, const int64_t *rowptr, const int32_t *col, const double *val, const double *x, double *y,
                 int nthreads)
{
#pragma omp parallel for num_threads(nthreads) schedule(static)
    for (int64_t i = 0; i < n; ++i) {
        double acc = 0.0;
        for (int64_t k = rowptr[i]; k < rowptr[i + 1]; ++k) acc += val[k] * x[col[k]];
        y[i] = acc;
    }
}

/* sum over unique DOFs of x*y, left to right: OwnerOverlapCopyCommunication::dot with the owner
 * mask applied, /root/reference/extern/dune-istl/dune/istl/owneroverlapcopy.hh:398-416 */
double oracle_dot(int64_t n, const double *x, const double *y)
{
    double s = 0.0;
    for (int64_t i = 0; i < n; ++i) s += x[i] * y[i];
    return s;
}

/* One subdomain solve x = A_i^{-1} b with a supernodal LL^T factor in cholmod_factor layout.
 * Dune::Cholmod::apply (/root/reference/extern/dune-istl/dune/istl/cholmod.hh:263-306) calls
 * cholmod_l_solve(CHOLMOD_A): permute, forward solve with L, backward solve with L^T, un-permute.
 * The supernodal sweeps restate the published algorithm of CHOLMOD's cholmod_super_lsolve /
 * cholmod_super_ltsolve for one right-hand side (dense triangular solve on the diagonal block,
 * dense matrix-vector update with the rows below). */
void oracle_super_solve(int32_t n, int32_t nsuper, const int32_t *super, const int64_t *pi, const int64_t *px,
                        const int32_t *s, const double *x, const int32_t *perm, const double *b, double *out,
                        double *work /* n */)
{
    double *y = work;
    for (int32_t k = 0; k < n; ++k) y[k] = b[perm[k]];
    for (int32_t J = 0; J < nsuper; ++J) {
        const int32_t f = super[J], c = super[J + 1] - f;
        const int64_t m = pi[J + 1] - pi[J];
        const double *P = x + px[J];
        const int32_t *rows = s + pi[J];
        for (int32_t j = 0; j < c; ++j) {
            const double yj = y[f + j] / P[j + (int64_t)j * m];
            y[f + j] = yj;
            const double *colj = P + (int64_t)j * m;
            for (int32_t i = j + 1; i < c; ++i) y[f + i] -= colj[i] * yj;
            for (int64_t i = c; i < m; ++i) y[rows[i]] -= colj[i] * yj;
        }
    }
    for (int32_t J = nsuper - 1; J >= 0; --J) {
        const int32_t f = super[J], c = super[J + 1] - f;
        const int64_t m = pi[J + 1] - pi[J];
        const double *P = x + px[J];
        const int32_t *rows = s + pi[J];
        for (int32_t j = c - 1; j >= 0; --j) {
            const double *colj = P + (int64_t)j * m;
            double acc = y[f + j];
            for (int64_t i = c; i < m; ++i) acc -= colj[i] * y[rows[i]];
            for (int32_t i = j + 1; i < c; ++i) acc -= colj[i] * y[f + i];
            y[f + j] = acc / colj[j];
        }
    }
    for (int32_t k = 0; k < n; ++k) out[perm[k]] = y[k];
}

/* One subdomain solve with a supernodal LU factor (non-symmetric A_i): Dune::UMFPack::apply
 * (/root/reference/extern/dune-istl/dune/istl/umfpack.hh, selected at /root/reference/dune/ddm/schwarz.hh:85-92) calls
 * umfpack_dl_solve(UMFPACK_A): permute, forward solve with unit-lower L, backward solve with U, un-permute.  xl holds
 * the L panels, xu the panels of U^T in the same layout (diagonal block = U_JJ^T, lower triangular). */
void oracle_super_solve_lu(int32_t n, int32_t nsuper, const int32_t *super, const int64_t *pi, const int64_t *px,
                           const int32_t *s, const double *xl, const double *xu, const int32_t *perm, const double *b,
                           double *out, double *work /* n */)
{
    double *y = work;
    for (int32_t k = 0; k < n; ++k) y[k] = b[perm[k]];
    for (int32_t J = 0; J < nsuper; ++J) {
        const int32_t f = super[J], c = super[J + 1] - f;
        const int64_t m = pi[J + 1] - pi[J];
        const double *P = xl + px[J];
        const int32_t *rows = s + pi[J];
        for (int32_t j = 0; j < c; ++j) {
            const double yj = y[f + j]; /* unit diagonal */
            const double *colj = P + (int64_t)j * m;
            for (int32_t i = j + 1; i < c; ++i) y[f + i] -= colj[i] * yj;
            for (int64_t i = c; i < m; ++i) y[rows[i]] -= colj[i] * yj;
        }
    }
    for (int32_t J = nsuper - 1; J >= 0; --J) {
        const int32_t f = super[J], c = super[J + 1] - f;
        const int64_t m = pi[J + 1] - pi[J];
        const double *P = xu + px[J]; /* column j of the panel = row j of U */
        const int32_t *rows = s + pi[J];
        for (int32_t j = c - 1; j >= 0; --j) {
            const double *colj = P + (int64_t)j * m;
            double acc = y[f + j];
            for (int64_t i = c; i < m; ++i) acc -= colj[i] * y[rows[i]];
            for (int32_t i = j + 1; i < c; ++i) acc -= colj[i] * y[f + i];
            y[f + j] = acc / colj[j];
        }
    }
    for (int32_t k = 0; k < n; ++k) out[perm[k]] = y[k];
}

/* SchwarzPreconditioner::apply, /root/reference/dune/ddm/schwarz.hh:115-149.
 * restricted != 0: x = sum_i R_i^T D_i A_i^{-1} R_i d ; else x = sum_i R_i^T A_i^{-1} R_i d.
 * The halo sums of addOwnerCopyToOwnerCopy are taken in ascending rank order. */
void oracle_schwarz_apply(const oracle_setup *S, int restricted, const double *d, double *x, int nthreads)
{
    const int64_t tot = S->sub_ptr[S->nsub];
    double *xo = (double *)malloc(sizeof(double) * (size_t)tot);
#pragma omp parallel num_threads(nthreads)
    {
        double *dl = NULL, *wk = NULL;
        int64_t cap = 0;
#pragma omp for schedule(dynamic, 1)
        for (int32_t i = 0; i < S->nsub; ++i) {
            const int64_t o = S->sub_ptr[i], ni = S->sub_ptr[i + 1] - o;
            if (ni > cap) { free(dl); free(wk); cap = ni; dl = malloc(sizeof(double) * cap); wk = malloc(sizeof(double) * cap); }
            for (int64_t k = 0; k < ni; ++k) dl[k] = d[S->sub_dofs[o + k]]; /* restrict + copyOwnerToAll */
            if (S->xu && S->xu[i])
                oracle_super_solve_lu((int32_t)ni, S->nsuper[i], S->super[i], S->pi[i], S->px[i], S->s[i], S->x[i], S->xu[i],
                                      S->perm[i], dl, xo + o, wk);
            else
                oracle_super_solve((int32_t)ni, S->nsuper[i], S->super[i], S->pi[i], S->px[i], S->s[i], S->x[i], S->perm[i], dl,
                                   xo + o, wk);
            if (restricted && S->pou)
                for (int64_t k = 0; k < ni; ++k) xo[o + k] *= S->pou[o + k];
        }
        free(dl); free(wk);
    }
    memset(x, 0, sizeof(double) * (size_t)S->n);
    for (int32_t i = 0; i < S->nsub; ++i)
        for (int64_t k = S->sub_ptr[i]; k < S->sub_ptr[i + 1]; ++k) x[S->sub_dofs[k]] += xo[k];
    free(xo);
}

/* GalerkinPreconditioner::apply, /root/reference/dune/ddm/galerkin_preconditioner.hh:151-194:
 * restrict, k_i dot products per rank, gather, solve with A0 (here: dense Cholesky, the
 * reference uses a sparse direct solver on rank 0), scatter, prolongate, halo sum. */
void oracle_coarse_apply(const oracle_setup *S, const double *d, double *x, int nthreads)
{
    const int32_t N0 = S->k_off[S->nsub];
    double *d0 = (double *)calloc((size_t)N0 + 1, sizeof(double));
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1)
    for (int32_t i = 0; i < S->nsub; ++i) {
        const int64_t o = S->sub_ptr[i], ni = S->sub_ptr[i + 1] - o;
        const int32_t ki = S->k_off[i + 1] - S->k_off[i];
        const double *Zi = S->Z + S->z_ptr[i];
        for (int32_t k = 0; k < ki; ++k) {
            double acc = 0.0;
            for (int64_t j = 0; j < ni; ++j) acc += Zi[(int64_t)k * ni + j] * d[S->sub_dofs[o + j]];
            d0[S->k_off[i] + k] = acc;
        }
    }
    /* A0 y = d0: dense LU with row swaps for a non-symmetric coarse matrix, else L L^T y = d0 */
    const double *L = S->A0chol;
    if (S->A0lu) {
        const double *M = S->A0lu;
        for (int32_t i = 0; i < N0; ++i) {
            const int32_t p = S->A0piv[i];
            if (p != i) { const double t = d0[i]; d0[i] = d0[p]; d0[p] = t; }
        }
        for (int32_t i = 0; i < N0; ++i) {
            double acc = d0[i];
            for (int32_t j = 0; j < i; ++j) acc -= M[(int64_t)i * N0 + j] * d0[j];
            d0[i] = acc;
        }
        for (int32_t i = N0 - 1; i >= 0; --i) {
            double acc = d0[i];
            for (int32_t j = i + 1; j < N0; ++j) acc -= M[(int64_t)i * N0 + j] * d0[j];
            d0[i] = acc / M[(int64_t)i * N0 + i];
        }
    } else {
    for (int32_t i = 0; i < N0; ++i) {
        double acc = d0[i];
        for (int32_t j = 0; j < i; ++j) acc -= L[(int64_t)i * N0 + j] * d0[j];
        d0[i] = acc / L[(int64_t)i * N0 + i];
    }
    for (int32_t i = N0 - 1; i >= 0; --i) {
        double acc = d0[i];
        for (int32_t j = i + 1; j < N0; ++j) acc -= L[(int64_t)j * N0 + i] * d0[j];
        d0[i] = acc / L[(int64_t)i * N0 + i];
    }
    }
    memset(x, 0, sizeof(double) * (size_t)S->n);
    const int64_t tot = S->sub_ptr[S->nsub];
    double *xo = (double *)malloc(sizeof(double) * (size_t)tot);
#pragma omp parallel for num_threads(nthreads) schedule(dynamic, 1)
    for (int32_t i = 0; i < S->nsub; ++i) {
        const int64_t o = S->sub_ptr[i], ni = S->sub_ptr[i + 1] - o;
        const int32_t ki = S->k_off[i + 1] - S->k_off[i];
        const double *Zi = S->Z + S->z_ptr[i];
        for (int64_t j = 0; j < ni; ++j) xo[o + j] = 0.0;
        for (int32_t k = 0; k < ki; ++k) {
            const double c = d0[S->k_off[i] + k];
            for (int64_t j = 0; j < ni; ++j) xo[o + j] += c * Zi[(int64_t)k * ni + j];
        }
    }
    for (int32_t i = 0; i < S->nsub; ++i)
        for (int64_t k = S->sub_ptr[i]; k < S->sub_ptr[i + 1]; ++k) x[S->sub_dofs[k]] += xo[k];
    free(xo);
    free(d0);
}

/* CombinedPreconditioner::apply, /root/reference/dune/ddm/combined_preconditioner.hh:127-163.
 * mode 0 additive: x = M1 d + M0 d; mode 1 multiplicative: x = M1 d; x += M0 (d - A x).
 * with_coarse == 0 gives the one-level method. */
void oracle_twolevel_apply(const oracle_setup *S, int restricted, int mode, int with_coarse, const double *d, double *x,
                           int nthreads)
{
    oracle_schwarz_apply(S, restricted, d, x, nthreads);
    if (!with_coarse) return;
    double *xn = (double *)malloc(sizeof(double) * (size_t)S->n);
    if (mode == 0) {
        oracle_coarse_apply(S, d, xn, nthreads);
    } else {
        double *dn = (double *)malloc(sizeof(double) * (size_t)S->n);
        oracle_spmv(S->n, S->rowptr, S->col, S->val, x, dn, nthreads);
        for (int64_t i = 0; i < S->n; ++i) dn[i] = d[i] - dn[i];
        oracle_coarse_apply(S, dn, xn, nthreads);
        free(dn);
    }
    for (int64_t i = 0; i < S->n; ++i) x[i] += xn[i];
    free(xn);
}

/* Dune::CGSolver::apply, /root/reference/extern/dune-istl/dune/istl/solvers.hh:279-340, with the
 * stopping rule of Iteration::step (solver.hh:457-480): def < def0*reduction || def < 1e-30.
 * b is overwritten by the defect, x holds the solution.  Returns the iteration count; *converged and
 * the defect history (maxit+1 entries) are filled. */
int oracle_pcg(const oracle_setup *S, int restricted, int mode, int with_coarse, double *x, double *b, double reduction,
               int maxit, double *history, int *converged, int nthreads)
{
    const int64_t n = S->n;
    double *p = malloc(sizeof(double) * n), *q = malloc(sizeof(double) * n);
    oracle_spmv(n, S->rowptr, S->col, S->val, x, q, nthreads); /* applyscaleadd(-1,x,b) */
    for (int64_t i = 0; i < n; ++i) b[i] -= q[i];
    double def = sqrt(oracle_dot(n, b, b)), def0 = def;
    history[0] = def;
    int i = 0;
    *converged = (def < def0 * reduction || def < 1e-30);
    if (!*converged) {
        oracle_twolevel_apply(S, restricted, mode, with_coarse, b, p, nthreads);
        double rholast = oracle_dot(n, p, b);
        for (i = 1; i <= maxit; ++i) {
            oracle_spmv(n, S->rowptr, S->col, S->val, p, q, nthreads);
            const double alpha = oracle_dot(n, p, q);
            const double lambda = (def == 0.0) ? 0.0 : rholast / alpha;
            for (int64_t k = 0; k < n; ++k) x[k] += lambda * p[k];
            for (int64_t k = 0; k < n; ++k) b[k] -= lambda * q[k];
            def = sqrt(oracle_dot(n, b, b));
            history[i] = def;
            *converged = (def < def0 * reduction || def < 1e-30);
            if (*converged) break;
            oracle_twolevel_apply(S, restricted, mode, with_coarse, b, q, nthreads);
            const double rho = oracle_dot(n, q, b);
            const double beta = (def == 0.0) ? 0.0 : rho / rholast;
            for (int64_t k = 0; k < n; ++k) p[k] = beta * p[k] + q[k];
            rholast = rho;
        }
        if (i > maxit) i = maxit;
    }
    free(p); free(q);
    return i;
}

int oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* Dune::RestartedGMResSolver::apply, /root/reference/extern/dune-istl/dune/istl/solvers.hh:923-1035 (left-preconditioned,
 * modified Gram-Schmidt, Givens rotations :1072-1111, update :1039-1059).  The defect is the PRECONDITIONED residual norm
 * |s[i+1]| (:948,1002); stopping rule as in oracle_pcg.  b is overwritten, x holds the solution.  Returns the iteration
 * count; history gets maxit+1 entries. */
static void gen_rot(double dx, double dy, double *cs, double *sn)
{
    const double eps = 1e-15, ndx = fabs(dx), ndy = fabs(dy);
    const double nmax = ndx > ndy ? ndx : ndy, nmin = ndx > ndy ? ndy : ndx;
    const double temp = nmin / nmax;
    if (ndy < eps) { *cs = 1.0; *sn = 0.0; }
    else if (ndx < eps) { *cs = 0.0; *sn = 1.0; }
    else if (ndy > ndx) { *cs = 1.0 / sqrt(1.0 + temp * temp) * temp; *sn = 1.0 / sqrt(1.0 + temp * temp) * dx * dy / ndx / ndy; }
    else { *cs = 1.0 / sqrt(1.0 + temp * temp); *sn = 1.0 / sqrt(1.0 + temp * temp) * (dy / dx); }
}
static void app_rot(double *dx, double *dy, double cs, double sn)
{
    const double t = cs * *dx + sn * *dy;
    *dy = -sn * *dx + cs * *dy;
    *dx = t;
}

int oracle_gmres(const oracle_setup *S, int restricted, int mode, int with_coarse, double *x, double *b, double reduction,
                 int restart, int maxit, double *history, int *converged, int nthreads)
{
    const int64_t n = S->n;
    const int m = restart;
    double *s = calloc(m + 1, sizeof(double)), *sn = calloc(m, sizeof(double)), *cs = calloc(m, sizeof(double));
    double *H = calloc((size_t)(m + 1) * m, sizeof(double)); /* H[k*m + i] */
    double *b2 = malloc(sizeof(double) * n), *w = malloc(sizeof(double) * n), *t = malloc(sizeof(double) * n);
    double **v = malloc(sizeof(double *) * (m + 1));
    for (int k = 0; k <= m; ++k) v[k] = malloc(sizeof(double) * n);
    memcpy(b2, b, sizeof(double) * n);
    oracle_spmv(n, S->rowptr, S->col, S->val, x, t, nthreads);
    for (int64_t q = 0; q < n; ++q) b[q] -= t[q];
    oracle_twolevel_apply(S, restricted, mode, with_coarse, b, v[0], nthreads);
    double norm = sqrt(oracle_dot(n, v[0], v[0]));
    const double def0 = norm;
    history[0] = norm;
    *converged = (norm < def0 * reduction || norm < 1e-30);
    int j = 1;
    while (j <= maxit && !*converged) {
        int i = 0;
        const double inv = (norm == 0.0) ? 0.0 : 1.0 / norm;
        for (int64_t q = 0; q < n; ++q) v[0][q] *= inv;
        s[0] = norm;
        for (i = 1; i < m + 1; ++i) s[i] = 0.0;
        for (i = 0; i < m && j <= maxit && !*converged; ++i, ++j) {
            oracle_spmv(n, S->rowptr, S->col, S->val, v[i], v[i + 1], nthreads);
            oracle_twolevel_apply(S, restricted, mode, with_coarse, v[i + 1], w, nthreads);
            for (int k = 0; k < i + 1; ++k) {
                const double h = oracle_dot(n, v[k], w);
                H[(size_t)k * m + i] = h;
                for (int64_t q = 0; q < n; ++q) w[q] -= h * v[k][q];
            }
            const double hn = sqrt(oracle_dot(n, w, w));
            H[(size_t)(i + 1) * m + i] = hn;
            const double sc = (norm == 0.0) ? 0.0 : 1.0 / hn;
            for (int64_t q = 0; q < n; ++q) v[i + 1][q] = w[q] * sc;
            for (int k = 0; k < i; ++k) app_rot(&H[(size_t)k * m + i], &H[(size_t)(k + 1) * m + i], cs[k], sn[k]);
            gen_rot(H[(size_t)i * m + i], H[(size_t)(i + 1) * m + i], &cs[i], &sn[i]);
            app_rot(&H[(size_t)i * m + i], &H[(size_t)(i + 1) * m + i], cs[i], sn[i]);
            app_rot(&s[i], &s[i + 1], cs[i], sn[i]);
            norm = fabs(s[i + 1]);
            history[j] = norm;
            *converged = (norm < def0 * reduction || norm < 1e-30);
        }
        /* update(w, i, H, s, v) */
        memset(w, 0, sizeof(double) * n);
        double *y = malloc(sizeof(double) * (m + 1));
        memcpy(y, s, sizeof(double) * (m + 1));
        for (int a = i - 1; a >= 0; --a) {
            double rhs = s[a];
            for (int c = a + 1; c < i; ++c) rhs -= H[(size_t)a * m + c] * y[c];
            y[a] = (rhs == 0.0) ? 0.0 : rhs / H[(size_t)a * m + a];
            for (int64_t q = 0; q < n; ++q) w[q] += y[a] * v[a][q];
        }
        free(y);
        for (int64_t q = 0; q < n; ++q) x[q] += w[q];
        if (!*converged && j < maxit) {
            memcpy(b, b2, sizeof(double) * n);
            oracle_spmv(n, S->rowptr, S->col, S->val, x, t, nthreads);
            for (int64_t q = 0; q < n; ++q) b[q] -= t[q];
            oracle_twolevel_apply(S, restricted, mode, with_coarse, b, v[0], nthreads);
            norm = sqrt(oracle_dot(n, v[0], v[0]));
        }
    }
    for (int k = 0; k <= m; ++k) free(v[k]);
    free(v); free(s); free(sn); free(cs); free(H); free(b2); free(w); free(t);
    return j - 1;
}
