/* CPU ORACLE -- TEST INFRASTRUCTURE, NOT THE PRODUCT.  Declarations of oracle.c (see its header comment). */
#ifndef DDM_ORACLE_H
#define DDM_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Setup products of all subdomains, flat (see oracle/oracle.py for how they are packed). */
typedef struct {
    int32_t nsub;
    int64_t n;              /* unique global DOFs */
    const int64_t *sub_ptr; /* nsub+1 */
    const int32_t *sub_dofs;/* R_i index lists */
    const double *pou;      /* D_i, same layout as sub_dofs */
    /* factors, one entry per subdomain */
    const int32_t *nsuper;
    const int32_t **super;
    const int64_t **pi, **px;
    const int32_t **s;
    const double **x;
    const int32_t **perm;
    /* coarse space */
    const int64_t *z_ptr;   /* nsub+1 offsets into Z (column-major n_i x k_i blocks) */
    const int32_t *k_off;   /* nsub+1 exclusive scan of k_i */
    const double *Z;
    const double *A0chol;   /* dense lower Cholesky factor of A0, row-major N0 x N0 */
    /* operator (for the multiplicative combine) */
    const int64_t *rowptr;
    const int32_t *col;
    const double *val;
    /* non-symmetric problems (NULL otherwise): panels of U^T per subdomain, same symbolic layout as x (which then holds
     * the unit-lower L), and the dense coarse matrix as LAPACK getrf output (row-major L\\U, 0-based row swaps) */
    const double **xu;
    const double *A0lu;
    const int32_t *A0piv;
} oracle_setup;

void oracle_spmv(int64_t n, const int64_t *rowptr, const int32_t *col, const double *val, const double *x, double *y,
                 int nthreads);
double oracle_dot(int64_t n, const double *x, const double *y);
void oracle_super_solve(int32_t n, int32_t nsuper, const int32_t *super, const int64_t *pi, const int64_t *px,
                        const int32_t *s, const double *x, const int32_t *perm, const double *b, double *out, double *work);
void oracle_super_solve_lu(int32_t n, int32_t nsuper, const int32_t *super, const int64_t *pi, const int64_t *px,
                           const int32_t *s, const double *xl, const double *xu, const int32_t *perm, const double *b,
                           double *out, double *work);
void oracle_schwarz_apply(const oracle_setup *S, int restricted, const double *d, double *x, int nthreads);
void oracle_coarse_apply(const oracle_setup *S, const double *d, double *x, int nthreads);
void oracle_twolevel_apply(const oracle_setup *S, int restricted, int mode, int with_coarse, const double *d, double *x,
                           int nthreads);
int oracle_pcg(const oracle_setup *S, int restricted, int mode, int with_coarse, double *x, double *b, double reduction,
               int maxit, double *history, int *converged, int nthreads);
int oracle_gmres(const oracle_setup *S, int restricted, int mode, int with_coarse, double *x, double *b, double reduction,
                 int restart, int maxit, double *history, int *converged, int nthreads);
int oracle_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
