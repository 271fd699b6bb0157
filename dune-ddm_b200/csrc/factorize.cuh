// Setup-time kernels: numeric multifrontal Cholesky of one subdomain matrix on the device, with a symbolic analysis
// (ordering, supernodes, row structures) handed over by the caller.
//
// Stands in for cholmod_l_factorize inside Dune::Cholmod::setMatrix (extern/dune-istl/dune/istl/cholmod.hh:469-472) once
// cholmod_l_analyze has run, and for the numeric re-factorisation of TwoLevelSchwarzSolver's update branch
// (dune/ddm/twolevel_schwarz.hh:99-104).  The product is the supernodal LL^T factor in cholmod_factor layout (column-major
// panels, dense diagonal block on top), left on the device for the tile conversion of ddm_upload_factor.
//
// Fronts are processed level by level of the supernodal elimination tree.  A front is [panel (m x cw) | update (nb x nb)]:
// the panel is the supernode's own storage in the factor, the update matrix lives in an arena until the parent has
// consumed it.  Narrow supernodes (cw <= 32) are done by one CTA each, many per launch; wide ones go through
// cuSOLVER/cuBLAS (potrf, trsm, syrk) one after the other (setup only, not on the apply path).
#pragma once
#include "kernels.cuh"

namespace ddm {

struct MfFront {
    int64_t px;         // panel offset in the factor array
    int64_t pi;         // row-list offset in s
    int64_t uoff;       // update matrix (nb x nb, column-major, lower triangle valid) in the arena; -1: none
    int32_t f, cw, m;   // first column, columns, rows of the panel (m = cw + nb)
    int32_t child_ptr, child_end;  // children in child_idx
    int32_t pad;
};

// Position of pivot index `row` in a supernode's row list: own columns first, then the ascending structure.
__device__ __forceinline__ int mf_front_pos(const int32_t *__restrict__ s, int64_t p0, int32_t f, int32_t cw, int32_t m, int32_t row)
{
    if (row < f + cw) return row - f;
    const int32_t *st = s + p0 + cw;
    int lo = 0, hi = m - cw - 1;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (st[mid] < row) lo = mid + 1;
        else hi = mid;
    }
    return cw + lo;
}

// Entries of A_i (CSR, the subdomain's own numbering, both triangles stored) -> panels of the permuted matrix.
// One warp per row.  err[0] is set when an entry falls outside the symbolic structure.
__global__ void __launch_bounds__(256) mf_scatter_kernel(int32_t n, const int64_t *__restrict__ rowptr, const int32_t *__restrict__ col,
                                                         const double *__restrict__ val, const int32_t *__restrict__ iperm,
                                                         const int32_t *__restrict__ col2super, const MfFront *__restrict__ fr,
                                                         const int32_t *__restrict__ s, double *__restrict__ raw, int *err)
{
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (w >= n) return;
    const int32_t pr = iperm[w];
    const MfFront F = fr[col2super[pr]];
    for (int64_t q = rowptr[w] + lane; q < rowptr[w + 1]; q += 32) {
        const int32_t pc = iperm[col[q]];
        if (pc < pr) continue;  // the transposed entry lands in this panel column
        const int pos = mf_front_pos(s, F.pi, F.f, F.cw, F.m, pc);
        if (pos >= F.m || s[F.pi + pos] != pc) {
            atomicExch(err, 1);
            continue;
        }
        raw[F.px + (int64_t)(pr - F.f) * F.m + pos] = val[q];
    }
}

// rel[q], q in the structure part of supernode J's row list: position of that row in the parent's row list.
__global__ void __launch_bounds__(128) mf_relidx_kernel(int32_t nsuper, const MfFront *__restrict__ fr, const int32_t *__restrict__ parent,
                                                        const int32_t *__restrict__ s, int32_t *__restrict__ rel)
{
    const int32_t J = blockIdx.x;
    if (J >= nsuper) return;
    const MfFront F = fr[J];
    const int32_t p = parent[J];
    if (p < 0) return;
    const MfFront Pf = fr[p];
    for (int q = F.cw + threadIdx.x; q < F.m; q += blockDim.x) rel[F.pi + q] = mf_front_pos(s, Pf.pi, Pf.f, Pf.cw, Pf.m, s[F.pi + q]);
}

// Extend-add of one child's update matrix into the parent's front, lower triangle; column b of the child per warp.
__device__ __forceinline__ void mf_extend_add_cols(const MfFront &Pf, const MfFront &Cf, const int32_t *__restrict__ rel,
                                                   double *__restrict__ raw, double *__restrict__ arena, int warp, int nwarps, int lane)
{
    const int nbC = Cf.m - Cf.cw;
    const int nbP = Pf.m - Pf.cw;
    const double *UC = arena + Cf.uoff;
    const int32_t *rc = rel + Cf.pi + Cf.cw;
    double *UP = Pf.uoff >= 0 ? arena + Pf.uoff : nullptr;
    for (int b = warp; b < nbC; b += nwarps) {
        const int rb = rc[b];
        double *dst = rb < Pf.cw ? raw + Pf.px + (int64_t)rb * Pf.m : UP + (int64_t)(rb - Pf.cw) * nbP - Pf.cw;
        const double *src = UC + (int64_t)b * nbC;
        for (int a = b + lane; a < nbC; a += 32) dst[rc[a]] += src[a];
    }
}

__global__ void __launch_bounds__(256) mf_extend_add_kernel(const MfFront *__restrict__ fr, int32_t P, int32_t C,
                                                            const int32_t *__restrict__ rel, double *__restrict__ raw,
                                                            double *__restrict__ arena)
{
    const int nw = (gridDim.x * blockDim.x) >> 5;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    mf_extend_add_cols(fr[P], fr[C], rel, raw, arena, w, nw, threadIdx.x & 31);
}

// devInfo of cusolverDnDpotrf -> the factorisation's error flag (checked once, at the end).
__global__ void mf_collect_info_kernel(const int *info, int *err)
{
    if (threadIdx.x == 0 && *info != 0) atomicExch(err, 2);
}

// One narrow front (cw <= 32) per CTA: zero its update matrix, add the children's, factor the pivot block, solve for
// the rows below, subtract the outer product from the update matrix.
__global__ void __launch_bounds__(256, 2) mf_small_fronts_kernel(const MfFront *__restrict__ fr, const int32_t *__restrict__ list,
                                                              const int32_t *__restrict__ child_idx, const int32_t *__restrict__ rel,
                                                              double *__restrict__ raw, double *__restrict__ arena, int *err)
{
    __shared__ double Ls[32][33];
    __shared__ double As[32][33], Bs[32][33];
    const int J = list[blockIdx.x];
    const MfFront F = fr[J];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cw = F.cw, m = F.m, nb = m - cw;
    double *pan = raw + F.px;
    double *U = F.uoff >= 0 ? arena + F.uoff : nullptr;
    if (U)
        for (int64_t q = tid; q < (int64_t)nb * nb; q += 256) U[q] = 0.0;
    __syncthreads();
    for (int ci = F.child_ptr; ci < F.child_end; ++ci) {
        mf_extend_add_cols(F, fr[child_idx[ci]], rel, raw, arena, warp, 8, lane);
        __syncthreads();
    }
    // pivot block: left-looking Cholesky by warp 0, one column at a time
    for (int q = tid; q < 32 * 32; q += 256) {
        const int i = q & 31, j = q >> 5;
        Ls[i][j] = (i < cw && j < cw && i >= j) ? pan[(int64_t)j * m + i] : 0.0;
    }
    __syncthreads();
    if (warp == 0) {
        for (int j = 0; j < cw; ++j) {
            double acc = Ls[lane][j];
            for (int k = 0; k < j; ++k) acc -= Ls[lane][k] * Ls[j][k];
            const double djj = __shfl_sync(kFull, acc, j);
            if (!(djj > 0.0) && lane == 0) atomicExch(err, 2);
            const double d = sqrt(djj);
            __syncwarp();
            if (lane == j) Ls[lane][j] = d;
            else if (lane > j) Ls[lane][j] = acc / d;
            __syncwarp();
        }
    }
    __syncthreads();
    for (int q = tid; q < cw * cw; q += 256) {
        const int i = q % cw, j = q / cw;
        pan[(int64_t)j * m + i] = Ls[i][j];  // strict upper part: zero
    }
    // rows below: x L11^T = a, one row per thread
    for (int i = tid; i < nb; i += 256) {
        double x[32];
        double *row = pan + cw + i;
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = (j < cw) ? row[(int64_t)j * m] : 0.0;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (j < cw) {
                double acc = x[j];
#pragma unroll
                for (int k = 0; k < j; ++k) acc -= x[k] * Ls[j][k];
                x[j] = acc / Ls[j][j];
                row[(int64_t)j * m] = x[j];
            }
        }
    }
    if (!U || nb == 0) return;
    __syncthreads();
    // update matrix: U(a, b) -= sum_k L21(a, k) L21(b, k), lower triangle, 32 x 32 tiles
    const double *L21 = pan + cw;
    const int nt = (nb + 31) / 32;
    const int ti = tid & 31, tj = (tid >> 5) * 4;
    for (int ta = 0; ta < nt; ++ta) {
        for (int q = tid; q < 32 * 32; q += 256) {
            const int i = q & 31, k = q >> 5;
            As[k][i] = (k < cw && ta * 32 + i < nb) ? L21[(int64_t)k * m + ta * 32 + i] : 0.0;
        }
        for (int tb = 0; tb <= ta; ++tb) {
            __syncthreads();
            for (int q = tid; q < 32 * 32; q += 256) {
                const int i = q & 31, k = q >> 5;
                Bs[k][i] = (k < cw && tb * 32 + i < nb) ? L21[(int64_t)k * m + tb * 32 + i] : 0.0;
            }
            __syncthreads();
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
            for (int k = 0; k < cw; ++k) {
                const double av = As[k][ti];
                a0 = fma(av, Bs[k][tj], a0);
                a1 = fma(av, Bs[k][tj + 1], a1);
                a2 = fma(av, Bs[k][tj + 2], a2);
                a3 = fma(av, Bs[k][tj + 3], a3);
            }
            const int a = ta * 32 + ti, b = tb * 32 + tj;
            if (a < nb) {
                double *u = U + a;
                if (b < nb && a >= b) u[(int64_t)b * nb] -= a0;
                if (b + 1 < nb && a >= b + 1) u[(int64_t)(b + 1) * nb] -= a1;
                if (b + 2 < nb && a >= b + 2) u[(int64_t)(b + 2) * nb] -= a2;
                if (b + 3 < nb && a >= b + 3) u[(int64_t)(b + 3) * nb] -= a3;
            }
        }
        __syncthreads();
    }
}

// Inverses of the narrow pivot blocks for the block-LDL^T form of the sweeps (trisolve.cuh): one warp per supernode.
// Linv = L_JJ^{-1} (cw x cw, column-major, lower), Dinv = Tinv^T Linv with T = U_JJ^T (LU) or T = L_JJ (Cholesky).
struct SmallInv {
    int64_t src;   // offset of the pivot block in the raw panel array (leading dimension m)
    int64_t linv, dinv, tinv;  // offsets in aux; tinv < 0: symmetric
    int32_t m, cw;
};

__device__ __forceinline__ bool mf_tri_inverse(double (*L)[33], double (*X)[33], int cw, int lane)
{
    // lane j owns column j of the inverse: X(i, j), i >= j, by forward substitution
    bool ok = true;
    for (int i = 0; i < cw; ++i) {
        const double dii = L[i][i];
        if (dii == 0.0 || !isfinite(dii)) ok = false;
        if (lane <= i && lane < cw) {
            double acc = 0.0;
            for (int k = lane; k < i; ++k) acc += L[i][k] * X[k][lane];
            X[i][lane] = lane == i ? 1.0 / dii : -acc / dii;
        }
        __syncwarp();
    }
    return ok;
}

__global__ void __launch_bounds__(32) mf_small_inverse_kernel(int32_t count, const SmallInv *__restrict__ jobs, const double *__restrict__ raw,
                                                              const double *__restrict__ rawU, double *__restrict__ aux, int lu, int *err)
{
    __shared__ double L[32][33], X[32][33], T[32][33];
    const int lane = threadIdx.x;
    const int job = blockIdx.x;
    if (job >= count) return;
    const SmallInv S = jobs[job];
    const int cw = S.cw;
    for (int j = 0; j < cw; ++j) {
        L[lane][j] = (lane < cw && lane >= j) ? raw[S.src + (int64_t)j * S.m + lane] : 0.0;
        X[lane][j] = 0.0;
    }
    __syncwarp();
    bool ok = mf_tri_inverse(L, X, cw, lane);
    if (!lu)
        for (int j = 0; j < cw; ++j)
            if (!(L[j][j] > 0.0)) ok = false;
    for (int j = 0; j < cw; ++j)
        if (lane < cw) aux[S.linv + (int64_t)j * cw + lane] = X[lane][j];
    if (lu) {
        for (int j = 0; j < cw; ++j) {
            L[lane][j] = (lane < cw && lane >= j) ? rawU[S.src + (int64_t)j * S.m + lane] : 0.0;
            T[lane][j] = 0.0;
        }
        __syncwarp();
        ok = mf_tri_inverse(L, T, cw, lane) && ok;
        for (int j = 0; j < cw; ++j)
            if (lane < cw) aux[S.tinv + (int64_t)j * cw + lane] = T[lane][j];
    }
    __syncwarp();
    double(*Ti)[33] = lu ? T : X;
    // Dinv(a, b) = sum_{k >= max(a, b)} Tinv(k, a) Linv(k, b); lane = a
    if (lane < cw)
        for (int b = 0; b < cw; ++b) {
            double acc = 0.0;
            for (int k = max(lane, b); k < cw; ++k) acc += Ti[k][lane] * X[k][b];
            aux[S.dinv + (int64_t)b * cw + lane] = acc;
        }
    if (!ok && lane == 0) atomicExch(err, 3);
}

}  // namespace ddm
