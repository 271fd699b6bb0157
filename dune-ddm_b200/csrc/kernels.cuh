// Device kernels of the solve phase (sm_100a).  Every kernel here is HBM-bandwidth bound; the
// layouts are chosen so that each warp-wide access is a contiguous 256-byte (or wider) segment.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ddm {

constexpr int kWarp = 32;
constexpr unsigned kFull = 0xffffffffu;

// ------------------------------------------------------------------------------------------------
// K1  CSR SpMV, scalar FP64.  Replaces BCRSMatrix::mv / usmv
// (extern/dune-istl/dune/istl/bcrsmatrix.hh:1641-1661, 1713-1730).
// G lanes cooperate on one row (G chosen from the mean row length at upload); the G-lane groups of
// a warp read neighbouring rows, so value/column streams stay contiguous across the warp.
// mode 0: y = A x      mode 1: y += alpha * A x
// ------------------------------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(256) spmv_csr_kernel(int64_t n, const int32_t *__restrict__ rowptr,
                                                       const int32_t *__restrict__ col,
                                                       const double *__restrict__ val,
                                                       const double *__restrict__ x, double *__restrict__ y,
                                                       double alpha, int mode)
{
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t row = gtid / G;
    const int lane = (int)(gtid % G);
    double acc = 0.0;
    if (row < n) {
        const int32_t b = rowptr[row], e = rowptr[row + 1];
        for (int32_t k = b + lane; k < e; k += G) acc = fma(val[k], __ldg(x + col[k]), acc);
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o, G);
    if (row < n && lane == 0) {
        if (mode == 0) y[row] = acc;
        else y[row] += alpha * acc;
    }
}

// K1s  the same product from a sliced-ELL copy of the matrix (slices of 32 rows, entries of a slice stored column by column:
// entry k of row r sits at ptr[slice] + k*32 + lane).  One thread per row: every load of the value / column streams is one
// contiguous 256 / 128 byte segment per warp, all loads of a row are independent (no row-pointer indirection, no cross-lane
// reduction), and the row sum runs in ascending column order -- the order of BCRSMatrix::mv (bcrsmatrix.hh:1652-1660).  Used
// when padding the slices to their longest row costs < 10 % (structured and quasi-uniform meshes); padded entries carry
// value 0 and the row's own column.
__global__ void __launch_bounds__(256) spmv_sell_kernel(int64_t n, const int64_t *__restrict__ ptr, const int32_t *__restrict__ col,
                                                        const double *__restrict__ val, const double *__restrict__ x,
                                                        double *__restrict__ y, double alpha, int mode)
{
    const int64_t row = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t slice = row >> 5;
    if (slice * 32 >= n) return;
    const int64_t b = ptr[slice] + (threadIdx.x & 31);
    const int w = (int)((ptr[slice + 1] - ptr[slice]) >> 5);
    double acc = 0.0;
    int k = 0;
    for (; k + 4 <= w; k += 4) {
        const int64_t o = b + (int64_t)k * 32;
        const int32_t c0 = __ldcs(col + o), c1 = __ldcs(col + o + 32), c2 = __ldcs(col + o + 64), c3 = __ldcs(col + o + 96);
        const double v0 = __ldcs(val + o), v1 = __ldcs(val + o + 32), v2 = __ldcs(val + o + 64), v3 = __ldcs(val + o + 96);
        acc = fma(v0, __ldg(x + c0), acc);
        acc = fma(v1, __ldg(x + c1), acc);
        acc = fma(v2, __ldg(x + c2), acc);
        acc = fma(v3, __ldg(x + c3), acc);
    }
    for (; k < w; ++k) {
        const int64_t o = b + (int64_t)k * 32;
        acc = fma(__ldcs(val + o), __ldg(x + __ldcs(col + o)), acc);
    }
    if (row < n) {
        if (mode == 0) y[row] = acc;
        else y[row] += alpha * acc;
    }
}

// K1b  block-CSR SpMV with 3x3 blocks (BCRSMatrix<FieldMatrix<double,3,3>>, linear elasticity): same entry points as K1.
// The 9*nblk values of a block row are one contiguous stream; G lanes walk it element by element (fully coalesced: lane t
// reads value e = 9*first + t, t + G, ...), element e belongs to block e/9, row (e%9)/3, column e%3 of that block.  Each
// lane keeps three partial sums (one per block row), reduced across the G lanes at the end.  Vectors are node-major
// (BlockVector<FieldVector<double,3>>).
template <int G>
__global__ void __launch_bounds__(256) spmv_bcsr3_kernel(int64_t nb, const int32_t *__restrict__ rowptr,
                                                         const int32_t *__restrict__ col, const double *__restrict__ val,
                                                         const double *__restrict__ x, double *__restrict__ y, double alpha,
                                                         int mode)
{
    const int64_t gtid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t row = gtid / G;
    const int lane = (int)(gtid % G);
    double a0 = 0.0, a1 = 0.0, a2 = 0.0;
    if (row < nb) {
        const int32_t b = rowptr[row], e9 = 9 * rowptr[row + 1];
        for (int32_t e = 9 * b + lane; e < e9; e += G) {
            const int32_t blk = e / 9, rem = e - 9 * blk, rr = rem / 3, cc = rem - 3 * rr;
            const double p = val[e] * __ldg(x + 3 * (int64_t)col[blk] + cc);
            a0 += rr == 0 ? p : 0.0;
            a1 += rr == 1 ? p : 0.0;
            a2 += rr == 2 ? p : 0.0;
        }
    }
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) {
        a0 += __shfl_xor_sync(kFull, a0, o, G);
        a1 += __shfl_xor_sync(kFull, a1, o, G);
        a2 += __shfl_xor_sync(kFull, a2, o, G);
    }
    if (row < nb && lane < 3) {
        const double a = lane == 0 ? a0 : lane == 1 ? a1 : a2;
        double *yp = y + 3 * row + lane;
        if (mode == 0) *yp = a;
        else *yp += alpha * a;
    }
}

// ------------------------------------------------------------------------------------------------
// K8/K9  vector updates and deterministic two-stage reductions (Krylov recurrences of
// extern/dune-istl/dune/istl/solvers.hh:312-335; masked dot of owneroverlapcopy.hh:398-444).
// ------------------------------------------------------------------------------------------------
__global__ void axpy_kernel(int64_t n, double a, const double *__restrict__ x, double *__restrict__ y)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = fma(a, x[i], y[i]);
}
__global__ void scale_kernel(int64_t n, double a, double *__restrict__ y)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] *= a;
}
__global__ void fill_kernel(int64_t n, double a, double *__restrict__ y)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = a;
}
// y = x + a*y
__global__ void xpay_kernel(int64_t n, double a, const double *__restrict__ x, double *__restrict__ y)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = fma(a, y[i], x[i]);
}
// y = x + (num/den)*y with the scalars on the device (p = q + beta p, beta = rho / rholast)
__global__ void xpay_dev_kernel(int64_t n, const double *__restrict__ num, const double *__restrict__ den, int zero,
                                const double *__restrict__ x, double *__restrict__ y)
{
    const double a = zero ? 0.0 : (*num) / (*den);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = fma(a, y[i], x[i]);
}

__device__ __forceinline__ double block_sum_256(double v, double *sm)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sm[w] = v;
    __syncthreads();
    double t = 0.0;
    if (w == 0) {
        t = (l < (int)(blockDim.x >> 5)) ? sm[l] : 0.0;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(kFull, t, o);
    }
    __syncthreads();
    return t; // valid in thread 0
}

// stage 1: per-block partial of sum x*y (fixed grid => fixed summation tree => reproducible)
__global__ void __launch_bounds__(256) dot_partial_kernel(int64_t n, const double *__restrict__ x,
                                                          const double *__restrict__ y, double *__restrict__ partial)
{
    __shared__ double sm[8];
    double acc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        acc = fma(x[i], y[i], acc);
    const double t = block_sum_256(acc, sm);
    if (threadIdx.x == 0) partial[blockIdx.x] = t;
}
// stage 2: one block sums the partials in a fixed order
__global__ void __launch_bounds__(256) reduce_final_kernel(int np, const double *__restrict__ partial,
                                                           double *__restrict__ out)
{
    __shared__ double sm[8];
    double acc = 0.0;
    for (int i = threadIdx.x; i < np; i += blockDim.x) acc += partial[i];
    const double t = block_sum_256(acc, sm);
    if (threadIdx.x == 0) *out = t;
}
// fused CG update (solvers.hh:318-322): lambda = rholast/alpha; x += lambda p; r -= lambda q; partial |r|^2
__global__ void __launch_bounds__(256) cg_update_kernel(int64_t n, const double *__restrict__ rholast,
                                                        const double *__restrict__ alpha, int zero,
                                                        const double *__restrict__ p, const double *__restrict__ q,
                                                        double *__restrict__ x, double *__restrict__ r,
                                                        double *__restrict__ partial)
{
    __shared__ double sm[8];
    const double lambda = zero ? 0.0 : (*rholast) / (*alpha);
    double acc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        x[i] = fma(lambda, p[i], x[i]);
        const double ri = fma(-lambda, q[i], r[i]);
        r[i] = ri;
        acc = fma(ri, ri, acc);
    }
    const double t = block_sum_256(acc, sm);
    if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

// modified Gram-Schmidt step of RestartedGMResSolver (solvers.hh:975-981), two passes fused into one:
//   w -= h_prev * v_prev  (the axpy of the previous k, scalar on the device), then the partial of  v_k . w
// (v_k == nullptr: partial of |w|^2, the norm that closes the Arnoldi step, :982).
__global__ void __launch_bounds__(256) mgs_kernel(int64_t n, const double *__restrict__ hprev, const double *__restrict__ vprev,
                                                  const double *__restrict__ vk, double *__restrict__ w,
                                                  double *__restrict__ partial)
{
    __shared__ double sm[8];
    const double h = vprev ? *hprev : 0.0;
    double acc = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double wi = w[i];
        if (vprev) {
            wi = fma(-h, vprev[i], wi);
            w[i] = wi;
        }
        acc = fma(vk ? vk[i] : wi, wi, acc);
    }
    const double t = block_sum_256(acc, sm);
    if (threadIdx.x == 0) partial[blockIdx.x] = t;
}
// dst = src * (1 / sqrt(nrm2)) (v[i+1] = w; v[i+1] *= 1/H[i+1][i], solvers.hh:988-991); zero != 0 gives dst = 0
__global__ void normalize_kernel(int64_t n, const double *__restrict__ nrm2, int zero, const double *__restrict__ src,
                                 double *__restrict__ dst)
{
    const double a = zero ? 0.0 : 1.0 / sqrt(*nrm2);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = src[i] * a;
}

// ------------------------------------------------------------------------------------------------
// K2 + K5  restriction of the defect to all hosted subdomains, fused with the coarse restriction.
// schwarz.hh:121-125 / galerkin_preconditioner.hh:159-167: W[p] = d[gidx[p]] (pivot order), and for the
// segment's subdomain the partial sums  sum_p Z[p,k] d_p  for its k basis vectors.
// One block per segment (<= kSeg consecutive positions of one subdomain).
// ------------------------------------------------------------------------------------------------
constexpr int kSeg = 2048;
constexpr int kKmax = 16; // basis vectors handled per pass

struct Segment {
    int32_t sub;      // hosted subdomain
    int32_t start;    // first position (absolute offset into W)
    int32_t len;      // number of positions
    int32_t local0;   // start - base of the subdomain
};

__global__ void __launch_bounds__(256, 4) restrict_kernel(const Segment *__restrict__ segs, const int32_t *__restrict__ gidx,
                                                          const double *__restrict__ d, double *__restrict__ W,
                                                          double *__restrict__ W2, int with_coarse,
                                                          const int64_t *__restrict__ zoff, const int32_t *__restrict__ sub_n,
                                                          const int32_t *__restrict__ sub_k, const double *__restrict__ Zp,
                                                          double *__restrict__ partial /* [nseg][kpad] */, int kpad)
{
    __shared__ double sm[8][kKmax];
    const Segment sg = segs[blockIdx.x];
    const int k = with_coarse ? sub_k[sg.sub] : 0;
    const int32_t ni = sub_n[sg.sub];
    const double *Zi = with_coarse ? Zp + zoff[sg.sub] + sg.local0 : nullptr;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int k0 = 0; k0 < (k > 0 ? k : 1); k0 += kKmax) {
        double acc[kKmax];
#pragma unroll
        for (int j = 0; j < kKmax; ++j) acc[j] = 0.0;
        for (int t = threadIdx.x; t < sg.len; t += blockDim.x) {
            const int p = sg.start + t;
            const double v = __ldg(d + gidx[p]);
            if (k0 == 0) {
                W[p] = v;
                if (W2) W2[p] = 0.0;
            }
            // the basis loads of this position are issued eight at a time before their first use (the compiler otherwise
            // pairs each load with its FMA and keeps two in flight); two batches keep the kernel at three blocks per SM
#pragma unroll
            for (int h = 0; h < kKmax; h += 8) {
                double z[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) z[j] = (k0 + h + j < k) ? __ldcs(Zi + (int64_t)(k0 + h + j) * ni + t) : 0.0;
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[h + j] = fma(z[j], v, acc[h + j]);
            }
        }
        if (k > 0) {
            // fixed summation tree: lanes by butterfly, then the eight warps in ascending order; ONE barrier pair for all
            // basis vectors of the batch
#pragma unroll
            for (int j = 0; j < kKmax; ++j) {
                double s = acc[j];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(kFull, s, o);
                if (lane == 0) sm[warp][j] = s;
            }
            __syncthreads();
            if (threadIdx.x < kKmax && k0 + (int)threadIdx.x < k) {
                double s = 0.0;
#pragma unroll
                for (int w = 0; w < 8; ++w) s += sm[w][threadIdx.x];
                partial[(int64_t)blockIdx.x * kpad + k0 + threadIdx.x] = s;
            }
            __syncthreads();
        }
    }
}

// second stage of K5: d0[koff[i]+k] = sum over the segments of subdomain i, in segment order
__global__ void coarse_restrict_final_kernel(int nsub, const int32_t *__restrict__ seg_ptr, const int32_t *__restrict__ sub_k,
                                             const int32_t *__restrict__ koff, const double *__restrict__ partial, int kpad,
                                             double *__restrict__ d0)
{
    const int i = blockIdx.x;
    if (i >= nsub) return;
    for (int k = threadIdx.x; k < sub_k[i]; k += blockDim.x) {
        double acc = 0.0;
        for (int s = seg_ptr[i]; s < seg_ptr[i + 1]; ++s) acc += partial[(int64_t)s * kpad + k];
        d0[koff[i] + k] = acc;
    }
}

// K6  dense coarse solve as matrix-vector products with the explicit inverse (plus one refinement step
// driven from the host side): out = M v, M row-major N x N, one warp per row.
__global__ void __launch_bounds__(256) dense_gemv_kernel(int N, const double *__restrict__ M, const double *__restrict__ v,
                                                         double *out, const double *addto, double sign, int tri)
{
    // tri 0: full rows; 1: lower triangle (columns <= row); 2: upper triangle (columns >= row, start aligned down to 32)
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (row >= N) return;
    const double *Mr = M + (int64_t)row * N;
    const int j0 = tri == 2 ? (row & ~31) : 0, j1 = tri == 1 ? row + 1 : N;
    double acc = 0.0;
    for (int j = j0 + lane; j < j1; j += 32) acc = fma(Mr[j], __ldg(v + j), acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(kFull, acc, o);
    if (lane == 0) out[row] = addto ? addto[row] + sign * acc : sign * acc;
}

// ------------------------------------------------------------------------------------------------
// K4 + K7  prolongation.  Stage a (position-wise, coalesced):
//   X[p] <- (pou ? pou[p] : 1) * X[p] * fine  +  sum_k Z[p,k] y[koff+k]
// (schwarz.hh:139-141 and galerkin_preconditioner.hh:186-188), stage b (DOF-wise gather, no atomics,
// fixed order): x[v] (+)= sum over the positions that map to v  (the halo sums of schwarz.hh:138/142
// and galerkin_preconditioner.hh:190 for all subdomains hosted on this GPU).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) prolong_scale_kernel(const Segment *__restrict__ segs, double *__restrict__ X,
                                                            const double *__restrict__ pouP, int fine, int with_coarse,
                                                            const int64_t *__restrict__ zoff, const int32_t *__restrict__ sub_n,
                                                            const int32_t *__restrict__ sub_k, const int32_t *__restrict__ koff,
                                                            const double *__restrict__ Zp, const double *__restrict__ y0)
{
    __shared__ double ys[64];
    const Segment sg = segs[blockIdx.x];
    const int k = with_coarse ? sub_k[sg.sub] : 0;
    const int32_t ni = sub_n[sg.sub];
    const double *Zi = with_coarse ? Zp + zoff[sg.sub] + sg.local0 : nullptr;
    for (int t = threadIdx.x; t < sg.len; t += blockDim.x) {
        const int p = sg.start + t;
        double v = 0.0;
        if (fine) v = pouP ? pouP[p] * X[p] : X[p];
        X[p] = v;
    }
    for (int k0 = 0; k0 < k; k0 += 64) {
        __syncthreads();
        if (threadIdx.x < 64 && k0 + threadIdx.x < k) ys[threadIdx.x] = y0[koff[sg.sub] + k0 + threadIdx.x];
        __syncthreads();
        const int kk = min(64, k - k0);
        for (int t = threadIdx.x; t < sg.len; t += blockDim.x) {
            double v = X[sg.start + t];
            for (int j = 0; j < kk; ++j) v = fma(Zi[(int64_t)(k0 + j) * ni + t], ys[j], v);
            X[sg.start + t] = v;
        }
    }
}

__global__ void __launch_bounds__(256) gather_sum_kernel(int64_t nloc, const int32_t *__restrict__ dptr,
                                                         const int32_t *__restrict__ dpos, const double *__restrict__ X,
                                                         double *__restrict__ x, int accumulate)
{
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= nloc) return;
    double acc = 0.0;
    for (int32_t q = dptr[v]; q < dptr[v + 1]; ++q) acc += X[dpos[q]];
    x[v] = accumulate ? x[v] + acc : acc;
}


// ------------------------------------------------------------------------------------------------
// Segmented vector operations over the subdomain ranges of a block-diagonal problem (all subdomains advance in lock
// step): the recurrences of the batched shift-invert Lanczos used for the GenEO eigenproblems (SURVEY 8f N3;
// replaces one Spectra SymGEigsShiftSolver run per rank, dune/ddm/eigensolvers/spectra.hh:112-215).
// Segments are the <= kSeg-long pieces of the subdomain ranges built at finalize.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) seg_dot_partial_kernel(const Segment *__restrict__ segs, const double *__restrict__ x,
                                                              const double *__restrict__ y, double *__restrict__ partial)
{
    __shared__ double sm[8];
    const Segment sg = segs[blockIdx.x];
    double acc = 0.0;
    for (int t = threadIdx.x; t < sg.len; t += blockDim.x) acc = fma(x[sg.start + t], y[sg.start + t], acc);
    const double r = block_sum_256(acc, sm);
    if (threadIdx.x == 0) partial[blockIdx.x] = r;
}
__global__ void seg_dot_final_kernel(int nsub, const int32_t *__restrict__ seg_ptr, const double *__restrict__ partial,
                                     double *__restrict__ out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nsub) return;
    double acc = 0.0;
    for (int s = seg_ptr[i]; s < seg_ptr[i + 1]; ++s) acc += partial[s];
    out[i] = acc;
}
// y[p] = beta[sub(p)] * y[p] + alpha[sub(p)] * x[p]   (beta == nullptr: 1)
__global__ void __launch_bounds__(256) seg_axpby_kernel(const Segment *__restrict__ segs, const double *__restrict__ alpha,
                                                        const double *__restrict__ x, const double *__restrict__ beta,
                                                        double *__restrict__ y)
{
    const Segment sg = segs[blockIdx.x];
    const double a = alpha[sg.sub], b = beta ? beta[sg.sub] : 1.0;
    for (int t = threadIdx.x; t < sg.len; t += blockDim.x) {
        const int p = sg.start + t;
        y[p] = fma(a, x[p], b * y[p]);
    }
}

} // namespace ddm
