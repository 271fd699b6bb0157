// The two sweeps of the batched subdomain solve (see trisolve.cuh for the design).
#pragma once

namespace ddm {

// forward sweep: WX[rows] += panel(rows, chunks) * y(chunks)   (the panel holds D^{-1} rows and -Lt rows)
// Per task the vector operand (<= kGroup chunks of y) and the row targets are staged once in shared memory, so the
// tile loop touches shared memory only; the next ticket and descriptor are fetched while the tiles are consumed.
//
// DET = true is the reproducible variant (ddm_set_deterministic): no floating-point atomics.  Every task stores its results
// into private slots of S (its row targets are slot numbers), and "assemble" tasks (ntiles == 0), one per 32-column chunk of
// every supernode that has children, add the slots that belong to their columns in ascending slot number -- a fixed order
// -- before the supernode's compute tasks read y.  The default variant (DET = false) is the code as it was.
template <bool DET>
__global__ void __launch_bounds__(kTriWarps * 32, 1) trisolve_fwd_kernel_t(const TriTask *__restrict__ tasks, int ntasks, double *WX,
                                                                            int *cnt, int *ticket, int xoff, double *__restrict__ S,
                                                                            const int32_t *__restrict__ lptr, const int32_t *__restrict__ lidx)
{
    extern __shared__ __align__(128) unsigned char tri_smem[];
    TriWarp w;
    tri_setup(w, tasks, ntasks, ticket, tri_smem);
    const int lane = w.lane;
    w.start();
    while (w.hasA) {
        const int ncc = w.A.ncc;
        w.prefetch_ticket();
        if (DET && w.A.ntiles == 0) {
            // assemble y for 32 columns: restricted defect + the descendants' contributions, in slot order
            if (w.A.wait_cnt >= 0) wait_counter(cnt + w.A.wait_cnt, w.A.wait_need, lane);
            w.prefetch_task();
            if (lane < w.A.w) {
                const int p = w.A.vec + lane;
                double y = __ldcg(WX + p);
                for (int q = lptr[p]; q < lptr[p + 1]; ++q) y += __ldcg(S + lidx[q]);
                WX[p] = y;
            }
            __threadfence();
            __syncwarp();
            if (lane == 0) atomicAdd(cnt + w.A.signal_cnt, 1);
            w.advance();
            continue;
        }
        if (ncc == 0) {
            // Symmetric task on the inverse diagonal block: tiles (r, c) of its lower triangle, fixed chunk c.  Each tile
            // is used twice: z_r += T y_c, and (off the diagonal) z_c += T^T y_r -- the upper triangle is never stored.
            const int nt = w.A.ntiles, wc = w.A.w;
            const int k0 = (int)(intptr_t)w.A.rowidx, nvalid = w.A.gstride;
            if (w.A.wait_cnt >= 0) wait_counter(cnt + w.A.wait_cnt, w.A.wait_need, lane);
            const double *yp = WX + w.A.vec + lane;
            const double yc = (lane < wc) ? __ldcg(yp) : 0.0;
            double yr[kGroup - 1];
#pragma unroll
            for (int t = 0; t < kGroup - 1; ++t) yr[t] = (t < nt && t * 32 + lane < nvalid) ? __ldcg(yp + (k0 + t) * 32) : 0.0;
            w.prefetch_task();
            w.vals[lane] = yc;
#pragma unroll
            for (int t = 0; t < kGroup - 1; ++t) w.vals[(1 + t) * 32 + lane] = yr[t];
            __syncwarp();
            double acc[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) acc[j] = 0.0;
            double *zr = WX + xoff + w.A.vec + lane;
            for (int ti = 0; ti < nt; ++ti) {
                const double xr = (k0 + ti == 0) ? 0.0 : w.vals[(1 + ti) * 32 + lane];  // diagonal tile: stored in full
                const double *T = w.acquire_tile() + lane;
                double a0 = 0.0, a1 = 0.0;
                if (wc == 32) {
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {
                        const double t0 = T[j * 32], t1 = T[(j + 1) * 32];
                        a0 = fma(t0, w.vals[j], a0);
                        a1 = fma(t1, w.vals[j + 1], a1);
                        acc[j] = fma(t0, xr, acc[j]);
                        acc[j + 1] = fma(t1, xr, acc[j + 1]);
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < wc) {
                            const double t0 = T[j * 32];
                            a0 = fma(t0, w.vals[j], a0);
                            acc[j] = fma(t0, xr, acc[j]);
                        }
                }
                w.release_tile();
                if (DET) {
                    if (ti * 32 + lane < nvalid) S[w.A.signal_cnt + ti * 32 + lane] = a0 + a1;  // symmetric tasks carry their slot base here
                } else if (ti * 32 + lane < nvalid)
                    atomicAdd(zr + (k0 + ti) * 32, a0 + a1);
            }
#pragma unroll
            for (int half = 16; half >= 1; half >>= 1) {
                const bool upper = (lane & half) != 0;
#pragma unroll
                for (int k = 0; k < half; ++k) {
                    const double send = upper ? acc[k] : acc[k + half];
                    const double keep = upper ? acc[k + half] : acc[k];
                    acc[k] = keep + __shfl_xor_sync(kFull, send, half);
                }
            }
            if (DET) {
                if (lane < wc) S[w.A.signal_cnt + nt * 32 + lane] = acc[0];
            } else if (lane < wc)
                atomicAdd(zr, acc[0]);
            w.advance();  // nobody waits for the diagonal block: no fence, no signal
            continue;
        }
        const int ng = w.A.ntiles / ncc;
        int tg[kGroup];  // row targets: static data, fetched before the dependency wait
#pragma unroll
        for (int g = 0; g < kGroup; ++g) tg[g] = (g < ng) ? __ldg(w.A.rowidx + g * 32 + lane) : -1;
        if (w.A.wait_cnt >= 0) wait_counter(cnt + w.A.wait_cnt, w.A.wait_need, lane);
        const double *yp = WX + w.A.vec + lane;
        double yv[kGroup];
#pragma unroll
        for (int cc = 0; cc < kGroup; ++cc) {
            const int wn = (cc == ncc - 1) ? w.A.wlast : w.A.w;
            yv[cc] = (cc < ncc && lane < wn) ? __ldcg(yp + 32 * cc) : 0.0;
        }
        w.prefetch_task();
#pragma unroll
        for (int g = 0; g < kGroup; ++g) {
            w.idx[g * 32 + lane] = tg[g];
            w.vals[g * 32 + lane] = yv[g];
        }
        __syncwarp();
        for (int g = 0; g < ng; ++g) {
            double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
            for (int cc = 0; cc < ncc; ++cc) {
                const int wc = (cc == ncc - 1) ? w.A.wlast : w.A.w;
                const double *T = w.acquire_tile() + lane;
                const double *Y = w.vals + cc * 32;
                if (wc == 32) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        a0 = fma(T[j * 32], Y[j], a0);
                        a1 = fma(T[(j + 1) * 32], Y[j + 1], a1);
                        a2 = fma(T[(j + 2) * 32], Y[j + 2], a2);
                        a3 = fma(T[(j + 3) * 32], Y[j + 3], a3);
                    }
                } else {
                    for (int j = 0; j < wc; ++j) a0 = fma(T[j * 32], Y[j], a0);
                }
                w.release_tile();
            }
            const int tgt = w.idx[g * 32 + lane];
            if (DET) {
                if (tgt >= 0) S[tgt] = (a0 + a1) + (a2 + a3);  // row targets are slot numbers
            } else if (tgt >= 0)
                atomicAdd(WX + tgt, (a0 + a1) + (a2 + a3));
        }
        if (w.A.signal_cnt >= 0) {  // somebody waits for this task: publish the accumulations, then count
            __threadfence();
            __syncwarp();
            if (lane == 0) atomicAdd(cnt + w.A.signal_cnt, 1);
        }
        w.advance();
    }
}

// backward sweep: WX[cols] += panel(rows, cols)^T * x(rows)   (the panel holds -Lt)
// The row lists (static) are fetched before the dependency wait, all x gathers of the task are issued together
// after it and staged in shared memory.
// DET: compute tasks store their 32 partial sums into the slots at S[gstride ..]; assemble tasks (ntiles == 0), one per column
// chunk of every supernode, form x = (z slots of the forward sweep) + (slots of this supernode's backward tasks), in slot order.
template <bool DET>
__global__ void __launch_bounds__(kTriWarps * 32, 1) trisolve_bwd_kernel_t(const TriTask *__restrict__ tasks, int ntasks, double *WX,
                                                                            int *cnt, int *ticket, int xoff, double *__restrict__ S,
                                                                            const int32_t *__restrict__ lptr, const int32_t *__restrict__ lidx)
{
    extern __shared__ __align__(128) unsigned char tri_smem[];
    TriWarp w;
    tri_setup(w, tasks, ntasks, ticket, tri_smem);
    const int lane = w.lane;
    w.start();
    while (w.hasA) {
        const int nt = w.A.ntiles;
        const int wc = w.A.w;
        w.prefetch_ticket();
        if (DET && nt == 0) {
            if (w.A.wait_cnt >= 0) wait_counter(cnt + w.A.wait_cnt, w.A.wait_need, lane);
            w.prefetch_task();
            if (lane < wc) {
                const int p = w.A.vec + lane - xoff;  // list index = position inside X
                double x = 0.0;
                for (int q = lptr[p]; q < lptr[p + 1]; ++q) x += __ldcg(S + lidx[q]);
                WX[w.A.vec + lane] = x;
            }
            __threadfence();
            __syncwarp();
            if (lane == 0) atomicAdd(cnt + w.A.signal_cnt, 1);
            w.advance();
            continue;
        }
        int rr[kGroup];
#pragma unroll
        for (int t = 0; t < kGroup; ++t) rr[t] = (t < nt) ? __ldg(w.A.rowidx + t * 32 + lane) : -1;
        if (w.A.wait_cnt >= 0) wait_counter(cnt + w.A.wait_cnt, w.A.wait_need, lane);
        double xv[kGroup];
#pragma unroll
        for (int t = 0; t < kGroup; ++t) xv[t] = (rr[t] >= 0) ? __ldcg(WX + xoff + rr[t]) : 0.0;  // row lists hold W offsets
        w.prefetch_task();
#pragma unroll
        for (int t = 0; t < kGroup; ++t) w.vals[t * 32 + lane] = xv[t];
        __syncwarp();
        double acc[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = 0.0;
        for (int ti = 0; ti < nt; ++ti) {
            const double xr = w.vals[ti * 32 + lane];
            const double *T = w.acquire_tile() + lane;
            if (wc == 32) {
#pragma unroll
                for (int j = 0; j < 32; ++j) acc[j] = fma(T[j * 32], xr, acc[j]);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                    if (j < wc) acc[j] = fma(T[j * 32], xr, acc[j]);
            }
            w.release_tile();
        }
        // transpose-reduce: afterwards lane L holds the sum over lanes of acc[L]
#pragma unroll
        for (int half = 16; half >= 1; half >>= 1) {
            const bool upper = (lane & half) != 0;
#pragma unroll
            for (int k = 0; k < half; ++k) {
                const double send = upper ? acc[k] : acc[k + half];
                const double keep = upper ? acc[k + half] : acc[k];
                acc[k] = keep + __shfl_xor_sync(kFull, send, half);
            }
        }
        if (DET) {
            if (lane < wc) S[w.A.gstride + lane] = acc[0];
        } else if (lane < wc)
            atomicAdd(WX + w.A.vec + lane, acc[0]);
        if (w.A.signal_cnt >= 0) {  // somebody waits for this task: publish the accumulations, then count
            __threadfence();
            __syncwarp();
            if (lane == 0) atomicAdd(cnt + w.A.signal_cnt, 1);
        }
        w.advance();
    }
}

// the two instantiations of each sweep under the names the host code launches
#define trisolve_fwd_kernel trisolve_fwd_kernel_t<false>
#define trisolve_bwd_kernel trisolve_bwd_kernel_t<false>
#define trisolve_fwd_det_kernel trisolve_fwd_kernel_t<true>
#define trisolve_bwd_det_kernel trisolve_bwd_kernel_t<true>

// ---- setup-time conversion: column-major sources -> tiles --------------------------------------------------
// kind 0: tile = -(rows of the raw panel) * Linv  (Linv: w x w lower, column-major; small supernodes, w <= 32)
// kind 1: tile = copy of a column-major block (results of the cuBLAS path for wide supernodes, inverse diagonal blocks)
struct ConvTile {
    int64_t src;    // offset of element (first row of the tile, first column of the chunk) in its source array
    int64_t dst;    // offset in the tile array
    int64_t linv;   // kind 0: offset of Linv in the aux array
    int32_t ld;     // leading dimension of the source
    int32_t nrows;  // valid rows in this tile (<= 32)
    int16_t w;      // columns of the tile
    int16_t kind;
    int32_t srcsel;  // 0: raw panel array, 1: aux array
};

__global__ void __launch_bounds__(256) convert_tiles_kernel(int64_t ntiles, const ConvTile *__restrict__ tiles,
                                                            const double *__restrict__ raw, const double *__restrict__ aux,
                                                            double *__restrict__ out)
{
    const int64_t t = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (t >= ntiles) return;
    const ConvTile ct = tiles[t];
    const int w = ct.w;
    const double *src = (ct.srcsel ? aux : raw) + ct.src;
    if (ct.kind == 1) {
        for (int j = 0; j < w; ++j) out[ct.dst + (int64_t)j * 32 + lane] = (lane < ct.nrows) ? src[(int64_t)j * ct.ld + lane] : 0.0;
        return;
    }
    double p[32];
#pragma unroll
    for (int k = 0; k < 32; ++k) p[k] = (k < w && lane < ct.nrows) ? src[(int64_t)k * ct.ld + lane] : 0.0;
    const double *Li = aux + ct.linv;
    for (int j = 0; j < w; ++j) {
        double acc = 0.0;
#pragma unroll
        for (int k = 0; k < 32; ++k)
            if (k >= j && k < w) acc = fma(p[k], __ldg(Li + (int64_t)j * w + k), acc);
        out[ct.dst + (int64_t)j * 32 + lane] = -acc;
    }
}

}  // namespace ddm
