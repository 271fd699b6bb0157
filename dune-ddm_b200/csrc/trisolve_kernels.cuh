// The two sweeps of the batched subdomain solve (see trisolve.cuh for the design).
#pragma once

namespace ddm {

// forward sweep: WX[rows] += panel(rows, chunks) * y(chunks)   (the panel holds D^{-1} rows and -Lt rows)
__global__ void __launch_bounds__(kTriWarps * 32, 1) trisolve_fwd_kernel(const TriTask *__restrict__ tasks, int ntasks, double *WX,
                                                                          int *cnt, int *ticket)
{
    extern __shared__ __align__(128) unsigned char tri_smem[];
    TriWarp w;
    tri_setup(w, tasks, ntasks, ticket, tri_smem);
    const int lane = w.lane;
    w.start();
    while (w.hasA) {
        const int ncc = w.A.ncc, ng = w.A.ntiles / ncc;
        int tgt = __ldg(w.A.rowidx + lane);
        if (w.A.wait_cnt >= 0) wait_counter(cnt + w.A.wait_cnt, w.A.wait_need, lane);
        const double *yp = WX + w.A.vec + lane;
        double ynext = (lane < (ncc == 1 ? w.A.wlast : w.A.w)) ? __ldcg(yp) : 0.0;
        for (int g = 0; g < ng; ++g) {
            double acc = 0.0;
            for (int cc = 0; cc < ncc; ++cc) {
                const double yl = ynext;
                {   // y of the next tile (wraps to the first chunk for the next row tile)
                    const int cn = (cc + 1 == ncc) ? 0 : cc + 1;
                    const int wn = (cn == ncc - 1) ? w.A.wlast : w.A.w;
                    ynext = (lane < wn) ? __ldcg(yp + 32 * cn) : 0.0;
                }
                const int wc = (cc == ncc - 1) ? w.A.wlast : w.A.w;
                const double *T = w.acquire_tile() + lane;
                if (wc == 32) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) acc = fma(T[j * 32], __shfl_sync(kFull, yl, j), acc);
                } else {
#pragma unroll 4
                    for (int j = 0; j < wc; ++j) acc = fma(T[j * 32], __shfl_sync(kFull, yl, j), acc);
                }
                w.release_tile();
            }
            const int tnext = (g + 1 < ng) ? __ldg(w.A.rowidx + (g + 1) * 32 + lane) : -1;
            if (tgt >= 0) atomicAdd(WX + tgt, acc);
            tgt = tnext;
        }
        __threadfence();
        __syncwarp();
        if (lane == 0 && w.A.signal_cnt >= 0) atomicAdd(cnt + w.A.signal_cnt, 1);
        w.advance();
    }
}

// backward sweep: WX[cols] += panel(rows, cols)^T * x(rows)   (the panel holds -Lt)
__global__ void __launch_bounds__(kTriWarps * 32, 1) trisolve_bwd_kernel(const TriTask *__restrict__ tasks, int ntasks, double *WX,
                                                                          int *cnt, int *ticket, int xoff)
{
    extern __shared__ __align__(128) unsigned char tri_smem[];
    TriWarp w;
    tri_setup(w, tasks, ntasks, ticket, tri_smem);
    const int lane = w.lane;
    w.start();
    while (w.hasA) {
        const int nt = w.A.ntiles;
        const int wc = w.A.w;
        int rnext = __ldg(w.A.rowidx + lane);
        if (w.A.wait_cnt >= 0) wait_counter(cnt + w.A.wait_cnt, w.A.wait_need, lane);
        double xnext = (rnext >= 0) ? __ldcg(WX + xoff + rnext) : 0.0;  // row lists hold W offsets; x lives xoff further
        double acc[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = 0.0;
        for (int ti = 0; ti < nt; ++ti) {
            const double xr = xnext;
            if (ti + 1 < nt) {
                rnext = __ldg(w.A.rowidx + (ti + 1) * 32 + lane);
                xnext = (rnext >= 0) ? __ldcg(WX + xoff + rnext) : 0.0;
            }
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
        if (lane < wc) atomicAdd(WX + w.A.vec + lane, acc[0]);
        __threadfence();
        __syncwarp();
        if (lane == 0 && w.A.signal_cnt >= 0) atomicAdd(cnt + w.A.signal_cnt, 1);
        w.advance();
    }
}

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
