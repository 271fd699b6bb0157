// K3  batched subdomain solves  x_i = A_i^{-1} d_i  for all hosted subdomains, two persistent kernels (sm_100a).
//
// Replaces Dune::Cholmod::apply -> cholmod_l_solve (extern/dune-istl/dune/istl/cholmod.hh:263-306).
//
// Formulation.  At upload the supernodal LL^T factor is re-expressed, supernode by supernode, as a block LDL^T:
//     A = Lt D Lt^T,   Lt unit block lower triangular with  Lt_below(J) = L_below(J) L_JJ^{-1},   D_J = L_JJ L_JJ^T,
// and D_J is kept as its explicit inverse.  No triangular solve is left, and nothing inside a supernode is sequential:
//     forward  (tree order)    y_below -= Lt_below(J) y_J ;   z_J = D_J^{-1} y_J
//     backward (reverse order) x_J = z_J - Lt_below(J)^T x_below
// Each sweep reads every stored entry once: forward = the stacked panel [D_J^{-1}; -Lt_below(J)] times y_J, backward =
// the transposed below-panel times x_below.  Dependencies exist only between a supernode and its children in the
// elimination tree, so the critical path is the tree depth, not the number of columns.
//
// Layout.  A supernode with cw columns, Rd = ceil(cw/32) diagonal row tiles and Rb = ceil(nb/32) below row tiles is
// stored as (Rd+Rb) row tiles of 32 rows x cw columns; inside a row tile the columns are cut into chunks of 32, every
// chunk (a "tile": 32 rows x w<=32 columns) is column-major, so lane i owns row i and one column is 256 contiguous
// bytes.  Tile (r, c) starts at  base + r*32*cw + c*1024  doubles.  Tiles are moved by TMA bulk copies
// (cp.async.bulk, global -> shared, completion on an mbarrier) through a per-warp ring of kStages tiles, so the
// tile stream never waits for registers and runs ahead of the dependency waits.
//
// Scheduling.  Work is cut into tasks (forward: a block of row tiles x column chunks, up to kGroup tiles; backward: one column chunk
// x up to kGroup below row tiles), listed in elimination-tree level order over ALL subdomains.  Warps take tasks from
// a ticket counter and synchronise point-to-point through per-supernode counters; a task only ever waits for tasks
// with smaller ticket numbers, so the co-resident (cooperatively launched) grid cannot deadlock.
#pragma once
#include "kernels.cuh"

namespace ddm {

constexpr int kStages = 2;          // tiles in flight per warp
constexpr int kTriWarps = 11;       // warps per CTA (one CTA per SM): 11 x (2 x 8 KB ring + 3 KB operands) = 209 KB of shared memory
constexpr int kGroup = 8;           // tiles per task
constexpr int kTileDoubles = 1024;  // 32 x 32

struct alignas(16) TriTask {
    const double *tiles;    // first tile of the task
    const int32_t *rowidx;  // fwd: 32 target offsets in WX (-1 = padding); bwd: 32 offsets per tile, consecutive tiles
    int32_t vec;            // fwd: offset in WX of y for the first column chunk; bwd: offset in WX of the 32 output columns
    int32_t stride;         // doubles between consecutive tiles of one group
    int16_t ntiles;         // = groups x ncc
    int16_t ncc;            // tiles per group (fwd: column chunks per row tile; bwd: all tiles form one group)
    int16_t w;              // columns of every tile but the last of a group
    int16_t wlast;          // columns of the last tile of a group (fwd: trailing chunk of the supernode)
    int32_t wait_cnt;       // counter index to wait on (-1: none)
    int32_t wait_need;
    int32_t signal_cnt;     // counter to add 1 to on completion (-1: none)
    int32_t gstride;        // doubles between consecutive groups (fwd: row tiles)
};
static_assert(sizeof(TriTask) == 48, "TriTask layout");

// ---- PTX helpers: mbarrier + bulk async copy (TMA, non-tensor form) -----------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "WAIT_LOOP:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra.uni WAIT_DONE;\n"
                 "bra.uni WAIT_LOOP;\n"
                 "WAIT_DONE:\n"
                 "}\n" ::"r"(smem_u32(bar)), "r"(parity)
                 : "memory");
}
__device__ __forceinline__ void tma_load_1d(void *dst_smem, const void *src_gmem, unsigned bytes, uint64_t *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ int ld_acquire(const int *p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int ld_relaxed(const int *p)
{
    int v;
    asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// One lane polls with relaxed loads; a single acquire closes the wait (an acquire per poll invalidates L1 each time).
__device__ __forceinline__ void wait_counter(const int *cnt, int need, int lane)
{
    if (need > 0) {
        if (lane == 0) {
            while (ld_relaxed(cnt) < need) __nanosleep(20);
#ifndef DDM_RELAXED_WAIT
            (void)ld_acquire(cnt);
#endif
        }
        __syncwarp();
    }
}

}  // namespace ddm

#include "trisolve_stream.cuh"
#include "trisolve_kernels.cuh"
