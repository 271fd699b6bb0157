// The per-warp tile stream of the triangular-solve kernels (see trisolve.cuh for the design).
// A warp holds two tickets (the task it computes and the next one) and keeps kStages tile copies in flight over the
// concatenation of their tiles: the factor stream is static data, so it may run ahead of the dependency waits.
#pragma once

namespace ddm {

struct TriWarp {
    const TriTask *tasks;
    int ntasks;
    int *ticket;
    double *ring;    // kStages tiles of this warp in shared memory
    uint64_t *bars;  // kStages mbarriers of this warp
    double *vals;    // kGroup*32 doubles of this warp: the vector operand of the task (y chunks / gathered x)
    int32_t *idx;    // kGroup*32 ints of this warp: row targets of the task
    int lane;
    TriTask A, B, C;
    bool hasA, hasB;
    int tickC;       // lane 0: ticket requested ahead of time (-1: none pending)
    int issA, issB;
    unsigned head, tail;

    __device__ __forceinline__ int take_ticket()
    {
        int t = 0;
        if (lane == 0) t = atomicAdd(ticket, 1);
        return __shfl_sync(kFull, t, 0);
    }
    __device__ __forceinline__ static TriTask load_task(const TriTask *p)
    {
        TriTask t;
        const int4 *s = reinterpret_cast<const int4 *>(p);
        int4 *d = reinterpret_cast<int4 *>(&t);
        d[0] = __ldg(s);
        d[1] = __ldg(s + 1);
        d[2] = __ldg(s + 2);
        return t;
    }
    __device__ __forceinline__ void issue(const TriTask &tk, int ti)
    {
        if (lane == 0) {
            const unsigned slot = head % kStages;
            const int g = tk.ncc > 0 ? ti / tk.ncc : 0, cc = ti - g * tk.ncc;  // ncc == 0: one group (symmetric task)
            const int wc = (tk.ncc > 0 && cc == tk.ncc - 1) ? tk.wlast : tk.w;
            const unsigned bytes = 256u * (unsigned)wc;
            // no proxy fence: the slot's last generic-proxy reads completed before the __syncwarp in release_tile()
            mbar_expect_tx(bars + slot, bytes);
            tma_load_1d(ring + (size_t)slot * kTileDoubles, tk.tiles + (size_t)g * tk.gstride + (size_t)cc * tk.stride, bytes, bars + slot);
        }
        ++head;
    }
    __device__ __forceinline__ void topup()
    {
        while (head - tail < (unsigned)kStages) {
            if (hasA && issA < A.ntiles) issue(A, issA++);
            else if (hasB && issB < B.ntiles) issue(B, issB++);
            else break;
        }
    }
    // Requests the ticket after B (atomic in flight while the current task computes) ...
    __device__ __forceinline__ void prefetch_ticket()
    {
        tickC = -1;
        if (hasB && lane == 0) tickC = atomicAdd(ticket, 1);
    }
    // ... and later its descriptor (loads in flight until advance() needs them)
    __device__ __forceinline__ void prefetch_task()
    {
        tickC = __shfl_sync(kFull, tickC, 0);
        if (tickC >= 0 && tickC < ntasks) C = load_task(tasks + tickC);
    }
    __device__ __forceinline__ void start()
    {
        tickC = -1;
        head = tail = 0;
        issA = issB = 0;
        int t = take_ticket();
        hasA = t < ntasks;
        if (hasA) A = load_task(tasks + t);
        hasB = false;
        if (hasA) {
            t = take_ticket();
            hasB = t < ntasks;
            if (hasB) B = load_task(tasks + t);
        }
        topup();
    }
    __device__ __forceinline__ void advance()
    {
        const bool hadB = hasB;
        A = B;
        hasA = hasB;
        issA = issB;
        issB = 0;
        hasB = hadB && tickC >= 0 && tickC < ntasks;
        if (hasB) B = C;
        tickC = -1;
        topup();
    }
    // waits for the next tile of A; returns its shared-memory address
    __device__ __forceinline__ const double *acquire_tile()
    {
        const unsigned slot = tail % kStages;
        mbar_wait(bars + slot, (tail / kStages) & 1u);
        return ring + (size_t)slot * kTileDoubles;
    }
    __device__ __forceinline__ void release_tile()
    {
        __syncwarp();
        ++tail;
        topup();
    }
};

constexpr size_t kTriRingBytes = (size_t)kTriWarps * kStages * kTileDoubles * sizeof(double);
constexpr size_t kTriBarBytes = (size_t)kTriWarps * kStages * sizeof(uint64_t);
constexpr size_t kTriValBytes = (size_t)kTriWarps * kGroup * 32 * sizeof(double);
constexpr size_t kTriIdxBytes = (size_t)kTriWarps * kGroup * 32 * sizeof(int32_t);
constexpr size_t kTriSmemBytes = kTriRingBytes + kTriBarBytes + kTriValBytes + kTriIdxBytes;

__device__ __forceinline__ void tri_setup(TriWarp &w, const TriTask *tasks, int ntasks, int *ticket, unsigned char *smem)
{
    const int warp = threadIdx.x >> 5;
    w.lane = threadIdx.x & 31;
    w.tasks = tasks;
    w.ntasks = ntasks;
    w.ticket = ticket;
    w.ring = reinterpret_cast<double *>(smem) + (size_t)warp * kStages * kTileDoubles;
    w.bars = reinterpret_cast<uint64_t *>(smem + kTriRingBytes) + warp * kStages;
    w.vals = reinterpret_cast<double *>(smem + kTriRingBytes + kTriBarBytes) + warp * kGroup * 32;
    w.idx = reinterpret_cast<int32_t *>(smem + kTriRingBytes + kTriBarBytes + kTriValBytes) + warp * kGroup * 32;
    if (w.lane == 0) {
        for (int s = 0; s < kStages; ++s) mbar_init(w.bars + s, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
}

}  // namespace ddm
