// stats.cu -- horizon snapshot of the episode-statistics accumulator.
//
// Every step kernel adds its per-CTA partial sums into one of IGE_STATS_ROWS striped 128-B rows.  Once per
// horizon the rows are folded into the <= 16-float vector that the NCCL all-reduce carries
// (utils/rlgames_utils.py:89-107: one process per GPU, per-rank statistics) and re-zeroed, so a horizon's
// numbers never mix with the next one's and the fp32 counters stay far below 2^24.
//
// The accumulator is double-buffered by the host (distributed.StatsReducer): horizon h adds into rows[h & 1]
// while this kernel -- a forked branch of horizon h+1's CUDA graph -- folds rows[h & 1 ^ 1]; it is therefore
// never on the step's critical path and never races with the step kernels' atomics.
//
// ige_stats_exchange is the multi-GPU form: the same fold fused with a sum all-reduce over NVLink peer memory.
#include "ige_common.cuh"

namespace {

static_assert(IGE_STATS_ROWS >= IGE_MAX_RANKS, "stats_exchange_kernel: one thread per (rank, column)");

__global__ void __launch_bounds__(IGE_STATS_ROWS * IGE_NUM_STATS)
stats_snapshot_kernel(float *__restrict__ rows, float *__restrict__ out, const int zero_rows, const int accumulate) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    __shared__ float s[IGE_STATS_ROWS][IGE_NUM_STATS + 1];
    const int r = threadIdx.x / IGE_NUM_STATS, c = threadIdx.x % IGE_NUM_STATS;
    s[r][c] = __ldcg(rows + r * IGE_STATS_STRIDE + c);   // atomics land in L2
    __syncthreads();
    if (zero_rows) rows[r * IGE_STATS_STRIDE + c] = 0.0f;
    if (r == 0) {
        float acc = 0.0f;   // fixed order: the snapshot is deterministic for a given set of row values
#pragma unroll
        for (int k = 0; k < IGE_STATS_ROWS; ++k) acc += s[k][c];
        out[c] = accumulate ? out[c] + acc : acc;
    }
}

// ---------------------------------------------------------------------------------------------
// fold + all-reduce over peer memory (see include/ige_b200.h).  thread t: column c = t % 16, rank slot s = t / 16.
// ---------------------------------------------------------------------------------------------
constexpr int kSlotWords = IGE_NUM_STATS;                       // 64-bit words per (parity, rank) slot
__device__ __forceinline__ void st_sys_u64(uint64_t *p, uint64_t v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_sys_u64(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t global_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

__global__ void __launch_bounds__(IGE_STATS_ROWS * IGE_NUM_STATS) stats_exchange_kernel(const __grid_constant__ IgeStatsExchangeArgs a) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    __shared__ float s[IGE_STATS_ROWS][IGE_NUM_STATS + 1];
    __shared__ float s_vec[IGE_NUM_STATS];
    __shared__ float s_peer[IGE_MAX_RANKS][IGE_NUM_STATS];
    __shared__ int s_missing;
    const int tid = threadIdx.x, r = tid / IGE_NUM_STATS, c = tid % IGE_NUM_STATS;
    // ---- fold (every thread owns one (row, column) of the accumulator) ----
    s[r][c] = __ldcg(a.stats + r * IGE_STATS_STRIDE + c);
    if (tid == 0) s_missing = 0;
    __syncthreads();
    if (a.zero_stats) a.stats[r * IGE_STATS_STRIDE + c] = 0.0f;
    if (r == 0) {
        float acc = 0.0f;
#pragma unroll
        for (int k = 0; k < IGE_STATS_ROWS; ++k) acc += s[k][c];
        s_vec[c] = acc;
        if (a.local) a.local[c] = acc;
    }
    __syncthreads();
    const uint64_t epoch = *a.epoch + 1ull;                       // 1, 2, ...: zero-filled memory never matches
    const uint32_t tag = (uint32_t)epoch;
    const size_t parity_off = (size_t)(epoch & 1ull) * a.world * kSlotWords;
    // ---- push: slot [parity][my rank] of EVERY rank's mailbox (r = destination rank) ----
    if (r < a.world)
        st_sys_u64(a.mailbox[r] + parity_off + (size_t)a.rank * kSlotWords + c,
                   ((uint64_t)tag << 32) | (uint64_t)__float_as_uint(s_vec[c]));
    // ---- wait: slot [parity][r] of the LOCAL mailbox (r = source rank) ----
    if (r < a.world) {
        const uint64_t *src = a.mailbox[a.rank] + parity_off + (size_t)r * kSlotWords + c;
        const uint64_t limit = (uint64_t)(a.timeout_ms ? a.timeout_ms : 2000u) * 1000000ull;
        uint64_t v = ld_sys_u64(src), t0 = 0;
        while ((uint32_t)(v >> 32) != tag) {
            if (t0 == 0) t0 = global_ns();
            else if (global_ns() - t0 > limit) { atomicMax(&s_missing, r + 1); v = 0; break; }
            __nanosleep(200);
            v = ld_sys_u64(src);
        }
        s_peer[r][c] = __uint_as_float((uint32_t)v);
    }
    __syncthreads();
    if (r == 0) {
        float acc = 0.0f;                                         // rank order: the same bits on every rank
        for (int k = 0; k < a.world; ++k) acc += s_peer[k][c];
        a.out[c] = acc;
    }
    if (tid == 0) {
        *a.epoch = epoch;
        if (s_missing && a.error) *a.error = s_missing;
    }
}

}  // namespace

extern "C" uint64_t ige_mailbox_bytes(int32_t world) { return (uint64_t)2 * (world > 0 ? world : 1) * kSlotWords * sizeof(uint64_t); }

extern "C" int ige_mailbox_create(int32_t world, void **mailbox, unsigned char handle[IGE_IPC_HANDLE_BYTES]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == IGE_IPC_HANDLE_BYTES, "IPC handle size");
    if (!mailbox || !handle) return ige_fail(IGE_ENULL, "ige_mailbox_create: NULL argument");
    if (world < 1 || world > IGE_MAX_RANKS) return ige_fail(IGE_ESIZE, "ige_mailbox_create: world=%d outside [1, %d]", world, IGE_MAX_RANKS);
    void *p = nullptr;
    cudaError_t e = cudaMalloc(&p, ige_mailbox_bytes(world));
    if (e == cudaSuccess) e = cudaMemset(p, 0, ige_mailbox_bytes(world));
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) {
        if (p) cudaFree(p);
        cudaGetLastError();
        return ige_fail((int)e, "ige_mailbox_create: %s", cudaGetErrorString(e));
    }
    for (int i = 0; i < IGE_IPC_HANDLE_BYTES; ++i) handle[i] = reinterpret_cast<unsigned char *>(&h)[i];
    *mailbox = p;
    return IGE_OK;
}
extern "C" int ige_mailbox_open(const unsigned char handle[IGE_IPC_HANDLE_BYTES], void **mailbox) {
    if (!mailbox || !handle) return ige_fail(IGE_ENULL, "ige_mailbox_open: NULL argument");
    cudaIpcMemHandle_t h;
    for (int i = 0; i < IGE_IPC_HANDLE_BYTES; ++i) reinterpret_cast<unsigned char *>(&h)[i] = handle[i];
    void *p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return ige_fail((int)e, "ige_mailbox_open: %s", cudaGetErrorString(e));
    }
    *mailbox = p;
    return IGE_OK;
}
extern "C" int ige_mailbox_close(void *peer_mailbox) {
    cudaError_t e = cudaIpcCloseMemHandle(peer_mailbox);
    if (e != cudaSuccess) { cudaGetLastError(); return ige_fail((int)e, "ige_mailbox_close: %s", cudaGetErrorString(e)); }
    return IGE_OK;
}
extern "C" int ige_mailbox_destroy(void *own_mailbox) {
    cudaError_t e = cudaFree(own_mailbox);
    if (e != cudaSuccess) { cudaGetLastError(); return ige_fail((int)e, "ige_mailbox_destroy: %s", cudaGetErrorString(e)); }
    return IGE_OK;
}

extern "C" int ige_stats_exchange(const IgeStatsExchangeArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    IGE_REQUIRE_PTR(a->stats);
    IGE_REQUIRE_PTR(a->out);
    IGE_OPTIONAL_PTR(a->local);
    if (!a->epoch) return ige_fail(IGE_ENULL, "epoch is NULL");
    if (a->world < 1 || a->world > IGE_MAX_RANKS || a->rank < 0 || a->rank >= a->world)
        return ige_fail(IGE_ESIZE, "ige_stats_exchange: rank=%d world=%d invalid (max %d ranks)", a->rank, a->world, IGE_MAX_RANKS);
    for (int r = 0; r < a->world; ++r)
        if (!a->mailbox[r] || (reinterpret_cast<uintptr_t>(a->mailbox[r]) & 7u))
            return ige_fail(IGE_ENULL, "ige_stats_exchange: mailbox[%d] is NULL / misaligned", r);
ige_launch(stats_exchange_kernel, 1, IGE_STATS_ROWS * IGE_NUM_STATS, 0, (cudaStream_t)stream, *a);
    return ige_check_launch("ige_stats_exchange");
}

extern "C" int ige_stats_snapshot(const IgeStatsSnapshotArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    IGE_REQUIRE_PTR(a->stats);
    IGE_REQUIRE_PTR(a->out);
ige_launch(stats_snapshot_kernel, 1, IGE_STATS_ROWS * IGE_NUM_STATS, 0, (cudaStream_t)stream, a->stats, a->out, a->zero_stats,
                                                                                           a->accumulate);
    return ige_check_launch("ige_stats_snapshot");
}
