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
#include "ige_common.cuh"

namespace {

__global__ void __launch_bounds__(IGE_STATS_ROWS * IGE_NUM_STATS)
stats_snapshot_kernel(float *__restrict__ rows, float *__restrict__ out, const int zero_rows, const int accumulate) {
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

}  // namespace

extern "C" int ige_stats_snapshot(const IgeStatsSnapshotArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    IGE_REQUIRE_PTR(a->stats);
    IGE_REQUIRE_PTR(a->out);
    stats_snapshot_kernel<<<1, IGE_STATS_ROWS * IGE_NUM_STATS, 0, (cudaStream_t)stream>>>(a->stats, a->out, a->zero_stats,
                                                                                           a->accumulate);
    return ige_check_launch("ige_stats_snapshot");
}
