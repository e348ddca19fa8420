// task_common.cuh -- thread-per-env skeleton of the smaller single-agent tasks.
//
// These tasks move a few hundred bytes per env with a handful of flops; what they share is the reset plumbing:
// reset_buf.nonzero() -> ascending int32 id list for gym.set_*_tensor_indexed, per-env reset noise (Philox in
// production, injected uniforms by reset rank for parity with the reference's torch.rand stream), a step word that
// advances once per launch.  The skeleton gives every task that, with the same protocol as the big kernels
// (ige::list_publish_slot / list_scanner_cta): workers publish ballots, one extra CTA writes the ids.
#pragma once

#include "ige_common.cuh"

namespace ige {

constexpr int kTaskThreads = 128;

// uniform `j` of torch.rand draw `draw` (a (k, cols) tensor, rows by reset rank) for env `env`
__device__ __forceinline__ float task_uniform(const IgeTaskRand &r, uint64_t step, int draw, int env, int cols, int j) {
    if (r.mode == IGE_RAND_PHILOX) {
        uint32_t c[4] = {(uint32_t)env, ((uint32_t)draw << 16) | ((uint32_t)j >> 2), (uint32_t)step, (uint32_t)(step >> 32)};
        philox4x32_10(c, (uint32_t)r.seed, (uint32_t)(r.seed >> 32));
        return (float)(c[j & 3] >> 8) * (1.0f / 16777216.0f);
    }
    return __ldg(r.u[draw] + (size_t)__ldg(r.rank + env) * cols + j);
}
// same, with an explicit rank array (tasks with a second id list, e.g. Ingenuity's target re-draws)
__device__ __forceinline__ float task_uniform_ranked(const IgeTaskRand &r, const int32_t *rank, uint64_t step, int draw,
                                                     int env, int cols, int j) {
    if (r.mode == IGE_RAND_PHILOX) {
        uint32_t c[4] = {(uint32_t)env, ((uint32_t)draw << 16) | ((uint32_t)j >> 2), (uint32_t)step, (uint32_t)(step >> 32)};
        philox4x32_10(c, (uint32_t)r.seed, (uint32_t)(r.seed >> 32));
        return (float)(c[j & 3] >> 8) * (1.0f / 16777216.0f);
    }
    return __ldg(r.u[draw] + (size_t)__ldg(rank + env) * cols + j);
}
// torch_rand_float (utils/torch_jit_utils.py:216-218): (upper - lower) * rand + lower with the Python-double span and
// lower bound rounded to fp32 where they meet the tensor
__device__ __forceinline__ float rand_span(double lower, double upper, float u) {
    return (float)(upper - lower) * u + (float)lower;
}

// Body of a resetting kernel.  `env(e, flagged, step)` does the per-env work of the launch for env e (flagged: its
// reset flag was set at entry); the skeleton owns the flag loads, the reset list and the step word.
//   reset_buf[N] int64 in; list optional.  Grid = `workers` CTAs that stride over 128-env chunks (+ one scanner CTA
//   when the list is requested, ige::list_scanner_cta: the workers only publish their warps' ballots).
//   `tail(total)`: run by every thread of the scanner CTA once the list is complete (total = number of resets) --
//   for the rare cross-env side effect (Ingenuity re-sets the rotor speed of ALL envs when any env resets).
struct NoTail { __device__ void operator()(int) const {} };
//   CHUNK_ANY: env() additionally receives "some env of this CTA's current 128-env chunk is flagged" (a CTA-wide OR).
template <bool CHUNK_ANY = false, class EnvFn, class TailFn = NoTail>
__device__ __forceinline__ void task_reset_skeleton(int num_envs, int workers, const int64_t *reset_buf,
                                                    const IgeResetList &L, const IgeTaskRand &rand, EnvFn env,
                                                    TailFn tail = TailFn()) {
    if ((int)blockIdx.x >= workers) {
        pdl_trigger();
        tail(list_scanner_cta<kTaskThreads>(L, (num_envs + 31) / 32, rand.step));
        return;
    }
    __shared__ uint64_t s_step;
    __shared__ uint32_t s_tag;
    const int tid = threadIdx.x, lane = tid & 31;
    const bool emit = L.scratch != nullptr;
    if (tid == 0) {
        s_step = (rand.mode == IGE_RAND_PHILOX && rand.step) ? *reinterpret_cast<const volatile uint64_t *>(rand.step) : 0ull;
        s_tag = emit ? list_epoch_tag(L) : 0u;
    }
    __syncthreads();   // epoch and step word were READ before this CTA's first publish
    const uint64_t step = s_step;
    const uint32_t tag = s_tag;
    for (int e0 = blockIdx.x * kTaskThreads; e0 < num_envs; e0 += workers * kTaskThreads) {
        const int e = e0 + tid;
        const bool f = e < num_envs && reset_buf[e] != 0;
        const uint32_t m = __ballot_sync(0xffffffffu, f);
        if (emit && lane == 0 && e < num_envs) list_publish_slot(L, tag, e >> 5, m);
        if constexpr (CHUNK_ANY) {
            const bool any = __syncthreads_or(f ? 1 : 0) != 0;
            if (e < num_envs) env(e, f, step, any);
        } else {
            if (e < num_envs) env(e, f, step);
        }
    }
    pdl_trigger();   // main loop done: the next kernel's grid may be scheduled behind this one's tail
}

}  // namespace ige

// host: grid of the skeleton -- workers <= resident CTAs - 1 (occupancy query cached per kernel and device)
template <class K>
static inline int ige_task_grid(K kernel, int num_envs, int *workers, const char *name) {
    static int cap[IGE_MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= IGE_MAX_DEVICES) return ige_fail(IGE_EPARAM, "%s: device ordinal %d not supported", name, dev);
    if (cap[dev] == 0) {
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, ige::kTaskThreads, 0);
        if (e != cudaSuccess || per_sm < 1 || sms < 1)
            return ige_fail(e != cudaSuccess ? (int)e : IGE_EPARAM, "%s: occupancy query failed", name);
        cap[dev] = sms * per_sm > 1 ? sms * per_sm - 1 : 1;
    }
    const int chunks = (num_envs + ige::kTaskThreads - 1) / ige::kTaskThreads;
    const int per = (chunks + cap[dev] - 1) / cap[dev];
    *workers = per > 0 ? (chunks + per - 1) / per : 1;
    if (*workers < 1) *workers = 1;
    return IGE_OK;
}
