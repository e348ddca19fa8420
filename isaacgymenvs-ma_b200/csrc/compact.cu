// compact.cu -- reset_buf.nonzero() as a single-pass, order-preserving stream compaction.
//
// Replaces `env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()` (tasks/ant.py:291,
// tasks/cartpole.py:168, tasks/humanoid.py:291) and, with group = A, the multi-agent AND filter
// `_agent_ids_to_env_ids` (tasks/franka_reach_MA.py:875-884): env e is emitted iff all `group`
// consecutive agent flags are set.  Output: ascending int32 ids (what gym.set_*_indexed wants),
// their count, and each env's exclusive rank (the row of its reset noise in the injected-uniform
// parity mode).  Integer work, bit-exact by construction.
//
// One launch, chained-scan with decoupled look-back: chunk order is taken from an atomic ticket so
// a waiting CTA's predecessors are always already running; the last CTA to finish zeroes the
// scratch again, so the launch is re-playable inside a CUDA graph with no memset node.
#include "ige_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kItems = 8;
constexpr int kChunk = kThreads * kItems;
constexpr uint32_t kFlagAgg = 1u << 30, kFlagPrefix = 2u << 30, kValMask = (1u << 30) - 1;

__global__ void __launch_bounds__(kThreads) compact_resets_kernel(IgeCompactArgs a, int num_chunks) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    uint32_t *scratch = a.scratch;
    __shared__ int s_chunk;
    __shared__ int s_warp[kThreads / 32];
    __shared__ int s_prefix;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_chunk = (int)atomicAdd(&scratch[0], 1u);
    __syncthreads();
    const int chunk = s_chunk;
    const int n = a.num_rows / a.group;
    const int base = chunk * kChunk + tid * kItems;

    // per-thread flags for kItems consecutive envs
    uint32_t bits = 0;
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const int e = base + i;
        if (e < n) {
            bool all = true;
            for (int g = 0; g < a.group; ++g) all &= (__ldg(a.reset_buf + (size_t)e * a.group + g) != 0);
            bits |= (all ? 1u : 0u) << i;
        }
    }
    const int mine = __popc(bits);
    // block exclusive scan of `mine`
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int warp_off = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) {
        const int c = s_warp[w];
        if (w < warp) warp_off += c;
        total += c;
    }
    const int local_excl = warp_off + incl - mine;

    // decoupled look-back (thread 0)
    if (tid == 0) {
        uint32_t prefix = 0;
        if (chunk > 0) {
            atomicExch(&scratch[2 + chunk], kFlagAgg | (uint32_t)total);
            int j = chunk - 1;
            while (true) {
                uint32_t s;
                do { s = atomicAdd(&scratch[2 + j], 0u); } while ((s >> 30) == 0u);
                prefix += s & kValMask;
                if ((s >> 30) == 2u) break;
                --j;
            }
        }
        atomicExch(&scratch[2 + chunk], kFlagPrefix | (prefix + (uint32_t)total));
        s_prefix = (int)prefix;
        if (chunk == num_chunks - 1) *a.count = (int)prefix + total;
    }
    __syncthreads();
    int rank = s_prefix + local_excl;
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        const int e = base + i;
        if (e < n) {
            if (a.reset_rank) a.reset_rank[e] = rank;
            if ((bits >> i) & 1u) a.reset_ids[rank++] = e;
        }
    }
    // retire: the last CTA restores the scratch to zero for the next launch
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        const uint32_t d = atomicAdd(&scratch[1], 1u);
        s_chunk = (d == (uint32_t)num_chunks - 1u);
    }
    __syncthreads();
    if (s_chunk) {
        for (int i = tid; i < num_chunks; i += kThreads) scratch[2 + i] = 0u;
        if (tid == 0) {
            scratch[0] = 0u;
            scratch[1] = 0u;
        }
    }
}

}  // namespace

// scratch words needed for `num_rows / group` envs
extern "C" int ige_compact_scratch_words(int32_t num_envs) {
    const int chunks = num_envs > 0 ? (num_envs + kChunk - 1) / kChunk : 1;   // an empty env set still launches one chunk
    return 2 + chunks;
}

extern "C" int ige_compact_resets(const IgeCompactArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    if (a->num_rows < 0 || a->group < 1 || a->num_rows % a->group != 0)
        return ige_fail(IGE_ESIZE, "num_rows=%d group=%d invalid", a->num_rows, a->group);
    if (!a->reset_buf || !a->reset_ids || !a->count || !a->scratch)
        return ige_fail(IGE_ENULL, "reset_buf / reset_ids / count / scratch must be non-NULL");
    const int n = a->num_rows / a->group;
    if (n > (1 << 30) - 1) return ige_fail(IGE_ESIZE, "too many envs");
    const int num_chunks = n == 0 ? 1 : (n + kChunk - 1) / kChunk;
ige_launch(compact_resets_kernel, num_chunks, kThreads, 0, (cudaStream_t)stream, *a, num_chunks);
    return ige_check_launch("ige_compact_resets");
}

extern "C" int ige_reset_list_scratch_words(int32_t num_envs) { return ige::reset_list_scratch_words(num_envs); }
