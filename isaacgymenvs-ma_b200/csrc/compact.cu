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
#include <mutex>

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

// ---------------------------------------------------------------------------------------------
// ige_compact_resets_static -- the same compaction for callers that have no way to hand us scratch words.
//
// Who that is: a plug-in running inside a host process whose device memory belongs to somebody else.  A simulator
// that sized its GPU heap at start-up, a trainer that gave the whole card to a caching allocator with a hard cap, or a
// binding written in a language with no CUDA runtime of its own only owns the tensors it was handed: the flags going
// in and the id list coming out.  It cannot call cudaMalloc for a few look-back words (no budget left, no runtime
// handle, or a control loop that cannot afford cudaMalloc's device-wide synchronisation), and ige_compact_resets
// insists on a caller-owned scratch buffer.  This variant keeps the look-back words in the library image instead:
// a file-scope __device__ array, present once on every GPU that loads the library, from load until the process exits
// (kStaticWords * 4 = 8200 bytes).  The price is a size limit (IGE_COMPACT_STATIC_MAX_ENVS envs) and that every call
// on a device goes through the one array: two calls whose kernels can be on the GPU at the same time (different
// streams, different host threads) corrupt each other's prefix sums.  Serialise them, or use ige_compact_resets.
//
// The array's C++ name is only a handle on the host side; the kernel needs the device address behind it, which
// cudaGetSymbolAddress returns for the CURRENT device.  It is looked up the first time the entry point runs on a
// device and remembered per device ordinal.  That first call is therefore a set-up call (it may make the runtime
// load the module): make it once per device, outside any stream capture, before the calls that matter.
// Nothing here calls cudaMalloc / cudaFree or any library routine that allocates.
// ---------------------------------------------------------------------------------------------
namespace {
constexpr int kStaticMaxEnvs = IGE_COMPACT_STATIC_MAX_ENVS;
constexpr int kStaticMaxChunks = (kStaticMaxEnvs + kChunk - 1) / kChunk;
constexpr int kStaticWords = 2 + kStaticMaxChunks;   // ticket, retirement counter, one look-back slot per chunk
constexpr int kStaticMaxDevices = 64;
}  // namespace

static __device__ __align__(256) uint32_t g_compact_static_scratch[kStaticWords];

static std::mutex g_static_scratch_mu;
static uint32_t *g_static_scratch_addr[kStaticMaxDevices] = {};   // device address of the array, per device ordinal

extern "C" int ige_compact_resets_static(const IgeCompactArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    if (a->num_rows < 0 || a->group < 1 || a->num_rows % a->group != 0)
        return ige_fail(IGE_ESIZE, "num_rows=%d group=%d invalid", a->num_rows, a->group);
    if (!a->reset_buf || !a->reset_ids || !a->count)
        return ige_fail(IGE_ENULL, "reset_buf / reset_ids / count must be non-NULL");
    const int n = a->num_rows / a->group;
    if (n > kStaticMaxEnvs)
        return ige_fail(IGE_ESIZE,
                        "ige_compact_resets_static: %d envs exceed the built-in scratch (IGE_COMPACT_STATIC_MAX_ENVS = %d); "
                        "use ige_compact_resets with caller-owned scratch",
                        n, kStaticMaxEnvs);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { cudaGetLastError(); return ige_fail((int)e, "cudaGetDevice: %s", cudaGetErrorString(e)); }
    if (dev < 0 || dev >= kStaticMaxDevices)
        return ige_fail(IGE_EPARAM, "ige_compact_resets_static: device ordinal %d outside [0, %d)", dev, kStaticMaxDevices);
    uint32_t *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_static_scratch_mu);
        scratch = g_static_scratch_addr[dev];
        if (!scratch) {   // first call on this device: resolve the symbol while this device is current
            e = cudaGetSymbolAddress((void **)&scratch, g_compact_static_scratch);
            if (e != cudaSuccess || !scratch) {
                cudaGetLastError();
                return ige_fail(e != cudaSuccess ? (int)e : IGE_EPARAM, "cudaGetSymbolAddress(compact scratch) on device %d: %s",
                                dev, cudaGetErrorString(e));
            }
            g_static_scratch_addr[dev] = scratch;
        }
    }
    const int num_chunks = n == 0 ? 1 : (n + kChunk - 1) / kChunk;
    // The array still holds whatever the previous call left.  The kernel hands it back zeroed, but a launch that was
    // cut short would not have; the words this launch reads must be zero, so clear exactly those, in stream order.
    e = cudaMemsetAsync(scratch, 0, sizeof(uint32_t) * (size_t)(2 + num_chunks), (cudaStream_t)stream);
    if (e != cudaSuccess) { cudaGetLastError(); return ige_fail((int)e, "cudaMemsetAsync(compact scratch): %s", cudaGetErrorString(e)); }
    IgeCompactArgs k = *a;
    k.scratch = scratch;   // whatever the caller put there is ignored
    // plain stream-ordered launch (no programmatic dependent launch): the kernel must not start before the clear above
    ige_launch_opt(false, compact_resets_kernel, num_chunks, kThreads, 0, (cudaStream_t)stream, k, num_chunks);
    return ige_check_launch("ige_compact_resets_static");
}
