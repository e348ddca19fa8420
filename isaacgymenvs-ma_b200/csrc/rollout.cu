// rollout.cu -- agent-side per-step bookkeeping of A2CAgentMA.play_steps.
//
// ige_ma_rollout_step replaces learning/A2CAgent_MA.py:35-56 with ONE launch: rewards shaper,
// time-out value bootstrap, experience-buffer rows, current_* accumulate and done masking,
// env_done_indices = dones.nonzero()[::num_agents] (ascending ROW ids) and the three AverageMeter
// updates -- sync-free (the reference reads len(env_done_indices) on the host).
//
// HBM-bound, 45 B/row (R rewards 4 + dones 8 + time_outs 1 + values 4 + current_* 12; W mb_rewards 4
// + dones_out 1 + current_* 12 minus nothing re-read) -> thread owns 4 consecutive rows so every stream
// moves as 16-B (4-B for the byte streams) vectors.  The `[::A]` stride is over the LIST of done rows, so
// a row is kept iff its rank in nonzero() is a multiple of A: single-pass chained scan with decoupled,
// warp-wide look-back over chunk counts; all element-wise stores are issued before
// the look-back so they overlap it.  Meter sums: per-CTA fp32 partials -> fp64 atomics; the CTA that
// retires last applies the AverageMeter recurrence and restores the scratch (graph re-playable).
//
// ige_ma_discount_values: rl_games' A2CBase.discount_values (GAE) as called from A2CAgent_MA.py:60-67.
#include "ige_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kItems = 4;  // rows per thread: one 16-B vector per float stream, two per int64 stream, 4 B per byte stream
constexpr int kChunk = kThreads * kItems;
constexpr uint32_t kFlagAgg = 1u << 30, kFlagPrefix = 2u << 30, kValMask = (1u << 30) - 1;

__device__ __forceinline__ void load4(const float *p, float (&v)[kItems]) {
    const float4 a = __ldg(reinterpret_cast<const float4 *>(p));
    v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w;
}
__device__ __forceinline__ void load4_rw(const float *p, float (&v)[kItems]) {  // buffers this kernel rewrites
    const float4 a = *reinterpret_cast<const float4 *>(p);
    v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w;
}
__device__ __forceinline__ void store4(float *p, const float (&v)[kItems]) {
    *reinterpret_cast<float4 *>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ uint32_t poll(const uint32_t *p) { return *reinterpret_cast<const volatile uint32_t *>(p); }

// Chunk c is CTA c (CTAs are dispatched in block-index order, so a CTA's predecessors are always running or done --
// the same assumption as the reset list of the locomotion kernel); slots: scratch[2 + c], retirement counter: scratch[1].
__global__ void __launch_bounds__(kThreads) ma_rollout_step_kernel(IgeRolloutStepArgs a, int num_chunks, int vec, int use_ticket) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    uint32_t *scratch = a.scratch;
    __shared__ int s_last;
    __shared__ int s_warp[kThreads / 32];
    __shared__ int s_prefix;
    __shared__ float s_sum[kThreads / 32][3];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n = a.num_rows;
    // meter state: fetched now by the three threads that may need it at the very end (hides a cold DRAM round trip)
    float mt_mean = 0.0f;
    int mt_size = 0;
    if (tid < 3) mt_mean = a.meter_mean[tid], mt_size = a.meter_size[tid];
    // The look-back below only ever waits for a chunk with a LOWER index, so chunk indices must be handed out in an order
    // that guarantees those chunks are running or done, whatever order the hardware dispatches CTAs in: when the grid fits
    // on the device at once (the launcher checks) chunk = block index is safe and free; otherwise chunks are taken from an
    // atomic ticket (one L2 round trip per CTA, hidden by the oversubscribed grid).
    __shared__ int s_chunk;
    if (use_ticket) {
        if (tid == 0) s_chunk = (int)atomicAdd(&scratch[0], 1u);
        __syncthreads();
    }
    for (int chunk = use_ticket ? s_chunk : (int)blockIdx.x; chunk < num_chunks; chunk += use_ticket ? num_chunks : (int)gridDim.x) {
    const int base = chunk * kChunk + tid * kItems;
    const bool full = vec && base + kItems <= n;  // vec: every stream is 16-B aligned (host-checked)
    const bool boot = a.value_bootstrap != 0;

    float rew[kItems], val[kItems], cr[kItems], cs[kItems], cl[kItems];
    long long dn[kItems];
    uint8_t to[kItems];
    if (full) {
        load4(a.rewards + base, rew);
        load4_rw(a.current_rewards + base, cr);
        load4_rw(a.current_shaped_rewards + base, cs);
        load4_rw(a.current_lengths + base, cl);
#pragma unroll
        for (int i = 0; i < kItems; i += 2) {
            const longlong2 d = __ldg(reinterpret_cast<const longlong2 *>(a.dones + base + i));
            dn[i] = d.x, dn[i + 1] = d.y;
        }
        if (boot) {
            load4(a.values + base, val);
            const uint32_t t = __ldg(reinterpret_cast<const uint32_t *>(a.time_outs + base));
#pragma unroll
            for (int i = 0; i < kItems; ++i) to[i] = (t >> (8 * i)) & 0xffu;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kItems; ++i) {
            const int r = base + i;
            const bool ok = r < n;
            rew[i] = ok ? __ldg(a.rewards + r) : 0.0f;
            cr[i] = ok ? a.current_rewards[r] : 0.0f;
            cs[i] = ok ? a.current_shaped_rewards[r] : 0.0f;
            cl[i] = ok ? a.current_lengths[r] : 0.0f;
            dn[i] = ok ? __ldg(a.dones + r) : 0;
            val[i] = (ok && boot) ? __ldg(a.values + r) : 0.0f;
            to[i] = (ok && boot) ? __ldg(a.time_outs + r) : 0;
        }
    }

    // element-wise part (:34-42, :54-56); the pre-mask current_* stay in registers for the meters (:47-49)
    float shaped[kItems], m_cr[kItems], m_cs[kItems], m_cl[kItems];
    uint32_t bits = 0;
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        float s = rew[i] + a.shaper_shift;  // DefaultRewardsShaper.__call__
        s = s * a.shaper_scale;
        s = fminf(fmaxf(s, a.shaper_min), a.shaper_max);
        if (boot) {  // :36-37  shaped += gamma * values * time_outs.float()
            const float gv = a.gamma * val[i];
            s = s + gv * (float)to[i];
        }
        shaped[i] = s;
        cr[i] = cr[i] + rew[i];  // :41-43
        cs[i] = cs[i] + s;
        cl[i] = cl[i] + 1.0f;
        const float nd = 1.0f - (float)dn[i];  // :52
        m_cr[i] = cr[i] * nd;                  // :54-56
        m_cs[i] = cs[i] * nd;
        m_cl[i] = cl[i] * nd;
        bits |= (dn[i] != 0 ? 1u : 0u) << i;
    }
    if (full) {
        store4(a.mb_rewards + base, shaped);
        store4(a.current_rewards + base, m_cr);
        store4(a.current_shaped_rewards + base, m_cs);
        store4(a.current_lengths + base, m_cl);
        if (a.dones_out) {
            uint32_t o = 0u;
#pragma unroll
            for (int i = 0; i < kItems; ++i) o |= (uint32_t)(uint8_t)dn[i] << (8 * i);
            *reinterpret_cast<uint32_t *>(a.dones_out + base) = o;
        }
    } else {
#pragma unroll
        for (int i = 0; i < kItems; ++i) {
            const int r = base + i;
            if (r < n) {
                a.mb_rewards[r] = shaped[i];
                a.current_rewards[r] = m_cr[i];
                a.current_shaped_rewards[r] = m_cs[i];
                a.current_lengths[r] = m_cl[i];
                if (a.dones_out) a.dones_out[r] = (uint8_t)dn[i];
            }
        }
    }

    // rank of every done row in dones.nonzero(): block scan + decoupled look-back over chunk counts
    const int mine = __popc(bits);
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
    if (warp == 0) {  // warp-wide look-back: lane l polls predecessor (window start - l)
        uint32_t prefix = 0;
        if (chunk > 0) {
            if (lane == 0) atomicExch(&scratch[2 + chunk], kFlagAgg | (uint32_t)total);
            for (int j0 = chunk - 1;; j0 -= 32) {
                const int j = j0 - lane;
                uint32_t s = kFlagPrefix;  // before chunk 0: an inclusive prefix of zero
                if (j >= 0)
                    do { s = poll(&scratch[2 + j]); } while ((s >> 30) == 0u);
                const unsigned pm = __ballot_sync(0xffffffffu, (s >> 30) == 2u);
                const int first = pm ? __ffs(pm) - 1 : 32;
                uint32_t v = lane <= first ? (s & kValMask) : 0u;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                prefix += v;
                if (pm) break;
            }
        }
        if (lane == 0) {
            atomicExch(&scratch[2 + chunk], kFlagPrefix | (prefix + (uint32_t)total));
            s_prefix = (int)prefix;
            if (chunk == num_chunks - 1)  // :45  len(all_done_indices[::A]) = ceil(total / A)
                *a.done_count = ((int)prefix + total + a.num_agents - 1) / a.num_agents;
        }
    }
    __syncthreads();
    int rank = s_prefix + local_excl;
    float p0 = 0.0f, p1 = 0.0f, p2 = 0.0f;
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
        if ((bits >> i) & 1u) {
            if (rank % a.num_agents == 0) {  // :45 all_done_indices[::num_agents]
                a.done_rows[rank / a.num_agents] = base + i;
                p0 += cr[i], p1 += cs[i], p2 += cl[i];
            }
            ++rank;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        p0 += __shfl_xor_sync(0xffffffffu, p0, o);
        p1 += __shfl_xor_sync(0xffffffffu, p1, o);
        p2 += __shfl_xor_sync(0xffffffffu, p2, o);
    }
    if (lane == 0) s_sum[warp][0] = p0, s_sum[warp][1] = p1, s_sum[warp][2] = p2;
    __syncthreads();
    if (tid < 3 && total > 0) {  // three lanes, one fp64 atomic each
        double t = 0.0;
        for (int w = 0; w < kThreads / 32; ++w) t += s_sum[w][tid];
        atomicAdd(&a.reduce_scratch[tid], t);
    }
    if (tid < 32) {
        __threadfence();
        __syncwarp();
        if (tid == 0) s_last = (atomicAdd(&scratch[1], 1u) == (uint32_t)num_chunks - 1u);
    }
    __syncthreads();
    if (s_last) {  // last CTA to retire: AverageMeter.update x3 (:47-49), restore the scratch
        __threadfence();
        if (tid < 3) {
            const int k = __ldcg(a.done_count);
            const double sum = __ldcg(&a.reduce_scratch[tid]);
            if (k > 0) {
                const float new_mean = (float)sum / (float)k;  // torch.mean(values.float(), dim=0)
                const int size = min(k, a.meter_max_size);     // np.clip(size, 0, max_size)
                const int old_size = min(a.meter_max_size - size, mt_size);
                const int size_sum = old_size + size;
                a.meter_size[tid] = size_sum;
                const float x = mt_mean * (float)old_size, y = new_mean * (float)size;
                a.meter_mean[tid] = (x + y) / (float)size_sum;
            }
            a.reduce_scratch[tid] = 0.0;
        }
        for (int i = tid; i < num_chunks; i += kThreads) scratch[2 + i] = 0u;
        if (tid == 0) { scratch[0] = 0u; scratch[1] = 0u; }
    }
    __syncthreads();   // the shared scratch is reused by the next chunk
    }
}

// thread per row, backwards over the horizon; every (t, rows) slab is read/written coalesced
__global__ void __launch_bounds__(256) ma_discount_values_kernel(IgeDiscountArgs a) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= a.num_rows) return;
    const size_t n = (size_t)a.num_rows;
    float lastgaelam = 0.0f;
    float nextvalues = __ldg(a.last_values + r);
    float nextnonterminal = 1.0f - (float)__ldg(a.dones + r);
    for (int t = a.horizon - 1; t >= 0; --t) {
        const size_t i = (size_t)t * n + r;
        const float v = __ldg(a.mb_values + i), rw = __ldg(a.mb_rewards + i);
        const float d = (float)__ldg(a.mb_dones + i);
        const float gv = a.gamma * nextvalues;       // delta = r + gamma * nextvalues * nextnonterminal - v
        const float delta = (rw + gv * nextnonterminal) - v;
        const float gl = a.gamma_tau * nextnonterminal;  // delta + gamma * tau * nextnonterminal * lastgaelam
        lastgaelam = delta + gl * lastgaelam;
        if (a.mb_advs) a.mb_advs[i] = lastgaelam;
        a.mb_returns[i] = lastgaelam + v;
        nextvalues = v;
        nextnonterminal = 1.0f - d;
    }
}

}  // namespace

// scratch words for `num_rows` rows: retirement counter + one look-back slot per chunk
extern "C" int ige_rollout_scratch_words(int32_t num_rows) { return 4 + (num_rows + kChunk - 1) / kChunk; }

extern "C" int ige_ma_rollout_step(const IgeRolloutStepArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    if (a->num_rows < 0 || a->num_rows > (1 << 30) - 1) return ige_fail(IGE_ESIZE, "num_rows=%d out of range", a->num_rows);
    if (a->num_agents < 1) return ige_fail(IGE_ESIZE, "num_agents=%d must be >= 1", a->num_agents);
    if (a->meter_max_size < 1) return ige_fail(IGE_ESIZE, "meter_max_size=%d must be >= 1", a->meter_max_size);
    if (!a->rewards || !a->dones || !a->mb_rewards || !a->current_rewards || !a->current_shaped_rewards || !a->current_lengths)
        return ige_fail(IGE_ENULL, "rewards / dones / mb_rewards / current_* must be non-NULL");
    if (a->value_bootstrap && (!a->time_outs || !a->values))
        return ige_fail(IGE_ENULL, "value_bootstrap needs time_outs and values");
    // the 16-B vector path needs every stream 16-B aligned (row slices of (horizon, rows) buffers are not, when
    // rows % 4 != 0); otherwise the same kernel takes its scalar path.  Natural alignment is always required.
    uintptr_t all = 0, nat = 0;
    const void *f32s[] = {a->rewards, a->mb_rewards, a->current_rewards, a->current_shaped_rewards, a->current_lengths,
                          a->value_bootstrap ? a->values : nullptr};
    for (const void *q : f32s) all |= reinterpret_cast<uintptr_t>(q), nat |= reinterpret_cast<uintptr_t>(q) & 3u;
    all |= reinterpret_cast<uintptr_t>(a->dones);
    nat |= reinterpret_cast<uintptr_t>(a->dones) & 7u;
    if (nat) return ige_fail(IGE_EALIGN, "a float / int64 stream is not naturally aligned");
    int vec = (all & 15u) == 0;
    if (a->dones_out && (reinterpret_cast<uintptr_t>(a->dones_out) & 3u)) vec = 0;
    if (a->value_bootstrap && (reinterpret_cast<uintptr_t>(a->time_outs) & 3u)) vec = 0;
    if (!a->done_rows || !a->done_count || !a->meter_mean || !a->meter_size || !a->reduce_scratch || !a->scratch)
        return ige_fail(IGE_ENULL, "done_rows / done_count / meter_* / reduce_scratch / scratch must be non-NULL");
    const int num_chunks = a->num_rows == 0 ? 1 : (a->num_rows + kChunk - 1) / kChunk;
    static int cap[IGE_MAX_DEVICES];   // resident CTAs per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= IGE_MAX_DEVICES) return ige_fail(IGE_EPARAM, "device ordinal %d not supported", dev);
    if (cap[dev] == 0) {
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ma_rollout_step_kernel, kThreads, 0);
        if (e != cudaSuccess || per_sm < 1 || sms < 1) return ige_fail(e != cudaSuccess ? (int)e : IGE_EPARAM, "ige_ma_rollout_step: occupancy query failed");
        cap[dev] = sms * per_sm;
    }
    ige_launch(ma_rollout_step_kernel, num_chunks, kThreads, 0, (cudaStream_t)stream, *a, num_chunks, vec,
               num_chunks > cap[dev] ? 1 : 0);
    return ige_check_launch("ige_ma_rollout_step");
}

extern "C" int ige_ma_discount_values(const IgeDiscountArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    if (a->num_rows < 0 || a->horizon < 0) return ige_fail(IGE_ESIZE, "num_rows=%d horizon=%d invalid", a->num_rows, a->horizon);
    if (!a->dones || !a->last_values || !a->mb_dones || !a->mb_values || !a->mb_rewards || !a->mb_returns)
        return ige_fail(IGE_ENULL, "dones / last_values / mb_* must be non-NULL");
    if (a->num_rows == 0 || a->horizon == 0) return IGE_OK;
ige_launch(ma_discount_values_kernel, (a->num_rows + 255) / 256, 256, 0, (cudaStream_t)stream, *a);
    return ige_check_launch("ige_ma_discount_values");
}
