// anymal.cu -- Anymal.post_physics_step as ONE fused kernel (first of the remaining
// single-agent tasks).  Replaces tasks/anymal.py:231-240 (progress += 1, reset_buf.nonzero(), reset_idx :277-304,
// compute_anymal_observations :355-386, compute_anymal_reward :311-351) and the step() epilogue
// (tasks/base/vec_task.py:396-404).
//
// HBM-bound: 544 B/env (R root 52 + dof_state 96 + torques 48 + actions 48 + contact rows 60 + commands 12 + flags 16;
// W obs 192 + rew 4 + flags 16), ~250 flops.  Thread per env; the 96-B DOF row, torque / action rows and the 192-B
// observation row move as 16-B vectors (rows are 16-B multiples).  The reset list (== reset_buf.nonzero(), ascending)
// and each reset env's row in the injected uniforms come from a chained scan: chunk = CTA (128 envs), warp-wide
// decoupled look-back over chunk counts, all loads issued before the look-back.  The CTA that retires last restores
// the scratch and advances the Philox step, so a launch is re-playable inside a CUDA graph.
#include "ige_common.cuh"

namespace {

constexpr int kThreads = 128;
constexpr uint32_t kFlagAgg = 1u << 30, kFlagPrefix = 2u << 30, kValMask = (1u << 30) - 1;

__device__ __forceinline__ uint32_t poll(const uint32_t *p) { return *reinterpret_cast<const volatile uint32_t *>(p); }

// torch_rand_float (utils/torch_jit_utils.py:216-218): (upper - lower) * rand + lower, scalars rounded to fp32
__device__ __forceinline__ float rand_float(float span, float lower, float u) { return span * u + lower; }

// 4 uniforms of draw `draw`, columns 4*grp .. 4*grp+3 (same counter layout as ige::philox_uniform / the oracle)
__device__ __forceinline__ void philox4(const IgeAnymalParams &p, uint64_t step, uint32_t row, uint32_t draw, uint32_t grp,
                                        float (&out)[4]) {
    uint32_t c[4] = {row, (draw << 16) | grp, (uint32_t)step, (uint32_t)(step >> 32)};
    ige::philox4x32_10(c, (uint32_t)p.seed, (uint32_t)(p.seed >> 32));
#pragma unroll
    for (int i = 0; i < 4; ++i) out[i] = (float)(c[i] >> 8) * (1.0f / 16777216.0f);
}

__global__ void __launch_bounds__(kThreads) anymal_post_physics_kernel(IgeAnymalParams p, IgeAnymalBuffers b, int num_chunks, int use_ticket) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    __shared__ int s_warp[kThreads / 32];
    __shared__ int s_prefix, s_last;
    uint32_t *scratch = b.scratch;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint64_t step = (p.rand_mode == IGE_RAND_PHILOX && p.step) ? *reinterpret_cast<const volatile uint64_t *>(p.step) : 0ull;
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
    const int e = chunk * kThreads + tid;
    const bool ok = e < p.num_envs;

    // ---- every load whose address depends only on the env index, up front ----
    long long reset_in = 0, progress = 0;
    float root[13], ds[24], tq[12], act[12], cmd[3], cfb[3], cfk[IGE_ANYMAL_MAX_KNEES][3];
    if (ok) {
        reset_in = __ldg(b.reset_buf + e);
        progress = b.progress_buf[e] + 1;  // :232
#pragma unroll
        for (int i = 0; i < 13; ++i) root[i] = __ldg(b.root_states + (size_t)e * 13 + i);
#pragma unroll
        for (int i = 0; i < 6; ++i) {
            const float4 v = reinterpret_cast<const float4 *>(b.dof_state + (size_t)e * 24)[i];
            ds[4 * i] = v.x, ds[4 * i + 1] = v.y, ds[4 * i + 2] = v.z, ds[4 * i + 3] = v.w;
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const float4 t = __ldg(reinterpret_cast<const float4 *>(b.torques + (size_t)e * 12) + i);
            tq[4 * i] = t.x, tq[4 * i + 1] = t.y, tq[4 * i + 2] = t.z, tq[4 * i + 3] = t.w;
            const float4 a = __ldg(reinterpret_cast<const float4 *>(b.actions + (size_t)e * 12) + i);
            act[4 * i] = a.x, act[4 * i + 1] = a.y, act[4 * i + 2] = a.z, act[4 * i + 3] = a.w;
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) cmd[i] = b.commands[(size_t)e * 3 + i];
        const float *cf = b.contact_forces + (size_t)e * p.num_bodies * 3;
#pragma unroll
        for (int i = 0; i < 3; ++i) cfb[i] = __ldg(cf + p.base_index * 3 + i);
#pragma unroll
        for (int k = 0; k < IGE_ANYMAL_MAX_KNEES; ++k)
            if (k < p.num_knees) {
#pragma unroll
                for (int i = 0; i < 3; ++i) cfk[k][i] = __ldg(cf + p.knee_indices[k] * 3 + i);
            }
    }

    // ---- rank of this env in reset_buf.nonzero(): ballots + chained scan over chunk counts ----
    const bool flagged = ok && reset_in != 0;
    const unsigned bal = __ballot_sync(0xffffffffu, flagged);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    int warp_off = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) {
        const int c = s_warp[w];
        if (w < warp) warp_off += c;
        total += c;
    }
    if (warp == 0) {
        uint32_t prefix = 0;
        if (chunk > 0) {
            if (lane == 0) atomicExch(&scratch[2 + chunk], kFlagAgg | (uint32_t)total);
            for (int j0 = chunk - 1;; j0 -= 32) {
                const int j = j0 - lane;
                uint32_t s = kFlagPrefix;
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
            if (chunk == num_chunks - 1) *b.reset_count = (int)prefix + total;
        }
    }
    __syncthreads();
    const int rank = s_prefix + warp_off + __popc(bal & ((1u << lane) - 1u));

    if (ok) {
        // ---- reset_idx :277-304 ----
        if (flagged) {
            float u0[12], u1[12], uc[3];
            if (p.rand_mode == IGE_RAND_PHILOX) {
#pragma unroll
                for (int g = 0; g < 3; ++g) {
                    float o[4];
                    philox4(p, step, (uint32_t)e, 0u, (uint32_t)g, o);
#pragma unroll
                    for (int i = 0; i < 4; ++i) u0[4 * g + i] = o[i];
                    philox4(p, step, (uint32_t)e, 1u, (uint32_t)g, o);
#pragma unroll
                    for (int i = 0; i < 4; ++i) u1[4 * g + i] = o[i];
                }
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    float o[4];
                    philox4(p, step, (uint32_t)e, (uint32_t)(2 + d), 0u, o);
                    uc[d] = o[0];
                }
            } else {
#pragma unroll
                for (int j = 0; j < 12; ++j) {
                    u0[j] = __ldg(p.u[0] + (size_t)rank * 12 + j);
                    u1[j] = __ldg(p.u[1] + (size_t)rank * 12 + j);
                }
#pragma unroll
                for (int d = 0; d < 3; ++d) uc[d] = __ldg(p.u[2 + d] + rank);
            }
#pragma unroll
            for (int j = 0; j < 12; ++j) {
                ds[2 * j] = p.default_dof_pos[j] * rand_float(1.0f, 0.5f, u0[j]);            // :282, :285
                ds[2 * j + 1] = rand_float((float)(0.1 - (-0.1)), (float)(-0.1), u1[j]);     // :283, :286
            }
            const float *rg[3] = {p.command_x_range, p.command_y_range, p.command_yaw_range};
#pragma unroll
            for (int d = 0; d < 3; ++d)                                                      // :300-302
                cmd[d] = rand_float((float)((double)rg[d][1] - (double)rg[d][0]), rg[d][0], uc[d]);
#pragma unroll
            for (int i = 0; i < 6; ++i)
                reinterpret_cast<float4 *>(b.dof_state + (size_t)e * 24)[i] =
                    make_float4(ds[4 * i], ds[4 * i + 1], ds[4 * i + 2], ds[4 * i + 3]);
#pragma unroll
            for (int d = 0; d < 3; ++d) b.commands[(size_t)e * 3 + d] = cmd[d];
            progress = 0;  // :303
            b.reset_ids[rank] = e;
        }
        // ---- compute_anymal_observations :355-386 ----
        const float grav[3] = {0.0f, 0.0f, -1.0f};
        float blv[3], bav[3], pg[3], obs[48];
        ige::quat_rotate<true>(root + 3, root + 7, blv);
        ige::quat_rotate<true>(root + 3, root + 10, bav);
        ige::quat_rotate<false>(root + 3, grav, pg);
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            obs[i] = blv[i] * p.lin_vel_scale;
            obs[3 + i] = bav[i] * p.ang_vel_scale;
            obs[6 + i] = pg[i];
            obs[9 + i] = cmd[i] * (i < 2 ? p.lin_vel_scale : p.ang_vel_scale);
        }
#pragma unroll
        for (int j = 0; j < 12; ++j) {
            obs[12 + j] = (ds[2 * j] - p.default_dof_pos[j]) * p.dof_pos_scale;
            obs[24 + j] = ds[2 * j + 1] * p.dof_vel_scale;
            obs[36 + j] = act[j];
        }
#pragma unroll
        for (int i = 0; i < 12; ++i)
            reinterpret_cast<float4 *>(b.obs_buf + (size_t)e * 48)[i] =
                make_float4(obs[4 * i], obs[4 * i + 1], obs[4 * i + 2], obs[4 * i + 3]);
        if (b.obs_clamped) {  // vec_task.py:404
#pragma unroll
            for (int i = 0; i < 12; ++i) {
                float4 v;
                v.x = fminf(fmaxf(obs[4 * i], -p.clip_obs), p.clip_obs);
                v.y = fminf(fmaxf(obs[4 * i + 1], -p.clip_obs), p.clip_obs);
                v.z = fminf(fmaxf(obs[4 * i + 2], -p.clip_obs), p.clip_obs);
                v.w = fminf(fmaxf(obs[4 * i + 3], -p.clip_obs), p.clip_obs);
                reinterpret_cast<float4 *>(b.obs_clamped + (size_t)e * 48)[i] = v;
            }
        }
        // ---- compute_anymal_reward :311-351 ----
        const float d0 = cmd[0] - blv[0], d1 = cmd[1] - blv[1], d2 = cmd[2] - bav[2];
        const float lin_err = d0 * d0 + d1 * d1;  // torch.sum over 2 elements
        const float ang_err = d2 * d2;
        const float r_lin = expf(-lin_err / 0.25f) * p.rew_lin_vel_xy;
        const float r_ang = expf(-ang_err / 0.25f) * p.rew_ang_vel_z;
        const float tsum = ige::row_sum<12>([&](int j) { return tq[j] * tq[j]; });
        const float r_tq = tsum * p.rew_torque;
        float total_rew = (r_lin + r_ang) + r_tq;
        total_rew = fmaxf(total_rew, 0.0f);
        bool reset = ige::norm3(cfb[0], cfb[1], cfb[2]) > 1.0f;
#pragma unroll
        for (int k = 0; k < IGE_ANYMAL_MAX_KNEES; ++k)
            if (k < p.num_knees) reset |= ige::norm3(cfk[k][0], cfk[k][1], cfk[k][2]) > 1.0f;
        const bool time_out = (float)progress >= p.max_episode_length - 1.0f;  // :348
        reset |= time_out;
        b.rew_buf[e] = total_rew;
        b.reset_buf[e] = reset ? 1 : 0;
        b.progress_buf[e] = progress;
        if (b.timeout_buf) b.timeout_buf[e] = (uint8_t)(time_out && reset);  // vec_task.py:396
    }

    // ---- retire: the last CTA restores the scratch and advances the Philox step ----
    __syncthreads();
    if (tid == 0) {
        __threadfence();
        s_last = (atomicAdd(&scratch[1], 1u) == (uint32_t)num_chunks - 1u);
    }
    __syncthreads();
    if (s_last) {
        for (int i = tid; i < num_chunks; i += kThreads) scratch[2 + i] = 0u;
        if (tid == 0) {
            scratch[0] = 0u;
            scratch[1] = 0u;
            if (p.rand_mode == IGE_RAND_PHILOX && p.step) *p.step = step + 1;
        }
    }
    __syncthreads();   // s_warp / s_prefix / s_last are reused by the next chunk
    }
}

}  // namespace

extern "C" int ige_anymal_scratch_words(int32_t num_envs) { return 4 + (num_envs + kThreads - 1) / kThreads; }

extern "C" int ige_anymal_post_physics(const IgeAnymalParams *p, const IgeAnymalBuffers *b, void *stream) {
    if (!p || !b) return ige_fail(IGE_ENULL, "params / buffers is NULL");
    if (p->num_envs < 0 || p->num_envs > (1 << 30) - 1) return ige_fail(IGE_ESIZE, "num_envs=%d out of range", p->num_envs);
    if (p->num_knees < 0 || p->num_knees > IGE_ANYMAL_MAX_KNEES) return ige_fail(IGE_ESIZE, "num_knees=%d out of range", p->num_knees);
    if (p->base_index < 0 || p->base_index >= p->num_bodies) return ige_fail(IGE_ESIZE, "base_index out of range");
    for (int k = 0; k < p->num_knees; ++k)
        if (p->knee_indices[k] < 0 || p->knee_indices[k] >= p->num_bodies) return ige_fail(IGE_ESIZE, "knee index out of range");
    if (p->rand_mode != IGE_RAND_INJECTED && p->rand_mode != IGE_RAND_PHILOX)
        return ige_fail(IGE_EPARAM, "rand_mode must be IGE_RAND_INJECTED or IGE_RAND_PHILOX");
    IGE_REQUIRE_PTR(b->dof_state);
    IGE_REQUIRE_PTR(b->torques);
    IGE_REQUIRE_PTR(b->actions);
    IGE_REQUIRE_PTR(b->obs_buf);
    IGE_OPTIONAL_PTR(b->obs_clamped);
    if (!b->root_states || !b->contact_forces || !b->commands || !b->rew_buf || !b->reset_buf || !b->progress_buf ||
        !b->reset_ids || !b->reset_count || !b->scratch)
        return ige_fail(IGE_ENULL, "root_states / contact_forces / commands / rew / flags / reset list / scratch must be non-NULL");
    const int num_chunks = p->num_envs == 0 ? 1 : (p->num_envs + kThreads - 1) / kThreads;
    static int cap[IGE_MAX_DEVICES];   // resident CTAs per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= IGE_MAX_DEVICES) return ige_fail(IGE_EPARAM, "device ordinal %d not supported", dev);
    if (cap[dev] == 0) {
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, anymal_post_physics_kernel, kThreads, 0);
        if (e != cudaSuccess || per_sm < 1 || sms < 1) return ige_fail(e != cudaSuccess ? (int)e : IGE_EPARAM, "ige_anymal_post_physics: occupancy query failed");
        cap[dev] = sms * per_sm;
    }
    ige_launch(anymal_post_physics_kernel, num_chunks, kThreads, 0, (cudaStream_t)stream, *p, *b, num_chunks,
               num_chunks > cap[dev] ? 1 : 0);
    return ige_check_launch("ige_anymal_post_physics");
}
