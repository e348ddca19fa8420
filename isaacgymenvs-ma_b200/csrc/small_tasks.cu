// small_tasks.cu -- BallBalance, Quadcopter and Ingenuity for sm_100a.
//
// Replaces (reference, isaacgymenvs/tasks/):
//   ball_balance.py  pre_physics_step :410-430 (reset_idx :360-408), post_physics_step :432-441
//                    (compute_observations :330-345, compute_bbot_reward :460-476)
//   quadcopter.py    pre_physics_step :301-331 (reset_idx :280-299), post_physics_step :333-341
//                    (compute_observations :362-373, compute_quadcopter_reward :391-425)
//   ingenuity.py     pre_physics_step :322-364 (set_targets :283-294, reset_idx :296-320), post_physics_step
//                    :366-374 (compute_observations :396-401, compute_ingenuity_reward :420-455)
// plus the step() epilogue (tasks/base/vec_task.py:396-404).  These tasks reset in PRE-physics position, so each has
// two entry points; every kernel is one launch, thread per env, persistent grid (task_common.cuh).  HBM-bound with
// 100-300 B per env and no reuse: rows are read with the widest aligned vectors their odd sizes allow.
#include "task_common.cuh"

namespace {

using ige::kTaskThreads;
constexpr float kPiF = 3.14159274101257324219f;   // float32(math.pi)

__device__ __forceinline__ float clamp_obs(float x, float c) { return fminf(fmaxf(x, -c), c); }
__device__ __forceinline__ float tensor_clampf(float t, float lo, float hi) { return fmaxf(fminf(t, hi), lo); }  // torch_jit_utils.py:228-230

// ------------------------------------------------------------------------------------------------
// BallBalance
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTaskThreads) ball_balance_pre_kernel(const __grid_constant__ IgeBallBalanceArgs a,
                                                                        const int workers) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    ige::task_reset_skeleton(a.num_envs, workers, a.reset_buf, a.list, a.rand, [&](int e, bool flagged, uint64_t step) {
        float tg[6], act[3];
#pragma unroll
        for (int j = 0; j < 6; ++j) tg[j] = a.dof_position_targets[(size_t)e * 6 + j];
#pragma unroll
        for (int j = 0; j < 3; ++j) act[j] = __ldg(a.actions + (size_t)e * 3 + j);
        if (flagged) {  // reset_idx :360-408
            float u[5];
#pragma unroll
            for (int d = 0; d < 5; ++d) u[d] = ige::task_uniform(a.rand, step, d, e, 1, 0);
            const float dists = ige::rand_span(0.001, 0.5, u[0]);
            const float angle = ige::rand_span(-3.14159265358979323846, 3.14159265358979323846, u[1]);  // torch_random_dir_2 :222-225
            const float dc = cosf(angle), dsn = sinf(angle);
            const float hx = dists * dc, hy = dists * dsn;
            const float speedscales = (dists - 0.001f) / (float)(0.5 - 0.001);
            const float hspeeds = ige::rand_span(0.0, 5.0, u[2]);
            const float hvx = (-speedscales * hspeeds) * dc, hvy = (-speedscales * hspeeds) * dsn;
            const float vspeed = -ige::rand_span(5.0, 5.0, u[3]);
            const float height = ige::rand_span(1.0, 2.0, u[4]);
            float *root = a.root_states + (size_t)e * 26;
            const float *init = a.initial_root_states + (size_t)e * 26;
#pragma unroll
            for (int i = 0; i < 13; ++i) root[i] = __ldg(init + i);    // the balance bot's row
            float *ball = root + 13;
            ball[0] = hx; ball[1] = hy; ball[2] = height;
            ball[3] = 0.f; ball[4] = 0.f; ball[5] = 0.f; ball[6] = 1.f;
            ball[7] = hvx; ball[8] = hvy; ball[9] = vspeed;
            ball[10] = 0.f; ball[11] = 0.f; ball[12] = 0.f;
#pragma unroll
            for (int i = 0; i < 3; ++i)
                reinterpret_cast<float4 *>(a.dof_states + (size_t)e * 12)[i] =
                    __ldg(reinterpret_cast<const float4 *>(a.initial_dof_states + (size_t)e * 12) + i);
            a.reset_buf[e] = 0;
            a.progress_buf[e] = 0;
        }
        // position targets :419-427 (actuated DOFs 1, 3, 5)
#pragma unroll
        for (int k = 0; k < 3; ++k) tg[2 * k + 1] += a.target_step * act[k];
#pragma unroll
        for (int j = 0; j < 6; ++j) {
            tg[j] = tensor_clampf(tg[j], a.dof_lower[j], a.dof_upper[j]);
            if (flagged) tg[j] = 0.0f;
            a.dof_position_targets[(size_t)e * 6 + j] = tg[j];
        }
    });
}

__global__ void __launch_bounds__(kTaskThreads) ball_balance_post_kernel(const __grid_constant__ IgeBallBalanceArgs a) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < a.num_envs; e += gridDim.x * blockDim.x) {
        const long long progress = a.progress_buf[e] + 1;   // :434
        long long reset = a.reset_buf[e];
        float dof[12], ball[10], sens[18];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const float4 v = reinterpret_cast<const float4 *>(a.dof_states + (size_t)e * 12)[i];
            dof[4 * i] = v.x; dof[4 * i + 1] = v.y; dof[4 * i + 2] = v.z; dof[4 * i + 3] = v.w;
        }
        const float *br = a.root_states + (size_t)e * 26 + 13;
#pragma unroll
        for (int i = 0; i < 10; ++i) ball[i] = br[i];
#pragma unroll
        for (int i = 0; i < 9; ++i) {
            const float2 v = __ldg(reinterpret_cast<const float2 *>(a.sensor_tensor + (size_t)e * 18) + i);
            sens[2 * i] = v.x; sens[2 * i + 1] = v.y;
        }
        float obs[24];
#pragma unroll
        for (int k = 0; k < 3; ++k) {  // compute_observations :330-345
            obs[k] = dof[2 * (2 * k + 1)];
            obs[3 + k] = dof[2 * (2 * k + 1) + 1];
            obs[6 + k] = ball[k];
            obs[9 + k] = ball[7 + k];
            obs[12 + k] = sens[6 * k] / 20.0f;
            obs[15 + k] = sens[6 * k + 3] / 20.0f;
            obs[18 + k] = sens[6 * k + 4] / 20.0f;
            obs[21 + k] = sens[6 * k + 5] / 20.0f;
        }
#pragma unroll
        for (int i = 0; i < 6; ++i)
            reinterpret_cast<float4 *>(a.obs_buf + (size_t)e * 24)[i] = make_float4(obs[4 * i], obs[4 * i + 1], obs[4 * i + 2], obs[4 * i + 3]);
        if (a.obs_clamped) {
#pragma unroll
            for (int i = 0; i < 6; ++i)
                reinterpret_cast<float4 *>(a.obs_clamped + (size_t)e * 24)[i] =
                    make_float4(clamp_obs(obs[4 * i], a.clip_obs), clamp_obs(obs[4 * i + 1], a.clip_obs),
                                clamp_obs(obs[4 * i + 2], a.clip_obs), clamp_obs(obs[4 * i + 3], a.clip_obs));
        }
        // compute_bbot_reward :460-476
        const float zz = ball[2] - 0.7f;
        const float ball_dist = __fsqrt_rn((ball[0] * ball[0] + zz * zz) + ball[1] * ball[1]);
        const float ball_speed = __fsqrt_rn((ball[7] * ball[7] + ball[8] * ball[8]) + ball[9] * ball[9]);
        const float pos_reward = 1.0f / (1.0f + ball_dist);
        const float speed_reward = 1.0f / (1.0f + ball_speed);
        const bool late = (float)progress >= a.max_episode_length - 1.0f;
        if (late) reset = 1;
        if (ball[2] < (float)((double)a.ball_radius * 1.5)) reset = 1;
        a.rew_buf[e] = pos_reward * speed_reward;
        a.reset_buf[e] = reset;
        a.progress_buf[e] = progress;
        if (a.timeout_buf) a.timeout_buf[e] = (late && reset != 0) ? 1 : 0;   // vec_task.py:396
    }
}

// ------------------------------------------------------------------------------------------------
// Quadcopter
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTaskThreads) quadcopter_pre_kernel(const __grid_constant__ IgeQuadcopterArgs a,
                                                                      const int workers) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    ige::task_reset_skeleton(a.num_envs, workers, a.reset_buf, a.list, a.rand, [&](int e, bool flagged, uint64_t step) {
        float tg[8], th[4], act[12], pos[8];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const float4 v = reinterpret_cast<const float4 *>(a.dof_position_targets + (size_t)e * 8)[i];
            tg[4 * i] = v.x; tg[4 * i + 1] = v.y; tg[4 * i + 2] = v.z; tg[4 * i + 3] = v.w;
        }
        {
            const float4 v = *reinterpret_cast<const float4 *>(a.thrusts + (size_t)e * 4);
            th[0] = v.x; th[1] = v.y; th[2] = v.z; th[3] = v.w;
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const float4 v = __ldg(reinterpret_cast<const float4 *>(a.actions + (size_t)e * 12) + i);
            act[4 * i] = v.x; act[4 * i + 1] = v.y; act[4 * i + 2] = v.z; act[4 * i + 3] = v.w;
        }
        if (flagged) {  // reset_idx :280-299
            float root[13];
#pragma unroll
            for (int i = 0; i < 13; ++i) root[i] = __ldg(a.initial_root_states + (size_t)e * 13 + i);
            root[0] += ige::rand_span(-1.5, 1.5, ige::task_uniform(a.rand, step, 0, e, 1, 0));
            root[1] += ige::rand_span(-1.5, 1.5, ige::task_uniform(a.rand, step, 1, e, 1, 0));
            root[2] += ige::rand_span(-0.2, 1.5, ige::task_uniform(a.rand, step, 2, e, 1, 0));
#pragma unroll
            for (int i = 0; i < 13; ++i) a.root_states[(size_t)e * 13 + i] = root[i];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                pos[j] = ige::rand_span(-0.2, 0.2, ige::task_uniform(a.rand, step, 3, e, 8, j));
                reinterpret_cast<float2 *>(a.dof_states)[(size_t)e * 8 + j] = make_float2(pos[j], 0.0f);
            }
            a.reset_buf[e] = 0;
            a.progress_buf[e] = 0;
        }
        // DOF position targets, thrusts, rotor forces :308-320
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            tg[j] += a.dof_target_step * act[j];
            tg[j] = tensor_clampf(tg[j], a.dof_lower[j], a.dof_upper[j]);
            if (flagged) tg[j] = pos[j];   // :325
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            th[k] += a.thrust_step * act[8 + k];
            th[k] = tensor_clampf(th[k], a.thrust_lower[k], a.thrust_upper[k]);
        }
        float *f = a.forces + (size_t)e * a.num_bodies * 3;
        if (flagged) {  // :323-324
#pragma unroll
            for (int k = 0; k < 4; ++k) th[k] = 0.0f;
            for (int i = 0; i < a.num_bodies * 3; ++i) f[i] = 0.0f;
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) f[a.rotor_bodies[k] * 3 + 2] = th[k];
        }
#pragma unroll
        for (int i = 0; i < 2; ++i)
            reinterpret_cast<float4 *>(a.dof_position_targets + (size_t)e * 8)[i] = make_float4(tg[4 * i], tg[4 * i + 1], tg[4 * i + 2], tg[4 * i + 3]);
        *reinterpret_cast<float4 *>(a.thrusts + (size_t)e * 4) = make_float4(th[0], th[1], th[2], th[3]);
    });
}

// shared by Quadcopter and Ingenuity: reward terms of compute_quadcopter_reward :391-425 / compute_ingenuity_reward :420-455
__device__ __forceinline__ float copter_reward(const float *root, float target_dist, float up_scale) {
    const float pos_reward = 1.0f / (1.0f + target_dist * target_dist);
    const float zaxis[3] = {0.0f, 0.0f, 1.0f};
    float ups[3];
    ige::quat_rotate<false>(root + 3, zaxis, ups);   // quat_axis(root_quats, 2), torch_jit_utils.py:280-284
    const float tiltage = fabsf(1.0f - ups[2]);
    const float up_reward = up_scale / (1.0f + tiltage * tiltage);
    const float spinnage = fabsf(root[12]);
    const float spinnage_reward = 1.0f / (1.0f + spinnage * spinnage);
    return pos_reward + pos_reward * (up_reward + spinnage_reward);
}

__global__ void __launch_bounds__(kTaskThreads) quadcopter_post_kernel(const __grid_constant__ IgeQuadcopterArgs a) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < a.num_envs; e += gridDim.x * blockDim.x) {
        const long long progress = a.progress_buf[e] + 1;   // :335
        float root[13], obs[21];
#pragma unroll
        for (int i = 0; i < 13; ++i) root[i] = a.root_states[(size_t)e * 13 + i];
        obs[0] = (0.0f - root[0]) / 3.0f;   // compute_observations :362-373
        obs[1] = (0.0f - root[1]) / 3.0f;
        obs[2] = (1.0f - root[2]) / 3.0f;
#pragma unroll
        for (int i = 0; i < 4; ++i) obs[3 + i] = root[3 + i];
#pragma unroll
        for (int i = 0; i < 3; ++i) { obs[7 + i] = root[7 + i] / 2.0f; obs[10 + i] = root[10 + i] / kPiF; }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float4 v = reinterpret_cast<const float4 *>(a.dof_states + (size_t)e * 16)[i];
            obs[13 + 2 * i] = v.x; obs[13 + 2 * i + 1] = v.z;
        }
        float *o = a.obs_buf + (size_t)e * 21;
#pragma unroll
        for (int i = 0; i < 21; ++i) o[i] = obs[i];
        if (a.obs_clamped) {
#pragma unroll
            for (int i = 0; i < 21; ++i) a.obs_clamped[(size_t)e * 21 + i] = clamp_obs(obs[i], a.clip_obs);
        }
        const float dz = 1.0f - root[2];
        const float target_dist = __fsqrt_rn((root[0] * root[0] + root[1] * root[1]) + dz * dz);
        const float reward = copter_reward(root, target_dist, 1.0f);
        const bool die = target_dist > 3.0f || root[2] < 0.3f;
        const bool late = (float)progress >= a.max_episode_length - 1.0f;
        const long long reset = (late || die) ? 1 : 0;
        a.rew_buf[e] = reward;
        a.reset_buf[e] = reset;
        a.progress_buf[e] = progress;
        if (a.timeout_buf) a.timeout_buf[e] = (late && reset != 0) ? 1 : 0;
    }
}

// ------------------------------------------------------------------------------------------------
// Ingenuity
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ingenuity_set_target(const IgeIngenuityArgs &a, int e, float ux, float uy, float uz, float *marker) {
    const float tx = ux * 10.0f - 5.0f, ty = uy * 10.0f - 5.0f, tz = uz + 1.0f;   // set_targets :283-294
    a.target_root_positions[(size_t)e * 3 + 0] = tx;
    a.target_root_positions[(size_t)e * 3 + 1] = ty;
    a.target_root_positions[(size_t)e * 3 + 2] = tz;
    marker[0] = tx; marker[1] = ty; marker[2] = tz + 0.4f;
}

__global__ void __launch_bounds__(kTaskThreads) ingenuity_pre_kernel(const __grid_constant__ IgeIngenuityArgs a,
                                                                     const int workers) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    ige::task_reset_skeleton<true>(
        a.num_envs, workers, a.reset_buf, a.list, a.rand,
        [&](int e, bool flagged, uint64_t step, bool chunk_any) {
            // :299-300 re-sets the rotor speeds of ALL envs whenever ANY env resets.  A chunk that holds a flagged env
            // knows the answer and writes its own envs here; the scanner CTA (below) covers the chunks that hold none.
            if (chunk_any) {
                a.dof_states[(size_t)e * 8 + 3] = -50.0f;
                a.dof_states[(size_t)e * 8 + 7] = 50.0f;
            }
            float *copter = a.root_tensor + (size_t)e * 26, *marker = copter + 13;
            const bool retarget = a.progress_buf[e] % 500 == 0;   // :325 (progress is read before reset_idx zeroes it)
            if (retarget)
                ingenuity_set_target(a, e, ige::task_uniform_ranked(a.rand, a.target_rank, step, 0, e, 2, 0),
                                     ige::task_uniform_ranked(a.rand, a.target_rank, step, 0, e, 2, 1),
                                     ige::task_uniform_ranked(a.rand, a.target_rank, step, 1, e, 1, 0), marker);
            if (flagged) {   // reset_idx :296-320 (its nested set_targets draws first)
                ingenuity_set_target(a, e, ige::task_uniform(a.rand, step, 2, e, 2, 0), ige::task_uniform(a.rand, step, 2, e, 2, 1),
                                     ige::task_uniform(a.rand, step, 3, e, 1, 0), marker);
                float root[13];
#pragma unroll
                for (int i = 0; i < 13; ++i) root[i] = __ldg(a.initial_root_states + (size_t)e * 13 + i);
                root[0] += ige::rand_span(-1.5, 1.5, ige::task_uniform(a.rand, step, 4, e, 1, 0));
                root[1] += ige::rand_span(-1.5, 1.5, ige::task_uniform(a.rand, step, 5, e, 1, 0));
                root[2] += ige::rand_span(-0.2, 1.5, ige::task_uniform(a.rand, step, 6, e, 1, 0));
#pragma unroll
                for (int i = 0; i < 13; ++i) copter[i] = root[i];
                a.reset_buf[e] = 0;
                a.progress_buf[e] = 0;
            }
            if (a.actor_dirty) {   // the sorted unique() of :341, as flags over the (N, 2) actor rows
                a.actor_dirty[(size_t)e * 2] = flagged ? 1 : 0;
                a.actor_dirty[(size_t)e * 2 + 1] = (flagged || retarget) ? 1 : 0;
            }
            // rotor thrusts :345-360
            float act[6], th[6];
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                const float2 v = __ldg(reinterpret_cast<const float2 *>(a.actions + (size_t)e * 6) + i);
                act[2 * i] = v.x; act[2 * i + 1] = v.y;
            }
#pragma unroll
            for (int r = 0; r < 2; ++r) {
                const float vert = fminf(fmaxf(act[3 * r + 2] * 2000.0f, -a.thrust_upper_limit), a.thrust_upper_limit);
                th[3 * r + 2] = a.dt * vert;
                th[3 * r] = th[3 * r + 2] * fminf(fmaxf(act[3 * r], -a.thrust_lateral_component), a.thrust_lateral_component);
                th[3 * r + 1] = th[3 * r + 2] * fminf(fmaxf(act[3 * r + 1], -a.thrust_lateral_component), a.thrust_lateral_component);
            }
            float *f = a.forces + (size_t)e * a.num_bodies * 3;
            if (flagged) {
#pragma unroll
                for (int i = 0; i < 6; ++i) th[i] = 0.0f;
                for (int i = 0; i < a.num_bodies * 3; ++i) f[i] = 0.0f;
            } else {
#pragma unroll
                for (int i = 0; i < 3; ++i) { f[3 + i] = th[i]; f[9 + i] = th[3 + i]; }   // bodies 1 and 3
            }
#pragma unroll
            for (int i = 0; i < 3; ++i)
                reinterpret_cast<float2 *>(a.thrusts + (size_t)e * 6)[i] = make_float2(th[2 * i], th[2 * i + 1]);
        },
        [&](int total) {   // :299-300 for the 128-env chunks without a flagged env (their 4 slots are all zero)
            if (total == 0) return;
            const uint64_t *slots = reinterpret_cast<const uint64_t *>(a.list.scratch + ige::kListMaskBase);
            const int tiles = (a.num_envs + 31) / 32, chunks = (a.num_envs + kTaskThreads - 1) / kTaskThreads;
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
            // warp w looks at 32 chunks at a time (lane = chunk: its four slots, independent loads), then writes the
            // envs of the few empty ones together
            for (int c0 = warp * 32; c0 < chunks; c0 += kTaskThreads) {
                const int c = c0 + lane;
                uint32_t m = 1u;
                if (c < chunks) {
                    m = 0u;
                    for (int t = 4 * c; t < 4 * c + 4 && t < tiles; ++t) m |= (uint32_t)ige::ld_relaxed_gpu_u64(slots + t);
                }
                for (uint32_t em = __ballot_sync(0xffffffffu, m == 0u); em; em &= em - 1) {
                    const int base = (c0 + __ffs(em) - 1) * kTaskThreads;
                    for (int e = base + lane; e < base + kTaskThreads && e < a.num_envs; e += 32) {
                        a.dof_states[(size_t)e * 8 + 3] = -50.0f;
                        a.dof_states[(size_t)e * 8 + 7] = 50.0f;
                    }
                }
            }
        });
}

__global__ void __launch_bounds__(kTaskThreads) ingenuity_post_kernel(const __grid_constant__ IgeIngenuityArgs a) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < a.num_envs; e += gridDim.x * blockDim.x) {
        const long long progress = a.progress_buf[e] + 1;   // :368
        float root[13], tgt[3], obs[13];
#pragma unroll
        for (int i = 0; i < 13; ++i) root[i] = a.root_tensor[(size_t)e * 26 + i];
#pragma unroll
        for (int i = 0; i < 3; ++i) tgt[i] = a.target_root_positions[(size_t)e * 3 + i];
#pragma unroll
        for (int i = 0; i < 3; ++i) obs[i] = (tgt[i] - root[i]) / 3.0f;   // compute_observations :396-401
#pragma unroll
        for (int i = 0; i < 4; ++i) obs[3 + i] = root[3 + i];
#pragma unroll
        for (int i = 0; i < 3; ++i) { obs[7 + i] = root[7 + i] / 2.0f; obs[10 + i] = root[10 + i] / kPiF; }
#pragma unroll
        for (int i = 0; i < 13; ++i) a.obs_buf[(size_t)e * 13 + i] = obs[i];
        if (a.obs_clamped) {
#pragma unroll
            for (int i = 0; i < 13; ++i) a.obs_clamped[(size_t)e * 13 + i] = clamp_obs(obs[i], a.clip_obs);
        }
        const float dx = tgt[0] - root[0], dy = tgt[1] - root[1], dz = tgt[2] - root[2];
        const float target_dist = __fsqrt_rn((dx * dx + dy * dy) + dz * dz);
        const float reward = copter_reward(root, target_dist, 5.0f);
        const bool die = target_dist > 8.0f || root[2] < 0.5f;
        const bool late = (float)progress >= a.max_episode_length - 1.0f;
        const long long reset = (late || die) ? 1 : 0;
        a.rew_buf[e] = reward;
        a.reset_buf[e] = reset;
        a.progress_buf[e] = progress;
        if (a.timeout_buf) a.timeout_buf[e] = (late && reset != 0) ? 1 : 0;
    }
}

template <class Args>
int check_common(const Args *a, const char *name) {
    if (!a) return ige_fail(IGE_ENULL, "%s: args is NULL", name);
    if (a->num_envs < 0) return ige_fail(IGE_ESIZE, "%s: num_envs < 0", name);
    if (!a->reset_buf || !a->progress_buf) return ige_fail(IGE_ENULL, "%s: reset_buf / progress_buf is NULL", name);
    if (a->rand.mode != IGE_RAND_INJECTED && a->rand.mode != IGE_RAND_PHILOX) return ige_fail(IGE_EPARAM, "%s: rand.mode invalid", name);
    if (a->rand.mode == IGE_RAND_INJECTED && !a->rand.rank) return ige_fail(IGE_ENULL, "%s: rand.rank is NULL (INJECTED mode)", name);
    return ige_check_list(a->list);
}
int post_grid(int n) {   // grid-stride kernels: at most 16 CTAs per SM
    const int sms = ige_current_device_sms();
    const int g = (n + kTaskThreads - 1) / kTaskThreads;
    return g < 1 ? 1 : (g > sms * 16 ? sms * 16 : g);
}
// workers + one scanner CTA when the reset list is requested
template <class K, class Args>
int launch_pre(K kernel, const Args *a, void *stream, const char *name) {
    int workers = 0;
    if (int rc = ige_task_grid(kernel, a->num_envs, &workers, name)) return rc;
ige_launch(kernel, workers + (a->list.scratch ? 1 : 0), kTaskThreads, 0, (cudaStream_t)stream, *a, workers);
    return ige_check_launch(name);
}

}  // namespace

extern "C" int ige_ball_balance_pre_physics(const IgeBallBalanceArgs *a, void *stream) {
    if (int rc = check_common(a, "ige_ball_balance_pre_physics")) return rc;
    IGE_REQUIRE_PTR(a->root_states);
    IGE_REQUIRE_PTR(a->dof_states);
    IGE_REQUIRE_PTR(a->initial_root_states);
    IGE_REQUIRE_PTR(a->initial_dof_states);
    IGE_REQUIRE_PTR(a->actions);
    IGE_REQUIRE_PTR(a->dof_position_targets);
    if (a->num_envs == 0) return IGE_OK;
    return launch_pre(ball_balance_pre_kernel, a, stream, "ige_ball_balance_pre_physics");
}
extern "C" int ige_ball_balance_post_physics(const IgeBallBalanceArgs *a, void *stream) {
    if (int rc = check_common(a, "ige_ball_balance_post_physics")) return rc;
    IGE_REQUIRE_PTR(a->root_states);
    IGE_REQUIRE_PTR(a->dof_states);
    IGE_REQUIRE_PTR(a->sensor_tensor);
    IGE_REQUIRE_PTR(a->obs_buf);
    IGE_REQUIRE_PTR(a->rew_buf);
    IGE_OPTIONAL_PTR(a->obs_clamped);
    if (a->num_envs == 0) return IGE_OK;
ige_launch(ball_balance_post_kernel, post_grid(a->num_envs), kTaskThreads, 0, (cudaStream_t)stream, *a);
    return ige_check_launch("ige_ball_balance_post_physics");
}

extern "C" int ige_quadcopter_pre_physics(const IgeQuadcopterArgs *a, void *stream) {
    if (int rc = check_common(a, "ige_quadcopter_pre_physics")) return rc;
    IGE_REQUIRE_PTR(a->root_states);
    IGE_REQUIRE_PTR(a->dof_states);
    IGE_REQUIRE_PTR(a->initial_root_states);
    IGE_REQUIRE_PTR(a->actions);
    IGE_REQUIRE_PTR(a->dof_position_targets);
    IGE_REQUIRE_PTR(a->thrusts);
    IGE_REQUIRE_PTR(a->forces);
    if (a->num_bodies < 1) return ige_fail(IGE_ESIZE, "num_bodies < 1");
    for (int k = 0; k < 4; ++k)
        if (a->rotor_bodies[k] < 0 || a->rotor_bodies[k] >= a->num_bodies) return ige_fail(IGE_ESIZE, "rotor body index out of range");
    if (a->num_envs == 0) return IGE_OK;
    return launch_pre(quadcopter_pre_kernel, a, stream, "ige_quadcopter_pre_physics");
}
extern "C" int ige_quadcopter_post_physics(const IgeQuadcopterArgs *a, void *stream) {
    if (int rc = check_common(a, "ige_quadcopter_post_physics")) return rc;
    IGE_REQUIRE_PTR(a->root_states);
    IGE_REQUIRE_PTR(a->dof_states);
    IGE_REQUIRE_PTR(a->obs_buf);
    IGE_REQUIRE_PTR(a->rew_buf);
    IGE_OPTIONAL_PTR(a->obs_clamped);
    if (a->num_envs == 0) return IGE_OK;
ige_launch(quadcopter_post_kernel, post_grid(a->num_envs), kTaskThreads, 0, (cudaStream_t)stream, *a);
    return ige_check_launch("ige_quadcopter_post_physics");
}

extern "C" int ige_ingenuity_pre_physics(const IgeIngenuityArgs *a, void *stream) {
    if (int rc = check_common(a, "ige_ingenuity_pre_physics")) return rc;
    IGE_REQUIRE_PTR(a->root_tensor);
    IGE_REQUIRE_PTR(a->dof_states);
    IGE_REQUIRE_PTR(a->initial_root_states);
    IGE_REQUIRE_PTR(a->actions);
    IGE_REQUIRE_PTR(a->target_root_positions);
    IGE_REQUIRE_PTR(a->thrusts);
    IGE_REQUIRE_PTR(a->forces);
    IGE_OPTIONAL_PTR(a->actor_dirty);
    if (a->rand.mode == IGE_RAND_INJECTED && !a->target_rank) return ige_fail(IGE_ENULL, "target_rank is NULL (INJECTED mode)");
    if (!a->list.scratch)   // the rotor speeds of all envs depend on whether ANY env resets: the scanner CTA knows
        return ige_fail(IGE_ENULL, "ige_ingenuity_pre_physics needs the reset list (list.scratch)");
    if (a->num_bodies < 4) return ige_fail(IGE_ESIZE, "num_bodies < 4");
    if (a->num_envs == 0) return IGE_OK;
    return launch_pre(ingenuity_pre_kernel, a, stream, "ige_ingenuity_pre_physics");
}
extern "C" int ige_ingenuity_post_physics(const IgeIngenuityArgs *a, void *stream) {
    if (int rc = check_common(a, "ige_ingenuity_post_physics")) return rc;
    IGE_REQUIRE_PTR(a->root_tensor);
    IGE_REQUIRE_PTR(a->target_root_positions);
    IGE_REQUIRE_PTR(a->obs_buf);
    IGE_REQUIRE_PTR(a->rew_buf);
    IGE_OPTIONAL_PTR(a->obs_clamped);
    if (a->num_envs == 0) return IGE_OK;
ige_launch(ingenuity_post_kernel, post_grid(a->num_envs), kTaskThreads, 0, (cudaStream_t)stream, *a);
    return ige_check_launch("ige_ingenuity_post_physics");
}
