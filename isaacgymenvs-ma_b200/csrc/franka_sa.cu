// franka_sa.cu -- the single-agent Franka tasks for sm_100a: FrankaReach, FrankaCubeStack2,
// FrankaCubeStack (one kernel, variant = template parameter) and FrankaCabinet.
//
// Replaces (reference, isaacgymenvs/tasks/):
//   franka_reach.py        post_physics_step :695-704, reset_idx :479-535, _reset_init_cube_state :538-632,
//                          FSM / _update_states :410-458, compute_observations :468-477, compute_franka_reward :877-928
//   franka_cube_stack2.py  post_physics_step :707-760, reset_idx :501-558, _reset_init_cube_state :557-647,
//                          FSM / _update_states :408-466, compute_observations :486-499, compute_franka_reward :774-850
//   franka_cube_stack.py   post_physics_step :657-665, reset_idx :458-514, _reset_init_cube_state :516-599,
//                          _update_states :416-433, compute_observations :444-456, compute_franka_reward :698-747
//   franka_cabinet.py      pre_physics_step :417-425, post_physics_step :427-436, reset_idx :385-415,
//                          compute_observations :352-383 (compute_grasp_transforms :554-567), compute_franka_reward :475-551
// plus the step() epilogue (tasks/base/vec_task.py:396-404).  Every entry point is ONE launch: thread per env over a
// persistent grid (task_common.cuh), the reset of a flagged env done by its own thread before the env's observation
// is formed (the reference resets first, then observes the re-drawn cube), the reset id list by the scanner CTA.
// HBM-bound: 150-400 B per env, no reuse.
#include "task_common.cuh"

namespace {

using ige::kTaskThreads;

__device__ __forceinline__ float clamp_obs(float x, float c) { return fminf(fmaxf(x, -c), c); }
__device__ __forceinline__ float tensor_clampf(float t, float lo, float hi) { return fmaxf(fminf(t, hi), lo); }  // torch_jit_utils.py:228-230

// per-CTA partial sums -> reduce_scratch; the CTA that retires last turns them into means and restores the scratch
template <int NC>
__device__ __forceinline__ void retire_component_means(float (&acc)[NC], float *scratch, float *out, int workers, int num_envs) {
    __shared__ float s_part[NC][kTaskThreads / 32];
    __shared__ bool s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
        const float v = ige::warp_sum(acc[c]);
        if (lane == 0) s_part[c][warp] = v;
    }
    __syncthreads();
    if (tid < NC) {
        float v = 0.0f;
#pragma unroll
        for (int w = 0; w < kTaskThreads / 32; ++w) v += s_part[tid][w];
        atomicAdd(scratch + tid, v);
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        const unsigned prev = atomicAdd(reinterpret_cast<unsigned *>(scratch + NC), 1u);
        s_last = prev == (unsigned)workers - 1u;
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        if (tid < NC) out[tid] = atomicExch(scratch + tid, 0.0f) / (float)num_envs;
        if (tid == 0) *reinterpret_cast<unsigned *>(scratch + NC) = 0u;
    }
}

// axisangle2quat((0, 0, az)) then quat_mul(., (0,0,0,1)): franka_reach.py:18-49, :626-629
__device__ __forceinline__ void yaw_quat(float az, float *q) {
    const float angle = ige::norm3(0.0f, 0.0f, az);
    float aa[4] = {0.0f, 0.0f, 0.0f, 1.0f};
    if (angle > 1e-6f) {
        const float s = sinf(angle / 2.0f);
        aa[0] = 0.0f * s / angle; aa[1] = 0.0f * s / angle; aa[2] = az * s / angle;
        aa[3] = cosf(angle / 2.0f);
    }
    const float ident[4] = {0.0f, 0.0f, 0.0f, 1.0f};
    ige::quat_mul(aa, ident, q);
}

template <int V>
__device__ __forceinline__ float sa_uniform(const IgeFrankaSAArgs &a, uint64_t step, int which, int e, int cols, int j) {
    return ige::task_uniform(a.rand, step, a.draw_slot[which], e, cols, j);
}

// _reset_init_cube_state for one env; which: 0 = cubeB, 1 = cubeA
template <int V>
__device__ __forceinline__ void sample_cube(const IgeFrankaSAArgs &a, uint64_t step, int e, int which, const float *other_xy,
                                            float *row) {
    const int base = which ? 3 : 0;
#pragma unroll
    for (int i = 0; i < 13; ++i) row[i] = 0.0f;
    row[2] = which ? a.cubeA_z_base : a.cubeB_z_base;
    if (a.cube_z_range != 0.0f) row[2] = row[2] + sa_uniform<V>(a, step, base + 0, e, 1, 0) * a.cube_z_range;
    row[6] = 1.0f;
    float ux = sa_uniform<V>(a, step, base + 1, e, 2, 0), uy = sa_uniform<V>(a, step, base + 1, e, 2, 1);
    row[0] = a.cube_xy_center[0] + a.cube_xy_mult * (ux - 0.5f);
    row[1] = a.cube_xy_center[1] + a.cube_xy_mult * (uy - 0.5f);
    if (V == IGE_FRANKA_SA_STACK && which && a.rand.mode == IGE_RAND_PHILOX) {
        // check_valid (franka_cube_stack.py:567-585): re-sample while the cubes would touch; INJECTED mode is handed the
        // accepted sample
        for (int t = 1; t < 100 && ige::norm2(row[0] - other_xy[0], row[1] - other_xy[1]) < a.min_dist; ++t) {
            ux = ige::task_uniform(a.rand, step, 8 + t, e, 2, 0);
            uy = ige::task_uniform(a.rand, step, 8 + t, e, 2, 1);
            row[0] = a.cube_xy_center[0] + a.cube_xy_mult * (ux - 0.5f);
            row[1] = a.cube_xy_center[1] + a.cube_xy_mult * (uy - 0.5f);
        }
    }
    if (a.cube_yaw_mult > 0.0f) yaw_quat(a.cube_yaw_mult * (sa_uniform<V>(a, step, base + 2, e, 1, 0) - 0.5f), row + 3);
}

template <int V>
__global__ void __launch_bounds__(kTaskThreads) franka_sa_kernel(const __grid_constant__ IgeFrankaSAArgs a, const int workers) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    constexpr int NOBS = V == IGE_FRANKA_SA_REACH ? 14 : (V == IGE_FRANKA_SA_STACK2 ? 21 : 19);
    constexpr int NC = IGE_FRANKA_SA_COMPONENTS;
    float comp[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) comp[c] = 0.0f;
    ige::task_reset_skeleton(a.num_envs, workers, a.reset_buf, a.list, a.rand, [&](int e, bool flagged, uint64_t step) {
        long long progress = a.progress_buf[e] + 1;
        long long reset = flagged ? 1 : 0;
        // env-indexed loads that do not depend on the reset: gripper site (+ finger tips), action, issued first
        const float *rb = a.rigid_body_state + (size_t)e * a.num_bodies * 13;
        float eef[7], lf[3], rf[3];
#pragma unroll
        for (int i = 0; i < 7; ++i) eef[i] = __ldg(rb + (size_t)a.eef_body * 13 + i);
        if (V == IGE_FRANKA_SA_STACK) {
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                lf[i] = __ldg(rb + (size_t)a.lf_body * 13 + i);
                rf[i] = __ldg(rb + (size_t)a.rf_body * 13 + i);
            }
        }
        const float agrip = __ldg(a.actions + (size_t)e * 7 + 6);
        float *cubeA_row = a.root_state + ((size_t)e * a.num_actors + a.cubeA_actor) * 13;
        float *cubeB_row = a.root_state + ((size_t)e * a.num_actors + a.cubeB_actor) * 13;
        float *dof = a.dof_state + (size_t)e * 18;
        float cubeA[7], cubeB[3] = {0.0f, 0.0f, 0.0f}, qg[2];
        if (flagged) {  // reset_idx: cube B first, then A, then the arm
            float row[13], bxy[2] = {0.0f, 0.0f};
            if (V != IGE_FRANKA_SA_REACH) {
                sample_cube<V>(a, step, e, 0, bxy, row);
#pragma unroll
                for (int i = 0; i < 13; ++i) {
                    a.init_cubeB_state[(size_t)e * 13 + i] = row[i];
                    cubeB_row[i] = row[i];
                }
                bxy[0] = row[0]; bxy[1] = row[1];
                cubeB[0] = row[0]; cubeB[1] = row[1]; cubeB[2] = row[2];
            }
            sample_cube<V>(a, step, e, 1, bxy, row);
#pragma unroll
            for (int i = 0; i < 13; ++i) {
                a.init_cubeA_state[(size_t)e * 13 + i] = row[i];
                cubeA_row[i] = row[i];
            }
#pragma unroll
            for (int i = 0; i < 7; ++i) cubeA[i] = row[i];
#pragma unroll
            for (int j = 0; j < 9; ++j) {
                float v = a.default_dof_pos[j] + a.dof_noise_mult * (sa_uniform<V>(a, step, 6, e, 9, j) - 0.5f);
                v = tensor_clampf(v, a.dof_lower[j], a.dof_upper[j]);
                if (j >= 7) v = a.default_dof_pos[j];
                reinterpret_cast<float2 *>(dof)[j] = make_float2(v, 0.0f);
                a.pos_control[(size_t)e * 9 + j] = v;
                a.effort_control[(size_t)e * 9 + j] = 0.0f;
                if (j >= 7) qg[j - 7] = v;
            }
            progress = 0; reset = 0;
        } else {
#pragma unroll
            for (int i = 0; i < 7; ++i) cubeA[i] = cubeA_row[i];
            if (V != IGE_FRANKA_SA_REACH) {
#pragma unroll
                for (int i = 0; i < 3; ++i) cubeB[i] = cubeB_row[i];
            }
            qg[0] = dof[14]; qg[1] = dof[16];
        }
        // _update_states
        const float rel[3] = {cubeA[0] - eef[0], cubeA[1] - eef[1], cubeA[2] - eef[2]};
        const float a2b[3] = {cubeB[0] - cubeA[0], cubeB[1] - cubeA[1], cubeB[2] - cubeA[2]};
        const float height = cubeA[2] - a.table_height;
        const float d = ige::norm3(rel[0], rel[1], rel[2]);
        float obs[NOBS];
        float reward;
        if (V == IGE_FRANKA_SA_STACK) {
            int k = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) obs[k++] = cubeA[3 + i];
#pragma unroll
            for (int i = 0; i < 3; ++i) obs[k++] = cubeA[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) obs[k++] = a2b[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) obs[k++] = eef[i];
#pragma unroll
            for (int i = 0; i < 4; ++i) obs[k++] = eef[3 + i];
            obs[k++] = qg[0]; obs[k++] = qg[1];
            // compute_franka_reward franka_cube_stack.py:698-747
            const float target_height = a.cubeB_size + a.cubeA_size / 2.0f;
            const float d_lf = ige::norm3(cubeA[0] - lf[0], cubeA[1] - lf[1], cubeA[2] - lf[2]);
            const float d_rf = ige::norm3(cubeA[0] - rf[0], cubeA[1] - rf[1], cubeA[2] - rf[2]);
            float dist_reward = 1.0f - tanhf(10.0f * ((d + d_lf) + d_rf) / 3.0f);
            const bool lifted = (height - a.cubeA_size) > 0.04f;
            const float d_ab = ige::norm3(a2b[0] + 0.0f, a2b[1] + 0.0f, a2b[2] + (a.cubeA_size + a.cubeB_size) / 2.0f);
            const float align_reward = (1.0f - tanhf(10.0f * d_ab)) * (lifted ? 1.0f : 0.0f);
            dist_reward = fmaxf(dist_reward, align_reward);
            const bool stack = (ige::norm2(a2b[0], a2b[1]) < 0.02f) && (fabsf(height - target_height) < 0.02f) && (d > 0.04f);
            reward = stack ? a.r_stack_scale * 1.0f
                           : (a.r_dist_scale * dist_reward + a.r_lift_scale * (lifted ? 1.0f : 0.0f)) + a.r_align_scale * align_reward;
            if (stack) reset = 1;
        } else {
            // FSM franka_reach.py:410-430 / franka_cube_stack2.py:408-439
            const bool on = d <= (a.cubeA_size / 2.0f) * 0.9f;
            const bool closed = agrip < 0.0f;
            int fsm = 0;
            if (on) fsm = 1;
            if (on && closed) fsm = 2;
            if (V == IGE_FRANKA_SA_STACK2) {
                const bool lifted = height >= (a.cubeB_size + a.cubeA_size * 0.5f);
                if (on && closed && lifted) fsm = 3;
                const bool aligned = ige::norm2(a2b[0], a2b[1]) <= a.cubeA_size * 0.5f;
                if (aligned) fsm = 4;
                const bool super_close = ige::norm3(a2b[0], a2b[1], a2b[2]) < (a.cubeA_size * 0.7071f + a.cubeB_size * 0.5f);
                if (aligned && super_close) fsm = 5;
                if (aligned && super_close && !closed) fsm = 6;
            }
            int k = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) obs[k++] = eef[3 + i];
#pragma unroll
            for (int i = 0; i < 3; ++i) obs[k++] = eef[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) obs[k++] = rel[i];
            obs[k++] = height;
            if (V == IGE_FRANKA_SA_STACK2) {
#pragma unroll
                for (int i = 0; i < 4; ++i) obs[k++] = cubeA[3 + i];
#pragma unroll
                for (int i = 0; i < 3; ++i) obs[k++] = a2b[i];
            }
            obs[k++] = qg[0]; obs[k++] = qg[1];
            obs[k++] = (float)(1 << fsm);   // torch.pow(2.0, FSM)
            const float dist_reward = 1.0f / (0.5f + d * d) * 0.5f;
            const float sf = (float)fsm;
            if (V == IGE_FRANKA_SA_REACH) {  // franka_reach.py:877-928
                const float s0 = (dist_reward + tanhf(agrip * 3.0f)) / 2.0f;
                const float s1 = (dist_reward + tanhf(-agrip * 3.0f)) / 2.0f;
                const float s2 = expf(2.5f * height) - 1.0f;
                reward = fsm == 0 ? s0 : 0.0f;
                reward += s1;
                reward += fsm == 1 ? s1 : 0.0f;
                reward += fsm == 2 ? s2 : 0.0f;
                reward += sf;
                comp[0] += s0; comp[1] += s1; comp[2] += s2; comp[7] += sf; comp[8] += height;
            } else {  // franka_cube_stack2.py:774-850
                const float s0 = dist_reward;
                const float s1 = (dist_reward + fmaxf(tanhf(-agrip * 3.0f), 0.0f)) / 2.0f;
                const float s2 = fsm == 2 ? fminf(height / 0.095f, 1.0f) : 0.0f;
                const float off = (a.cubeA_size + a.cubeB_size) / 2.0f;
                const float tz = a2b[2] + off;
                const float tdist = ige::norm3(a2b[0], a2b[1], tz);
                const float tdist2 = ige::norm3(a2b[0], a2b[1], tz + off);   // the extras block adds the offset once more (:752-756)
                const float s3 = tanhf(5.0f * -tdist) + 1.0f;
                const float s4 = tanhf(6.0f * -tdist) + 1.0f;
                const float s5 = tanhf(agrip * 7.0f) + 1.0f;
                const float s6 = tanhf(7.0f * d);
                reward = fsm == 0 ? s0 : 0.0f;
                reward += fsm == 1 ? s1 : 0.0f;
                reward += fsm == 2 ? s2 : 0.0f;
                reward += fsm == 3 ? s3 : 0.0f;
                reward += fsm == 4 ? s4 : 0.0f;
                reward += fsm == 5 ? s5 : 0.0f;
                reward += fsm == 6 ? s6 + 10.0f : 0.0f;
                reward += sf;
                comp[0] += s0; comp[1] += s1; comp[2] += s2; comp[3] += s3; comp[4] += s4; comp[5] += s5; comp[6] += s6;
                comp[7] += sf; comp[8] += height; comp[9] += tdist2;
            }
            reward = fmaxf(reward, 0.0f);
        }
        const bool late = (float)progress >= a.max_episode_length - 1.0f;
        if (late) reset = 1;
        float *o = a.obs_buf + (size_t)e * NOBS;
#pragma unroll
        for (int i = 0; i < NOBS; ++i) o[i] = obs[i];
        if (a.obs_clamped) {
#pragma unroll
            for (int i = 0; i < NOBS; ++i) a.obs_clamped[(size_t)e * NOBS + i] = clamp_obs(obs[i], a.clip_obs);
        }
        a.rew_buf[e] = reward;
        a.reset_buf[e] = reset;
        a.progress_buf[e] = progress;
        if (a.timeout_buf) a.timeout_buf[e] = (late && reset != 0) ? 1 : 0;   // vec_task.py:396
    });
    if ((int)blockIdx.x < workers && a.reward_components && V != IGE_FRANKA_SA_STACK)
        retire_component_means<NC>(comp, a.reduce_scratch, a.reward_components, workers, a.num_envs);
}

// ------------------------------------------------------------------------------------------------
// FrankaCabinet
// ------------------------------------------------------------------------------------------------
// utils/torch_jit_utils.py:71-77 quat_apply
__device__ __forceinline__ void quat_apply3(const float *q, const float *b, float *out) {
    float t[3], c2[3];
    t[0] = (q[1] * b[2] - q[2] * b[1]) * 2.0f;
    t[1] = (q[2] * b[0] - q[0] * b[2]) * 2.0f;
    t[2] = (q[0] * b[1] - q[1] * b[0]) * 2.0f;
    c2[0] = q[1] * t[2] - q[2] * t[1];
    c2[1] = q[2] * t[0] - q[0] * t[2];
    c2[2] = q[0] * t[1] - q[1] * t[0];
#pragma unroll
    for (int i = 0; i < 3; ++i) out[i] = (b[i] + q[3] * t[i]) + c2[i];
}

__global__ void __launch_bounds__(kTaskThreads) franka_cabinet_pre_kernel(const __grid_constant__ IgeFrankaCabinetArgs a) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    const int total = a.num_envs * 9;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {   // :417-425
        const int e = i / 9, j = i - e * 9;
        float *t = a.franka_dof_targets + (size_t)e * a.num_dofs + j;
        const float v = *t + a.dof_speed_scales[j] * a.target_step * __ldg(a.actions + i) * a.action_scale;
        *t = tensor_clampf(v, a.dof_lower[j], a.dof_upper[j]);
    }
}

__global__ void __launch_bounds__(kTaskThreads) franka_cabinet_post_kernel(const __grid_constant__ IgeFrankaCabinetArgs a,
                                                                           const int workers) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    ige::task_reset_skeleton(a.num_envs, workers, a.reset_buf, a.list, a.rand, [&](int e, bool flagged, uint64_t step) {
        const int D = a.num_dofs;
        long long progress = a.progress_buf[e] + 1;
        long long reset = flagged ? 1 : 0;
        const float *rb = a.rigid_body_states + (size_t)e * a.num_bodies * 13;
        float hand[7], drawer[7], lf[3], rf[3], act[9];
#pragma unroll
        for (int i = 0; i < 7; ++i) {
            hand[i] = __ldg(rb + (size_t)a.hand_body * 13 + i);
            drawer[i] = __ldg(rb + (size_t)a.drawer_body * 13 + i);
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            lf[i] = __ldg(rb + (size_t)a.lfinger_body * 13 + i);
            rf[i] = __ldg(rb + (size_t)a.rfinger_body * 13 + i);
        }
#pragma unroll
        for (int j = 0; j < 9; ++j) act[j] = __ldg(a.actions + (size_t)e * 9 + j);
        float *dof = a.dof_state + (size_t)e * D * 2;
        float pos[9], vel[9], cab_pos, cab_vel;
        if (flagged) {  // reset_idx :385-415
#pragma unroll
            for (int j = 0; j < 9; ++j) {
                float v = a.default_dof_pos[j] + 0.25f * (ige::task_uniform(a.rand, step, 0, e, 9, j) - 0.5f);
                v = tensor_clampf(v, a.dof_lower[j], a.dof_upper[j]);
                pos[j] = v; vel[j] = 0.0f;
                reinterpret_cast<float2 *>(dof)[j] = make_float2(v, 0.0f);
                a.franka_dof_targets[(size_t)e * D + j] = v;
            }
            for (int j = 9; j < D; ++j) reinterpret_cast<float2 *>(dof)[j] = make_float2(0.0f, 0.0f);
            cab_pos = 0.0f; cab_vel = 0.0f;
            const int words = a.num_props * 13;
            float *dst = a.root_state_tensor + ((size_t)e * a.num_actors + 2) * 13;
            const float *src = a.default_prop_states + (size_t)e * words;
            for (int i = 0; i < words; ++i) dst[i] = __ldg(src + i);
            progress = 0; reset = 0;
        } else {
#pragma unroll
            for (int j = 0; j < 9; ++j) {
                const float2 v = reinterpret_cast<const float2 *>(dof)[j];
                pos[j] = v.x; vel[j] = v.y;
            }
            const float2 c = reinterpret_cast<const float2 *>(dof)[12];   // cabinet_dof_*[:, 3]
            cab_pos = c.x; cab_vel = c.y;
        }
        // compute_grasp_transforms :554-567 (tf_combine, torch_jit_utils.py:148-149)
        float fg_rot[4], fg_pos[3], dg_rot[4], dg_pos[3], t3[3];
        ige::quat_mul(hand + 3, a.franka_local_grasp_rot, fg_rot);
        quat_apply3(hand + 3, a.franka_local_grasp_pos, t3);
#pragma unroll
        for (int i = 0; i < 3; ++i) fg_pos[i] = t3[i] + hand[i];
        ige::quat_mul(drawer + 3, a.drawer_local_grasp_rot, dg_rot);
        quat_apply3(drawer + 3, a.drawer_local_grasp_pos, t3);
#pragma unroll
        for (int i = 0; i < 3; ++i) dg_pos[i] = t3[i] + drawer[i];
        if (a.franka_grasp_pos) { for (int i = 0; i < 3; ++i) a.franka_grasp_pos[(size_t)e * 3 + i] = fg_pos[i]; }
        if (a.franka_grasp_rot) { for (int i = 0; i < 4; ++i) a.franka_grasp_rot[(size_t)e * 4 + i] = fg_rot[i]; }
        if (a.drawer_grasp_pos) { for (int i = 0; i < 3; ++i) a.drawer_grasp_pos[(size_t)e * 3 + i] = dg_pos[i]; }
        if (a.drawer_grasp_rot) { for (int i = 0; i < 4; ++i) a.drawer_grasp_rot[(size_t)e * 4 + i] = dg_rot[i]; }
        float obs[23];
#pragma unroll
        for (int j = 0; j < 9; ++j) {   // :378-381
            obs[j] = 2.0f * (pos[j] - a.dof_lower[j]) / (a.dof_upper[j] - a.dof_lower[j]) - 1.0f;
            obs[9 + j] = vel[j] * a.dof_vel_scale;
        }
#pragma unroll
        for (int i = 0; i < 3; ++i) obs[18 + i] = dg_pos[i] - fg_pos[i];
        obs[21] = cab_pos; obs[22] = cab_vel;
        // compute_franka_reward :475-551
        const float d = ige::norm3(fg_pos[0] - dg_pos[0], fg_pos[1] - dg_pos[1], fg_pos[2] - dg_pos[2]);
        float dist_reward = 1.0f / (1.0f + d * d);
        dist_reward *= dist_reward;
        if (d <= 0.02f) dist_reward = dist_reward * 2.0f;
        float ax1[3], ax2[3], ax3[3], ax4[3];
        quat_apply3(fg_rot, a.gripper_forward_axis, ax1);
        quat_apply3(dg_rot, a.drawer_inward_axis, ax2);
        quat_apply3(fg_rot, a.gripper_up_axis, ax3);
        quat_apply3(dg_rot, a.drawer_up_axis, ax4);
        const float dot1 = ige::dot3(ax1, ax2), dot2 = ige::dot3(ax3, ax4);
        const float sg1 = (float)((dot1 > 0.0f) - (dot1 < 0.0f)), sg2 = (float)((dot2 > 0.0f) - (dot2 < 0.0f));
        const float rot_reward = 0.5f * (sg1 * (dot1 * dot1) + sg2 * (dot2 * dot2));
        const bool around = lf[2] > dg_pos[2] && rf[2] < dg_pos[2];
        const float around_handle_reward = around ? 0.0f + 0.5f : 0.0f;
        const float lfd = fabsf(lf[2] - dg_pos[2]), rfd = fabsf(rf[2] - dg_pos[2]);
        const float finger_dist_reward = around ? (0.04f - lfd) + (0.04f - rfd) : 0.0f;
        const float action_penalty = ige::row_sum<9>([&](int j) { return act[j] * act[j]; });
        const float open_reward = cab_pos * around_handle_reward + cab_pos;
        float rew = ((((a.dist_reward_scale * dist_reward + a.rot_reward_scale * rot_reward) +
                       a.around_handle_reward_scale * around_handle_reward) + a.open_reward_scale * open_reward) +
                     a.finger_dist_reward_scale * finger_dist_reward) - a.action_penalty_scale * action_penalty;
        if (cab_pos > 0.01f) rew = rew + 0.5f;
        if (cab_pos > 0.2f) rew = rew + around_handle_reward;
        if (cab_pos > 0.39f) rew = rew + (2.0f * around_handle_reward);
        if (lf[0] < dg_pos[0] - a.distX_offset) rew = -1.0f;
        if (rf[0] < dg_pos[0] - a.distX_offset) rew = -1.0f;
        if (cab_pos > 0.39f) reset = 1;
        const bool late = (float)progress >= a.max_episode_length - 1.0f;
        if (late) reset = 1;
        float *o = a.obs_buf + (size_t)e * 23;
#pragma unroll
        for (int i = 0; i < 23; ++i) o[i] = obs[i];
        if (a.obs_clamped) {
#pragma unroll
            for (int i = 0; i < 23; ++i) a.obs_clamped[(size_t)e * 23 + i] = clamp_obs(obs[i], a.clip_obs);
        }
        a.rew_buf[e] = rew;
        a.reset_buf[e] = reset;
        a.progress_buf[e] = progress;
        if (a.timeout_buf) a.timeout_buf[e] = (late && reset != 0) ? 1 : 0;
    });
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

template <class K, class Args>
int launch_workers(K kernel, const Args *a, void *stream, const char *name) {
    int workers = 0;
    if (int rc = ige_task_grid(kernel, a->num_envs, &workers, name)) return rc;
ige_launch(kernel, workers + (a->list.scratch ? 1 : 0), kTaskThreads, 0, (cudaStream_t)stream, *a, workers);
    return ige_check_launch(name);
}

}  // namespace

extern "C" int ige_franka_sa_post_physics(const IgeFrankaSAArgs *a, void *stream) {
    const char *name = "ige_franka_sa_post_physics";
    if (int rc = check_common(a, name)) return rc;
    if (a->variant < IGE_FRANKA_SA_REACH || a->variant > IGE_FRANKA_SA_STACK) return ige_fail(IGE_EPARAM, "%s: variant invalid", name);
    IGE_REQUIRE_PTR(a->rigid_body_state);
    IGE_REQUIRE_PTR(a->root_state);
    IGE_REQUIRE_PTR(a->dof_state);
    IGE_REQUIRE_PTR(a->actions);
    IGE_REQUIRE_PTR(a->pos_control);
    IGE_REQUIRE_PTR(a->effort_control);
    IGE_REQUIRE_PTR(a->init_cubeA_state);
    if (a->variant != IGE_FRANKA_SA_REACH) IGE_REQUIRE_PTR(a->init_cubeB_state);
    IGE_REQUIRE_PTR(a->obs_buf);
    IGE_REQUIRE_PTR(a->rew_buf);
    IGE_OPTIONAL_PTR(a->obs_clamped);
    if (a->reward_components && !a->reduce_scratch) return ige_fail(IGE_ENULL, "%s: reward_components needs reduce_scratch", name);
    if (a->num_bodies < 1 || a->num_actors < 1) return ige_fail(IGE_ESIZE, "%s: num_bodies / num_actors < 1", name);
    const int bodies[3] = {a->eef_body, a->lf_body, a->rf_body};
    for (int i = 0; i < (a->variant == IGE_FRANKA_SA_STACK ? 3 : 1); ++i)
        if (bodies[i] < 0 || bodies[i] >= a->num_bodies) return ige_fail(IGE_ESIZE, "%s: body handle out of range", name);
    if (a->cubeA_actor < 0 || a->cubeA_actor >= a->num_actors) return ige_fail(IGE_ESIZE, "%s: cubeA_actor out of range", name);
    if (a->variant != IGE_FRANKA_SA_REACH && (a->cubeB_actor < 0 || a->cubeB_actor >= a->num_actors))
        return ige_fail(IGE_ESIZE, "%s: cubeB_actor out of range", name);
    for (int i = 0; i < 7; ++i)
        if (a->draw_slot[i] >= 8) return ige_fail(IGE_EPARAM, "%s: draw_slot[%d] out of range", name, i);
    if (a->num_envs == 0) return IGE_OK;
    switch (a->variant) {
        case IGE_FRANKA_SA_REACH: return launch_workers(franka_sa_kernel<IGE_FRANKA_SA_REACH>, a, stream, name);
        case IGE_FRANKA_SA_STACK2: return launch_workers(franka_sa_kernel<IGE_FRANKA_SA_STACK2>, a, stream, name);
        default: return launch_workers(franka_sa_kernel<IGE_FRANKA_SA_STACK>, a, stream, name);
    }
}

static int cabinet_check(const IgeFrankaCabinetArgs *a, const char *name) {
    if (a->num_dofs < 13) return ige_fail(IGE_ESIZE, "%s: num_dofs < 13 (9 franka + 4 cabinet)", name);
    if (a->num_props < 0 || a->num_actors < 2 + a->num_props) return ige_fail(IGE_ESIZE, "%s: num_actors < 2 + num_props", name);
    const int bodies[4] = {a->hand_body, a->drawer_body, a->lfinger_body, a->rfinger_body};
    for (int i = 0; i < 4; ++i)
        if (bodies[i] < 0 || bodies[i] >= a->num_bodies) return ige_fail(IGE_ESIZE, "%s: body handle out of range", name);
    return IGE_OK;
}

extern "C" int ige_franka_cabinet_pre_physics(const IgeFrankaCabinetArgs *a, void *stream) {
    const char *name = "ige_franka_cabinet_pre_physics";
    if (!a) return ige_fail(IGE_ENULL, "%s: args is NULL", name);
    if (a->num_envs < 0) return ige_fail(IGE_ESIZE, "%s: num_envs < 0", name);
    if (a->num_dofs < 9) return ige_fail(IGE_ESIZE, "%s: num_dofs < 9", name);
    IGE_REQUIRE_PTR(a->actions);
    IGE_REQUIRE_PTR(a->franka_dof_targets);
    if (a->num_envs == 0) return IGE_OK;
    const int sms = ige_current_device_sms();
    int g = (a->num_envs * 9 + kTaskThreads - 1) / kTaskThreads;
    g = g > sms * 16 ? sms * 16 : g;
ige_launch(franka_cabinet_pre_kernel, g, kTaskThreads, 0, (cudaStream_t)stream, *a);
    return ige_check_launch(name);
}

extern "C" int ige_franka_cabinet_post_physics(const IgeFrankaCabinetArgs *a, void *stream) {
    const char *name = "ige_franka_cabinet_post_physics";
    if (int rc = check_common(a, name)) return rc;
    if (int rc = cabinet_check(a, name)) return rc;
    IGE_REQUIRE_PTR(a->rigid_body_states);
    IGE_REQUIRE_PTR(a->root_state_tensor);
    IGE_REQUIRE_PTR(a->dof_state);
    IGE_REQUIRE_PTR(a->actions);
    IGE_REQUIRE_PTR(a->franka_dof_targets);
    if (a->num_props > 0) IGE_REQUIRE_PTR(a->default_prop_states);
    IGE_REQUIRE_PTR(a->obs_buf);
    IGE_REQUIRE_PTR(a->rew_buf);
    IGE_OPTIONAL_PTR(a->obs_clamped);
    if (a->num_envs == 0) return IGE_OK;
    return launch_workers(franka_cabinet_post_kernel, a, stream, name);
}
