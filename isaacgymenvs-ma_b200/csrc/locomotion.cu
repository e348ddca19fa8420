// locomotion.cu -- fused Ant / Humanoid post_physics_step for sm_100a.
//
// Replaces (reference, isaacgymenvs/tasks/):
//   Ant.post_physics_step       ant.py:287-297       Humanoid.post_physics_step  humanoid.py:287-296
//   Ant.reset_idx               ant.py:252-279       Humanoid.reset_idx          humanoid.py:253-279
//   compute_ant_observations    ant.py:374-408       compute_humanoid_observations humanoid.py:378-413
//   compute_ant_reward          ant.py:325-371       compute_humanoid_reward     humanoid.py:323-375
//   + the VecTask.step epilogue base/vec_task.py:396,404 (time_outs, clamped obs) when requested.
//
// Data movement (HBM-bound path, DESIGN.md "Locomotion kernel"): the gymtorch tensors are arrays
// of structs with awkward row sizes (root 52 B, DOF 8 B, sensors 24 B), but a TILE of TE
// consecutive envs is one contiguous, 16-byte aligned span in every tensor.  Each CTA therefore
// moves its tile with a handful of TMA 1-D bulk copies (cp.async.bulk, completion on one
// mbarrier), computes out of shared memory, assembles the (TE x NUM_OBS) observation tile densely
// in shared memory and writes it back with one bulk store.  Nothing is read twice and every
// global access is a full-line burst.
//
// Thread roles inside a CTA (NT threads, TE envs):
//   all threads : reset scatter over (env, dof), the element-wise observation columns
//   threads < TE: per-env quaternion / heading / Euler math, the reward reductions, flags
#include "ige_common.cuh"

namespace {

template <int D, int NS, bool HUM>
struct Layout {
    static constexpr int NUM_OBS = 12 + 3 * D + NS + (HUM ? D : 0);
    static constexpr int O_DOFPOS = 12, O_DOFVEL = 12 + D, O_FORCE = 12 + 2 * D;
    static constexpr int O_SENS = 12 + 2 * D + (HUM ? D : 0), O_ACT = O_SENS + NS;
};

__device__ __forceinline__ float clampf(float x, float c) { return x < -c ? -c : (x > c ? c : x); }

template <typename T>
__device__ __forceinline__ void coop_load(T *dst, const T *src, int count) {
    for (int i = threadIdx.x; i < count; i += blockDim.x) dst[i] = src[i];
}
template <typename T>
__device__ __forceinline__ void coop_store(T *dst, const T *src, int count) {
    for (int i = threadIdx.x; i < count; i += blockDim.x) dst[i] = src[i];
}

template <int D, int NS, bool HUM, int TE, int NT>
__global__ void __launch_bounds__(NT)
locomotion_post_physics_kernel(const __grid_constant__ IgeLocomotionParams p,
                               const __grid_constant__ IgeLocomotionBuffers b, const int use_tma) {
    using L = Layout<D, NS, HUM>;
    constexpr int NUM_OBS = L::NUM_OBS;
    static_assert(TE % 16 == 0 && TE <= NT, "tile must be a multiple of 16 envs and fit the CTA");

    // ---- shared-memory tile (every array starts 16-B aligned: all sizes are multiples of 16 B) ----
    __shared__ __align__(128) float s_obs[TE * NUM_OBS];
    __shared__ __align__(16) float s_root[TE * 13];
    __shared__ __align__(16) float s_dof[TE * D * 2];
    __shared__ __align__(16) float s_sens[TE * NS];
    __shared__ __align__(16) float s_force[HUM ? TE * D : 4];
    __shared__ __align__(16) float s_act[TE * D];
    __shared__ __align__(16) float s_pot[TE];
    __shared__ __align__(16) float s_prev[TE];
    __shared__ __align__(16) float s_rew[TE];
    __shared__ __align__(16) float s_up[TE * 3];
    __shared__ __align__(16) float s_head[TE * 3];
    __shared__ __align__(16) long long s_reset[TE];
    __shared__ __align__(16) long long s_progress[TE];
    __shared__ __align__(16) long long s_randomize[TE];
    __shared__ __align__(8) uint64_t s_bar;
    __shared__ float s_red[6 * (NT / 32)];

    const int tid = threadIdx.x;
    const int e0 = blockIdx.x * TE;
    const int ne = min(TE, p.num_envs - e0);
    const bool bulk = use_tma && (ne % 4 == 0);
    const bool has_rnd = b.randomize_buf != nullptr;
    const uint64_t step = (p.rand.mode == IGE_RAND_PHILOX && p.rand.step) ? __ldg(p.rand.step) : 0ull;

    // ---- 1. tile load ----
    if (bulk) {
        if (tid == 0) {
            ige::mbar_init(&s_bar, 1);
            ige::fence_mbar_init();
        }
        __syncthreads();
        if (tid == 0) {
            const uint32_t une = (uint32_t)ne;
            uint32_t bytes = une * (13 * 4 + D * 8 + NS * 4 + D * 4 + 4 + 8 + 8);
            if (HUM) bytes += une * D * 4;
            if (has_rnd) bytes += une * 8;
            ige::mbar_expect_tx(&s_bar, bytes);
            ige::bulk_g2s(s_root, b.root_states + (size_t)e0 * 13, une * 52, &s_bar);
            ige::bulk_g2s(s_dof, b.dof_state + (size_t)e0 * D * 2, une * D * 8, &s_bar);
            ige::bulk_g2s(s_sens, b.vec_sensor_tensor + (size_t)e0 * NS, une * NS * 4, &s_bar);
            if (HUM) ige::bulk_g2s(s_force, b.dof_force_tensor + (size_t)e0 * D, une * D * 4, &s_bar);
            ige::bulk_g2s(s_act, b.actions + (size_t)e0 * D, une * D * 4, &s_bar);
            ige::bulk_g2s(s_pot, b.potentials + e0, une * 4, &s_bar);
            ige::bulk_g2s(s_reset, b.reset_buf + e0, une * 8, &s_bar);
            ige::bulk_g2s(s_progress, b.progress_buf + e0, une * 8, &s_bar);
            if (has_rnd) ige::bulk_g2s(s_randomize, b.randomize_buf + e0, une * 8, &s_bar);
        }
        ige::mbar_wait(&s_bar, 0);
    } else {
        coop_load(s_root, b.root_states + (size_t)e0 * 13, ne * 13);
        coop_load(s_dof, b.dof_state + (size_t)e0 * D * 2, ne * D * 2);
        coop_load(s_sens, b.vec_sensor_tensor + (size_t)e0 * NS, ne * NS);
        if (HUM) coop_load(s_force, b.dof_force_tensor + (size_t)e0 * D, ne * D);
        coop_load(s_act, b.actions + (size_t)e0 * D, ne * D);
        coop_load(s_pot, b.potentials + e0, ne);
        coop_load(s_reset, (const long long *)b.reset_buf + e0, ne);
        coop_load(s_progress, (const long long *)b.progress_buf + e0, ne);
        if (has_rnd) coop_load(s_randomize, (const long long *)b.randomize_buf + e0, ne);
        __syncthreads();
    }

    // ---- 2. reset_idx: scatter new DOF rows (ant.py:257-261), all threads over (env, dof) ----
    for (int i = tid; i < ne * D; i += NT) {
        const int le = i / D, j = i - le * D;
        if (s_reset[le] != 0) {
            const int e = e0 + le;
            // torch_rand_float(lo, hi): (hi - lo) * rand + lo (torch_jit_utils.py:215-218)
            const float pos = 0.4f * ige::reset_uniform(p.rand, step, 0, e, D, j) + (-0.2f);
            const float vel = 0.2f * ige::reset_uniform(p.rand, step, 1, e, D, j) + (-0.1f);
            float v = __ldg(b.initial_dof_pos + (size_t)e * D + j) + pos;
            v = fminf(v, p.dof_limits_upper[j]);  // tensor_clamp torch_jit_utils.py:228-230
            v = fmaxf(v, p.dof_limits_lower[j]);
            const float2 row = make_float2(v, vel);
            reinterpret_cast<float2 *>(s_dof)[i] = row;
            reinterpret_cast<float2 *>(b.dof_state)[(size_t)e0 * D + i] = row;  // ant.py:260-261
        }
    }
    __syncthreads();

    // ---- 3a. element-wise observation columns, all threads ----
    for (int i = tid; i < ne * D; i += NT) {
        const int le = i / D, j = i - le * D;
        const float2 pv = reinterpret_cast<const float2 *>(s_dof)[i];
        const float lo = p.dof_limits_lower[j], hi = p.dof_limits_upper[j];
        float *o = s_obs + le * NUM_OBS;
        o[L::O_DOFPOS + j] = (2.0f * pv.x - hi - lo) / (hi - lo);  // unscale torch_jit_utils.py:238-240
        o[L::O_DOFVEL + j] = pv.y * p.dof_vel_scale;
        if (HUM) o[L::O_FORCE + j] = s_force[i] * p.contact_force_scale;
        o[L::O_ACT + j] = s_act[i];
    }
    for (int i = tid; i < ne * NS; i += NT) {
        const int le = i / NS, j = i - le * NS;
        s_obs[le * NUM_OBS + L::O_SENS + j] = s_sens[i] * p.contact_force_scale;
    }

    // ---- 3b. per-env torso math, threads < ne ----
    float potentials = 0.f, prev_potentials = 0.f;
    long long progress = 0, reset = 0;
    if (tid < ne) {
        const int e = e0 + tid;
        const float *root = s_root + tid * 13;
        progress = s_progress[tid] + 1;  // ant.py:288
        reset = s_reset[tid];
        potentials = s_pot[tid];
        if (reset != 0) {  // ant.py:273-279
            const float ix = __ldg(b.initial_root_states + (size_t)e * 13 + 0);
            const float iy = __ldg(b.initial_root_states + (size_t)e * 13 + 1);
            const float nn = -ige::norm3(p.targets[0] - ix, p.targets[1] - iy, 0.0f);
            potentials = p.potentials_recip_mul ? nn * (1.0f / p.dt) : nn / p.dt;
            progress = 0;
            reset = 0;
        }
        // compute_*_observations (ant.py:382-391)
        const float to_target[3] = {p.targets[0] - root[0], p.targets[1] - root[1], 0.0f};
        prev_potentials = potentials;
        {
            const float nn = -ige::norm3(to_target[0], to_target[1], 0.0f);
            potentials = p.potentials_recip_mul ? nn * (1.0f / p.dt) : nn / p.dt;
        }
        // compute_heading_and_up torch_jit_utils.py:247-262
        float target_dirs[3], torso_quat[4], up_vec[3], heading_vec[3];
        ige::normalize3(to_target, target_dirs);
        ige::quat_mul(root + 3, p.inv_start_rot, torso_quat);
        ige::quat_rotate<false>(torso_quat, p.basis_vec1, up_vec);
        ige::quat_rotate<false>(torso_quat, p.basis_vec0, heading_vec);
        const float heading_proj = ige::dot3(heading_vec, target_dirs);
        // compute_rot torch_jit_utils.py:265-276
        float vel_loc[3], angvel_loc[3], roll, yaw;
        ige::quat_rotate<true>(torso_quat, root + 7, vel_loc);
        ige::quat_rotate<true>(torso_quat, root + 10, angvel_loc);
        ige::euler_roll_yaw(torso_quat, roll, yaw);
        const float walk_target_angle = atan2f(p.targets[2] - root[2], p.targets[0] - root[0]);
        const float angle_to_target = walk_target_angle - yaw;

        float *o = s_obs + tid * NUM_OBS;
        o[0] = root[2];
        o[1] = vel_loc[0]; o[2] = vel_loc[1]; o[3] = vel_loc[2];
        if (HUM) {
            o[4] = angvel_loc[0] * p.angular_velocity_scale;
            o[5] = angvel_loc[1] * p.angular_velocity_scale;
            o[6] = angvel_loc[2] * p.angular_velocity_scale;
            o[7] = ige::normalize_angle(yaw);  // humanoid.py:402-404
            o[8] = ige::normalize_angle(roll);
            o[9] = ige::normalize_angle(angle_to_target);
        } else {
            o[4] = angvel_loc[0]; o[5] = angvel_loc[1]; o[6] = angvel_loc[2];
            o[7] = yaw; o[8] = roll; o[9] = angle_to_target;  // ant.py:403
        }
        o[10] = up_vec[2];
        o[11] = heading_proj;
        s_up[tid * 3 + 0] = up_vec[0]; s_up[tid * 3 + 1] = up_vec[1]; s_up[tid * 3 + 2] = up_vec[2];
        s_head[tid * 3 + 0] = heading_vec[0]; s_head[tid * 3 + 1] = heading_vec[1]; s_head[tid * 3 + 2] = heading_vec[2];
    }
    __syncthreads();

    // ---- 4. reward + termination (ant.py:325-371, humanoid.py:323-375), threads < ne ----
    float st_n = 0.f, st_rew = 0.f, st_done = 0.f, st_to = 0.f, st_len = 0.f, st_obj = 0.f;
    if (tid < ne) {
        const float *o = s_obs + tid * NUM_OBS;
        const float heading_reward = o[11] > 0.8f ? p.heading_weight : p.heading_weight * o[11] / 0.8f;
        const float up_reward = o[10] > 0.93f ? 0.0f + p.up_weight : 0.0f;
        const float actions_cost = ige::row_sum<D>([&](int j) { const float a = o[L::O_ACT + j]; return a * a; });
        const float progress_reward = potentials - prev_potentials;
        float total;
        if (HUM) {
            const float dof_at_limit_cost = ige::row_sum<D>([&](int j) {
                const float a = fabsf(o[L::O_DOFPOS + j]);
                const float scaled = p.joints_at_limit_cost_scale * (a - 0.98f) / 0.02f;  // humanoid.py:356
                return (a > 0.98f ? 1.0f : 0.0f) * scaled * p.motor_effort_ratio[j];
            });
            const float electricity_cost = ige::row_sum<D>([&](int j) {
                return fabsf(o[L::O_ACT + j] * o[L::O_DOFVEL + j]) * p.motor_effort_ratio[j];
            });
            total = progress_reward + 2.0f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                    p.energy_cost_scale * electricity_cost - dof_at_limit_cost;
        } else {
            const float electricity_cost =
                ige::row_sum<D>([&](int j) { return fabsf(o[L::O_ACT + j] * o[L::O_DOFVEL + j]); });
            int at_limit = 0;
#pragma unroll
            for (int j = 0; j < D; ++j) at_limit += o[L::O_DOFPOS + j] > 0.99f;
            total = progress_reward + 0.5f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                    p.energy_cost_scale * electricity_cost - (float)at_limit * p.joints_at_limit_cost_scale;
        }
        const bool fallen = o[0] < p.termination_height;
        const bool late = (float)progress >= p.max_episode_length - 1.0f;
        if (fallen) total = p.death_cost;
        if (fallen || late) reset = 1;
        s_rew[tid] = total;
        s_pot[tid] = potentials;
        s_prev[tid] = prev_potentials;
        s_reset[tid] = reset;
        s_progress[tid] = progress;
        if (has_rnd) s_randomize[tid] += 1;  // ant.py:289
        const bool timeout = late && reset != 0;  // vec_task.py:396
        if (b.timeout_buf) b.timeout_buf[e0 + tid] = timeout ? 1 : 0;
        st_n = 1.f; st_rew = total; st_done = reset != 0 ? 1.f : 0.f; st_to = timeout ? 1.f : 0.f;
        st_len = reset != 0 ? (float)progress : 0.f; st_obj = s_root[tid * 13 + 7];
    }

    // ---- 5. tile store ----
    const bool clamp_copy = b.obs_clamped != nullptr && p.clip_obs < INFINITY;
    if (bulk) {
        ige::fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            const uint32_t une = (uint32_t)ne;
            ige::bulk_s2g(b.obs_buf + (size_t)e0 * NUM_OBS, s_obs, une * NUM_OBS * 4);
            if (b.obs_clamped && !clamp_copy) ige::bulk_s2g(b.obs_clamped + (size_t)e0 * NUM_OBS, s_obs, une * NUM_OBS * 4);
            ige::bulk_s2g(b.rew_buf + e0, s_rew, une * 4);
            ige::bulk_s2g(b.potentials + e0, s_pot, une * 4);
            ige::bulk_s2g(b.prev_potentials + e0, s_prev, une * 4);
            ige::bulk_s2g(b.up_vec + (size_t)e0 * 3, s_up, une * 12);
            ige::bulk_s2g(b.heading_vec + (size_t)e0 * 3, s_head, une * 12);
            ige::bulk_s2g(b.reset_buf + e0, s_reset, une * 8);
            ige::bulk_s2g(b.progress_buf + e0, s_progress, une * 8);
            if (has_rnd) ige::bulk_s2g(b.randomize_buf + e0, s_randomize, une * 8);
            ige::bulk_commit();
            ige::bulk_wait_read_all();
        }
        if (clamp_copy) {
            __syncthreads();  // thread 0 has drained the reads of s_obs
            for (int i = tid; i < ne * NUM_OBS; i += NT) s_obs[i] = clampf(s_obs[i], p.clip_obs);
            ige::fence_proxy_async();
            __syncthreads();
            if (tid == 0) {
                ige::bulk_s2g(b.obs_clamped + (size_t)e0 * NUM_OBS, s_obs, (uint32_t)ne * NUM_OBS * 4);
                ige::bulk_commit();
                ige::bulk_wait_read_all();
            }
        }
    } else {
        __syncthreads();
        coop_store(b.obs_buf + (size_t)e0 * NUM_OBS, s_obs, ne * NUM_OBS);
        if (b.obs_clamped)
            for (int i = tid; i < ne * NUM_OBS; i += NT) b.obs_clamped[(size_t)e0 * NUM_OBS + i] = clampf(s_obs[i], p.clip_obs);
        coop_store(b.rew_buf + e0, s_rew, ne);
        coop_store(b.potentials + e0, s_pot, ne);
        coop_store(b.prev_potentials + e0, s_prev, ne);
        coop_store(b.up_vec + (size_t)e0 * 3, s_up, ne * 3);
        coop_store(b.heading_vec + (size_t)e0 * 3, s_head, ne * 3);
        coop_store((long long *)b.reset_buf + e0, s_reset, ne);
        coop_store((long long *)b.progress_buf + e0, s_progress, ne);
        if (has_rnd) coop_store((long long *)b.randomize_buf + e0, s_randomize, ne);
    }

    // ---- 6. episode statistics (one atomic per slot per CTA) ----
    if (b.stats) {
        const float v[6] = {st_n, st_rew, st_done, st_to, st_len, st_obj};
        const int slot[6] = {IGE_STAT_STEPS, IGE_STAT_REWARD, IGE_STAT_DONE, IGE_STAT_TIMEOUT, IGE_STAT_DONE_LENGTH,
                             IGE_STAT_TASK0};
        ige::block_stats_add<6>(b.stats, v, slot, s_red);
    }
}

int g_use_tma = -1;
int use_tma() {
    if (g_use_tma < 0) {
        const char *e = getenv("IGE_NO_TMA");
        g_use_tma = (e && e[0] == '1') ? 0 : 1;
    }
    return g_use_tma;
}

int check(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, int d, int ns, bool hum) {
    if (!p || !b) return ige_fail(IGE_ENULL, "params / buffers is NULL");
    if (p->num_envs < 0) return ige_fail(IGE_ESIZE, "num_envs < 0");
    if (p->num_dof != d || p->num_sensor_floats != ns)
        return ige_fail(IGE_ESIZE, "expected num_dof=%d num_sensor_floats=%d, got %d / %d", d, ns, p->num_dof,
                        p->num_sensor_floats);
    IGE_REQUIRE_PTR(b->root_states);
    IGE_REQUIRE_PTR(b->dof_state);
    IGE_REQUIRE_PTR(b->vec_sensor_tensor);
    if (hum) IGE_REQUIRE_PTR(b->dof_force_tensor);
    IGE_REQUIRE_PTR(b->actions);
    IGE_REQUIRE_PTR(b->initial_root_states);
    IGE_REQUIRE_PTR(b->initial_dof_pos);
    IGE_REQUIRE_PTR(b->potentials);
    IGE_REQUIRE_PTR(b->prev_potentials);
    IGE_REQUIRE_PTR(b->up_vec);
    IGE_REQUIRE_PTR(b->heading_vec);
    IGE_REQUIRE_PTR(b->obs_buf);
    IGE_REQUIRE_PTR(b->rew_buf);
    IGE_REQUIRE_PTR(b->reset_buf);
    IGE_REQUIRE_PTR(b->progress_buf);
    IGE_OPTIONAL_PTR(b->randomize_buf);
    IGE_OPTIONAL_PTR(b->obs_clamped);
    return ige_check_rand(p->rand);
}

}  // namespace

extern "C" void ige_set_use_tma(int on) { g_use_tma = on ? 1 : 0; }

extern "C" int ige_ant_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, void *stream) {
    if (int rc = check(p, b, 8, 24, false)) return rc;
    if (p->num_envs == 0) return IGE_OK;
    constexpr int TE = 32, NT = 128;
    const int grid = (p->num_envs + TE - 1) / TE;
    locomotion_post_physics_kernel<8, 24, false, TE, NT><<<grid, NT, 0, (cudaStream_t)stream>>>(*p, *b, use_tma());
    return ige_check_launch("ige_ant_post_physics");
}

extern "C" int ige_humanoid_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, void *stream) {
    if (int rc = check(p, b, 21, 12, true)) return rc;
    if (p->num_envs == 0) return IGE_OK;
    constexpr int TE = 32, NT = 128;
    const int grid = (p->num_envs + TE - 1) / TE;
    locomotion_post_physics_kernel<21, 12, true, TE, NT><<<grid, NT, 0, (cudaStream_t)stream>>>(*p, *b, use_tma());
    return ige_check_launch("ige_humanoid_post_physics");
}
