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
// of structs with awkward row sizes (root 52 B, DOF 8 B, sensors 24 B), but a TILE of 32
// consecutive envs is one contiguous, 16-byte aligned span in every tensor.  Each CTA moves its
// tile with a handful of TMA 1-D bulk copies (cp.async.bulk, completion on one mbarrier),
// computes out of shared memory, assembles the (32 x NUM_OBS) observation tile densely in shared
// memory and writes it back with one bulk store.  Nothing is read twice; every global access is a
// full-line burst.
//
// Work split (warp-specialised; lane == env inside the tile for the torso roles).  Roles rotate
// with blockIdx so that every SM sub-partition sees the same instruction mix:
//   HEADING  : torso quat, up / heading vectors, heading projection, potentials (+ reset potentials)
//   VELOCITY : local linear / angular velocity (two inverse rotations), torso height
//   ANGLES   : Euler roll / yaw, walk-target angle (the three atan2f)
//   ELEM x k : the element-wise observation columns (unscale, scales, copies), float4 when the
//              layout allows (Ant), plus the per-DOF reward terms into a padded transposed scratch
// After one CTA barrier the warp that ran VELOCITY finishes the tile: reward reductions in the
// reference's summation order, termination / progress flags, time-outs, episode statistics, and
// issues the bulk stores.  The other warps retire immediately.
#include "ige_common.cuh"

namespace {

constexpr int TE = 32;  // envs per CTA tile == warp size (lane == env for the torso roles)

template <int D, int NS, bool HUM>
struct Layout {
    static constexpr int NUM_OBS = 12 + 3 * D + NS + (HUM ? D : 0);
    static constexpr int O_DOFPOS = 12, O_DOFVEL = 12 + D, O_FORCE = 12 + 2 * D;
    static constexpr int O_SENS = 12 + 2 * D + (HUM ? D : 0), O_ACT = O_SENS + NS;
    // float4 path: every element-wise segment starts 16-B aligned in the obs row and has a
    // multiple-of-4 length (Ant: 12, 20, 28, 52 with D = 8, NS = 24).
    static constexpr bool VEC4 = !HUM && (D % 4 == 0) && (NS % 4 == 0) && (NUM_OBS % 4 == 0);
};

__device__ __forceinline__ float clampf(float x, float c) { return x < -c ? -c : (x > c ? c : x); }

template <typename T>
__device__ __forceinline__ void coop_copy(T *dst, const T *src, int count) {
    for (int i = threadIdx.x; i < count; i += blockDim.x) dst[i] = src[i];
}

// atan2(sin x, cos x) (torch_jit_utils.py:126-128) for x in (-3*pi, 2*pi]: the exact wrap into
// (-pi, pi].  Differs from the libm composition by < 1e-6 abs (tests/test_gpu_parity.py bound).
__device__ __forceinline__ float wrap_pi(float x) {
    constexpr float kPi = 3.14159274101257324219f;
    if (x >= kPi) return x - ige::kTwoPi;
    if (x < -kPi) return x + ige::kTwoPi;
    return x;
}

enum Role { HEADING = 0, VELOCITY = 1, ANGLES = 2, ELEM0 = 3 };

template <int D, int NS, bool HUM, int NW_ELEM>
__global__ void __launch_bounds__(32 * (3 + NW_ELEM))
locomotion_post_physics_kernel(const __grid_constant__ IgeLocomotionParams p,
                               const __grid_constant__ IgeLocomotionBuffers b, const int use_tma) {
    using L = Layout<D, NS, HUM>;
    constexpr int NUM_OBS = L::NUM_OBS;
    constexpr int NW = 3 + NW_ELEM;
    constexpr int NT = 32 * NW;
    constexpr int TS = TE + 1;  // padded stride of the transposed reward-term scratch

    // ---- shared-memory tile (every bulk-copied array starts 16-B aligned, sizes are multiples of 16 B) ----
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
    __shared__ float s_t0[D * TS], s_t1[D * TS], s_t2[D * TS];  // reward terms [dof][env], padded
    __shared__ __align__(8) uint64_t s_bar;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int role = (warp + (int)blockIdx.x) % NW;
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
        coop_copy(s_root, b.root_states + (size_t)e0 * 13, ne * 13);
        coop_copy(s_dof, (const float *)b.dof_state + (size_t)e0 * D * 2, ne * D * 2);
        coop_copy(s_sens, b.vec_sensor_tensor + (size_t)e0 * NS, ne * NS);
        if (HUM) coop_copy(s_force, b.dof_force_tensor + (size_t)e0 * D, ne * D);
        coop_copy(s_act, b.actions + (size_t)e0 * D, ne * D);
        coop_copy(s_pot, (const float *)b.potentials + e0, ne);
        coop_copy(s_reset, (const long long *)b.reset_buf + e0, ne);
        coop_copy(s_progress, (const long long *)b.progress_buf + e0, ne);
        if (has_rnd) coop_copy(s_randomize, (const long long *)b.randomize_buf + e0, ne);
        __syncthreads();
    }

    // ---- 2. reset_idx: new DOF rows for flagged envs (ant.py:257-261); skipped for clean tiles ----
    const int my_flag = (tid < ne) && (s_reset[tid] != 0);
    if (__syncthreads_or(my_flag)) {
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
    }

    // ---- 3. roles ----
    if (role == HEADING) {
        if (lane < ne) {
            const float *root = s_root + lane * 13;
            float potentials = s_pot[lane];
            if (s_reset[lane] != 0) {  // ant.py:273-276
                const float *init = b.initial_root_states + (size_t)(e0 + lane) * 13;
                const float nn = -ige::norm3(p.targets[0] - __ldg(init), p.targets[1] - __ldg(init + 1), 0.0f);
                potentials = p.potentials_recip_mul ? nn * (1.0f / p.dt) : nn / p.dt;
            }
            // compute_*_observations ant.py:384-391
            const float to_target[3] = {p.targets[0] - root[0], p.targets[1] - root[1], 0.0f};
            s_prev[lane] = potentials;
            float n = ige::norm3(to_target[0], to_target[1], 0.0f);
            s_pot[lane] = p.potentials_recip_mul ? (-n) * (1.0f / p.dt) : (-n) / p.dt;
            // compute_heading_and_up torch_jit_utils.py:247-262
            n = n < 1e-9f ? 1e-9f : n;  // normalize(), torch_jit_utils.py:65-67
            const float target_dirs[3] = {to_target[0] / n, to_target[1] / n, to_target[2] / n};
            float torso_quat[4], up_vec[3], heading_vec[3];
            ige::quat_mul(root + 3, p.inv_start_rot, torso_quat);
            ige::quat_rotate<false>(torso_quat, p.basis_vec1, up_vec);
            ige::quat_rotate<false>(torso_quat, p.basis_vec0, heading_vec);
            float *o = s_obs + lane * NUM_OBS;
            o[10] = up_vec[2];
            o[11] = ige::dot3(heading_vec, target_dirs);
            s_up[lane * 3 + 0] = up_vec[0]; s_up[lane * 3 + 1] = up_vec[1]; s_up[lane * 3 + 2] = up_vec[2];
            s_head[lane * 3 + 0] = heading_vec[0]; s_head[lane * 3 + 1] = heading_vec[1];
            s_head[lane * 3 + 2] = heading_vec[2];
        }
    } else if (role == VELOCITY) {
        if (lane < ne) {
            const float *root = s_root + lane * 13;
            float torso_quat[4], vel_loc[3], angvel_loc[3];
            ige::quat_mul(root + 3, p.inv_start_rot, torso_quat);
            ige::quat_rotate<true>(torso_quat, root + 7, vel_loc);  // compute_rot torch_jit_utils.py:265-267
            ige::quat_rotate<true>(torso_quat, root + 10, angvel_loc);
            float *o = s_obs + lane * NUM_OBS;
            o[0] = root[2];
            o[1] = vel_loc[0]; o[2] = vel_loc[1]; o[3] = vel_loc[2];
            if (HUM) {
                o[4] = angvel_loc[0] * p.angular_velocity_scale;
                o[5] = angvel_loc[1] * p.angular_velocity_scale;
                o[6] = angvel_loc[2] * p.angular_velocity_scale;
            } else {
                o[4] = angvel_loc[0]; o[5] = angvel_loc[1]; o[6] = angvel_loc[2];
            }
        }
    } else if (role == ANGLES) {
        if (lane < ne) {
            const float *root = s_root + lane * 13;
            float torso_quat[4], roll, yaw;
            ige::quat_mul(root + 3, p.inv_start_rot, torso_quat);
            ige::euler_roll_yaw(torso_quat, roll, yaw);  // torch_jit_utils.py:269
            const float walk_target_angle = atan2f(p.targets[2] - root[2], p.targets[0] - root[0]);
            const float angle_to_target = walk_target_angle - yaw;
            float *o = s_obs + lane * NUM_OBS;
            if (HUM) {  // humanoid.py:402-404 normalize_angle
                o[7] = wrap_pi(yaw); o[8] = wrap_pi(roll); o[9] = wrap_pi(angle_to_target);
            } else {  // ant.py:403
                o[7] = yaw; o[8] = roll; o[9] = angle_to_target;
            }
        }
    } else {
        // ---- element-wise columns + per-DOF reward terms ----
        const int share = role - ELEM0;
        constexpr int STRIDE = 32 * NW_ELEM;
        if (L::VEC4) {
            // DOF groups of 4: two float4 loads -> (pos4, vel4)
            for (int it = share * 32 + lane; it < ne * (D / 4); it += STRIDE) {
                const int le = it / (D / 4), g = it - le * (D / 4);
                const float4 a = reinterpret_cast<const float4 *>(s_dof)[it * 2];      // p0 v0 p1 v1
                const float4 c = reinterpret_cast<const float4 *>(s_dof)[it * 2 + 1];  // p2 v2 p3 v3
                const float4 act = reinterpret_cast<const float4 *>(s_act)[it];
                const float pos[4] = {a.x, a.z, c.x, c.z}, vel[4] = {a.y, a.w, c.y, c.w};
                const float av[4] = {act.x, act.y, act.z, act.w};
                float ps[4], vs[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const int j = g * 4 + k;
                    const float lo = p.dof_limits_lower[j], hi = p.dof_limits_upper[j];
                    ps[k] = (2.0f * pos[k] - hi - lo) / (hi - lo);  // unscale torch_jit_utils.py:238-240
                    vs[k] = vel[k] * p.dof_vel_scale;
                    s_t0[j * TS + le] = av[k] * av[k];              // actions ** 2        ant.py:353
                    s_t1[j * TS + le] = fabsf(av[k] * vs[k]);       // |a * dof_vel_obs|   ant.py:354
                    s_t2[j * TS + le] = ps[k] > 0.99f ? 1.0f : 0.0f;  // dof at limit      ant.py:355
                }
                float *o = s_obs + le * NUM_OBS;
                reinterpret_cast<float4 *>(o + L::O_DOFPOS)[g] = make_float4(ps[0], ps[1], ps[2], ps[3]);
                reinterpret_cast<float4 *>(o + L::O_DOFVEL)[g] = make_float4(vs[0], vs[1], vs[2], vs[3]);
                reinterpret_cast<float4 *>(o + L::O_ACT)[g] = act;
            }
            for (int it = share * 32 + lane; it < ne * (NS / 4); it += STRIDE) {
                const int le = it / (NS / 4), g = it - le * (NS / 4);
                float4 s = reinterpret_cast<const float4 *>(s_sens)[it];
                s.x *= p.contact_force_scale; s.y *= p.contact_force_scale;
                s.z *= p.contact_force_scale; s.w *= p.contact_force_scale;
                reinterpret_cast<float4 *>(s_obs + le * NUM_OBS + L::O_SENS)[g] = s;
            }
        } else {
            for (int it = share * 32 + lane; it < ne * D; it += STRIDE) {
                const int le = it / D, j = it - le * D;
                const float2 pv = reinterpret_cast<const float2 *>(s_dof)[it];
                const float lo = p.dof_limits_lower[j], hi = p.dof_limits_upper[j];
                const float ps = (2.0f * pv.x - hi - lo) / (hi - lo);
                const float vs = pv.y * p.dof_vel_scale;
                const float a = s_act[it];
                float *o = s_obs + le * NUM_OBS;
                o[L::O_DOFPOS + j] = ps;
                o[L::O_DOFVEL + j] = vs;
                o[L::O_ACT + j] = a;
                s_t0[j * TS + le] = a * a;
                if (HUM) {
                    o[L::O_FORCE + j] = s_force[it] * p.contact_force_scale;
                    const float ratio = p.motor_effort_ratio[j];
                    s_t1[j * TS + le] = fabsf(a * vs) * ratio;  // humanoid.py:359
                    const float ap = fabsf(ps);
                    const float scaled = p.joints_at_limit_cost_scale * (ap - 0.98f) / 0.02f;  // humanoid.py:356
                    s_t2[j * TS + le] = (ap > 0.98f ? 1.0f : 0.0f) * scaled * ratio;            // humanoid.py:357
                } else {
                    s_t1[j * TS + le] = fabsf(a * vs);
                    s_t2[j * TS + le] = ps > 0.99f ? 1.0f : 0.0f;
                }
            }
            for (int it = share * 32 + lane; it < ne * NS; it += STRIDE) {
                const int le = it / NS, j = it - le * NS;
                s_obs[le * NUM_OBS + L::O_SENS + j] = s_sens[it] * p.contact_force_scale;
            }
        }
    }
    ige::fence_proxy_async();  // obs tile -> visible to the bulk-store (async) proxy
    __syncthreads();

    // optional clamped second copy with a finite clip (vec_task.py:404); inf clip reuses the tile
    const bool clamp_copy = b.obs_clamped != nullptr && p.clip_obs < INFINITY;
    if (clamp_copy)
        for (int i = tid; i < ne * NUM_OBS; i += NT)
            b.obs_clamped[(size_t)e0 * NUM_OBS + i] = clampf(s_obs[i], p.clip_obs);

    // ---- 4. reward + termination (ant.py:325-371, humanoid.py:323-375): one warp, lane == env ----
    if (role == VELOCITY) {
        float st_rew = 0.f, st_len = 0.f, st_obj = 0.f;
        int st_done = 0, st_to = 0;
        if (lane < ne) {
            const float *o = s_obs + lane * NUM_OBS;
            const float z = s_root[lane * 13 + 2];
            long long progress = s_progress[lane] + 1;  // ant.py:288
            long long reset = s_reset[lane];
            if (reset != 0) { progress = 0; reset = 0; }  // ant.py:278-279
            const float heading_reward = o[11] > 0.8f ? p.heading_weight : p.heading_weight * o[11] / 0.8f;
            const float up_reward = o[10] > 0.93f ? 0.0f + p.up_weight : 0.0f;
            const float actions_cost = ige::row_sum<D>([&](int j) { return s_t0[j * TS + lane]; });
            const float electricity_cost = ige::row_sum<D>([&](int j) { return s_t1[j * TS + lane]; });
            const float progress_reward = s_pot[lane] - s_prev[lane];
            float total;
            if (HUM) {
                const float dof_at_limit_cost = ige::row_sum<D>([&](int j) { return s_t2[j * TS + lane]; });
                total = progress_reward + 2.0f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                        p.energy_cost_scale * electricity_cost - dof_at_limit_cost;
            } else {
                float at_limit = 0.0f;  // integer count <= 8: exact in fp32 (ant.py:355 sums int64)
#pragma unroll
                for (int j = 0; j < D; ++j) at_limit += s_t2[j * TS + lane];
                total = progress_reward + 0.5f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                        p.energy_cost_scale * electricity_cost - at_limit * p.joints_at_limit_cost_scale;
            }
            const bool fallen = z < p.termination_height;
            const bool late = (float)progress >= p.max_episode_length - 1.0f;
            if (fallen) total = p.death_cost;
            if (fallen || late) reset = 1;
            s_rew[lane] = total;
            s_reset[lane] = reset;
            s_progress[lane] = progress;
            if (has_rnd) s_randomize[lane] += 1;  // ant.py:289
            const bool timeout = late && reset != 0;  // vec_task.py:396
            if (b.timeout_buf) b.timeout_buf[e0 + lane] = timeout ? 1 : 0;
            st_rew = total; st_done = reset != 0; st_to = timeout;
            st_len = reset != 0 ? (float)progress : 0.f; st_obj = s_root[lane * 13 + 7];
        }
        if (b.stats) {  // one multi-lane atomic per CTA, spread over IGE_STATS_ROWS lines
            const float v[6] = {lane < ne ? 1.f : 0.f, st_rew, (float)st_done, (float)st_to, st_len, st_obj};
            const int slot[6] = {IGE_STAT_STEPS, IGE_STAT_REWARD, IGE_STAT_DONE, IGE_STAT_TIMEOUT,
                                 IGE_STAT_DONE_LENGTH, IGE_STAT_TASK0};
            ige::warp_stats_add<6>(b.stats, v, slot);
        }
        if (bulk) {
            ige::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                const uint32_t une = (uint32_t)ne;
                ige::bulk_s2g(b.obs_buf + (size_t)e0 * NUM_OBS, s_obs, une * NUM_OBS * 4);
                if (b.obs_clamped && !clamp_copy)
                    ige::bulk_s2g(b.obs_clamped + (size_t)e0 * NUM_OBS, s_obs, une * NUM_OBS * 4);
                ige::bulk_s2g(b.rew_buf + e0, s_rew, une * 4);
                ige::bulk_s2g(b.potentials + e0, s_pot, une * 4);
                ige::bulk_s2g(b.prev_potentials + e0, s_prev, une * 4);
                ige::bulk_s2g(b.up_vec + (size_t)e0 * 3, s_up, une * 12);
                ige::bulk_s2g(b.heading_vec + (size_t)e0 * 3, s_head, une * 12);
                ige::bulk_s2g(b.reset_buf + e0, s_reset, une * 8);
                ige::bulk_s2g(b.progress_buf + e0, s_progress, une * 8);
                if (has_rnd) ige::bulk_s2g(b.randomize_buf + e0, s_randomize, une * 8);
                ige::bulk_commit();
                ige::bulk_wait_read_all();  // smem must outlive the async reads
            }
        }
    }
    if (!bulk) {  // ragged tail tile / TMA disabled: plain coalesced stores by the whole CTA
        __syncthreads();
        coop_copy(b.obs_buf + (size_t)e0 * NUM_OBS, (const float *)s_obs, ne * NUM_OBS);
        if (b.obs_clamped && !clamp_copy) coop_copy(b.obs_clamped + (size_t)e0 * NUM_OBS, (const float *)s_obs, ne * NUM_OBS);
        coop_copy(b.rew_buf + e0, (const float *)s_rew, ne);
        coop_copy(b.potentials + e0, (const float *)s_pot, ne);
        coop_copy(b.prev_potentials + e0, (const float *)s_prev, ne);
        coop_copy(b.up_vec + (size_t)e0 * 3, (const float *)s_up, ne * 3);
        coop_copy(b.heading_vec + (size_t)e0 * 3, (const float *)s_head, ne * 3);
        coop_copy((long long *)b.reset_buf + e0, (const long long *)s_reset, ne);
        coop_copy((long long *)b.progress_buf + e0, (const long long *)s_progress, ne);
        if (has_rnd) coop_copy((long long *)b.randomize_buf + e0, (const long long *)s_randomize, ne);
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
    constexpr int NW_ELEM = 1;
    const int grid = (p->num_envs + TE - 1) / TE;
    locomotion_post_physics_kernel<8, 24, false, NW_ELEM>
        <<<grid, 32 * (3 + NW_ELEM), 0, (cudaStream_t)stream>>>(*p, *b, use_tma());
    return ige_check_launch("ige_ant_post_physics");
}

extern "C" int ige_humanoid_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, void *stream) {
    if (int rc = check(p, b, 21, 12, true)) return rc;
    if (p->num_envs == 0) return IGE_OK;
    constexpr int NW_ELEM = 3;
    const int grid = (p->num_envs + TE - 1) / TE;
    locomotion_post_physics_kernel<21, 12, true, NW_ELEM>
        <<<grid, 32 * (3 + NW_ELEM), 0, (cudaStream_t)stream>>>(*p, *b, use_tma());
    return ige_check_launch("ige_humanoid_post_physics");
}
