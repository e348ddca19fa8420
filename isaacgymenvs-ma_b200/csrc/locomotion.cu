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
// consecutive envs is one contiguous, 16-byte aligned span in every tensor.  A tile therefore
// moves with a handful of TMA 1-D bulk copies (cp.async.bulk, completion on an mbarrier), is
// processed out of shared memory, and its (32 x NUM_OBS) observation tile is assembled densely in
// shared memory and written back with one bulk store.  Nothing is read twice; every global access
// is a full-line burst.
//
// Main kernel: persistent + software-pipelined.  Each CTA walks tiles blockIdx, blockIdx + grid, ...
//   * input ring of IN_STAGES tile buffers, one "full" mbarrier per stage; loads for tile
//     it + IN_STAGES are issued as soon as tile `it` has been consumed, so HBM latency hides behind
//     IN_STAGES - 1 tiles of work;
//   * warp-specialised roles (lane == env for the torso roles), rotated with blockIdx so that every
//     SM sub-partition sees the same instruction mix:
//       HEADING  torso quat, up / heading vectors, heading projection, potentials (+ reset potentials)
//       VELOCITY local linear / angular velocity, torso height
//       ANGLES   Euler roll / yaw, walk-target angle (the three atan2f)
//       ELEM x k element-wise observation columns (float4 where the layout allows), the reset_idx
//                DOF scatter for flagged envs, per-DOF reward terms into a padded transposed scratch
//       REWARD   one tile behind the others: reward reductions in the reference's summation order,
//                termination / progress flags, time-outs, statistics; issues the bulk stores of
//                tile `it` and the bulk loads of tile it + IN_STAGES;
//   * output ring of OUT_STAGES tile buffers so stores of tile `it` overlap the math of `it + 1`;
//   * ONE __syncthreads per tile; episode statistics stay in registers until the CTA retires.
// A second, non-pipelined kernel (same math, plain coalesced copies) serves the ragged last tile
// (num_envs % 4 != 0) and the IGE_NO_TMA debug mode.
#include "ige_common.cuh"

namespace {

constexpr int TE = 32;       // envs per tile == warp size (lane == env for the torso roles)
constexpr int TS = TE + 1;   // padded stride of the transposed reward-term scratch

template <int D, int NS, bool HUM>
struct Layout {
    static constexpr int NUM_OBS = 12 + 3 * D + NS + (HUM ? D : 0);
    static constexpr int O_DOFPOS = 12, O_DOFVEL = 12 + D, O_FORCE = 12 + 2 * D;
    static constexpr int O_SENS = 12 + 2 * D + (HUM ? D : 0), O_ACT = O_SENS + NS;
    // float4 path: every element-wise segment starts 16-B aligned in the obs row and has a
    // multiple-of-4 length (Ant: 12, 20, 28, 52 with D = 8, NS = 24).
    static constexpr bool VEC4 = !HUM && (D % 4 == 0) && (NS % 4 == 0) && (NUM_OBS % 4 == 0);
};

// one tile of inputs as it lies in shared memory (each member a multiple of 16 B)
template <int D, int NS, bool HUM>
struct __align__(128) InTile {
    float root[TE * 13];
    float dof[TE * D * 2];
    float sens[TE * NS];
    float force[HUM ? TE * D : 4];
    float act[TE * D];
    float pot[TE];
    long long reset[TE];
    long long progress[TE];
    long long randomize[TE];
};
template <int NUM_OBS>
struct __align__(128) OutTile {
    float obs[TE * NUM_OBS];
    float rew[TE];
    float pot[TE];
    float prev[TE];
    float up[TE * 3];
    float head[TE * 3];
    long long reset[TE];
    long long progress[TE];
    long long randomize[TE];
};
template <int D>
struct Terms {
    float t0[D * TS], t1[D * TS], t2[D * TS];  // reward terms [dof][env], padded: conflict-free both ways
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

// atan2 results lie in [-pi, pi], so `% 2*pi` (torch_jit_utils.py:195) is exactly one conditional add.
__device__ __forceinline__ float mod_two_pi(float a) { return a < 0.0f ? a + ige::kTwoPi : a; }

__device__ __forceinline__ float neg_norm_over_dt(const IgeLocomotionParams &p, float n) {
    return p.potentials_recip_mul ? (-n) * (1.0f / p.dt) : (-n) / p.dt;
}

// ---------------------------------------------------------------------------------------------
// roles (lane == env); `in`/`out`/`terms` are shared-memory tiles, e0 the first env of the tile
// ---------------------------------------------------------------------------------------------
template <int D, int NS, bool HUM>
__device__ __forceinline__ void role_heading(const IgeLocomotionParams &p, const IgeLocomotionBuffers &b,
                                             const InTile<D, NS, HUM> &in,
                                             OutTile<Layout<D, NS, HUM>::NUM_OBS> &out, int lane, int e0) {
    constexpr int NUM_OBS = Layout<D, NS, HUM>::NUM_OBS;
    const float *root = in.root + lane * 13;
    float potentials = in.pot[lane];
    if (in.reset[lane] != 0) {  // ant.py:273-276
        const float *init = b.initial_root_states + (size_t)(e0 + lane) * 13;
        potentials = neg_norm_over_dt(p, ige::norm3(p.targets[0] - __ldg(init), p.targets[1] - __ldg(init + 1), 0.0f));
    }
    // compute_*_observations ant.py:384-391
    const float to_target[3] = {p.targets[0] - root[0], p.targets[1] - root[1], 0.0f};
    out.prev[lane] = potentials;
    float n = ige::norm3(to_target[0], to_target[1], 0.0f);
    out.pot[lane] = neg_norm_over_dt(p, n);
    // compute_heading_and_up torch_jit_utils.py:247-262
    n = n < 1e-9f ? 1e-9f : n;  // normalize(), torch_jit_utils.py:65-67
    const float target_dirs[3] = {to_target[0] / n, to_target[1] / n, to_target[2] / n};
    float torso_quat[4], up_vec[3], heading_vec[3];
    ige::quat_mul(root + 3, p.inv_start_rot, torso_quat);
    ige::quat_rotate<false>(torso_quat, p.basis_vec1, up_vec);
    ige::quat_rotate<false>(torso_quat, p.basis_vec0, heading_vec);
    float *o = out.obs + lane * NUM_OBS;
    o[10] = up_vec[2];
    o[11] = ige::dot3(heading_vec, target_dirs);
    out.up[lane * 3 + 0] = up_vec[0]; out.up[lane * 3 + 1] = up_vec[1]; out.up[lane * 3 + 2] = up_vec[2];
    out.head[lane * 3 + 0] = heading_vec[0]; out.head[lane * 3 + 1] = heading_vec[1];
    out.head[lane * 3 + 2] = heading_vec[2];
}

template <int D, int NS, bool HUM>
__device__ __forceinline__ void role_velocity(const IgeLocomotionParams &p, const InTile<D, NS, HUM> &in,
                                              OutTile<Layout<D, NS, HUM>::NUM_OBS> &out, int lane) {
    constexpr int NUM_OBS = Layout<D, NS, HUM>::NUM_OBS;
    const float *root = in.root + lane * 13;
    float torso_quat[4], vel_loc[3], angvel_loc[3];
    ige::quat_mul(root + 3, p.inv_start_rot, torso_quat);
    ige::quat_rotate<true>(torso_quat, root + 7, vel_loc);  // compute_rot torch_jit_utils.py:265-267
    ige::quat_rotate<true>(torso_quat, root + 10, angvel_loc);
    float *o = out.obs + lane * NUM_OBS;
    o[0] = root[2];
    o[1] = vel_loc[0]; o[2] = vel_loc[1]; o[3] = vel_loc[2];
    const float s = HUM ? p.angular_velocity_scale : 1.0f;  // x * 1.0f is exact
    o[4] = angvel_loc[0] * s; o[5] = angvel_loc[1] * s; o[6] = angvel_loc[2] * s;
}

template <int D, int NS, bool HUM>
__device__ __forceinline__ void role_angles(const IgeLocomotionParams &p, const InTile<D, NS, HUM> &in,
                                            OutTile<Layout<D, NS, HUM>::NUM_OBS> &out, int lane) {
    constexpr int NUM_OBS = Layout<D, NS, HUM>::NUM_OBS;
    const float *root = in.root + lane * 13;
    float q[4];
    ige::quat_mul(root + 3, p.inv_start_rot, q);
    // get_euler_xyz torch_jit_utils.py:175-195 (pitch is never consumed on this path)
    const float qx = q[0], qy = q[1], qz = q[2], qw = q[3];
    const float sinr_cosp = 2.0f * (qw * qx + qy * qz);
    const float cosr_cosp = qw * qw - qx * qx - qy * qy + qz * qz;
    const float siny_cosp = 2.0f * (qw * qz + qx * qy);
    const float cosy_cosp = qw * qw + qx * qx - qy * qy - qz * qz;
    const float roll = mod_two_pi(atan2f(sinr_cosp, cosr_cosp));
    const float yaw = mod_two_pi(atan2f(siny_cosp, cosy_cosp));
    const float walk_target_angle = atan2f(p.targets[2] - root[2], p.targets[0] - root[0]);  // :272-273
    const float angle_to_target = walk_target_angle - yaw;
    float *o = out.obs + lane * NUM_OBS;
    if (HUM) {  // humanoid.py:402-404 normalize_angle
        o[7] = wrap_pi(yaw); o[8] = wrap_pi(roll); o[9] = wrap_pi(angle_to_target);
    } else {  // ant.py:403
        o[7] = yaw; o[8] = roll; o[9] = angle_to_target;
    }
}

// reset_idx for one DOF (ant.py:257-261): returns the new (pos, vel) row
__device__ __forceinline__ float2 reset_dof(const IgeLocomotionParams &p, const IgeLocomotionBuffers &b, uint64_t step,
                                            int e, int d, int j) {
    // torch_rand_float(lo, hi): (hi - lo) * rand + lo (torch_jit_utils.py:215-218)
    const float pos = 0.4f * ige::reset_uniform(p.rand, step, 0, e, d, j) + (-0.2f);
    const float vel = 0.2f * ige::reset_uniform(p.rand, step, 1, e, d, j) + (-0.1f);
    float v = __ldg(b.initial_dof_pos + (size_t)e * d + j) + pos;
    v = fminf(v, p.dof_limits_upper[j]);  // tensor_clamp torch_jit_utils.py:228-230
    v = fmaxf(v, p.dof_limits_lower[j]);
    return make_float2(v, vel);
}

// element-wise columns + reset DOF scatter + per-DOF reward terms; `share` of `nshares` warps
template <int D, int NS, bool HUM>
__device__ __forceinline__ void role_elem(const IgeLocomotionParams &p, const IgeLocomotionBuffers &b,
                                          InTile<D, NS, HUM> &in, OutTile<Layout<D, NS, HUM>::NUM_OBS> &out,
                                          Terms<D> &tm, int lane, int share, int nshares, int ne, int e0,
                                          uint64_t step) {
    using L = Layout<D, NS, HUM>;
    constexpr int NUM_OBS = L::NUM_OBS;
    const int stride = 32 * nshares;
    if (L::VEC4) {
        for (int it = share * 32 + lane; it < ne * (D / 4); it += stride) {
            const int le = it / (D / 4), g = it - le * (D / 4);
            float4 a = reinterpret_cast<const float4 *>(in.dof)[it * 2];      // p0 v0 p1 v1
            float4 c = reinterpret_cast<const float4 *>(in.dof)[it * 2 + 1];  // p2 v2 p3 v3
            if (in.reset[le] != 0) {  // reset_idx: rewrite this env's DOF rows (global + local copy)
                const int e = e0 + le;
                const float2 r0 = reset_dof(p, b, step, e, D, g * 4 + 0), r1 = reset_dof(p, b, step, e, D, g * 4 + 1);
                const float2 r2 = reset_dof(p, b, step, e, D, g * 4 + 2), r3 = reset_dof(p, b, step, e, D, g * 4 + 3);
                a = make_float4(r0.x, r0.y, r1.x, r1.y);
                c = make_float4(r2.x, r2.y, r3.x, r3.y);
                float4 *row = reinterpret_cast<float4 *>(b.dof_state + ((size_t)e * D + g * 4) * 2);
                row[0] = a;
                row[1] = c;
            }
            const float4 act = reinterpret_cast<const float4 *>(in.act)[it];
            const float pos[4] = {a.x, a.z, c.x, c.z}, vel[4] = {a.y, a.w, c.y, c.w};
            const float av[4] = {act.x, act.y, act.z, act.w};
            float ps[4], vs[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int j = g * 4 + k;
                const float lo = p.dof_limits_lower[j], hi = p.dof_limits_upper[j];
                ps[k] = (2.0f * pos[k] - hi - lo) / (hi - lo);  // unscale torch_jit_utils.py:238-240
                vs[k] = vel[k] * p.dof_vel_scale;
                tm.t0[j * TS + le] = av[k] * av[k];                // actions ** 2        ant.py:353
                tm.t1[j * TS + le] = fabsf(av[k] * vs[k]);         // |a * dof_vel_obs|   ant.py:354
                tm.t2[j * TS + le] = ps[k] > 0.99f ? 1.0f : 0.0f;  // dof at limit        ant.py:355
            }
            float *o = out.obs + le * NUM_OBS;
            reinterpret_cast<float4 *>(o + L::O_DOFPOS)[g] = make_float4(ps[0], ps[1], ps[2], ps[3]);
            reinterpret_cast<float4 *>(o + L::O_DOFVEL)[g] = make_float4(vs[0], vs[1], vs[2], vs[3]);
            reinterpret_cast<float4 *>(o + L::O_ACT)[g] = act;
        }
        for (int it = share * 32 + lane; it < ne * (NS / 4); it += stride) {
            const int le = it / (NS / 4), g = it - le * (NS / 4);
            float4 s = reinterpret_cast<const float4 *>(in.sens)[it];
            s.x *= p.contact_force_scale; s.y *= p.contact_force_scale;
            s.z *= p.contact_force_scale; s.w *= p.contact_force_scale;
            reinterpret_cast<float4 *>(out.obs + le * NUM_OBS + L::O_SENS)[g] = s;
        }
    } else {
        for (int it = share * 32 + lane; it < ne * D; it += stride) {
            const int le = it / D, j = it - le * D;
            float2 pv = reinterpret_cast<const float2 *>(in.dof)[it];
            if (in.reset[le] != 0) {
                pv = reset_dof(p, b, step, e0 + le, D, j);
                reinterpret_cast<float2 *>(b.dof_state)[(size_t)e0 * D + it] = pv;
            }
            const float lo = p.dof_limits_lower[j], hi = p.dof_limits_upper[j];
            const float ps = (2.0f * pv.x - hi - lo) / (hi - lo);
            const float vs = pv.y * p.dof_vel_scale;
            const float a = in.act[it];
            float *o = out.obs + le * NUM_OBS;
            o[L::O_DOFPOS + j] = ps;
            o[L::O_DOFVEL + j] = vs;
            o[L::O_ACT + j] = a;
            tm.t0[j * TS + le] = a * a;
            if (HUM) {
                o[L::O_FORCE + j] = in.force[it] * p.contact_force_scale;
                const float ratio = p.motor_effort_ratio[j];
                tm.t1[j * TS + le] = fabsf(a * vs) * ratio;  // humanoid.py:359
                const float ap = fabsf(ps);
                const float scaled = p.joints_at_limit_cost_scale * (ap - 0.98f) / 0.02f;  // humanoid.py:356
                tm.t2[j * TS + le] = (ap > 0.98f ? 1.0f : 0.0f) * scaled * ratio;           // humanoid.py:357
            } else {
                tm.t1[j * TS + le] = fabsf(a * vs);
                tm.t2[j * TS + le] = ps > 0.99f ? 1.0f : 0.0f;
            }
        }
        for (int it = share * 32 + lane; it < ne * NS; it += stride) {
            const int le = it / NS, j = it - le * NS;
            out.obs[le * NUM_OBS + L::O_SENS + j] = in.sens[it] * p.contact_force_scale;
        }
    }
}

struct StatAcc {
    float rew = 0.f, len = 0.f, obj = 0.f, n = 0.f, done = 0.f, to = 0.f;
};

// reward + termination for env `lane` of a tile (ant.py:325-371, humanoid.py:323-375, vec_task.py:396)
template <int D, int NS, bool HUM>
__device__ __forceinline__ void role_reward(const IgeLocomotionParams &p, const IgeLocomotionBuffers &b,
                                            const InTile<D, NS, HUM> &in, OutTile<Layout<D, NS, HUM>::NUM_OBS> &out,
                                            const Terms<D> &tm, int lane, int e0, bool has_rnd, StatAcc &st) {
    constexpr int NUM_OBS = Layout<D, NS, HUM>::NUM_OBS;
    const float *o = out.obs + lane * NUM_OBS;
    const float z = o[0];
    long long progress = in.progress[lane] + 1;  // ant.py:288
    long long reset = in.reset[lane];
    if (reset != 0) { progress = 0; reset = 0; }  // ant.py:278-279
    const float heading_reward = o[11] > 0.8f ? p.heading_weight : p.heading_weight * o[11] / 0.8f;
    const float up_reward = o[10] > 0.93f ? 0.0f + p.up_weight : 0.0f;
    const float actions_cost = ige::row_sum<D>([&](int j) { return tm.t0[j * TS + lane]; });
    const float electricity_cost = ige::row_sum<D>([&](int j) { return tm.t1[j * TS + lane]; });
    const float progress_reward = out.pot[lane] - out.prev[lane];
    float total;
    if (HUM) {
        const float dof_at_limit_cost = ige::row_sum<D>([&](int j) { return tm.t2[j * TS + lane]; });
        total = progress_reward + 2.0f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                p.energy_cost_scale * electricity_cost - dof_at_limit_cost;
    } else {
        float at_limit = 0.0f;  // integer count <= 8: exact in fp32 (ant.py:355 sums int64)
#pragma unroll
        for (int j = 0; j < D; ++j) at_limit += tm.t2[j * TS + lane];
        total = progress_reward + 0.5f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                p.energy_cost_scale * electricity_cost - at_limit * p.joints_at_limit_cost_scale;
    }
    const bool fallen = z < p.termination_height;
    const bool late = (float)progress >= p.max_episode_length - 1.0f;
    if (fallen) total = p.death_cost;
    if (fallen || late) reset = 1;
    out.rew[lane] = total;
    out.reset[lane] = reset;
    out.progress[lane] = progress;
    if (has_rnd) out.randomize[lane] = in.randomize[lane] + 1;  // ant.py:289
    const bool timeout = late && reset != 0;  // vec_task.py:396
    if (b.timeout_buf) b.timeout_buf[e0 + lane] = timeout ? 1 : 0;
    st.n += 1.f; st.rew += total; st.done += reset != 0 ? 1.f : 0.f; st.to += timeout ? 1.f : 0.f;
    st.len += reset != 0 ? (float)progress : 0.f; st.obj += in.root[lane * 13 + 7];
}

__device__ __forceinline__ void flush_stats(float *stats, const StatAcc &st) {
    const float v[6] = {st.n, st.rew, st.done, st.to, st.len, st.obj};
    const int slot[6] = {IGE_STAT_STEPS, IGE_STAT_REWARD, IGE_STAT_DONE, IGE_STAT_TIMEOUT, IGE_STAT_DONE_LENGTH,
                         IGE_STAT_TASK0};
    ige::warp_stats_add<6>(stats, v, slot);
}

template <int D, int NS, bool HUM>
__device__ __forceinline__ void issue_tile_loads(const IgeLocomotionBuffers &b, InTile<D, NS, HUM> &in, uint64_t *bar,
                                                 int e0, int ne, bool has_rnd) {
    const uint32_t une = (uint32_t)ne;
    uint32_t bytes = une * (13 * 4 + D * 8 + NS * 4 + D * 4 + 4 + 8 + 8);
    if (HUM) bytes += une * D * 4;
    if (has_rnd) bytes += une * 8;
    ige::mbar_expect_tx(bar, bytes);
    ige::bulk_g2s(in.root, b.root_states + (size_t)e0 * 13, une * 52, bar);
    ige::bulk_g2s(in.dof, b.dof_state + (size_t)e0 * D * 2, une * D * 8, bar);
    ige::bulk_g2s(in.sens, b.vec_sensor_tensor + (size_t)e0 * NS, une * NS * 4, bar);
    if (HUM) ige::bulk_g2s(in.force, b.dof_force_tensor + (size_t)e0 * D, une * D * 4, bar);
    ige::bulk_g2s(in.act, b.actions + (size_t)e0 * D, une * D * 4, bar);
    ige::bulk_g2s(in.pot, b.potentials + e0, une * 4, bar);
    ige::bulk_g2s(in.reset, b.reset_buf + e0, une * 8, bar);
    ige::bulk_g2s(in.progress, b.progress_buf + e0, une * 8, bar);
    if (has_rnd) ige::bulk_g2s(in.randomize, b.randomize_buf + e0, une * 8, bar);
}

template <int NUM_OBS>
__device__ __forceinline__ void issue_tile_stores(const IgeLocomotionBuffers &b, OutTile<NUM_OBS> &out, int e0, int ne,
                                                  bool has_rnd, bool second_obs_copy) {
    const uint32_t une = (uint32_t)ne;
    ige::bulk_s2g(b.obs_buf + (size_t)e0 * NUM_OBS, out.obs, une * NUM_OBS * 4);
    if (second_obs_copy) ige::bulk_s2g(b.obs_clamped + (size_t)e0 * NUM_OBS, out.obs, une * NUM_OBS * 4);
    ige::bulk_s2g(b.rew_buf + e0, out.rew, une * 4);
    ige::bulk_s2g(b.potentials + e0, out.pot, une * 4);
    ige::bulk_s2g(b.prev_potentials + e0, out.prev, une * 4);
    ige::bulk_s2g(b.up_vec + (size_t)e0 * 3, out.up, une * 12);
    ige::bulk_s2g(b.heading_vec + (size_t)e0 * 3, out.head, une * 12);
    ige::bulk_s2g(b.reset_buf + e0, out.reset, une * 8);
    ige::bulk_s2g(b.progress_buf + e0, out.progress, une * 8);
    if (has_rnd) ige::bulk_s2g(b.randomize_buf + e0, out.randomize, une * 8);
    ige::bulk_commit();
}

template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

enum Role { HEADING = 0, VELOCITY = 1, ANGLES = 2, REWARD = 3, ELEM0 = 4 };

// ---------------------------------------------------------------------------------------------
// main kernel: persistent, pipelined, TMA in / TMA out.  Handles tiles [0, num_tiles); every tile
// has ne % 4 == 0 envs (only the last may be partial).
// ---------------------------------------------------------------------------------------------
template <int D, int NS, bool HUM, int NW_ELEM, int IN_STAGES, int OUT_STAGES>
__global__ void __launch_bounds__(32 * (4 + NW_ELEM))
locomotion_pipelined_kernel(const __grid_constant__ IgeLocomotionParams p,
                            const __grid_constant__ IgeLocomotionBuffers b, const int num_tiles) {
    using L = Layout<D, NS, HUM>;
    constexpr int NUM_OBS = L::NUM_OBS;
    constexpr int NW = 4 + NW_ELEM;
    using In = InTile<D, NS, HUM>;
    using Out = OutTile<NUM_OBS>;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    In *in = reinterpret_cast<In *>(smem_raw);
    Out *out = reinterpret_cast<Out *>(smem_raw + sizeof(In) * IN_STAGES);
    Terms<D> *terms = reinterpret_cast<Terms<D> *>(smem_raw + sizeof(In) * IN_STAGES + sizeof(Out) * OUT_STAGES);
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + sizeof(In) * IN_STAGES + sizeof(Out) * OUT_STAGES +
                                                  sizeof(Terms<D>) * 2);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int role = (warp + (int)blockIdx.x) % NW;
    const int first = blockIdx.x, stride = gridDim.x;
    const int my_tiles = (num_tiles - first + stride - 1) / stride;
    const bool has_rnd = b.randomize_buf != nullptr;
    const bool clamp_copy = b.obs_clamped != nullptr && p.clip_obs < INFINITY;
    const bool second_obs_copy = b.obs_clamped != nullptr && !clamp_copy;
    const uint64_t step = (p.rand.mode == IGE_RAND_PHILOX && p.rand.step) ? __ldg(p.rand.step) : 0ull;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < IN_STAGES; ++s) ige::mbar_init(&full[s], 1);
        ige::fence_mbar_init();
    }
    __syncthreads();

    if (role == REWARD) {
        if (lane == 0) {
            for (int k = 0; k < IN_STAGES && k < my_tiles; ++k) {
                const int e0 = (first + k * stride) * TE;
                issue_tile_loads<D, NS, HUM>(b, in[k], &full[k], e0, min(TE, p.num_envs - e0), has_rnd);
            }
        }
        StatAcc st;
        for (int it = 0; it < my_tiles; ++it) {
            const int s = it % IN_STAGES, o = it % OUT_STAGES;
            const int e0 = (first + it * stride) * TE;
            const int ne = min(TE, p.num_envs - e0);
            // out[(it + 1) % OUT_STAGES] is written by the role warps right after the next barrier:
            // its previous store (tile it + 1 - OUT_STAGES) must have finished reading shared memory.
            if (lane == 0) bulk_wait_read<OUT_STAGES - 2>();
            __syncthreads();  // barrier #it: the role warps have produced tile `it`
            ige::mbar_wait(&full[s], (it / IN_STAGES) & 1);  // already complete: orders the TMA writes for us
            if (lane < ne) role_reward<D, NS, HUM>(p, b, in[s], out[o], terms[it & 1], lane, e0, has_rnd, st);
            ige::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
                issue_tile_stores<NUM_OBS>(b, out[o], e0, ne, has_rnd, second_obs_copy);
                if (it + IN_STAGES < my_tiles) {
                    const int n0 = (first + (it + IN_STAGES) * stride) * TE;
                    issue_tile_loads<D, NS, HUM>(b, in[s], &full[s], n0, min(TE, p.num_envs - n0), has_rnd);
                }
            }
            __syncwarp();
        }
        if (lane == 0) bulk_wait_read<0>();  // shared memory must outlive the async reads
        if (b.stats) flush_stats(b.stats, st);
    } else {
        for (int it = 0; it < my_tiles; ++it) {
            const int s = it % IN_STAGES, o = it % OUT_STAGES;
            const int e0 = (first + it * stride) * TE;
            const int ne = min(TE, p.num_envs - e0);
            ige::mbar_wait(&full[s], (it / IN_STAGES) & 1);
            if (role == HEADING) {
                if (lane < ne) role_heading<D, NS, HUM>(p, b, in[s], out[o], lane, e0);
            } else if (role == VELOCITY) {
                if (lane < ne) role_velocity<D, NS, HUM>(p, in[s], out[o], lane);
            } else if (role == ANGLES) {
                if (lane < ne) role_angles<D, NS, HUM>(p, in[s], out[o], lane);
            } else {
                role_elem<D, NS, HUM>(p, b, in[s], out[o], terms[it & 1], lane, role - ELEM0, NW_ELEM, ne, e0, step);
            }
            ige::fence_proxy_async();  // obs tile -> visible to the bulk-store (async) proxy
            __syncthreads();           // barrier #it
            if (clamp_copy) {          // finite clip: second, clamped copy with plain coalesced stores
                // (reads out[o] concurrently with the bulk store; it is not rewritten before barrier
                //  #(it + OUT_STAGES - 1), and every role warp passes through here first)
                const int reward_warp = (REWARD + NW - (int)(blockIdx.x % NW)) % NW;
                for (int i = (warp > reward_warp ? tid - 32 : tid); i < ne * NUM_OBS; i += 32 * (NW - 1))
                    b.obs_clamped[(size_t)e0 * NUM_OBS + i] = clampf(out[o].obs[i], p.clip_obs);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// fallback kernel: one tile per CTA, plain coalesced copies (ragged last tile / IGE_NO_TMA)
// ---------------------------------------------------------------------------------------------
template <int D, int NS, bool HUM, int NW_ELEM>
__global__ void __launch_bounds__(32 * (4 + NW_ELEM))
locomotion_tile_kernel(const __grid_constant__ IgeLocomotionParams p, const __grid_constant__ IgeLocomotionBuffers b,
                       const int tile_offset) {
    using L = Layout<D, NS, HUM>;
    constexpr int NUM_OBS = L::NUM_OBS;
    constexpr int NW = 4 + NW_ELEM;
    __shared__ InTile<D, NS, HUM> in;
    __shared__ OutTile<NUM_OBS> out;
    __shared__ Terms<D> tm;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int role = (warp + (int)blockIdx.x) % NW;
    const int e0 = (blockIdx.x + tile_offset) * TE;
    const int ne = min(TE, p.num_envs - e0);
    const bool has_rnd = b.randomize_buf != nullptr;
    const uint64_t step = (p.rand.mode == IGE_RAND_PHILOX && p.rand.step) ? __ldg(p.rand.step) : 0ull;

    coop_copy(in.root, b.root_states + (size_t)e0 * 13, ne * 13);
    coop_copy(in.dof, (const float *)b.dof_state + (size_t)e0 * D * 2, ne * D * 2);
    coop_copy(in.sens, b.vec_sensor_tensor + (size_t)e0 * NS, ne * NS);
    if (HUM) coop_copy(in.force, b.dof_force_tensor + (size_t)e0 * D, ne * D);
    coop_copy(in.act, b.actions + (size_t)e0 * D, ne * D);
    coop_copy(in.pot, (const float *)b.potentials + e0, ne);
    coop_copy(in.reset, (const long long *)b.reset_buf + e0, ne);
    coop_copy(in.progress, (const long long *)b.progress_buf + e0, ne);
    if (has_rnd) coop_copy(in.randomize, (const long long *)b.randomize_buf + e0, ne);
    __syncthreads();

    if (role == HEADING) {
        if (lane < ne) role_heading<D, NS, HUM>(p, b, in, out, lane, e0);
    } else if (role == VELOCITY) {
        if (lane < ne) role_velocity<D, NS, HUM>(p, in, out, lane);
    } else if (role == ANGLES) {
        if (lane < ne) role_angles<D, NS, HUM>(p, in, out, lane);
    } else if (role >= ELEM0) {
        role_elem<D, NS, HUM>(p, b, in, out, tm, lane, role - ELEM0, NW_ELEM, ne, e0, step);
    }
    __syncthreads();
    if (role == REWARD) {
        StatAcc st;
        if (lane < ne) role_reward<D, NS, HUM>(p, b, in, out, tm, lane, e0, has_rnd, st);
        if (b.stats) flush_stats(b.stats, st);
    }
    __syncthreads();
    coop_copy(b.obs_buf + (size_t)e0 * NUM_OBS, (const float *)out.obs, ne * NUM_OBS);
    if (b.obs_clamped)
        for (int i = tid; i < ne * NUM_OBS; i += blockDim.x)
            b.obs_clamped[(size_t)e0 * NUM_OBS + i] = clampf(out.obs[i], p.clip_obs);
    coop_copy(b.rew_buf + e0, (const float *)out.rew, ne);
    coop_copy(b.potentials + e0, (const float *)out.pot, ne);
    coop_copy(b.prev_potentials + e0, (const float *)out.prev, ne);
    coop_copy(b.up_vec + (size_t)e0 * 3, (const float *)out.up, ne * 3);
    coop_copy(b.heading_vec + (size_t)e0 * 3, (const float *)out.head, ne * 3);
    coop_copy((long long *)b.reset_buf + e0, (const long long *)out.reset, ne);
    coop_copy((long long *)b.progress_buf + e0, (const long long *)out.progress, ne);
    if (has_rnd) coop_copy((long long *)b.randomize_buf + e0, (const long long *)out.randomize, ne);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
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

template <int D, int NS, bool HUM, int NW_ELEM, int IN_STAGES, int OUT_STAGES>
int launch(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, cudaStream_t stream, const char *name) {
    constexpr int NT = 32 * (4 + NW_ELEM);
    constexpr int NUM_OBS = Layout<D, NS, HUM>::NUM_OBS;
    constexpr size_t smem = sizeof(InTile<D, NS, HUM>) * IN_STAGES + sizeof(OutTile<NUM_OBS>) * OUT_STAGES +
                            sizeof(Terms<D>) * 2 + sizeof(uint64_t) * IN_STAGES;
    auto kern = locomotion_pipelined_kernel<D, NS, HUM, NW_ELEM, IN_STAGES, OUT_STAGES>;
    static int max_ctas = 0;  // resident CTAs on the current device (per template instance)
    if (max_ctas == 0) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return ige_fail((int)e, "%s: cudaFuncSetAttribute(smem=%zu): %s", name, smem, cudaGetErrorString(e));
        int dev = 0, sms = 0, per_sm = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NT, smem);
        if (e != cudaSuccess || per_sm < 1 || sms < 1)
            return ige_fail(e != cudaSuccess ? (int)e : IGE_EPARAM, "%s: occupancy query failed", name);
        max_ctas = sms * per_sm;
    }
    const int n = p->num_envs;
    const int full_tiles = n / TE, rem = n % TE;
    const bool tma = use_tma() != 0;
    int pipelined_tiles = 0, fallback_offset = 0, fallback_tiles = 0;
    if (tma) {
        pipelined_tiles = full_tiles + ((rem > 0 && rem % 4 == 0) ? 1 : 0);
        fallback_offset = pipelined_tiles;
        fallback_tiles = (rem > 0 && rem % 4 != 0) ? 1 : 0;
    } else {
        fallback_tiles = full_tiles + (rem > 0 ? 1 : 0);
    }
    if (pipelined_tiles > 0) {
        const int grid = pipelined_tiles < max_ctas ? pipelined_tiles : max_ctas;
        kern<<<grid, NT, smem, stream>>>(*p, *b, pipelined_tiles);
        if (int rc = ige_check_launch(name)) return rc;
    }
    if (fallback_tiles > 0) {
        locomotion_tile_kernel<D, NS, HUM, NW_ELEM><<<fallback_tiles, NT, 0, stream>>>(*p, *b, fallback_offset);
        if (int rc = ige_check_launch(name)) return rc;
    }
    return IGE_OK;
}

}  // namespace

extern "C" void ige_set_use_tma(int on) { g_use_tma = on ? 1 : 0; }

extern "C" int ige_ant_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, void *stream) {
    if (int rc = check(p, b, 8, 24, false)) return rc;
    if (p->num_envs == 0) return IGE_OK;
    return launch<8, 24, false, 1, 2, 3>(p, b, (cudaStream_t)stream, "ige_ant_post_physics");
}

extern "C" int ige_humanoid_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, void *stream) {
    if (int rc = check(p, b, 21, 12, true)) return rc;
    if (p->num_envs == 0) return IGE_OK;
    return launch<21, 12, true, 3, 2, 3>(p, b, (cudaStream_t)stream, "ige_humanoid_post_physics");
}
