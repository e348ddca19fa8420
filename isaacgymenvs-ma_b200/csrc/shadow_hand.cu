// shadow_hand.cu -- fused ShadowHand reset (pre-physics position) and post_physics_step for sm_100a.
//
// Replaces (reference, isaacgymenvs/tasks/shadow_hand.py):
//   reset section of pre_physics_step :670-682  (reset_target_pose :586-602, reset_idx :604-668,
//                                                randomize_rotation :803-806)
//   post_physics_step :710-715                   (compute_observations :437-584 -- four observation
//                                                types + asymmetric states_buf --, compute_hand_reward
//                                                :746-800 incl. the global consecutive_successes update)
//   + the VecTask.step epilogue base/vec_task.py:396-408 when requested.
//
// Post-physics kernel (HBM-bound, 1732 B/env for full_state): a CTA owns a tile of 32 consecutive
// envs.  Every observation column is a function of ONE input element, so the tile's (32 x num_obs)
// rows are assembled in shared memory by element-parallel loops that read the contiguous input
// spans (dof_state, dof_force, sensors, actions) with fully coalesced loads, gather the 5 fingertip
// rows per env (52-B segments) and are written back with one TMA bulk store (rows are 844 B: only a
// tile of >= 4 envs is 16-B aligned, which is exactly why the row-major tile goes through shared
// memory).  Warp 0 (lane == env) does the per-env object/goal quaternion math, the reward and the
// flags, and feeds the two global sums of compute_hand_reward into a last-CTA reduction.
#include "ige_common.cuh"

namespace {

constexpr int TE = 32;
constexpr int NT = 256;
constexpr int MAX_OBS = 211;
constexpr float kPiF = 3.14159265358979323846f;

enum Seg { SEG_DOF_POS, SEG_DOF_VEL, SEG_DOF_FORCE, SEG_OBJ_POSE, SEG_OBJ_POS, SEG_OBJ_LINVEL, SEG_OBJ_ANGVEL,
           SEG_GOAL_POSE, SEG_QUAT_DIFF, SEG_FT_STATE, SEG_FT_POS, SEG_SENSORS, SEG_ACTIONS, SEG_END };
struct SegEntry { int kind, off; };
// shadow_hand.py:473-584
__constant__ SegEntry kLayout[7][12] = {
    {{SEG_FT_POS, 0}, {SEG_OBJ_POS, 15}, {SEG_QUAT_DIFF, 18}, {SEG_ACTIONS, 22}, {SEG_END, 42}},
    {{SEG_DOF_POS, 0}, {SEG_OBJ_POSE, 24}, {SEG_GOAL_POSE, 31}, {SEG_QUAT_DIFF, 38}, {SEG_FT_POS, 42}, {SEG_ACTIONS, 57},
     {SEG_END, 77}},
    {{SEG_DOF_POS, 0}, {SEG_DOF_VEL, 24}, {SEG_OBJ_POSE, 48}, {SEG_OBJ_LINVEL, 55}, {SEG_OBJ_ANGVEL, 58},
     {SEG_GOAL_POSE, 61}, {SEG_QUAT_DIFF, 68}, {SEG_FT_STATE, 72}, {SEG_ACTIONS, 137}, {SEG_END, 157}},
    {{SEG_DOF_POS, 0}, {SEG_DOF_VEL, 24}, {SEG_DOF_FORCE, 48}, {SEG_OBJ_POSE, 72}, {SEG_OBJ_LINVEL, 79},
     {SEG_OBJ_ANGVEL, 82}, {SEG_GOAL_POSE, 85}, {SEG_QUAT_DIFF, 92}, {SEG_FT_STATE, 96}, {SEG_SENSORS, 161},
     {SEG_ACTIONS, 191}, {SEG_END, 211}},
    // AllegroHand (allegro_hand.py:439-489): 16 DOFs, 16 actions, no fingertip / sensor segments
    {{SEG_DOF_POS, 0}, {SEG_OBJ_POSE, 16}, {SEG_GOAL_POSE, 23}, {SEG_QUAT_DIFF, 30}, {SEG_ACTIONS, 34}, {SEG_END, 50}},
    {{SEG_DOF_POS, 0}, {SEG_DOF_VEL, 16}, {SEG_OBJ_POSE, 32}, {SEG_OBJ_LINVEL, 39}, {SEG_OBJ_ANGVEL, 42},
     {SEG_GOAL_POSE, 45}, {SEG_QUAT_DIFF, 52}, {SEG_ACTIONS, 56}, {SEG_END, 72}},
    {{SEG_DOF_POS, 0}, {SEG_DOF_VEL, 16}, {SEG_DOF_FORCE, 32}, {SEG_OBJ_POSE, 48}, {SEG_OBJ_LINVEL, 55},
     {SEG_OBJ_ANGVEL, 58}, {SEG_GOAL_POSE, 61}, {SEG_QUAT_DIFF, 68}, {SEG_ACTIONS, 72}, {SEG_END, 88}}};
__host__ __device__ constexpr int hand_num_obs(int type) {
    return type == 0 ? 42 : type == 1 ? 77 : type == 2 ? 157 : type == 3 ? 211 : type == 4 ? 50 : type == 5 ? 72 : 88;
}
// layout of the asymmetric states_buf: the hand's own full_state (shadow_hand.py:529-556, allegro_hand.py:452-469)
__host__ __device__ constexpr int hand_state_layout(int type) {
    return type >= IGE_HAND_OBS_ALLEGRO_FULL_NO_VEL ? IGE_HAND_OBS_ALLEGRO_FULL_STATE : IGE_HAND_OBS_FULL_STATE;
}
__host__ __device__ constexpr bool hand_is_allegro(int type) { return type >= IGE_HAND_OBS_ALLEGRO_FULL_NO_VEL; }

__device__ __forceinline__ float clampf(float x, float c) { return x < -c ? -c : (x > c ? c : x); }

// ---------------------------------------------------------------------------------------------
// reset (pre-physics): one thread per env; only flagged envs do work
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ float hand_uniform(const IgeHandResetRand &r, uint64_t step, int draw, int rank, int env,
                                              int cols, int j) {
    if (r.mode == IGE_RAND_PHILOX) {
        uint32_t c[4] = {(uint32_t)env, ((uint32_t)draw << 16) | ((uint32_t)j >> 2), (uint32_t)step, (uint32_t)(step >> 32)};
        ige::philox4x32_10(c, (uint32_t)r.seed, (uint32_t)(r.seed >> 32));
        return (float)(c[j & 3] >> 8) * (1.0f / 16777216.0f);
    }
    const float *u = draw == 0 ? r.uA : draw == 1 ? r.uB : draw == 2 ? r.uC : r.uD;
    return __ldg(u + (size_t)rank * cols + j);
}
__device__ __forceinline__ float sym_uniform(float u) { return 2.0f * u + (-1.0f); }  // torch_rand_float(-1, 1)

// quat_from_angle_axis (torch_jit_utils.py:118-123) for a unit coordinate axis, incl. quat_unit()
__device__ __forceinline__ void quat_from_angle_axis_unit(float angle, int axis, float *q) {
    const float theta = angle / 2.0f;
    const float s = sinf(theta), c = cosf(theta);
    float v[4] = {0.f, 0.f, 0.f, c};
    v[axis] = 1.0f * s;
    float n2 = __fmul_rn(v[0], v[0]);
    n2 = __fmaf_rn(v[1], v[1], n2); n2 = __fmaf_rn(v[2], v[2], n2); n2 = __fmaf_rn(v[3], v[3], n2);
    float n = __fsqrt_rn(n2);
    n = n < 1e-9f ? 1e-9f : n;
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] = v[i] / n;
}
__device__ __forceinline__ void randomize_rotation(float rand0, float rand1, float *out) {  // shadow_hand.py:803-806
    float qa[4], qb[4];
    quat_from_angle_axis_unit(rand0 * kPiF, 0, qa);
    quat_from_angle_axis_unit(rand1 * kPiF, 1, qb);
    ige::quat_mul(qa, qb, out);
}
__device__ __forceinline__ void reset_target(const IgeShadowHandParams &p, const IgeShadowHandBuffers &b, int e, float r0,
                                             float r1) {  // reset_target_pose shadow_hand.py:586-602
    float rot[4];
    randomize_rotation(r0, r1, rot);
    float *gs = b.goal_states + (size_t)e * 13;
    const float *gi = b.goal_init_state + (size_t)e * 13;
    const long long gidx = b.goal_object_indices[e];
    float *root = b.root_state_tensor + (size_t)gidx * 13;
#pragma unroll
    for (int i = 0; i < 3; ++i) { const float v = gi[i]; gs[i] = v; root[i] = v + p.goal_displacement[i]; }
#pragma unroll
    for (int i = 0; i < 4; ++i) { gs[3 + i] = rot[i]; root[3 + i] = rot[i]; }
#pragma unroll
    for (int i = 7; i < 13; ++i) root[i] = 0.0f;
    if (b.actor_dirty) b.actor_dirty[gidx] = 1;
    b.reset_goal_buf[e] = 0;
}

// reset_idx (shadow_hand.py:604-668) for ONE env, by a whole warp: the 2*d + 5 + 2 + 1 uniforms are drawn one
// (Philox call) per lane and parked in shared memory, then the rows are written item-parallel (DOF j by lane j,
// rb_forces strided, the object row / goal pose / force probability by one lane each).  A single thread doing
// this serially (~60 Philox calls) used to set the latency of the whole launch.
__device__ void hand_reset_env_coop(const IgeShadowHandParams &p, const IgeShadowHandBuffers &b, uint64_t step, int e, int rk,
                                    int lane, float *us) {
    const IgeHandResetRand &r = p.rand;
    const int d = p.num_hand_dofs, dpe = p.num_dofs_per_env, cols = 2 * d + 5;
    // us[0 .. cols) draw 1 | us[cols], us[cols + 1] draw 2 | us[cols + 2] draw 3
    for (int j = lane; j < cols + 3; j += 32) {
        float u;
        if (j < cols) u = hand_uniform(r, step, 1, rk, e, cols, j);
        else if (j < cols + 2) u = hand_uniform(r, step, 2, rk, e, 4, j - cols);
        else u = hand_uniform(r, step, 3, rk, e, 1, 0);
        us[j] = u;
    }
    __syncwarp();
    if (lane == 31) reset_target(p, b, e, sym_uniform(us[cols]), sym_uniform(us[cols + 1]));  // :613
    float *rbf = b.rb_forces + (size_t)e * p.num_bodies * 3;
    for (int i = lane; i < p.num_bodies * 3; i += 32) rbf[i] = 0.0f;  // :616
    if (lane == 30) {
        const long long oidx = b.object_indices[e];
        float *obj = b.root_state_tensor + (size_t)oidx * 13;
        const float *oi = b.object_init_state + (size_t)e * 13;
        obj[0] = oi[0] + p.reset_position_noise * sym_uniform(us[0]);  // :619-623
        obj[1] = oi[1] + p.reset_position_noise * sym_uniform(us[1]);
        obj[2] = oi[2] + p.reset_position_noise * sym_uniform(us[2]);
        float rot[4];
        randomize_rotation(sym_uniform(us[3]), sym_uniform(us[4]), rot);  // :625
#pragma unroll
        for (int i = 0; i < 4; ++i) obj[3 + i] = rot[i];
#pragma unroll
        for (int i = 7; i < 13; ++i) obj[i] = 0.0f;
        if (b.actor_dirty) b.actor_dirty[oidx] = 1;
    }
    if (lane == 29) {
        const float l0 = logf(p.force_prob_range[0]), l1 = logf(p.force_prob_range[1]);  // :644-645
        b.random_force_prob[e] = expf((l0 - l1) * us[cols + 2] + l1);
        b.progress_buf[e] = 0;  // :666-668
        b.reset_buf[e] = 0;
        b.successes[e] = 0.0f;
    }
    for (int j = lane; j < d; j += 32) {  // :647-657
        const float delta_max = p.dof_upper[j] - p.dof_default_pos[j];
        const float delta_min = p.dof_lower[j] - p.dof_default_pos[j];
        const float rj = sym_uniform(us[5 + j]);
        const float rv = sym_uniform(us[5 + d + j]);
        const float rand_delta = delta_min + (delta_max - delta_min) * 0.5f * (rj + 1.0f);
        const float pos = p.dof_reset_base[j] + p.reset_dof_pos_noise * rand_delta;
        const float vel = p.dof_default_vel[j] + p.reset_dof_vel_noise * rv;
        reinterpret_cast<float2 *>(b.dof_state)[(size_t)e * dpe + j] = make_float2(pos, vel);
        b.prev_targets[(size_t)e * dpe + j] = pos;
        b.cur_targets[(size_t)e * dpe + j] = pos;
    }
    __syncwarp();
}

__global__ void __launch_bounds__(128) shadow_hand_reset_kernel(const __grid_constant__ IgeShadowHandParams p,
                                                                const __grid_constant__ IgeShadowHandBuffers b) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    __shared__ float s_u[4][2 * IGE_MAX_DOF + 8];
    const int e = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31;
    const IgeHandResetRand &r = p.rand;
    const uint64_t step = (r.mode == IGE_RAND_PHILOX && r.step) ? *reinterpret_cast<const volatile uint64_t *>(r.step) : 0ull;
    bool env = false;
    int rk = 0;
    if (e < p.num_envs) {
        const bool goal = b.reset_goal_buf[e] != 0;
        env = b.reset_buf[e] != 0;
        if (env) rk = r.env_rank ? __ldg(r.env_rank + e) : 0;
        if (goal && !env) {  // :675-679 (for an env that also resets, reset_idx's own reset_target_pose :613 overwrites it)
            const int grk = r.goal_rank ? __ldg(r.goal_rank + e) : 0;
            reset_target(p, b, e, sym_uniform(hand_uniform(r, step, 0, grk, e, 4, 0)),
                         sym_uniform(hand_uniform(r, step, 0, grk, e, 4, 1)));
        }
    }
    const uint32_t mask = __ballot_sync(0xffffffffu, env);
    for (uint32_t m = mask; m; m &= m - 1) {  // the warp resets its flagged envs one after the other, together
        const int src = __ffs(m) - 1;
        hand_reset_env_coop(p, b, step, e - lane + src, __shfl_sync(0xffffffffu, rk, src), lane, s_u[threadIdx.x >> 5]);
    }
    if (b.list.scratch) {  // reset list == reset_buf.nonzero() at entry
        if (lane == 0 && e < p.num_envs) ige::publish_tile_mask(b.list, p.num_envs, e >> 5, mask);
        ige::finish_reset_list(b.list, p.num_envs, nullptr, lane == 0);
    }
}

// ---------------------------------------------------------------------------------------------
// post-physics
// ---------------------------------------------------------------------------------------------
// Fill the element-parallel segments of one (TE x num_obs) tile held in shared memory.
__device__ __forceinline__ void fill_tile(const IgeShadowHandParams &p, const IgeShadowHandBuffers &b, int layout, int nobs,
                                          float *tile, int e0, int ne) {
    const int d = p.num_hand_dofs, dpe = p.num_dofs_per_env, na = p.num_actions, tid = threadIdx.x;
    for (int s = 0; kLayout[layout][s].kind != SEG_END; ++s) {
        const int kind = kLayout[layout][s].kind, off = kLayout[layout][s].off;
        if (kind == SEG_DOF_POS || kind == SEG_DOF_VEL) {
            const bool pos = kind == SEG_DOF_POS;
            for (int i = tid; i < ne * d; i += NT) {
                const int le = i / d, j = i - le * d;
                const float2 pv = __ldg(reinterpret_cast<const float2 *>(b.dof_state) + (size_t)(e0 + le) * dpe + j);
                const float lo = p.dof_lower[j], hi = p.dof_upper[j];
                tile[le * nobs + off + j] = pos ? (2.0f * pv.x - hi - lo) / (hi - lo)   // unscale :499-500
                                                : p.vel_obs_scale * pv.y;                // :514
            }
        } else if (kind == SEG_DOF_FORCE) {
            for (int i = tid; i < ne * d; i += NT) {
                const int le = i / d, j = i - le * d;
                tile[le * nobs + off + j] = p.force_torque_obs_scale * __ldg(b.dof_force_tensor + (size_t)e0 * d + i);
            }
        } else if (kind == SEG_SENSORS) {
            for (int i = tid; i < ne * 30; i += NT) {
                const int le = i / 30, j = i - le * 30;
                tile[le * nobs + off + j] = p.force_torque_obs_scale * __ldg(b.vec_sensor_tensor + (size_t)e0 * 30 + i);
            }
        } else if (kind == SEG_ACTIONS) {
            for (int i = tid; i < ne * na; i += NT) {
                const int le = i / na, j = i - le * na;
                tile[le * nobs + off + j] = __ldg(b.actions + (size_t)e0 * na + i);
            }
        } else if (kind == SEG_FT_STATE) {
            for (int i = tid; i < ne * 65; i += NT) {
                const int le = i / 65, j = i - le * 65, f = j / 13, k = j - f * 13;
                tile[le * nobs + off + j] =
                    __ldg(b.rigid_body_states + ((size_t)(e0 + le) * p.num_bodies + p.fingertip_handles[f]) * 13 + k);
            }
        } else if (kind == SEG_FT_POS) {
            for (int i = tid; i < ne * 15; i += NT) {
                const int le = i / 15, j = i - le * 15, f = j / 3, k = j - f * 3;
                tile[le * nobs + off + j] =
                    __ldg(b.rigid_body_states + ((size_t)(e0 + le) * p.num_bodies + p.fingertip_handles[f]) * 13 + k);
            }
        }
    }
}
// The per-env segments (object / goal rows, quat diff), by the thread that owns env `le`.
__device__ __forceinline__ void fill_env_segments(const IgeShadowHandParams &p, int layout, int nobs, float *row,
                                                  const float *obj, const float *goal, const float *qd) {
    for (int s = 0; kLayout[layout][s].kind != SEG_END; ++s) {
        float *o = row + kLayout[layout][s].off;
        switch (kLayout[layout][s].kind) {
        case SEG_OBJ_POSE:
#pragma unroll
            for (int i = 0; i < 7; ++i) o[i] = obj[i];
            break;
        case SEG_OBJ_POS: o[0] = obj[0]; o[1] = obj[1]; o[2] = obj[2]; break;
        case SEG_OBJ_LINVEL: o[0] = obj[7]; o[1] = obj[8]; o[2] = obj[9]; break;
        case SEG_OBJ_ANGVEL:
            o[0] = p.vel_obs_scale * obj[10]; o[1] = p.vel_obs_scale * obj[11]; o[2] = p.vel_obs_scale * obj[12];
            break;
        case SEG_GOAL_POSE:
#pragma unroll
            for (int i = 0; i < 7; ++i) o[i] = goal[i];
            break;
        case SEG_QUAT_DIFF: o[0] = qd[0]; o[1] = qd[1]; o[2] = qd[2]; o[3] = qd[3]; break;
        default: break;
        }
    }
}

__device__ __forceinline__ void store_tile(float *dst, float *dst_clamped, const float *tile, int count, float clip,
                                           bool bulk) {
    // `count` floats, contiguous in global.  bulk: one TMA store by thread 0 (after fence + barrier).
    if (bulk) {
        if (threadIdx.x == 0) {
            ige::bulk_s2g(dst, tile, (uint32_t)count * 4);
            ige::bulk_commit();
        }
    } else {
        for (int i = threadIdx.x; i < count; i += NT) dst[i] = tile[i];
    }
    if (dst_clamped)  // vec_task.py:404 (clip_obs is finite for this task: a real second pass)
        for (int i = threadIdx.x; i < count; i += NT) dst_clamped[i] = clampf(tile[i], clip);
}

// Called by every thread of every CTA (of all launches of one step) after its partial sums were added to
// reduce_scratch: the CTA that arrives last (`expected_ctas` arrivals) finishes compute_hand_reward's global
// part -- consecutive_successes (:795-798) --, re-arms the scratch and advances the Philox step.
__device__ __forceinline__ void finish_hand_reduction(const IgeShadowHandParams &p, const IgeShadowHandBuffers &b,
                                                      int expected_ctas, int *s_last) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        *s_last = atomicAdd(b.reduce_scratch + 1, 1u) == (unsigned)expected_ctas - 1u;
    }
    __syncthreads();
    if (*s_last && threadIdx.x == 0) {
        __threadfence();
        const unsigned num_resets = *reinterpret_cast<volatile unsigned *>(b.reduce_scratch + 0);
        const float finished = *reinterpret_cast<volatile float *>(b.reduce_scratch + 2);
        if (num_resets > 0)
            b.consecutive_successes[0] = p.av_factor * finished / (float)num_resets +
                                         (float)(1.0 - (double)p.av_factor) * b.consecutive_successes[0];
        b.reduce_scratch[0] = 0u; b.reduce_scratch[1] = 0u; b.reduce_scratch[2] = 0u;
        if (b.list.scratch && p.rand.step) *p.rand.step += 1ull;
    }
}

__global__ void __launch_bounds__(NT) shadow_hand_post_physics_kernel(const __grid_constant__ IgeShadowHandParams p,
                                                                      const __grid_constant__ IgeShadowHandBuffers b,
                                                                      const int use_tma, const int tile0,
                                                                      const int expected_ctas) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    extern __shared__ __align__(128) float smem[];
    const int nobs = hand_num_obs(p.obs_type);
    const int slay = hand_state_layout(p.obs_type), nst = hand_num_obs(slay);
    float *obs_tile = smem;                                             // TE * nobs
    float *st_tile = smem + ((TE * nobs + 31) / 32) * 32;               // TE * (211 | 88) when asymmetric
    __shared__ int s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int e0 = (tile0 + blockIdx.x) * TE;
    const int ne = min(TE, p.num_envs - e0);
    const bool bulk = use_tma && (ne % 4 == 0);

    fill_tile(p, b, p.obs_type, nobs, obs_tile, e0, ne);
    if (p.asymmetric_obs) fill_tile(p, b, slay, nst, st_tile, e0, ne);

    // ---- warp 0: per-env object / goal math, reward, flags (compute_hand_reward :746-800) ----
    int my_resets = 0;
    float my_finished = 0.f, st_rew = 0.f, st_len = 0.f;
    int st_to = 0;
    if (warp == 0) {
        if (lane < ne) {
            const int e = e0 + lane;
            float obj[13], goal[7];
            const float *op = b.root_state_tensor + (size_t)b.object_indices[e] * 13;
            const float *gp = b.goal_states + (size_t)e * 13;
#pragma unroll
            for (int i = 0; i < 13; ++i) obj[i] = op[i];
#pragma unroll
            for (int i = 0; i < 7; ++i) goal[i] = gp[i];
            float conj[4], qd[4];
            ige::quat_conjugate(goal + 3, conj);
            ige::quat_mul(obj + 3, conj, qd);
            fill_env_segments(p, p.obs_type, nobs, obs_tile + lane * nobs, obj, goal, qd);
            if (p.asymmetric_obs) fill_env_segments(p, slay, nst, st_tile + lane * nst, obj, goal, qd);

            long long progress = b.progress_buf[e] + 1;  // :711
            if (b.randomize_buf) b.randomize_buf[e] += 1;
            const float tol = p.ignore_z_rot ? 2.0f * p.success_tolerance : p.success_tolerance;  // :758-759
            const float goal_dist = ige::norm3(obj[0] - goal[0], obj[1] - goal[1], obj[2] - goal[2]);
            const float qn = ige::norm3(qd[0], qd[1], qd[2]);
            const float rot_dist = 2.0f * asinf(qn > 1.0f ? 1.0f : qn);
            const float dist_rew = goal_dist * p.dist_reward_scale;
            const float rot_rew = 1.0f / (fabsf(rot_dist) + p.rot_eps) * p.rot_reward_scale;
            // torch.sum(actions ** 2, dim=-1) for 20 columns: 8 vector lanes over 16, tail 4, then lanes
            const float *act = b.actions + (size_t)e * p.num_actions;
            float lanes8[8];
#pragma unroll
            for (int l = 0; l < 8; ++l) lanes8[l] = 0.f;
            const int na = p.num_actions, nv = na / 8;
            for (int i = 0; i < nv; ++i)
#pragma unroll
                for (int l = 0; l < 8; ++l) { const float a = __ldg(act + i * 8 + l); lanes8[l] += a * a; }
            float action_penalty = 0.f;
            for (int k = nv * 8; k < na; ++k) { const float a = __ldg(act + k); action_penalty += a * a; }
            if (nv > 0)
#pragma unroll
                for (int l = 0; l < 8; ++l) action_penalty += lanes8[l];
            float reward = dist_rew + rot_rew + action_penalty * p.action_penalty_scale;
            const bool hit = fabsf(rot_dist) <= tol;
            const long long goal_resets = hit ? 1 : b.reset_goal_buf[e];
            const float successes = b.successes[e] + (float)goal_resets;
            if (goal_resets == 1) reward = reward + p.reach_goal_bonus;
            const bool fell = goal_dist >= p.fall_dist;
            if (fell) reward = reward + p.fall_penalty;
            long long resets = fell ? 1 : b.reset_buf[e];
            if (p.max_consecutive_successes > 0) {
                if (hit) progress = 0;
                if (successes >= (float)p.max_consecutive_successes) resets = 1;
            }
            const bool late = (float)progress >= p.max_episode_length - 1.0f;
            if (late) resets = 1;
            if (p.max_consecutive_successes > 0 && late) reward = reward + 0.5f * p.fall_penalty;
            b.rew_buf[e] = reward;
            b.reset_buf[e] = resets;
            b.reset_goal_buf[e] = goal_resets;
            b.progress_buf[e] = progress;
            b.successes[e] = successes;
            const bool timeout = late && resets != 0;  // vec_task.py:396
            if (b.timeout_buf) b.timeout_buf[e] = timeout ? 1 : 0;
            my_resets = (int)resets;
            my_finished = successes * (float)resets;
            st_rew = reward; st_to = timeout; st_len = resets != 0 ? (float)progress : 0.f;
        }
        // global sums of compute_hand_reward (:795-796): exact (integer-valued terms)
        const int tile_resets = __reduce_add_sync(0xffffffffu, my_resets);
        const float tile_finished = ige::warp_sum(my_finished);
        if (lane == 0) {
            atomicAdd(b.reduce_scratch + 0, (unsigned)tile_resets);
            atomicAdd(reinterpret_cast<float *>(b.reduce_scratch) + 2, tile_finished);
        }
        if (b.stats) {
            const float v[7] = {lane < ne ? 1.f : 0.f, st_rew, (float)(my_resets != 0), (float)st_to, st_len,
                                (float)my_resets, my_finished};
            const int slot[7] = {IGE_STAT_STEPS, IGE_STAT_REWARD, IGE_STAT_DONE, IGE_STAT_TIMEOUT, IGE_STAT_DONE_LENGTH,
                                 IGE_STAT_TASK0, IGE_STAT_TASK0 + 1};
            ige::warp_stats_add<7>(b.stats, v, slot);
        }
    }
    if (bulk) ige::fence_proxy_async();
    __syncthreads();

    // ---- tile stores ----
    store_tile(b.obs_buf + (size_t)e0 * nobs, b.obs_clamped ? b.obs_clamped + (size_t)e0 * nobs : nullptr, obs_tile,
               ne * nobs, p.clip_obs, bulk);
    if (p.asymmetric_obs)
        store_tile(b.states_buf + (size_t)e0 * nst, b.states_clamped ? b.states_clamped + (size_t)e0 * nst : nullptr,
                   st_tile, ne * nst, p.clip_obs, bulk);
    if (bulk && tid == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");

    finish_hand_reduction(p, b, expected_ctas, &s_last);
}

// ---------------------------------------------------------------------------------------------
// post-physics, standard hand (24 DOFs == DOFs per env, 20 actions, 5 fingertips x 6 sensor floats):
// persistent CTAs over 32-env tiles, every global load of a tile issued before the first use.
// ---------------------------------------------------------------------------------------------
// The generic kernel above walks the layout segment by segment (one DRAM round trip per segment,
// runtime divisions); here the element -> (env, column) maps are compile-time, each thread first
// issues ALL of its loads for the tile (3 float2 DOF rows, one float4 each of dof_force / sensors /
// actions, <= 9 fingertip floats; warp 0 also the object / goal rows and the flags of its env) and
// only then scatters them into the shared (32 x num_obs) tile, so a tile costs two dependent DRAM
// round trips (object_indices -> object row is the second) instead of ~12.  Per-CTA partial sums stay in
// registers across tiles: 3 global atomics per CTA, not per tile.
// AllegroHand (layouts 4-6) is the same kernel instantiated with 16 DOFs / 16 actions and no sensor / fingertip columns.
constexpr int FS = 30, FT_N = 5;
constexpr int kFastCtasPerSm = 4;
constexpr int FNT = 256;       // threads per CTA of the fast kernel
constexpr int CT = FNT - 32;   // element threads (warp 0 is the env warp)
__host__ __device__ constexpr int fast_dofs(int layout) { return hand_is_allegro(layout) ? 16 : 24; }
__host__ __device__ constexpr int fast_actions(int layout) { return hand_is_allegro(layout) ? 16 : 20; }
struct SegOff { int dof_pos, dof_vel, dof_force, ft_state, ft_pos, sensors, actions; };
// element-segment offsets of the layouts (shadow_hand.py:473-584, allegro_hand.py:439-489; same numbers as kLayout)
__host__ __device__ constexpr SegOff seg_offsets(int layout) {
    return layout == 0 ? SegOff{-1, -1, -1, -1, 0, -1, 22}
         : layout == 1 ? SegOff{0, -1, -1, -1, 42, -1, 57}
         : layout == 2 ? SegOff{0, 24, -1, 72, -1, -1, 137}
         : layout == 3 ? SegOff{0, 24, 48, 96, -1, 161, 191}
         : layout == 4 ? SegOff{0, -1, -1, -1, -1, -1, 34}
         : layout == 5 ? SegOff{0, 16, -1, -1, -1, -1, 56}
                       : SegOff{0, 16, 32, -1, -1, -1, 72};
}

template <int LAYOUT, bool ASYM>
__global__ void __launch_bounds__(FNT, kFastCtasPerSm)
shadow_hand_post_physics_fast_kernel(const __grid_constant__ IgeShadowHandParams p,
                                     const __grid_constant__ IgeShadowHandBuffers b, const int num_tiles,
                                     const int expected_ctas) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    extern __shared__ __align__(128) float smem[];
    constexpr int nobs = hand_num_obs(LAYOUT);
    constexpr int FD = fast_dofs(LAYOUT), FA = fast_actions(LAYOUT);
    constexpr int SLAY = hand_state_layout(LAYOUT), NST = hand_num_obs(SLAY);   // states_buf: 211 | 88 wide
    // column-owner mapping: SUB threads share a column and split the 32 envs of the tile between them
    constexpr int DSUB = CT / FD, SSUB = CT / FS;
    constexpr int KD = (TE + DSUB - 1) / DSUB, KS = (TE + SSUB - 1) / SSUB;   // loads per thread: DOF rows, sensor floats
    static_assert(TE * FD / 4 <= CT && TE * FA / 4 <= CT && FD % 4 == 0 && FA % 4 == 0,
                  "fast kernel: one vector per thread for force / actions");
    float *obs_tile = smem;                                   // TE * nobs
    float *st_tile = smem + ((TE * nobs + 31) / 32) * 32;     // TE * NST when asymmetric
    __shared__ int s_last;
    __shared__ float s_lo[FD], s_hi[FD], s_rspan[FD];
    __shared__ int s_ft[FT_N];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr bool asym = ASYM;
    constexpr SegOff oo = seg_offsets(LAYOUT);
    constexpr SegOff so = seg_offsets(SLAY);   // states_buf
    constexpr bool need_dof = oo.dof_pos >= 0 || oo.dof_vel >= 0 || asym;
    constexpr bool need_force = oo.dof_force >= 0 || asym;
    constexpr bool need_sens = oo.sensors >= 0 || (asym && so.sensors >= 0);
    constexpr bool need_ft = oo.ft_state >= 0 || oo.ft_pos >= 0 || (asym && so.ft_state >= 0);
    constexpr bool ft_full = oo.ft_state >= 0 || (asym && so.ft_state >= 0);
    constexpr int FTW = ft_full ? 65 : 15, FTC = ft_full ? 13 : 3;   // fingertip block: 5 x (13 | 3) floats per env
    constexpr int FSUB = CT / FTW, KF = (TE + FSUB - 1) / FSUB;
    if (tid < FD) {
        const float lo = p.dof_lower[tid], hi = p.dof_upper[tid];
        s_lo[tid] = lo; s_hi[tid] = hi; s_rspan[tid] = 1.0f / (hi - lo);
    }
    if (need_ft && tid < FT_N) s_ft[tid] = p.fingertip_handles[tid];
    __syncthreads();

    int acc_resets = 0;                      // warp 0, lane == env: partial sums over this CTA's tiles
    float acc_finished = 0.f, acc_n = 0.f, acc_rew = 0.f, acc_done = 0.f, acc_to = 0.f, acc_len = 0.f;

    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int e0 = tile * TE;
        const int e = e0 + lane;             // warp 0 only
        float obj[13], goal[7], qd[4];
        long long progress = 0, resets_in = 0, goal_in = 0;
        float succ_in = 0.f;
        if (warp == 0) {
            // ---- env warp (lane == env): object / goal rows, flags, quaternion difference ----
            const long long oidx = b.object_indices[e];
            const float *gp = b.goal_states + (size_t)e * 13;
#pragma unroll
            for (int i = 0; i < 7; ++i) goal[i] = gp[i];
            progress = b.progress_buf[e];
            resets_in = b.reset_buf[e];
            goal_in = b.reset_goal_buf[e];
            succ_in = b.successes[e];
            const float *op = b.root_state_tensor + (size_t)oidx * 13;
#pragma unroll
            for (int i = 0; i < 13; ++i) obj[i] = op[i];
            if (b.randomize_buf) b.randomize_buf[e] += 1;
            float conj[4];
            ige::quat_conjugate(goal + 3, conj);
            ige::quat_mul(obj + 3, conj, qd);
            fill_env_segments(p, LAYOUT, nobs, obs_tile + lane * nobs, obj, goal, qd);
            if (asym) fill_env_segments(p, SLAY, NST, st_tile + lane * NST, obj, goal, qd);
        } else {
            // ---- element warps: every load of the tile first, then the scatter into the row-major tile(s).
            //      Each thread owns ONE column of a segment and walks the envs with a fixed stride, so addresses
            //      and tile offsets are a base + one add per element (no per-element divisions) ----
            const int ct = tid - 32;
            float2 dv[KD];
            float sv[KS], ft[KF];
            float4 fo = make_float4(0.f, 0.f, 0.f, 0.f), ac = fo;
            // DOF rows: column j = ct % 24, envs sub, sub + DSUB, ...
            const int dj = ct % FD, dsub = ct / FD;
            // sensors: column ct % 30; fingertips: column ct % FTW of the (5 x 13 | 5 x 3) block
            const int sj = ct % FS, ssub = ct / FS;
            const int fc = ct % FTW, fsub = ct / FTW;
            if (need_dof && dsub < DSUB) {
                const float2 *src = reinterpret_cast<const float2 *>(b.dof_state) + (size_t)e0 * FD + dj;
#pragma unroll
                for (int k = 0; k < KD; ++k) { const int le = dsub + k * DSUB; if (le < TE) dv[k] = __ldg(src + le * FD); }
            }
            if (need_force && ct < TE * FD / 4) fo = __ldg(reinterpret_cast<const float4 *>(b.dof_force_tensor + (size_t)e0 * FD) + ct);
            if (need_sens && ssub < SSUB) {
                const float *src = b.vec_sensor_tensor + (size_t)e0 * FS + sj;
#pragma unroll
                for (int k = 0; k < KS; ++k) { const int le = ssub + k * SSUB; if (le < TE) sv[k] = __ldg(src + le * FS); }
            }
            if (ct < TE * FA / 4) ac = __ldg(reinterpret_cast<const float4 *>(b.actions + (size_t)e0 * FA) + ct);
            if (need_ft && fsub < FSUB) {
                const int f = fc / FTC, kk = fc - f * FTC;
                const int estride = p.num_bodies * 13;
                const float *src = b.rigid_body_states + (size_t)e0 * estride + s_ft[f] * 13 + kk;
#pragma unroll
                for (int k = 0; k < KF; ++k) { const int le = fsub + k * FSUB; if (le < TE) ft[k] = __ldg(src + (size_t)le * estride); }
            }
            if (need_dof && dsub < DSUB) {
                const float lo = s_lo[dj], hi = s_hi[dj], rs = s_rspan[dj];
#pragma unroll
                for (int k = 0; k < KD; ++k) {
                    const int le = dsub + k * DSUB;
                    if (le < TE) {
                        // unscale :499-500; the divisor is loop-invariant per column -> exact reciprocal division
                        const float up = ige::div_by(2.0f * dv[k].x - hi - lo, hi - lo, rs);
                        const float uv = p.vel_obs_scale * dv[k].y;
                        if (oo.dof_pos >= 0) obs_tile[le * nobs + oo.dof_pos + dj] = up;
                        if (oo.dof_vel >= 0) obs_tile[le * nobs + oo.dof_vel + dj] = uv;
                        if (asym) { st_tile[le * NST + so.dof_pos + dj] = up; st_tile[le * NST + so.dof_vel + dj] = uv; }
                    }
                }
            }
            if (need_force && ct < TE * FD / 4) {
                const float v[4] = {fo.x, fo.y, fo.z, fo.w};
                const int i = ct * 4, le = i / FD, j = i - le * FD;   // FD % 4 == 0: one env per vector
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const float x = p.force_torque_obs_scale * v[q];
                    if (oo.dof_force >= 0) obs_tile[le * nobs + oo.dof_force + j + q] = x;
                    if (asym) st_tile[le * NST + so.dof_force + j + q] = x;
                }
            }
            if (need_sens && ssub < SSUB) {
#pragma unroll
                for (int k = 0; k < KS; ++k) {
                    const int le = ssub + k * SSUB;
                    if (le < TE) {
                        const float x = p.force_torque_obs_scale * sv[k];
                        if (oo.sensors >= 0) obs_tile[le * nobs + oo.sensors + sj] = x;
                        if (asym && so.sensors >= 0) st_tile[le * NST + so.sensors + sj] = x;
                    }
                }
            }
            if (ct < TE * FA / 4) {
                const float v[4] = {ac.x, ac.y, ac.z, ac.w};
                const int i = ct * 4, le = i / FA, j = i - le * FA;
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    obs_tile[le * nobs + oo.actions + j + q] = v[q];
                    if (asym) st_tile[le * NST + so.actions + j + q] = v[q];
                }
            }
            if (need_ft && fsub < FSUB) {
                const int f = fc / FTC, kk = fc - f * FTC;
#pragma unroll
                for (int k = 0; k < KF; ++k) {
                    const int le = fsub + k * FSUB;
                    if (le < TE) {
                        if (ft_full) {
                            if (oo.ft_state >= 0) obs_tile[le * nobs + oo.ft_state + fc] = ft[k];
                            if (oo.ft_pos >= 0 && kk < 3) obs_tile[le * nobs + oo.ft_pos + f * 3 + kk] = ft[k];
                            if (asym && so.ft_state >= 0) st_tile[le * NST + so.ft_state + fc] = ft[k];
                        } else {
                            obs_tile[le * nobs + oo.ft_pos + fc] = ft[k];
                        }
                    }
                }
            }
        }
        ige::fence_proxy_async();
        __syncthreads();
        // ---- tile out (warp 1 issues the bulk stores; warp 0 goes on with the reward) ----
        if (tid == 32) {
            ige::bulk_s2g(b.obs_buf + (size_t)e0 * nobs, obs_tile, (uint32_t)(TE * nobs) * 4);
            if (asym) ige::bulk_s2g(b.states_buf + (size_t)e0 * NST, st_tile, (uint32_t)(TE * NST) * 4);
            ige::bulk_commit();
        }
        if (warp == 0) {  // compute_hand_reward :746-800
            progress += 1;  // :711
            const float tol = p.ignore_z_rot ? 2.0f * p.success_tolerance : p.success_tolerance;  // :758-759
            const float goal_dist = ige::norm3(obj[0] - goal[0], obj[1] - goal[1], obj[2] - goal[2]);
            const float qn = ige::norm3(qd[0], qd[1], qd[2]);
            const float rot_dist = 2.0f * asinf(qn > 1.0f ? 1.0f : qn);
            const float dist_rew = goal_dist * p.dist_reward_scale;
            const float rot_rew = 1.0f / (fabsf(rot_dist) + p.rot_eps) * p.rot_reward_scale;
            const float *arow = obs_tile + lane * nobs + oo.actions;
            const float action_penalty = ige::row_sum<FA>([&](int k) { const float a = arow[k]; return a * a; });
            float reward = dist_rew + rot_rew + action_penalty * p.action_penalty_scale;
            const bool hit = fabsf(rot_dist) <= tol;
            const long long goal_resets = hit ? 1 : goal_in;
            const float successes = succ_in + (float)goal_resets;
            if (goal_resets == 1) reward = reward + p.reach_goal_bonus;
            const bool fell = goal_dist >= p.fall_dist;
            if (fell) reward = reward + p.fall_penalty;
            long long resets = fell ? 1 : resets_in;
            if (p.max_consecutive_successes > 0) {
                if (hit) progress = 0;
                if (successes >= (float)p.max_consecutive_successes) resets = 1;
            }
            const bool late = (float)progress >= p.max_episode_length - 1.0f;
            if (late) resets = 1;
            if (p.max_consecutive_successes > 0 && late) reward = reward + 0.5f * p.fall_penalty;
            b.rew_buf[e] = reward;
            b.reset_buf[e] = resets;
            b.reset_goal_buf[e] = goal_resets;
            b.progress_buf[e] = progress;
            b.successes[e] = successes;
            const bool timeout = late && resets != 0;  // vec_task.py:396
            if (b.timeout_buf) b.timeout_buf[e] = timeout ? 1 : 0;
            acc_resets += (int)resets;
            acc_finished += successes * (float)resets;
            acc_n += 1.f; acc_rew += reward; acc_done += resets != 0 ? 1.f : 0.f; acc_to += timeout ? 1.f : 0.f;
            acc_len += resets != 0 ? (float)progress : 0.f;
        }
        if (b.obs_clamped) {  // vec_task.py:404
            float *dc = b.obs_clamped + (size_t)e0 * nobs;
            for (int i = tid; i < TE * nobs; i += FNT) dc[i] = clampf(obs_tile[i], p.clip_obs);
        }
        if (asym && b.states_clamped) {
            float *dc = b.states_clamped + (size_t)e0 * NST;
            for (int i = tid; i < TE * NST; i += FNT) dc[i] = clampf(st_tile[i], p.clip_obs);
        }
        if (tid == 32) ige::bulk_wait_read_all();
        __syncthreads();  // the tile may be overwritten
    }

    ige::pdl_trigger();   // PDL: main loop done
    // ---- per-CTA sums -> global; last CTA: consecutive_successes (:798), re-arm, advance Philox ----
    if (warp == 0) {
        const int cta_resets = __reduce_add_sync(0xffffffffu, acc_resets);
        const float cta_finished = ige::warp_sum(acc_finished);   // integer-valued terms: exact in any order
        if (lane == 0) {
            atomicAdd(b.reduce_scratch + 0, (unsigned)cta_resets);
            atomicAdd(reinterpret_cast<float *>(b.reduce_scratch) + 2, cta_finished);
        }
        if (b.stats) {
            const float v[7] = {acc_n, acc_rew, acc_done, acc_to, acc_len, (float)acc_resets, acc_finished};
            const int slot[7] = {IGE_STAT_STEPS, IGE_STAT_REWARD, IGE_STAT_DONE, IGE_STAT_TIMEOUT, IGE_STAT_DONE_LENGTH,
                                 IGE_STAT_TASK0, IGE_STAT_TASK0 + 1};
            ige::warp_stats_add<7>(b.stats, v, slot);
        }
    }
    finish_hand_reduction(p, b, expected_ctas, &s_last);
}

typedef void (*HandFastKernel)(const IgeShadowHandParams, const IgeShadowHandBuffers, const int, const int);
HandFastKernel fast_kernel(int layout, bool asym) {
    static const HandFastKernel table[14] = {
        shadow_hand_post_physics_fast_kernel<0, false>, shadow_hand_post_physics_fast_kernel<0, true>,
        shadow_hand_post_physics_fast_kernel<1, false>, shadow_hand_post_physics_fast_kernel<1, true>,
        shadow_hand_post_physics_fast_kernel<2, false>, shadow_hand_post_physics_fast_kernel<2, true>,
        shadow_hand_post_physics_fast_kernel<3, false>, shadow_hand_post_physics_fast_kernel<3, true>,
        shadow_hand_post_physics_fast_kernel<4, false>, shadow_hand_post_physics_fast_kernel<4, true>,   // AllegroHand
        shadow_hand_post_physics_fast_kernel<5, false>, shadow_hand_post_physics_fast_kernel<5, true>,
        shadow_hand_post_physics_fast_kernel<6, false>, shadow_hand_post_physics_fast_kernel<6, true>};
    return table[layout * 2 + (asym ? 1 : 0)];
}

int check(const IgeShadowHandParams *p, const IgeShadowHandBuffers *b) {
    if (!p || !b) return ige_fail(IGE_ENULL, "params / buffers is NULL");
    if (p->num_envs < 0) return ige_fail(IGE_ESIZE, "num_envs < 0");
    if (p->num_hand_dofs < 1 || p->num_hand_dofs > IGE_MAX_DOF || p->num_dofs_per_env < p->num_hand_dofs)
        return ige_fail(IGE_ESIZE, "num_hand_dofs=%d num_dofs_per_env=%d invalid", p->num_hand_dofs, p->num_dofs_per_env);
    if (p->num_actions < 1 || p->num_actions > IGE_MAX_DOF) return ige_fail(IGE_ESIZE, "num_actions=%d invalid", p->num_actions);
    if (p->obs_type < 0 || p->obs_type > IGE_HAND_OBS_ALLEGRO_FULL_STATE) return ige_fail(IGE_EPARAM, "obs_type=%d invalid", p->obs_type);
    for (int f = 0; f < IGE_HAND_NUM_FINGERTIPS && !hand_is_allegro(p->obs_type); ++f)
        if (p->fingertip_handles[f] < 0 || p->fingertip_handles[f] >= p->num_bodies)
            return ige_fail(IGE_ESIZE, "fingertip_handles[%d]=%d outside [0, %d)", f, p->fingertip_handles[f], p->num_bodies);
    IGE_REQUIRE_PTR(b->dof_state);
    IGE_REQUIRE_PTR(b->root_state_tensor);
    IGE_REQUIRE_PTR(b->goal_states);
    IGE_REQUIRE_PTR(b->reset_buf);
    IGE_REQUIRE_PTR(b->reset_goal_buf);
    IGE_REQUIRE_PTR(b->progress_buf);
    IGE_REQUIRE_PTR(b->successes);
    if (!b->object_indices || !b->goal_object_indices) return ige_fail(IGE_ENULL, "object / goal index tensors are NULL");
    return ige_check_list(b->list);
}

}  // namespace

extern "C" int ige_shadow_hand_reset(const IgeShadowHandParams *p, const IgeShadowHandBuffers *b, void *stream) {
    if (int rc = check(p, b)) return rc;
    IGE_REQUIRE_PTR(b->goal_init_state);
    IGE_REQUIRE_PTR(b->object_init_state);
    IGE_REQUIRE_PTR(b->prev_targets);
    IGE_REQUIRE_PTR(b->cur_targets);
    IGE_REQUIRE_PTR(b->rb_forces);
    IGE_REQUIRE_PTR(b->random_force_prob);
    if (p->rand.mode == IGE_RAND_INJECTED && (!p->rand.goal_rank || !p->rand.env_rank))
        return ige_fail(IGE_ENULL, "rand.goal_rank / rand.env_rank is NULL (INJECTED mode)");
    if (p->rand.mode != IGE_RAND_INJECTED && p->rand.mode != IGE_RAND_PHILOX) return ige_fail(IGE_EPARAM, "rand.mode invalid");
    if (p->num_envs == 0) return IGE_OK;
    const int grid = (p->num_envs + 127) / 128;
ige_launch(shadow_hand_reset_kernel, grid, 128, 0, (cudaStream_t)stream, *p, *b);
    return ige_check_launch("ige_shadow_hand_reset");
}

extern "C" int ige_shadow_hand_post_physics(const IgeShadowHandParams *p, const IgeShadowHandBuffers *b, void *stream) {
    if (int rc = check(p, b)) return rc;
    const bool allegro = hand_is_allegro(p->obs_type);
    if (!allegro) IGE_REQUIRE_PTR(b->rigid_body_states);
    IGE_REQUIRE_PTR(b->actions);
    IGE_REQUIRE_PTR(b->obs_buf);
    IGE_REQUIRE_PTR(b->rew_buf);
    IGE_REQUIRE_PTR(b->consecutive_successes);
    if (!b->reduce_scratch) return ige_fail(IGE_ENULL, "reduce_scratch is NULL");
    const bool need_full = p->obs_type == IGE_HAND_OBS_FULL_STATE || p->obs_type == IGE_HAND_OBS_ALLEGRO_FULL_STATE || p->asymmetric_obs;
    if (need_full) {
        if (!allegro) IGE_REQUIRE_PTR(b->vec_sensor_tensor);
        IGE_REQUIRE_PTR(b->dof_force_tensor);
    }
    if (p->asymmetric_obs) IGE_REQUIRE_PTR(b->states_buf);
    IGE_OPTIONAL_PTR(b->obs_clamped);
    if (p->num_envs == 0) return IGE_OK;
    const int nobs = hand_num_obs(p->obs_type);
    const int nst = hand_num_obs(hand_state_layout(p->obs_type));
    const size_t smem = (size_t)(((TE * nobs + 31) / 32) * 32 + (p->asymmetric_obs ? TE * nst : 0)) * sizeof(float);
    // function attributes are per device (context): one flag per ordinal, not one per process
    int attr_dev = -1;
    ige_current_device_sms(&attr_dev);
    if (attr_dev < 0) return ige_fail(IGE_EPARAM, "ige_shadow_hand_post_physics: device ordinal not supported");
    static bool attr_set_of[IGE_MAX_DEVICES], fast_attr_set_of[IGE_MAX_DEVICES];
    bool &attr_set = attr_set_of[attr_dev], &fast_attr_set = fast_attr_set_of[attr_dev];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(shadow_hand_post_physics_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)((TE * MAX_OBS + 32 + TE * MAX_OBS) * sizeof(float)));
        if (e != cudaSuccess) return ige_fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        attr_set = true;
    }
    if (!fast_attr_set) {
        for (int i = 0; i < 14; ++i) {
            cudaError_t e = cudaFuncSetAttribute(fast_kernel(i >> 1, i & 1), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)((TE * MAX_OBS + 32 + TE * MAX_OBS) * sizeof(float)));
            if (e != cudaSuccess) return ige_fail((int)e, "cudaFuncSetAttribute: %s", cudaGetErrorString(e));
        }
        fast_attr_set = true;
    }
    const int tiles = (p->num_envs + TE - 1) / TE;
    cudaStream_t s = (cudaStream_t)stream;
    const int fd = fast_dofs(p->obs_type), fa = fast_actions(p->obs_type);
    const bool standard = p->num_hand_dofs == fd && p->num_dofs_per_env == fd && p->num_actions == fa && ige_use_tma() &&
                          ige_fast_paths();
    if (standard) {
        // full tiles: persistent fast kernel; a ragged last tile (num_envs % 32) goes through the generic one
        const int full = p->num_envs / TE, ragged = tiles - full;
        const int sms = ige_current_device_sms();
        const int grid = full < sms * kFastCtasPerSm ? full : sms * kFastCtasPerSm;
        if (full > 0) {
ige_launch(fast_kernel(p->obs_type, p->asymmetric_obs != 0), grid, FNT, smem, s, *p, *b, full, grid + ragged);
            if (int rc = ige_check_launch("ige_shadow_hand_post_physics(fast)")) return rc;
        }
        if (ragged == 0) return IGE_OK;
ige_launch(shadow_hand_post_physics_kernel, ragged, NT, smem, s, *p, *b, ige_use_tma(), full, grid + ragged);
    } else {
ige_launch(shadow_hand_post_physics_kernel, tiles, NT, smem, s, *p, *b, ige_use_tma(), 0, tiles);
    }
    return ige_check_launch("ige_shadow_hand_post_physics");
}
