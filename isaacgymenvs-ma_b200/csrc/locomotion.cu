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

#ifndef IGE_STAGE_ARGS
#define IGE_STAGE_ARGS 0
#endif
#ifndef IGE_ARG_PREFETCH
#define IGE_ARG_PREFETCH 0
#endif
#ifndef IGE_EXP_NOCOMPUTE
#define IGE_EXP_NOCOMPUTE 0
#endif
#ifndef IGE_EXP_NOREWARD
#define IGE_EXP_NOREWARD 0
#endif

namespace {

#ifdef IGE_ROLE_CLOCKS  // debug build only (tools/role_clocks.py): where does each role warp spend its cycles
__device__ unsigned long long g_role_clk[8][8];
__device__ __forceinline__ long long clk_now() {
    long long t;
    asm volatile("mov.u64 %0, %%clock64;" : "=l"(t)::"memory");
    return t;
}
#define CLK(var) const long long var = clk_now()
#define CLK_ADD(acc, a, b) acc += (b) - (a)
#define CLK_FLUSH(role, slot, v) do { if (lane == 0) atomicAdd(&g_role_clk[role][slot], (unsigned long long)(v)); } while (0)
#else
#define CLK(var)
#define CLK_ADD(acc, a, b)
#define CLK_FLUSH(role, slot, v)
#endif

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

// one tile of the large inputs as it lies in shared memory (each member a multiple of 16 B)
template <int D, int NS, bool HUM>
struct __align__(128) InTile {
    float root[TE * 13];
    float dof[TE * D * 2];
    float sens[TE * NS];
    float force[HUM ? TE * D : 4];
    float act[TE * D];
};
template <int NUM_OBS>
struct __align__(128) OutTile {
    float obs[TE * NUM_OBS];
};
// per-tile hand-off from the role warps to the REWARD warp
template <int D>
struct Handoff {
    float pot[TE], prev[TE], obj[TE];
    float t2[D * TS];  // joints-at-limit reward term [dof][env] (the one with a division), padded: conflict-free
};

enum Role { HEADING = 0, VELOCITY = 1, ANGLES = 2, REWARD = 3, ELEM0 = 4 };

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
// roles (lane == env `e` of the tile); `root` points at the env's 13-float row in shared memory
// ---------------------------------------------------------------------------------------------
// HEADING: potentials (+ reset potentials), up / heading vectors, projections.  Writes the final
// potentials / prev_potentials / up_vec / heading_vec rows straight to global (coalesced per warp).
// HEADING's per-env global outputs.  They are written AFTER the tile barrier (store_heading): the
// fence.proxy.async every producer executes before the barrier would otherwise wait for these stores.
struct HeadingOut {
    float prev, pot, up[3], hd[3];
};
__device__ __forceinline__ void store_heading(const IgeLocomotionBuffers &b, int e, const HeadingOut &r) {
    b.prev_potentials[e] = r.prev;
    b.potentials[e] = r.pot;
    float *up = b.up_vec + (size_t)e * 3, *hd = b.heading_vec + (size_t)e * 3;
    up[0] = r.up[0]; up[1] = r.up[1]; up[2] = r.up[2];
    hd[0] = r.hd[0]; hd[1] = r.hd[1]; hd[2] = r.hd[2];
}

template <int D, int NUM_OBS>
__device__ __forceinline__ void role_heading(const IgeLocomotionParams &p, const float *root, bool was_reset,
                                             float init_x, float init_y, float potentials, float *obs_row,
                                             Handoff<D> &h, int lane, HeadingOut &r) {
    if (was_reset)  // ant.py:273-276; (init_x, init_y) = initial_root_states[e, 0:2], prefetched a tile ahead
        potentials = neg_norm_over_dt(p, ige::norm3(p.targets[0] - init_x, p.targets[1] - init_y, 0.0f));
    // compute_*_observations ant.py:384-391
    const float to_target[3] = {p.targets[0] - root[0], p.targets[1] - root[1], 0.0f};
    float n = ige::norm3(to_target[0], to_target[1], 0.0f);
    const float new_pot = neg_norm_over_dt(p, n);
    h.prev[lane] = potentials;
    h.pot[lane] = new_pot;
    r.prev = potentials;
    r.pot = new_pot;
    // compute_heading_and_up torch_jit_utils.py:247-262
    n = n < 1e-9f ? 1e-9f : n;  // normalize(), torch_jit_utils.py:65-67
    const float target_dirs[3] = {to_target[0] / n, to_target[1] / n, to_target[2] / n};
    float torso_quat[4], up_vec[3], heading_vec[3];
    ige::quat_mul(root + 3, p.inv_start_rot, torso_quat);
    ige::quat_rotate<false>(torso_quat, p.basis_vec1, up_vec);
    ige::quat_rotate<false>(torso_quat, p.basis_vec0, heading_vec);
    obs_row[10] = up_vec[2];
    obs_row[11] = ige::dot3(heading_vec, target_dirs);
#pragma unroll
    for (int i = 0; i < 3; ++i) { r.up[i] = up_vec[i]; r.hd[i] = heading_vec[i]; }
}

template <int D, bool HUM>
__device__ __forceinline__ void role_velocity(const IgeLocomotionParams &p, const float *root, float *obs_row,
                                              Handoff<D> &h, int lane) {
    float torso_quat[4], vel_loc[3], angvel_loc[3];
    ige::quat_mul(root + 3, p.inv_start_rot, torso_quat);
    ige::quat_rotate<true>(torso_quat, root + 7, vel_loc);  // compute_rot torch_jit_utils.py:265-267
    ige::quat_rotate<true>(torso_quat, root + 10, angvel_loc);
    obs_row[0] = root[2];
    obs_row[1] = vel_loc[0]; obs_row[2] = vel_loc[1]; obs_row[3] = vel_loc[2];
    const float s = HUM ? p.angular_velocity_scale : 1.0f;  // x * 1.0f is exact
    obs_row[4] = angvel_loc[0] * s; obs_row[5] = angvel_loc[1] * s; obs_row[6] = angvel_loc[2] * s;
    h.obj[lane] = root[7];  // true_objective statistic (ant.py:245-250)
}

template <bool HUM>
__device__ __forceinline__ void role_angles(const IgeLocomotionParams &p, const float *root, float *obs_row) {
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
    if (HUM) {  // humanoid.py:402-404 normalize_angle
        obs_row[7] = wrap_pi(yaw); obs_row[8] = wrap_pi(roll); obs_row[9] = wrap_pi(angle_to_target);
    } else {  // ant.py:403
        obs_row[7] = yaw; obs_row[8] = roll; obs_row[9] = angle_to_target;
    }
}

// reset_idx noise for DOF j of env e (ant.py:257-258): torch_rand_float(lo, hi) = (hi - lo) * rand + lo
__device__ __forceinline__ float2 finish_reset_dof(float init_pos, float lo, float hi, float u_pos, float u_vel) {
    const float pos = 0.4f * u_pos + (-0.2f);
    const float vel = 0.2f * u_vel + (-0.1f);
    float v = init_pos + pos;
    v = fminf(v, hi);  // tensor_clamp torch_jit_utils.py:228-230
    v = fmaxf(v, lo);
    return make_float2(v, vel);  // ant.py:260-261
}
// four consecutive uniforms (columns 4g .. 4g+3) of draw `which` for env row `e`: ONE Philox call
__device__ __forceinline__ void reset_uniform4(const IgeResetRand &r, uint64_t step, int which, int e, int d, int g,
                                               float (&u)[4]) {
    if (r.mode == IGE_RAND_PHILOX) {
        uint32_t c[4] = {(uint32_t)e, ((uint32_t)which << 16) | (uint32_t)g, (uint32_t)step, (uint32_t)(step >> 32)};
        ige::philox4x32_10(c, (uint32_t)r.seed, (uint32_t)(r.seed >> 32));
#pragma unroll
        for (int k = 0; k < 4; ++k) u[k] = (float)(c[k] >> 8) * (1.0f / 16777216.0f);
    } else {
        const float *src = (which ? r.u1 : r.u0) + (size_t)__ldg(r.reset_rank + e) * d + g * 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) u[k] = __ldg(src + k);
    }
}

// Per-lane constants of an ELEM lane.  A lane owns FIXED DOF column(s) -- one float4 group (VEC4 layouts)
// or one column -- for the envs g, g + GROUPS, ... of every tile, so the limits, the hoisted reciprocal of
// the unscale divisor and the motor-effort ratio live in registers for the whole kernel.
template <int D, int NS, bool HUM, int NW_ELEM>
struct ElemLane {
    static constexpr bool VEC4 = Layout<D, NS, HUM>::VEC4;
    static constexpr int LANES = 32 * NW_ELEM;
    static constexpr int G = D / 4;                              // float4 groups per env row (VEC4)
    static constexpr int GROUPS = VEC4 ? LANES / G : LANES / D;  // envs per pass over the tile
    static constexpr int K = VEC4 ? 4 : 1;                       // DOF columns per lane
    static constexpr int P = (TE + GROUPS - 1) / GROUPS;         // passes over a tile
    static_assert(!VEC4 || LANES % G == 0, "VEC4 mapping needs LANES % (D / 4) == 0");
    bool active;
    int g, j;  // first env of the lane, first DOF column
    float lo[K], hi[K], den[K], rden[K], ratio[K];
    __device__ __forceinline__ void init(const IgeLocomotionParams &p, int L) {
        if (VEC4) {
            active = true; g = L / G; j = (L - g * G) * 4;
        } else {
            active = L < GROUPS * D; g = active ? L / D : 0; j = active ? L - g * D : 0;
        }
#pragma unroll
        for (int k = 0; k < K; ++k) {
            lo[k] = p.dof_limits_lower[j + k]; hi[k] = p.dof_limits_upper[j + k];
            den[k] = hi[k] - lo[k];
            rden[k] = 1.0f / den[k];
            ratio[k] = HUM ? p.motor_effort_ratio[j + k] : 1.0f;
        }
    }
};

// Reset-path inputs of one tile (reset_idx reads them for the few flagged envs: scattered rows, a full DRAM
// latency each), staged ONE TILE AHEAD in shared memory with cp.async (LDGSTS): no register holds an
// in-flight load, and the consumer -- the very thread that issued the copy -- only does a wait_group.
//   xy[le * 2 + {0, 1}] = initial_root_states[e, 0:2]      (HEADING, ant.py:273)
//   pos[le * D + j]     = initial_dof_pos[e, j]            (ELEM, ant.py:260; same [TE][D] layout as global)
template <int D>
struct ResetStage {
    float xy[TE * 2];
    float pos[TE * D];
};
__device__ __forceinline__ void cp_async4(float *dst_smem, const float *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(ige::smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async16(float *dst_smem, const float *src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(ige::smem_u32(dst_smem)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int D, int NS, bool HUM, int NW_ELEM>
__device__ __forceinline__ void prefetch_reset(const IgeLocomotionBuffers &b, const ElemLane<D, NS, HUM, NW_ELEM> &c,
                                               int role, int lane, int e0, uint32_t mask, ResetStage<D> &st) {
    using EL = ElemLane<D, NS, HUM, NW_ELEM>;
    if (mask != 0u) {
        if (role == HEADING) {
            if ((mask >> lane) & 1u) {
                const float *init = b.initial_root_states + (size_t)(e0 + lane) * 13;
                cp_async4(st.xy + lane * 2, init);
                cp_async4(st.xy + lane * 2 + 1, init + 1);
            }
        } else if (role >= ELEM0 && c.active) {
#pragma unroll
            for (int k = 0; k < EL::P; ++k) {
                const int le = k * EL::GROUPS + c.g;
                if (le < TE && ((mask >> le) & 1u)) {
                    const float *src = b.initial_dof_pos + (size_t)(e0 + le) * D + c.j;
                    if (EL::VEC4) cp_async16(st.pos + le * D + c.j, src);
                    else cp_async4(st.pos + le * D + c.j, src);
                }
            }
        }
    }
    cp_async_commit();  // one group per tile, empty or not: wait_group<1> then always means "the previous tile's"
}

// ELEM: DOF observation columns (unscaled position, scaled velocity; VEC4 layouts: actions too), the
// reset_idx DOF scatter for flagged envs and the joints-at-limit reward term.  `my_flag` is the reset flag
// of env (e0 + lane); items fetch their env's flag with a shuffle.
template <int D, int NS, bool HUM, int NW_ELEM>
__device__ __forceinline__ void role_elem(const IgeLocomotionParams &p, const IgeLocomotionBuffers &b,
                                          const InTile<D, NS, HUM> &in, float *obs, Handoff<D> &h,
                                          const ElemLane<D, NS, HUM, NW_ELEM> &c, int ne, int e0, uint64_t step,
                                          uint32_t mask, const float *init_pos /* [TE][D] rows of this tile */) {
    using L = Layout<D, NS, HUM>;
    using EL = ElemLane<D, NS, HUM, NW_ELEM>;
    constexpr int NUM_OBS = L::NUM_OBS;
    constexpr float kInv002 = 1.0f / 0.02f;
    if (!c.active) return;
#pragma unroll
    for (int k = 0; k < EL::P; ++k) {
        const int le = k * EL::GROUPS + c.g;
        if (le >= ne) continue;
        const bool was_reset = (mask >> le) & 1u;
        float *o = obs + le * NUM_OBS;
        if (L::VEC4) {
            const int g = c.j >> 2, it = le * EL::G + g;
            float4 a = reinterpret_cast<const float4 *>(in.dof)[it * 2];       // p0 v0 p1 v1
            float4 d = reinterpret_cast<const float4 *>(in.dof)[it * 2 + 1];   // p2 v2 p3 v3
            if (was_reset) {  // reset_idx: rewrite this env's DOF rows (global copy + the values used below)
                const int e = e0 + le;
                float up[4], uv[4];
                const float4 ip = *reinterpret_cast<const float4 *>(init_pos + le * D + c.j);
                reset_uniform4(p.rand, step, 0, e, D, g, up);
                reset_uniform4(p.rand, step, 1, e, D, g, uv);
                const float2 r0 = finish_reset_dof(ip.x, c.lo[0], c.hi[0], up[0], uv[0]);
                const float2 r1 = finish_reset_dof(ip.y, c.lo[1], c.hi[1], up[1], uv[1]);
                const float2 r2 = finish_reset_dof(ip.z, c.lo[2], c.hi[2], up[2], uv[2]);
                const float2 r3 = finish_reset_dof(ip.w, c.lo[3], c.hi[3], up[3], uv[3]);
                a = make_float4(r0.x, r0.y, r1.x, r1.y);
                d = make_float4(r2.x, r2.y, r3.x, r3.y);
                float4 *row = reinterpret_cast<float4 *>(b.dof_state + ((size_t)e * D + c.j) * 2);
                row[0] = a;
                row[1] = d;
            }
            const float4 act = reinterpret_cast<const float4 *>(in.act)[it];
            const float pos[4] = {a.x, a.z, d.x, d.z}, vel[4] = {a.y, a.w, d.y, d.w};
            float ps[4], vs[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                ps[i] = ige::div_by(2.0f * pos[i] - c.hi[i] - c.lo[i], c.den[i], c.rden[i]);  // unscale torch_jit_utils.py:238-240
                vs[i] = vel[i] * p.dof_vel_scale;
                h.t2[(c.j + i) * TS + le] = ps[i] > 0.99f ? 1.0f : 0.0f;  // dof at limit        ant.py:355
            }
            reinterpret_cast<float4 *>(o + L::O_DOFPOS)[g] = make_float4(ps[0], ps[1], ps[2], ps[3]);
            reinterpret_cast<float4 *>(o + L::O_DOFVEL)[g] = make_float4(vs[0], vs[1], vs[2], vs[3]);
            reinterpret_cast<float4 *>(o + L::O_ACT)[g] = act;
        } else {
            const int j = c.j, it = le * D + j;
            float2 pv = reinterpret_cast<const float2 *>(in.dof)[it];
            if (was_reset) {
                const int e = e0 + le;
                pv = finish_reset_dof(init_pos[it], c.lo[0], c.hi[0], ige::reset_uniform(p.rand, step, 0, e, D, j),
                                      ige::reset_uniform(p.rand, step, 1, e, D, j));
                reinterpret_cast<float2 *>(b.dof_state)[(size_t)e0 * D + it] = pv;
            }
            const float ps = ige::div_by(2.0f * pv.x - c.hi[0] - c.lo[0], c.den[0], c.rden[0]);  // unscale :238-240
            o[L::O_DOFPOS + j] = ps;
            o[L::O_DOFVEL + j] = pv.y * p.dof_vel_scale;
            if (HUM) {
                const float ap = fabsf(ps);
                const float scaled = ige::div_by(p.joints_at_limit_cost_scale * (ap - 0.98f), 0.02f, kInv002);  // humanoid.py:356
                h.t2[j * TS + le] = (ap > 0.98f ? 1.0f : 0.0f) * scaled * c.ratio[0];                            // humanoid.py:357
            } else {
                h.t2[j * TS + le] = ps > 0.99f ? 1.0f : 0.0f;
            }
        }
    }
}

// STREAM: the observation columns that are a copy / scaling of ONE input element (force sensors, DOF
// forces, and -- scalar layouts -- the actions).  Shared by the three torso warps, which run their slice
// after their own short role, so the ELEM warps keep only the division-heavy DOF columns.
template <int D, int NS, bool HUM>
__device__ __forceinline__ void stream_columns(const IgeLocomotionParams &p, const InTile<D, NS, HUM> &in, float *obs,
                                               int part, int nparts, int lane, int ne) {
    using L = Layout<D, NS, HUM>;
    constexpr int NUM_OBS = L::NUM_OBS;
    const int first = part * 32 + lane, stride = 32 * nparts;
    if (L::VEC4) {
        constexpr int GS = NS / 4;
        for (int it = first; it < ne * GS; it += stride) {
            const int le = it / GS, g = it - le * GS;
            float4 s = reinterpret_cast<const float4 *>(in.sens)[it];
            s.x *= p.contact_force_scale; s.y *= p.contact_force_scale;
            s.z *= p.contact_force_scale; s.w *= p.contact_force_scale;
            reinterpret_cast<float4 *>(obs + le * NUM_OBS + L::O_SENS)[g] = s;
        }
    } else {
        for (int it = first; it < ne * D; it += stride) {
            const int le = it / D, j = it - le * D;
            float *o = obs + le * NUM_OBS;
            o[L::O_ACT + j] = in.act[it];
            if (HUM) o[L::O_FORCE + j] = in.force[it] * p.contact_force_scale;
        }
        for (int it = first; it < ne * NS; it += stride) {
            const int le = it / NS, cix = it - le * NS;
            obs[le * NUM_OBS + L::O_SENS + cix] = in.sens[it] * p.contact_force_scale;
        }
    }
}

struct StatAcc {
    float rew = 0.f, len = 0.f, obj = 0.f, n = 0.f, done = 0.f, to = 0.f;
};

// REWARD: reward + termination for env e (ant.py:325-371, humanoid.py:323-375, vec_task.py:396).
// Writes rew / reset / progress / randomize / time_out straight to global (coalesced per warp).
template <int D, int NS, bool HUM>
__device__ __forceinline__ void role_reward(const IgeLocomotionParams &p, const IgeLocomotionBuffers &b,
                                            const float *obs_row, const Handoff<D> &h, int lane, int e,
                                            long long reset, long long progress, long long randomize, StatAcc &st) {
    const float z = obs_row[0];
    progress += 1;  // ant.py:288
    if (reset != 0) { progress = 0; reset = 0; }  // ant.py:278-279
    const float heading_reward = obs_row[11] > 0.8f ? p.heading_weight : p.heading_weight * obs_row[11] / 0.8f;
    const float up_reward = obs_row[10] > 0.93f ? 0.0f + p.up_weight : 0.0f;
    using L = Layout<D, NS, HUM>;
    const float actions_cost = ige::row_sum<D>([&](int j) { const float a = obs_row[L::O_ACT + j]; return a * a; });
    const float electricity_cost = ige::row_sum<D>([&](int j) {
        const float t = fabsf(obs_row[L::O_ACT + j] * obs_row[L::O_DOFVEL + j]);   // ant.py:354 / humanoid.py:359
        return HUM ? t * p.motor_effort_ratio[j] : t;
    });
    const float progress_reward = h.pot[lane] - h.prev[lane];
    float total;
    if (HUM) {
        const float dof_at_limit_cost = ige::row_sum<D>([&](int j) { return h.t2[j * TS + lane]; });
        total = progress_reward + 2.0f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                p.energy_cost_scale * electricity_cost - dof_at_limit_cost;
    } else {
        float at_limit = 0.0f;  // integer count <= 8: exact in fp32 (ant.py:355 sums int64)
#pragma unroll
        for (int j = 0; j < D; ++j) at_limit += h.t2[j * TS + lane];
        total = progress_reward + 0.5f + up_reward + heading_reward - p.actions_cost_scale * actions_cost -
                p.energy_cost_scale * electricity_cost - at_limit * p.joints_at_limit_cost_scale;
    }
    const bool fallen = z < p.termination_height;
    const bool late = (float)progress >= p.max_episode_length - 1.0f;
    if (fallen) total = p.death_cost;
    if (fallen || late) reset = 1;
    b.rew_buf[e] = total;
    b.reset_buf[e] = reset;
    b.progress_buf[e] = progress;
    if (b.randomize_buf) b.randomize_buf[e] = randomize + 1;  // ant.py:289
    const bool timeout = late && reset != 0;  // vec_task.py:396
    if (b.timeout_buf) b.timeout_buf[e] = timeout ? 1 : 0;
    st.n += 1.f; st.rew += total; st.done += reset != 0 ? 1.f : 0.f; st.to += timeout ? 1.f : 0.f;
    st.len += reset != 0 ? (float)progress : 0.f; st.obj += h.obj[lane];
}

__device__ __forceinline__ void flush_stats(float *stats, const StatAcc &st) {
    const float v[6] = {st.n, st.rew, st.done, st.to, st.len, st.obj};
    const int slot[6] = {IGE_STAT_STEPS, IGE_STAT_REWARD, IGE_STAT_DONE, IGE_STAT_TIMEOUT, IGE_STAT_DONE_LENGTH,
                         IGE_STAT_TASK0};
    ige::warp_stats_add<6>(stats, v, slot);
}

template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}

enum EmitMode { EMIT_NONE = 0, EMIT_MASKS = 1, EMIT_APPEND = 2 };

// Bulk loads of one tile, split over the lane 0s of the four role warps so they issue in parallel.
// HEADING also arms the barrier (expect_tx may legally trail the first complete_tx: the phase
// cannot flip before the single expected arrival).
template <int D, int NS, bool HUM>
__device__ __forceinline__ void issue_tile_loads(int role, const IgeLocomotionBuffers &b, InTile<D, NS, HUM> &in,
                                                 uint64_t *bar, int e0, int ne) {
    const uint32_t une = (uint32_t)ne;
    if (role == HEADING) {
        uint32_t bytes = une * (13 * 4 + D * 8 + NS * 4 + D * 4);
        if (HUM) bytes += une * D * 4;
        ige::mbar_expect_tx(bar, bytes);
        ige::bulk_g2s(in.root, b.root_states + (size_t)e0 * 13, une * 52, bar);
    } else if (role == VELOCITY) {
        ige::bulk_g2s(in.dof, b.dof_state + (size_t)e0 * D * 2, une * D * 8, bar);
    } else if (role == ANGLES) {
        ige::bulk_g2s(in.sens, b.vec_sensor_tensor + (size_t)e0 * NS, une * NS * 4, bar);
        if (HUM) ige::bulk_g2s(in.force, b.dof_force_tensor + (size_t)e0 * D, une * D * 4, bar);
    } else if (role == ELEM0) {
        ige::bulk_g2s(in.act, b.actions + (size_t)e0 * D, une * D * 4, bar);
    }
}

// ---------------------------------------------------------------------------------------------
// main kernel: persistent, pipelined, TMA in / TMA out.  Handles tiles [0, num_tiles); every tile
// has ne % 4 == 0 envs (only the last may be partial).
// ---------------------------------------------------------------------------------------------
template <int D, int NS, bool HUM, int NW_ELEM, int IN_STAGES>
__global__ void __launch_bounds__(32 * (4 + NW_ELEM), HUM ? 3 : 6)  // shared memory allows 6 (Ant) / 3 (Humanoid) CTAs per SM
locomotion_pipelined_kernel(const __grid_constant__ IgeLocomotionParams p_in,
                            const __grid_constant__ IgeLocomotionBuffers b_in, const int num_tiles,
                            const int workers) {
    ige::pdl_trigger();   // PDL (ige_common.cuh): the next step's grid may be scheduled behind this one
    const IgeLocomotionParams &p = p_in;
    const IgeLocomotionBuffers &b = b_in;
    using L = Layout<D, NS, HUM>;
    constexpr int NUM_OBS = L::NUM_OBS;
    constexpr int NW = 4 + NW_ELEM;
    using In = InTile<D, NS, HUM>;
    using Out = OutTile<NUM_OBS>;

#ifdef IGE_ROLE_CLOCKS
    __shared__ long long dbg_arrive[2][8];
#endif
    extern __shared__ __align__(128) unsigned char smem_raw[];
    In *in = reinterpret_cast<In *>(smem_raw);
    Out *out = reinterpret_cast<Out *>(smem_raw + sizeof(In) * IN_STAGES);
    constexpr size_t kBarBytes = (sizeof(uint64_t) * IN_STAGES + 15) / 16 * 16;
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + sizeof(In) * IN_STAGES + sizeof(Out) * 2);
    ResetStage<D> *stage = reinterpret_cast<ResetStage<D> *>(smem_raw + sizeof(In) * IN_STAGES + sizeof(Out) * 2 + kBarBytes);
    Handoff<D> *hand = reinterpret_cast<Handoff<D> *>(stage + 2);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int role = (warp + (int)blockIdx.x) % NW;
    // Reset index list: the workers only publish their tiles' entry-flag ballots; the extra last CTA turns them
    // into the ascending id list (ige::list_scanner_cta) -- no worker ever waits for another CTA.
    if ((int)blockIdx.x >= workers) {
        ige::pdl_wait();
        ige::list_scanner_cta<32 * NW>(b_in.list, num_tiles, p_in.rand.step);
        return;
    }
#ifdef IGE_SCAN_DEBUG
    if (threadIdx.x == 0) atomicMin(&ige::g_scan_dbg[0][0], ige::gtime());
#endif
    // strided assignment: worker c owns tiles c, c + workers, ... -- the slots of one scanner round are published
    // at about the same time
    const int my_tiles = (num_tiles - (int)blockIdx.x + workers - 1) / workers;

    CLK(kstart);
    // Philox step word: read by ONE thread (lane 0 of the REWARD warp, which also publishes the CTA's reset-list
    // slots: its load has COMPLETED before its first publish, and the scanner CTA advances the word only after it
    // has seen every slot) and handed to the ELEM warps through shared memory + an mbarrier.  The tile loads below
    // are issued without waiting for it: only the reset path consumes it, one DRAM round trip later.
    __shared__ uint64_t s_step;
    __shared__ __align__(8) uint64_t s_step_bar;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < IN_STAGES; ++s) ige::mbar_init(&full[s], 1);
        ige::mbar_init(&s_step_bar, 1);
        ige::fence_mbar_init();
    }
    __syncthreads();   // mbarriers are visible (no global memory touched yet)
    ige::pdl_wait();   // everything above overlapped the previous kernel's tail; global memory from here on

    const bool clamp_copy = b.obs_clamped != nullptr && p.clip_obs < INFINITY;
    const bool second_obs_copy = b.obs_clamped != nullptr && !clamp_copy;
    auto tile_e0 = [&](int it) { return ((int)blockIdx.x + it * workers) * TE; };
    auto tile_ne = [&](int it) { return min(TE, p.num_envs - tile_e0(it)); };
    // prologue: the first IN_STAGES tiles
    if (lane == 0 && role != REWARD)
        for (int k = 0; k < IN_STAGES && k < my_tiles; ++k)
            issue_tile_loads<D, NS, HUM>(role, b, in[k], &full[k], tile_e0(k), tile_ne(k));

    if (role == REWARD) {
        // small per-env inputs are prefetched one tile ahead into registers (coalesced 8-B loads)
        long long n_reset = 0, n_progress = 0, n_rand = 0;
        if (lane < tile_ne(0)) {
            const int e = tile_e0(0) + lane;
            n_reset = b.reset_buf[e]; n_progress = b.progress_buf[e];
            if (b.randomize_buf) n_rand = b.randomize_buf[e];
        }
        // reset index list: tile 0's ballot goes out as soon as its flags are here, tile it + 1's at the end of
        // iteration it (its flags were prefetched at the top of the iteration)
        const bool emit = b.list.scratch != nullptr;
        uint32_t ltag = 0u;
        if (emit) ltag = ige::list_epoch_tag(b.list);
        if (lane == 0) {   // (both words were READ -- their values used -- before the publish below)
            s_step = ige::load_step(p_in.rand);
            ige::mbar_arrive(&s_step_bar);
        }
        if (emit) {
            const uint32_t mask = __ballot_sync(0xffffffffu, n_reset != 0);
            if (lane == 0) ige::list_publish_slot(b.list, ltag, tile_e0(0) / TE, mask);
        }
        StatAcc st;
        long long k_work = 0, k_bar = 0;
        (void)k_work; (void)k_bar;
        CLK(rp);
        for (int it = 0; it < my_tiles; ++it) {
            const long long reset = n_reset, progress = n_progress, randomize = n_rand;
            if (it + 1 < my_tiles && lane < tile_ne(it + 1)) {
                const int e = tile_e0(it + 1) + lane;
                n_reset = b.reset_buf[e]; n_progress = b.progress_buf[e];
                if (b.randomize_buf) n_rand = b.randomize_buf[e];
            } else {
                n_reset = 0;  // lanes past the end of a partial tile must not leak stale flags
            }
            CLK(r0);
            __syncthreads();  // barrier #it: the role warps have produced tile `it`
            CLK(r1);
#ifdef IGE_ROLE_CLOCKS
            if (lane == 0) {  // rows 5 / 6 of g_role_clk: [role] = times last to arrive / sum of (last - own arrival)
                long long last = 0; int who = 0;
                for (int r = 0; r < NW; ++r) { if (r == REWARD) continue; const long long t = dbg_arrive[it & 1][r]; if (t > last) { last = t; who = r; } }
                atomicAdd(&g_role_clk[5][min(who, 4)], 1ull);
                for (int r = 0; r < NW; ++r) { if (r == REWARD) continue; atomicAdd(&g_role_clk[6][min(r, 4)], (unsigned long long)(last - dbg_arrive[it & 1][r])); }
                atomicAdd(&g_role_clk[5][REWARD], (unsigned long long)(r0 > last ? 1 : 0));          // REWARD arrived last
                atomicAdd(&g_role_clk[6][REWARD], (unsigned long long)(r0 > last ? r0 - last : 0));  // ... by this much
                atomicAdd(&g_role_clk[5][7], (unsigned long long)(last > r0 ? last - r0 : 0));       // REWARD idle before the barrier
            }
#endif
#if !IGE_EXP_NOREWARD
            if (lane < tile_ne(it))
                role_reward<D, NS, HUM>(p, b, out[it & 1].obs + lane * NUM_OBS, hand[it & 1], lane, tile_e0(it) + lane,
                                    reset, progress, randomize, st);
#else
            if (lane < tile_ne(it)) { b.rew_buf[tile_e0(it) + lane] = (float)(reset + progress + randomize); }
#endif
            if (emit && it + 1 < my_tiles) {
                const uint32_t mask = __ballot_sync(0xffffffffu, n_reset != 0);
                if (lane == 0) ige::list_publish_slot(b.list, ltag, tile_e0(it + 1) / TE, mask);
            }
            CLK(r2);
            CLK_ADD(k_bar, r0, r1); CLK_ADD(k_work, r1, r2);
        }
        CLK(r3);
        if (b.stats) flush_stats(b.stats, st);
        CLK(r4);
        CLK_FLUSH(REWARD, 1, k_work); CLK_FLUSH(REWARD, 2, k_bar); CLK_FLUSH(REWARD, 3, r4 - r3);
        CLK_FLUSH(REWARD, 0, rp - kstart); CLK_FLUSH(REWARD, 4, r4 - kstart); CLK_FLUSH(REWARD, 7, my_tiles);
    } else {
        // Look-ahead of the small per-env inputs: reset flags two tiles ahead, the reset-path rows
        // (prefetch_reset, cp.async into stage[]) and HEADING's potentials one tile ahead.
        using EL = ElemLane<D, NS, HUM, NW_ELEM>;
        EL elem;
        elem.init(p, role >= ELEM0 ? (role - ELEM0) * 32 + lane : 0);
        auto load_flag = [&](int t) -> long long {
            return (role != VELOCITY && role != ANGLES && t < my_tiles && lane < tile_ne(t)) ? b.reset_buf[tile_e0(t) + lane]
                                                                                             : 0ll;
        };
        long long f_next = load_flag(1);
        float n_pot = 0.f;
        if (role == HEADING && lane < tile_ne(0)) n_pot = b.potentials[tile_e0(0) + lane];
        uint32_t m_cur = __ballot_sync(0xffffffffu, load_flag(0) != 0), m_next = 0u;
        prefetch_reset<D, NS, HUM, NW_ELEM>(b, elem, role, lane, tile_e0(0), m_cur, stage[0]);
        if (role >= ELEM0) {   // the Philox step word (reset path only): handed over by the REWARD warp, long before tile 0 lands
            ige::mbar_wait(&s_step_bar, 0);
        }
        long long k_load = 0, k_work = 0, k_bar = 0, k_issue = 0;
        (void)k_load; (void)k_work; (void)k_bar; (void)k_issue;
        CLK(cbeg);
        for (int it = 0; it < my_tiles; ++it) {
            const int s = it % IN_STAGES, o = it & 1;
            const int e0 = tile_e0(it), ne = tile_ne(it);
            const bool was_reset = (m_cur >> lane) & 1u;
            const float pot_in = n_pot;
            m_next = __ballot_sync(0xffffffffu, f_next != 0);  // tile it + 1 (flags loaded one iteration ago)
            f_next = load_flag(it + 2);
            if (role == HEADING && it + 1 < my_tiles && lane < tile_ne(it + 1)) n_pot = b.potentials[tile_e0(it + 1) + lane];
            // m_next == 0 past the last tile; stage[(it + 1) & 1] was last read two tiles ago by this same thread
            prefetch_reset<D, NS, HUM, NW_ELEM>(b, elem, role, lane, tile_e0(it + 1), m_next, stage[(it + 1) & 1]);
            HeadingOut hout;
            CLK(c0);
            ige::mbar_wait(&full[s], (it / IN_STAGES) & 1);
            cp_async_wait<1>();  // my staged reset rows of tile `it` (the group of tile it + 1 may still fly)
            CLK(c1);
            float *obs_row = out[o].obs + lane * NUM_OBS;
            const float *root = in[s].root + lane * 13;
#if !IGE_EXP_NOCOMPUTE   // (timing experiment: the data-movement skeleton alone)
            if (role == HEADING) {
                if (lane < ne)
                    role_heading<D, NUM_OBS>(p, root, was_reset, stage[o].xy[lane * 2], stage[o].xy[lane * 2 + 1], pot_in,
                                             obs_row, hand[o], lane, hout);
            } else if (role == VELOCITY) {
                if (lane < ne) role_velocity<D, HUM>(p, root, obs_row, hand[o], lane);
            } else if (role == ANGLES) {
                if (lane < ne) role_angles<HUM>(p, root, obs_row);
            } else {
                role_elem<D, NS, HUM, NW_ELEM>(p, b, in[s], out[o].obs, hand[o], elem, ne, e0,
                                               *reinterpret_cast<const volatile uint64_t *>(&s_step), m_cur, stage[o].pos);
            }
            if (role < REWARD) stream_columns<D, NS, HUM>(p, in[s], out[o].obs, role, 3, lane, ne);
#else
            hout.prev = hout.pot = root[0]; hout.up[0] = hout.up[1] = hout.up[2] = hout.hd[0] = hout.hd[1] = hout.hd[2] = pot_in;
            (void)obs_row; (void)was_reset;
#endif
            ige::fence_proxy_async();  // obs tile -> visible to the bulk-store (async) proxy
            // out[o ^ 1] is rewritten right after this barrier: its store (tile it - 1, issued by this
            // same thread one iteration ago) must have finished reading shared memory
            CLK(c2);
#ifdef IGE_ROLE_CLOCKS
            if (lane == 0) dbg_arrive[o][min(role, 4) + (role > ELEM0 ? role - ELEM0 : 0)] = c2;
#endif
            if (role == ELEM0 && lane == 0) bulk_wait_read<0>();
            __syncthreads();  // barrier #it
            CLK(c3);
            CLK_ADD(k_load, c0, c1); CLK_ADD(k_work, c1, c2); CLK_ADD(k_bar, c2, c3);
            if (role == HEADING && lane < ne) store_heading(b, e0 + lane, hout);
            if (lane == 0) {
                if (role == ELEM0) {  // tile `it` is complete in out[o]: one bulk store (two with the epilogue copy)
                    ige::bulk_s2g(b.obs_buf + (size_t)e0 * NUM_OBS, out[o].obs, (uint32_t)ne * NUM_OBS * 4);
                    if (second_obs_copy)
                        ige::bulk_s2g(b.obs_clamped + (size_t)e0 * NUM_OBS, out[o].obs, (uint32_t)ne * NUM_OBS * 4);
                    ige::bulk_commit();
                }
                if (it + IN_STAGES < my_tiles)  // in[s] is free again: refill it
                    issue_tile_loads<D, NS, HUM>(role, b, in[s], &full[s], tile_e0(it + IN_STAGES),
                                                 tile_ne(it + IN_STAGES));
            }
            if (clamp_copy) {  // finite clip: second, clamped copy with plain coalesced stores
                const int reward_warp = (REWARD + NW - (int)(blockIdx.x % NW)) % NW;
                for (int i = (warp > reward_warp ? tid - 32 : tid); i < ne * NUM_OBS; i += 32 * (NW - 1))
                    b.obs_clamped[(size_t)e0 * NUM_OBS + i] = clampf(out[o].obs[i], p.clip_obs);
            }
            m_cur = m_next;
            CLK(c6);
            CLK_ADD(k_issue, c3, c6);
        }
        CLK(c4);
        if (role == ELEM0 && lane == 0) bulk_wait_read<0>();  // shared memory must outlive the async reads
        CLK(c5);
        CLK_FLUSH(min(role, 4), 0, k_load); CLK_FLUSH(min(role, 4), 1, k_work); CLK_FLUSH(min(role, 4), 2, k_bar);
        CLK_FLUSH(min(role, 4), 3, c5 - c4); CLK_FLUSH(min(role, 4), 7, my_tiles);
        CLK_FLUSH(min(role, 4), 4, k_issue); CLK_FLUSH(min(role, 4), 5, c5 - cbeg); CLK_FLUSH(min(role, 4), 6, cbeg - kstart);
    }
#ifdef IGE_SCAN_DEBUG
    if ((threadIdx.x & 31) == 0) atomicMax(&ige::g_scan_dbg[0][1], ige::gtime());
#endif
}

// ---------------------------------------------------------------------------------------------
// fallback kernel: one tile per CTA, plain coalesced copies (ragged last tile / IGE_NO_TMA)
// ---------------------------------------------------------------------------------------------
template <int D, int NS, bool HUM, int NW_ELEM>
__global__ void __launch_bounds__(32 * (4 + NW_ELEM))
locomotion_tile_kernel(const __grid_constant__ IgeLocomotionParams p, const __grid_constant__ IgeLocomotionBuffers b,
                       const int tile_offset, const int emit_mode, const int step_bias) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    using L = Layout<D, NS, HUM>;
    constexpr int NUM_OBS = L::NUM_OBS;
    constexpr int NW = 4 + NW_ELEM;
    __shared__ InTile<D, NS, HUM> in;
    __shared__ OutTile<NUM_OBS> out;
    __shared__ Handoff<D> hand;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int role = (warp + (int)blockIdx.x) % NW;
    const int e0 = (blockIdx.x + tile_offset) * TE;
    const int ne = min(TE, p.num_envs - e0);
    // step_bias = 1: this launch follows the pipelined kernel of the SAME step, whose scanner CTA has already advanced
    // the Philox step word
    const uint64_t step = ige::load_step(p.rand) - (uint64_t)step_bias;

    coop_copy(in.root, b.root_states + (size_t)e0 * 13, ne * 13);
    coop_copy(in.dof, (const float *)b.dof_state + (size_t)e0 * D * 2, ne * D * 2);
    coop_copy(in.sens, b.vec_sensor_tensor + (size_t)e0 * NS, ne * NS);
    if (HUM) coop_copy(in.force, b.dof_force_tensor + (size_t)e0 * D, ne * D);
    coop_copy(in.act, b.actions + (size_t)e0 * D, ne * D);
    long long reset = 0, progress = 0, randomize = 0;
    float pot_in = 0.f;
    if (lane < ne) {
        reset = b.reset_buf[e0 + lane];
        if (role == HEADING) pot_in = b.potentials[e0 + lane];
        if (role == REWARD) {
            progress = b.progress_buf[e0 + lane];
            if (b.randomize_buf) randomize = b.randomize_buf[e0 + lane];
        }
    }
    __syncthreads();

    float *obs_row = out.obs + lane * NUM_OBS;
    const float *root = in.root + lane * 13;
    using EL = ElemLane<D, NS, HUM, NW_ELEM>;
    EL elem;
    elem.init(p, role >= ELEM0 ? (role - ELEM0) * 32 + lane : 0);
    const uint32_t mask = __ballot_sync(0xffffffffu, reset != 0);
    float init_x = 0.f, init_y = 0.f;
    if (role == HEADING && reset != 0) {
        init_x = __ldg(b.initial_root_states + (size_t)(e0 + lane) * 13);
        init_y = __ldg(b.initial_root_states + (size_t)(e0 + lane) * 13 + 1);
    }
    if (role == HEADING) {
        if (lane < ne) {
            HeadingOut hout;
            role_heading<D, NUM_OBS>(p, root, reset != 0, init_x, init_y, pot_in, obs_row, hand, lane, hout);
            store_heading(b, e0 + lane, hout);
        }
    } else if (role == VELOCITY) {
        if (lane < ne) role_velocity<D, HUM>(p, root, obs_row, hand, lane);
    } else if (role == ANGLES) {
        if (lane < ne) role_angles<HUM>(p, root, obs_row);
    } else if (role >= ELEM0) {
        role_elem<D, NS, HUM, NW_ELEM>(p, b, in, out.obs, hand, elem, ne, e0, step, mask, b.initial_dof_pos + (size_t)e0 * D);
    }
    if (role < REWARD) stream_columns<D, NS, HUM>(p, in, out.obs, role, 3, lane, ne);
    __syncthreads();
    if (role == REWARD) {
        if (b.list.scratch && emit_mode == EMIT_MASKS) {  // sole kernel of the step: masks, expanded below
            const uint32_t mask = __ballot_sync(0xffffffffu, reset != 0);
            if (lane == 0) ige::publish_tile_mask(b.list, p.num_envs, blockIdx.x + tile_offset, mask);
        }
        if (b.list.scratch && emit_mode == EMIT_APPEND) {  // the one ragged tile after the pipelined kernel
            const uint32_t mask = __ballot_sync(0xffffffffu, reset != 0);
            const int base = *b.list.count;
            __syncwarp();
            if (reset != 0) b.list.ids[base + __popc(mask & ((1u << lane) - 1u))] = e0 + lane;
            if (lane == 0) *b.list.count = base + __popc(mask);
        }
        StatAcc st;
        if (lane < ne) role_reward<D, NS, HUM>(p, b, obs_row, hand, lane, e0 + lane, reset, progress, randomize, st);
        if (b.stats) flush_stats(b.stats, st);
    }
    coop_copy(b.obs_buf + (size_t)e0 * NUM_OBS, (const float *)out.obs, ne * NUM_OBS);
    if (b.obs_clamped)
        for (int i = tid; i < ne * NUM_OBS; i += blockDim.x)
            b.obs_clamped[(size_t)e0 * NUM_OBS + i] = clampf(out.obs[i], p.clip_obs);
    if (b.list.scratch && emit_mode == EMIT_MASKS)
        ige::finish_reset_list(b.list, p.num_envs, p.rand.step, role == REWARD && lane == 0);
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
    if (int rc = ige_check_list(b->list)) return rc;
    return ige_check_rand(p->rand);
}

template <int D, int NS, bool HUM, int NW_ELEM, int IN_STAGES>
int launch(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, cudaStream_t stream, const char *name) {
    constexpr int NT = 32 * (4 + NW_ELEM);
    constexpr int NUM_OBS = Layout<D, NS, HUM>::NUM_OBS;
    constexpr size_t smem = sizeof(InTile<D, NS, HUM>) * IN_STAGES + sizeof(OutTile<NUM_OBS>) * 2 +
                            (sizeof(uint64_t) * IN_STAGES + 15) / 16 * 16 + sizeof(ResetStage<D>) * 2 +
                            sizeof(Handoff<D>) * 2;
    auto kern = locomotion_pipelined_kernel<D, NS, HUM, NW_ELEM, IN_STAGES>;
    static int cap[IGE_MAX_DEVICES];  // resident CTAs per device (per template instance)
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= IGE_MAX_DEVICES) return ige_fail(IGE_EPARAM, "%s: device ordinal %d not supported", name, dev);
    if (cap[dev] == 0) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return ige_fail((int)e, "%s: cudaFuncSetAttribute(smem=%zu): %s", name, smem, cudaGetErrorString(e));
        int sms = 0, per_sm = 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NT, smem);
        if (e != cudaSuccess || per_sm < 1 || sms < 1)
            return ige_fail(e != cudaSuccess ? (int)e : IGE_EPARAM, "%s: occupancy query failed", name);
        cap[dev] = sms * per_sm;
    }
    const int max_ctas = cap[dev];
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
        // workers (+ the scanner CTA of the reset list) <= resident CTAs, balanced tile counts
        const int scanner = b->list.scratch ? 1 : 0;
        const int per_cta = (pipelined_tiles + (max_ctas - scanner) - 1) / (max_ctas - scanner);
        const int workers = (pipelined_tiles + per_cta - 1) / per_cta;
ige_launch(kern, workers + scanner, NT, smem, stream, *p, *b, pipelined_tiles, workers);
        if (int rc = ige_check_launch(name)) return rc;
    }
    if (fallback_tiles > 0) {  // TMA mode: at most the one ragged tile, whose ids come last -> append
ige_launch(locomotion_tile_kernel<D, NS, HUM, NW_ELEM>, fallback_tiles, NT, 0, stream, *p, *b, fallback_offset,
                   pipelined_tiles == 0 ? (int)EMIT_MASKS : (int)EMIT_APPEND,
                   (pipelined_tiles > 0 && b->list.scratch && p->rand.mode == IGE_RAND_PHILOX && p->rand.step) ? 1 : 0);
        if (int rc = ige_check_launch(name)) return rc;
    }
    return IGE_OK;
}

}  // namespace

#ifdef IGE_ROLE_CLOCKS
extern "C" int ige_debug_role_clocks(unsigned long long *out, int reset) {
    cudaDeviceSynchronize();
    if (out) cudaMemcpyFromSymbol(out, g_role_clk, sizeof(unsigned long long) * 64);
    if (reset) { unsigned long long z[64] = {0}; cudaMemcpyToSymbol(g_role_clk, z, sizeof(z)); }
    return 0;
}
#endif

#ifdef IGE_SCAN_DEBUG
extern "C" int ige_debug_scan(unsigned long long *out, int reset) {
    cudaDeviceSynchronize();
    if (out) cudaMemcpyFromSymbol(out, ige::g_scan_dbg, sizeof(unsigned long long) * 80 * 8);
    if (reset) {
        static unsigned long long z[80 * 8];
        for (int i = 0; i < 80 * 8; ++i) z[i] = 0;
        z[0] = ~0ull;
        cudaMemcpyToSymbol(ige::g_scan_dbg, z, sizeof(z));
    }
    return 0;
}
#endif

extern "C" void ige_set_use_tma(int on) { g_use_tma = on ? 1 : 0; }
extern "C" int ige_use_tma() { return use_tma(); }

extern "C" int ige_ant_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, void *stream) {
    if (int rc = check(p, b, 8, 24, false)) return rc;
    if (p->num_envs == 0) return IGE_OK;
    return launch<8, 24, false, 1, 2>(p, b, (cudaStream_t)stream, "ige_ant_post_physics");
}

extern "C" int ige_humanoid_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, void *stream) {
    if (int rc = check(p, b, 21, 12, true)) return rc;
    if (p->num_envs == 0) return IGE_OK;
    return launch<21, 12, true, 3, 2>(p, b, (cudaStream_t)stream, "ige_humanoid_post_physics");
}
