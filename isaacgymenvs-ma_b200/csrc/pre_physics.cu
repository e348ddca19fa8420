// pre_physics.cu -- the action-scaling half of pre_physics_step, one launch each.
//
// Replaces (reference, isaacgymenvs/tasks/): the clamp of VecTask.step (base/vec_task.py:373) fused with
//   Ant.pre_physics_step       ant.py:281-285       forces = actions * joint_gears * power_scale
//   Humanoid.pre_physics_step  humanoid.py:281-285  forces = actions * motor_efforts * power_scale
//   ShadowHand.pre_physics_step shadow_hand.py:684-698 (absolute / relative position targets, moving average,
//                              limit clamp, prev_targets update) -- the reset section :670-682 is
//                              ige_shadow_hand_reset.
// Element-wise and HBM-bound (12 B / action element for the locomotion tasks): one thread per float4 of the
// (N, D) action matrix, grid-stride; the per-DOF constants come from the constant bank through a small
// shared-memory table (the column of a flat element is not warp-uniform).
#include "ige_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kCtasPerSm = 8;

__device__ __forceinline__ float clamp_sym(float x, float c) { return fminf(fmaxf(x, -c), c); }  // torch.clamp

__global__ void __launch_bounds__(kThreads) action_force_kernel(const __grid_constant__ IgeActionForceArgs a) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    __shared__ float s_gear[IGE_MAX_DOF];
    if (threadIdx.x < IGE_MAX_DOF) s_gear[threadIdx.x] = a.gears[threadIdx.x];
    __syncthreads();
    const int D = a.num_dof;
    const long long total = (long long)a.num_envs * D;
    const long long vec = total / 4;
    const float4 *in4 = reinterpret_cast<const float4 *>(a.actions_in);
    float4 *act4 = reinterpret_cast<float4 *>(a.actions);
    float4 *frc4 = reinterpret_cast<float4 *>(a.forces);
    for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < vec; i += (long long)gridDim.x * kThreads) {
        const float4 v = __ldg(in4 + i);
        float x[4] = {v.x, v.y, v.z, v.w}, f[4];
        int col = (int)((i * 4) % D);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            x[q] = clamp_sym(x[q], a.clip_actions);
            f[q] = x[q] * s_gear[col] * a.power_scale;   // (actions * gears) * power_scale, left to right
            col = col + 1 == D ? 0 : col + 1;
        }
        act4[i] = make_float4(x[0], x[1], x[2], x[3]);
        frc4[i] = make_float4(f[0], f[1], f[2], f[3]);
    }
    // tail (N * D not a multiple of 4)
    for (long long i = vec * 4 + (long long)blockIdx.x * kThreads + threadIdx.x; i < total; i += (long long)gridDim.x * kThreads) {
        const float x = clamp_sym(__ldg(a.actions_in + i), a.clip_actions);
        a.actions[i] = x;
        a.forces[i] = x * s_gear[(int)(i % D)] * a.power_scale;
    }
}

__global__ void __launch_bounds__(kThreads) hand_targets_kernel(const __grid_constant__ IgeHandTargetArgs a) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    __shared__ float s_lo[IGE_MAX_DOF], s_hi[IGE_MAX_DOF];
    __shared__ int s_dof[IGE_MAX_DOF];
    if (threadIdx.x < IGE_MAX_DOF) {
        s_lo[threadIdx.x] = a.lower[threadIdx.x];
        s_hi[threadIdx.x] = a.upper[threadIdx.x];
        s_dof[threadIdx.x] = a.actuated[threadIdx.x];
    }
    __syncthreads();
    const int NA = a.num_actions, DPE = a.num_dofs_per_env;
    const long long total = (long long)a.num_envs * NA;
    for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < total; i += (long long)gridDim.x * kThreads) {
        const long long e = i / NA;
        const int j = (int)(i - e * NA);
        const float x = clamp_sym(__ldg(a.actions_in + i), a.clip_actions);
        a.actions[i] = x;
        const size_t t = (size_t)e * DPE + s_dof[j];
        const float lo = s_lo[j], hi = s_hi[j], prev = a.prev_targets[t];
        float cur;
        if (a.use_relative_control) {               // shadow_hand.py:685-688
            cur = prev + a.speed_scale_dt * x;
        } else {                                    // :690-694: scale(), then the moving average
            cur = 0.5f * (x + 1.0f) * (hi - lo) + lo;
            cur = a.act_moving_average * cur + a.one_minus_mavg * prev;
        }
        cur = fmaxf(fminf(cur, hi), lo);            // tensor_clamp = max(min(t, upper), lower)
        a.cur_targets[t] = cur;
        a.prev_targets[t] = cur;                    // :697
    }
}

// the device helpers on their own (include/ige_b200.h, IgeHelperProbeArgs)
__global__ void __launch_bounds__(kThreads) helper_probe_kernel(const __grid_constant__ IgeHelperProbeArgs a) {
    ige::pdl_prologue();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += gridDim.x * blockDim.x) {
        float q[4], q2[4], v[3], o[IGE_PROBE_COLS];
#pragma unroll
        for (int k = 0; k < 4; ++k) { q[k] = a.q[(size_t)i * 4 + k]; q2[k] = a.q2[(size_t)i * 4 + k]; }
#pragma unroll
        for (int k = 0; k < 3; ++k) v[k] = a.v[(size_t)i * 3 + k];
        ige::quat_mul(q, q2, o);
        ige::quat_rotate<false>(q, v, o + 4);
        ige::quat_rotate<true>(q, v, o + 7);
        ige::normalize3(v, o + 10);
        ige::euler_roll_yaw(q, o[13], o[14]);
        o[15] = ige::normalize_angle(v[0]);
        ige::quat_conjugate(q, o + 16);
        o[20] = ige::norm3(v[0], v[1], v[2]);
        o[21] = ige::dot3(q, v);
        o[22] = (2.0f * v[0] - v[2] - v[1]) / (v[2] - v[1]);   // unscale(x, lower, upper), torch_jit_utils.py:238-240
        o[23] = ige::py_remainder(v[0], ige::kTwoPi);
#pragma unroll
        for (int k = 0; k < IGE_PROBE_COLS; ++k) a.out[(size_t)i * IGE_PROBE_COLS + k] = o[k];
    }
}

int grid_for(long long items) {
    const int sms = ige_current_device_sms();   // grid-stride cap: SM count x 8 CTAs
    const long long g = (items + kThreads - 1) / kThreads, cap = (long long)sms * kCtasPerSm;
    return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

extern "C" int ige_locomotion_pre_physics(const IgeActionForceArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    if (a->num_envs < 0 || a->num_dof < 1 || a->num_dof > IGE_MAX_DOF)
        return ige_fail(IGE_ESIZE, "num_envs=%d num_dof=%d invalid", a->num_envs, a->num_dof);
    IGE_REQUIRE_PTR(a->actions_in);
    IGE_REQUIRE_PTR(a->actions);
    IGE_REQUIRE_PTR(a->forces);
    if (a->actions_in == a->actions) return ige_fail(IGE_EPARAM, "actions_in must not alias actions");
    if (a->num_envs == 0) return IGE_OK;
ige_launch(action_force_kernel, grid_for((long long)a->num_envs * a->num_dof / 4), kThreads, 0, (cudaStream_t)stream, *a);
    return ige_check_launch("ige_locomotion_pre_physics");
}

extern "C" int ige_shadow_hand_pre_physics_targets(const IgeHandTargetArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    if (a->num_envs < 0 || a->num_actions < 1 || a->num_actions > IGE_MAX_DOF || a->num_dofs_per_env < a->num_actions)
        return ige_fail(IGE_ESIZE, "num_envs=%d num_actions=%d num_dofs_per_env=%d invalid", a->num_envs, a->num_actions,
                        a->num_dofs_per_env);
    for (int j = 0; j < a->num_actions; ++j)
        if (a->actuated[j] < 0 || a->actuated[j] >= a->num_dofs_per_env)
            return ige_fail(IGE_ESIZE, "actuated[%d]=%d outside [0, %d)", j, a->actuated[j], a->num_dofs_per_env);
    IGE_REQUIRE_PTR(a->actions_in);
    IGE_REQUIRE_PTR(a->actions);
    IGE_REQUIRE_PTR(a->prev_targets);
    IGE_REQUIRE_PTR(a->cur_targets);
    if (a->actions_in == a->actions) return ige_fail(IGE_EPARAM, "actions_in must not alias actions");
    if (a->num_envs == 0) return IGE_OK;
ige_launch(hand_targets_kernel, grid_for((long long)a->num_envs * a->num_actions), kThreads, 0, (cudaStream_t)stream, *a);
    return ige_check_launch("ige_shadow_hand_pre_physics_targets");
}

extern "C" int ige_debug_helper_probe(const IgeHelperProbeArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    if (a->n < 0) return ige_fail(IGE_ESIZE, "n < 0");
    if (!a->q || !a->q2 || !a->v || !a->out) return ige_fail(IGE_ENULL, "q / q2 / v / out must be non-NULL");
    if (a->n == 0) return IGE_OK;
    ige_launch(helper_probe_kernel, grid_for(a->n), kThreads, 0, (cudaStream_t)stream, *a);
    return ige_check_launch("ige_debug_helper_probe");
}
