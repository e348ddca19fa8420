// franka_osc.cu -- MA Franka pre_physics_step with the operational-space torque controller.
//
// Replaces Franka*MA.pre_physics_step + _compute_osc_torques (tasks/franka_reach_MA.py:770-818,
// franka_collect_MA.py:892-951, franka_combine_MA.py:910-969, franka_PPMA.py:897-956) and the action clamp of
// VecTask.step (vec_task.py:373): ~25 torch launches (two batched LAPACK inverses, eight bmm) -> ONE launch.
//
// Per agent row: R mm 196 + j_eef 168 + dof_state 72 + eef_vel 24 + actions 28, W actions 28 + arm torques 24 +
// gripper targets 8 = 548 B against ~2.4 kflop -> 4 flop/B, still under the fp32 ridge of a B200 (~10 flop/B):
// HBM-bound.  A warp owns 32 consecutive rows: their mm / j_eef / dof_state tiles are contiguous in memory and are
// staged into shared memory with coalesced 16-B loads; lane == row then pulls its matrices into registers
// (row strides 49 / 42 / 18 words: odd or 2-way conflicts) and does everything else in registers with fully
// unrolled loops: mm^-1 (7x7) and (J mm^-1 J^T)^-1 (6x6) by Gauss-Jordan without pivoting (both SPD).
#include "ige_common.cuh"

namespace {

constexpr int kWarps = 2;
constexpr int kTile = 32;
constexpr float kPi = 3.14159265358979323846f;

template <int N>
__device__ __forceinline__ void gj_inverse(float (&A)[N][N]) {  // in place; same op order as the oracle
#pragma unroll
    for (int k = 0; k < N; ++k) {
        const float p = 1.0f / A[k][k];
        A[k][k] = 1.0f;
#pragma unroll
        for (int j = 0; j < N; ++j) A[k][j] = A[k][j] * p;
#pragma unroll
        for (int i = 0; i < N; ++i) {
            if (i == k) continue;
            const float f = A[i][k];
            A[i][k] = 0.0f;
#pragma unroll
            for (int j = 0; j < N; ++j) A[i][j] = __fmaf_rn(-f, A[k][j], A[i][j]);
        }
    }
}

__device__ __forceinline__ float py_remainder(float a, float b) {  // torch.remainder
    float m = fmodf(a, b);
    if (m != 0.0f && ((b < 0.0f) != (m < 0.0f))) m = m + b;
    return m;
}

__device__ __forceinline__ void cp_async16(float *dst, const float *src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
}

// one warp starts the coalesced copy of `count` floats (16-B aligned global base) into shared memory: 16-B chunks
// asynchronously (LDGSTS), the tail of a ragged last tile synchronously
__device__ __forceinline__ void warp_copy_async(float *dst, const float *src, int count, int lane) {
    const int v4 = count >> 2;
    for (int i = lane; i < v4; i += 32) cp_async16(dst + 4 * i, src + 4 * i);
    for (int i = (v4 << 2) + lane; i < count; i += 32) dst[i] = __ldg(src + i);
}

// the row's small inputs straight from global: clamped actions (vec_task.py:373 + :933) and eef_vel
__device__ __forceinline__ void load_small(const IgeFrankaOscArgs &a, int r, bool ok, float (&act)[7], float (&ev)[6]) {
#pragma unroll
    for (int i = 0; i < 7; ++i) act[i] = 0.0f;
#pragma unroll
    for (int i = 0; i < 6; ++i) ev[i] = 0.0f;
    if (!ok) return;
    const int NA = a.num_actions;
#pragma unroll
    for (int i = 0; i < 7; ++i)
        if (i < NA) {
            const float x = __ldg(a.actions_in + (size_t)r * NA + i);
            act[i] = fminf(fmaxf(x, -a.clip_actions), a.clip_actions);
        }
    const int e = r / a.num_agents, g = r - e * a.num_agents;
    const float *p = a.rigid_body_state + ((size_t)e * a.num_bodies + a.eef_body_start + (size_t)g * a.eef_body_step) * 13 + 7;
#pragma unroll
    for (int i = 0; i < 6; ++i) ev[i] = __ldg(p + i);
}

// Software pipeline per warp: the tile's matrices move shared memory -> registers as soon as they land; the copy of
// the warp's NEXT tile (and its small per-row inputs) is then issued into the same shared buffer and flies during the
// ~2.6 k instructions of register arithmetic -- the register file is the second stage of the double buffer.
__global__ void __launch_bounds__(kWarps * 32) franka_osc_kernel(IgeFrankaOscArgs a, int rows) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    __shared__ __align__(16) float s_mm[kWarps][kTile * 49];
    __shared__ __align__(16) float s_j[kWarps][kTile * 42];
    __shared__ __align__(16) float s_ds[kWarps][kTile * 18];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int num_tiles = (rows + kTile - 1) / kTile;
    const int stride = gridDim.x * kWarps;
    int tile = blockIdx.x * kWarps + warp;
    float act_n[7], ev_n[6];
    auto issue = [&](int t) {
        const int r0 = t * kTile, cnt = min(kTile, rows - r0);
        warp_copy_async(s_mm[warp], a.mm + (size_t)r0 * 49, cnt * 49, lane);
        warp_copy_async(s_j[warp], a.j_eef + (size_t)r0 * 42, cnt * 42, lane);
        warp_copy_async(s_ds[warp], a.dof_state + (size_t)r0 * 18, cnt * 18, lane);
        asm volatile("cp.async.commit_group;" ::: "memory");
        load_small(a, r0 + lane, lane < cnt, act_n, ev_n);
    };
    if (tile < num_tiles) issue(tile);
    for (; tile < num_tiles; tile += stride) {
        const int r0 = tile * kTile;
        const int cnt = min(kTile, rows - r0);
        const int r = r0 + lane;
        const bool ok = lane < cnt;
        float act[7], ev[6], M[7][7], J[6][7], q[7], qd[7];
#pragma unroll
        for (int i = 0; i < 7; ++i) act[i] = act_n[i];
#pragma unroll
        for (int i = 0; i < 6; ++i) ev[i] = ev_n[i];
        asm volatile("cp.async.wait_all;" ::: "memory");
        __syncwarp();
        if (ok) {
            const float *mrow = s_mm[warp] + lane * 49, *jrow = s_j[warp] + lane * 42, *ds = s_ds[warp] + lane * 18;
#pragma unroll
            for (int i = 0; i < 7; ++i)
#pragma unroll
                for (int j = 0; j < 7; ++j) M[i][j] = mrow[i * 7 + j];
#pragma unroll
            for (int i = 0; i < 6; ++i)
#pragma unroll
                for (int j = 0; j < 7; ++j) J[i][j] = jrow[i * 7 + j];
#pragma unroll
            for (int k = 0; k < 7; ++k) q[k] = ds[2 * k], qd[k] = ds[2 * k + 1];
        }
        __syncwarp();
        if (tile + stride < num_tiles) issue(tile + stride);
        if (ok) {
            float v[6], un[7], w[7];
#pragma unroll
            for (int i = 0; i < 6; ++i) {  // :939, :913  kp * dpose - kd * eef_vel
                const float dpose = (act[i] * a.cmd_limit[i]) / a.action_scale;
                v[i] = a.kp[i] * dpose - a.kd[i] * ev[i];
            }
#pragma unroll
            for (int k = 0; k < 7; ++k) {  // :920
                const float d0 = (a.default_dof_pos[k] - q[k]) + kPi;
                const float wrap = py_remainder(d0, ige::kTwoPi) - kPi;
                un[k] = a.kd_null[k] * (-qd[k]) + a.kp_null[k] * wrap;
            }
#pragma unroll
            for (int i = 0; i < 7; ++i) {  // :922  u_null = mm @ u_null
                float s = 0.0f;
#pragma unroll
                for (int k = 0; k < 7; ++k) s = __fmaf_rn(M[i][k], un[k], s);
                w[i] = s;
            }
            gj_inverse<7>(M);  // :907
            float JM[6][7], E[6][6];
#pragma unroll
            for (int i = 0; i < 6; ++i)
#pragma unroll
                for (int j = 0; j < 7; ++j) {
                    float s = 0.0f;
#pragma unroll
                    for (int k = 0; k < 7; ++k) s = __fmaf_rn(J[i][k], M[k][j], s);
                    JM[i][j] = s;
                }
#pragma unroll
            for (int i = 0; i < 6; ++i)  // :908
#pragma unroll
                for (int j = 0; j < 6; ++j) {
                    float s = 0.0f;
#pragma unroll
                    for (int k = 0; k < 7; ++k) s = __fmaf_rn(JM[i][k], J[j][k], s);
                    E[i][j] = s;
                }
            gj_inverse<6>(E);  // :909 m_eef
            float t[6], y[6], z[6];
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                float s = 0.0f, s2 = 0.0f;
#pragma unroll
                for (int k = 0; k < 6; ++k) s = __fmaf_rn(E[i][k], v[k], s);
                t[i] = s;
#pragma unroll
                for (int k = 0; k < 7; ++k) s2 = __fmaf_rn(JM[i][k], w[k], s2);
                y[i] = s2;
            }
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                float s = 0.0f;
#pragma unroll
                for (int k = 0; k < 6; ++k) s = __fmaf_rn(E[i][k], y[k], s);
                z[i] = s;
            }
            float *eff = a.effort_control + (size_t)r * 9;
#pragma unroll
            for (int k = 0; k < 6; ++k) {  // :913 + :923, clamp :926, :944 only columns 0..5 are assigned
                float s = 0.0f, s2 = 0.0f;
#pragma unroll
                for (int i = 0; i < 6; ++i) s = __fmaf_rn(J[i][k], t[i], s);
#pragma unroll
                for (int i = 0; i < 6; ++i) s2 = __fmaf_rn(J[i][k], z[i], s2);
                const float x = s + (w[k] - s2);
                eff[k] = fminf(fmaxf(x, -a.effort_limits[k]), a.effort_limits[k]);
            }
            const int NA = a.num_actions;
#pragma unroll
            for (int i = 0; i < 7; ++i)
                if (i < NA) a.actions[(size_t)r * NA + i] = act[i];
            if (a.has_gripper) {  // :946-953
                const bool open = act[6] >= 0.0f;
                a.pos_control[(size_t)r * 9 + 7] = open ? a.gripper_upper[0] : a.gripper_lower[0];
                a.pos_control[(size_t)r * 9 + 8] = open ? a.gripper_upper[1] : a.gripper_lower[1];
            }
        }
        __syncwarp();
    }
}

}  // namespace

extern "C" int ige_franka_ma_pre_physics(const IgeFrankaOscArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    if (a->num_envs < 0 || a->num_agents < 1) return ige_fail(IGE_ESIZE, "num_envs=%d num_agents=%d invalid", a->num_envs, a->num_agents);
    if (a->num_actions < 6 || a->num_actions > 7) return ige_fail(IGE_ESIZE, "num_actions=%d must be 6 or 7", a->num_actions);
    if (a->has_gripper && a->num_actions != 7) return ige_fail(IGE_EPARAM, "has_gripper needs num_actions == 7");
    if (a->eef_body_start < 0 || a->eef_body_step < 0 ||
        a->eef_body_start + (a->num_agents - 1) * a->eef_body_step >= a->num_bodies)
        return ige_fail(IGE_ESIZE, "end-effector body index out of range");
    if (!(a->action_scale != 0.0f)) return ige_fail(IGE_EPARAM, "action_scale must be non-zero");
    IGE_REQUIRE_PTR(a->mm);
    IGE_REQUIRE_PTR(a->j_eef);
    IGE_REQUIRE_PTR(a->dof_state);
    if (!a->actions_in || !a->actions || !a->rigid_body_state || !a->effort_control)
        return ige_fail(IGE_ENULL, "actions_in / actions / rigid_body_state / effort_control must be non-NULL");
    if (a->has_gripper && !a->pos_control) return ige_fail(IGE_ENULL, "pos_control is NULL");
    const long long rows = (long long)a->num_envs * a->num_agents;
    if (rows > (1ll << 30)) return ige_fail(IGE_ESIZE, "too many agent rows");
    if (rows == 0) return IGE_OK;
    const int tiles = (int)((rows + kTile - 1) / kTile);
    static int resident = 0;  // persistent grid: exactly the CTAs that fit on the device at once
    if (resident == 0) {
        int dev = 0, sms = 148, per_sm = 4;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, franka_osc_kernel, kWarps * 32, 0) != cudaSuccess || per_sm < 1)
            per_sm = 4;
        resident = sms * per_sm;
    }
    const int grid = min((tiles + kWarps - 1) / kWarps, resident);
ige_launch(franka_osc_kernel, grid, kWarps * 32, 0, (cudaStream_t)stream, *a, (int)rows);
    return ige_check_launch("ige_franka_ma_pre_physics");
}
