// dr_noise.cu -- the observation / action noise of VecTask.step.
//
// Replaces  actions      = dr_randomizations['actions']['noise_lambda'](actions)            (vec_task.py:373-374)
//           self.obs_buf = dr_randomizations['observations']['noise_lambda'](self.obs_buf)   (vec_task.py:399-400)
// whose bodies are the closures apply_randomizations builds (vec_task.py:675-685 gaussian, :703-711 uniform):
//     corr  = params['corr'] (a randn_like drawn on the FIRST call and kept)            * corr_scale + corr_shift
//     noise = corr + draw * scale + shift          draw = randn_like (gaussian) | rand_like (uniform)
//     out   = tensor + noise  (additive)  |  tensor * noise  (scaling)
// with (corr_scale, corr_shift, scale, shift) = (var_corr, mu_corr, var, mu) or (hi_corr - lo_corr, lo_corr, hi - lo, lo),
// plus the clamp that follows in step(): clamp(actions, +-clip_actions) (:376) / clamp(obs_buf, +-clip_obs) (:404).
// One pass over the tensor: 16-B vector loads / stores, grid-stride, 3 reads + up to 2 writes per element.
//
// Draws: INJECTED = the caller passes the reference's own randn_like / rand_like tensors (parity: the arithmetic is
// op-by-op the reference's, -fmad=false, so results are bit-exact); PHILOX = drawn in the kernel from the counter
// (element / 4, stream, step) -- uniforms as in the reset noise, normals by Box-Muller on two of them.
#include "ige_common.cuh"

namespace {

constexpr int kThreads = 256;

__device__ __forceinline__ float noise_value(const IgeNoiseArgs &a, float x, float corr, float draw, float *clamped) {
    const float c = __fadd_rn(__fmul_rn(corr, a.corr_scale), a.corr_shift);   // corr * var_corr + mu_corr
    float n = __fadd_rn(c, __fmul_rn(draw, a.scale));                         // corr + draw * var
    n = __fadd_rn(n, a.shift);                                                //      ... + mu
    const float y = a.operation == IGE_NOISE_SCALING ? __fmul_rn(x, n) : __fadd_rn(x, n);
    // torch.clamp: min(max(y, -clip), clip); NaN propagates
    *clamped = (y != y) ? y : fminf(fmaxf(y, -a.clip), a.clip);
    return y;
}

// 4 uniforms in [0,1) of counter (block, stream, step)
__device__ __forceinline__ void philox_u4(const IgeNoiseArgs &a, uint64_t step, uint32_t block4, uint32_t stream, float u[4]) {
    uint32_t c[4] = {block4, stream, (uint32_t)step, (uint32_t)(step >> 32)};
    ige::philox4x32_10(c, (uint32_t)a.seed, (uint32_t)(a.seed >> 32));
#pragma unroll
    for (int k = 0; k < 4; ++k) u[k] = (float)(c[k] >> 8) * (1.0f / 16777216.0f);
}
// 4 standard normals: two Box-Muller pairs; (1 - u) keeps the logarithm's argument in (0, 1]
__device__ __forceinline__ void philox_n4(const IgeNoiseArgs &a, uint64_t step, uint32_t block4, uint32_t stream, float z[4]) {
    float u[4];
    philox_u4(a, step, block4, stream, u);
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const float r = sqrtf(__fmul_rn(-2.0f, logf(1.0f - u[2 * k])));
        const float t = __fmul_rn(ige::kTwoPi, u[2 * k + 1]);
        z[2 * k] = __fmul_rn(r, cosf(t));
        z[2 * k + 1] = __fmul_rn(r, sinf(t));
    }
}

__global__ void __launch_bounds__(kThreads) dr_noise_kernel(const __grid_constant__ IgeNoiseArgs a) {
    ige::pdl_prologue();
    const uint64_t step = (a.rand_mode == IGE_RAND_PHILOX && a.step) ? *a.step : 0ull;
    const int64_t quads = (a.count + 3) / 4;
    const bool vec = ((reinterpret_cast<uintptr_t>(a.x) | reinterpret_cast<uintptr_t>(a.y) |
                       reinterpret_cast<uintptr_t>(a.corr) | reinterpret_cast<uintptr_t>(a.draw) |
                       reinterpret_cast<uintptr_t>(a.y_clamped)) & 15) == 0;
    for (int64_t q = (int64_t)blockIdx.x * kThreads + threadIdx.x; q < quads; q += (int64_t)gridDim.x * kThreads) {
        const int64_t i0 = q * 4;
        const int m = a.count - i0 >= 4 ? 4 : (int)(a.count - i0);
        float x[4], corr[4], draw[4], y[4], yc[4];
        const bool full = vec && m == 4;
        if (full) {
            const float4 v = *reinterpret_cast<const float4 *>(a.x + i0);
            x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
        } else {
            for (int k = 0; k < 4; ++k) x[k] = k < m ? a.x[i0 + k] : 0.0f;
        }
        if (a.rand_mode == IGE_RAND_PHILOX) {
            if (a.init_corr) {                    // the reference's first call draws and keeps `corr` (:678-680)
                philox_n4(a, step, (uint32_t)q, 2u * a.stream + 1u, corr);
                for (int k = 0; k < m; ++k) a.corr[i0 + k] = corr[k];
            }
            if (a.distribution == IGE_NOISE_GAUSSIAN) philox_n4(a, step, (uint32_t)q, 2u * a.stream, draw);
            else philox_u4(a, step, (uint32_t)q, 2u * a.stream, draw);
        }
        if (!(a.rand_mode == IGE_RAND_PHILOX && a.init_corr)) {
            if (full) {
                const float4 v = *reinterpret_cast<const float4 *>(a.corr + i0);
                corr[0] = v.x; corr[1] = v.y; corr[2] = v.z; corr[3] = v.w;
            } else {
                for (int k = 0; k < 4; ++k) corr[k] = k < m ? a.corr[i0 + k] : 0.0f;
            }
        }
        if (a.rand_mode != IGE_RAND_PHILOX) {
            if (full) {
                const float4 v = __ldg(reinterpret_cast<const float4 *>(a.draw + i0));
                draw[0] = v.x; draw[1] = v.y; draw[2] = v.z; draw[3] = v.w;
            } else {
                for (int k = 0; k < 4; ++k) draw[k] = k < m ? a.draw[i0 + k] : 0.0f;
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) y[k] = noise_value(a, x[k], corr[k], draw[k], &yc[k]);
        if (full) {
            *reinterpret_cast<float4 *>(a.y + i0) = make_float4(y[0], y[1], y[2], y[3]);
            if (a.y_clamped) *reinterpret_cast<float4 *>(a.y_clamped + i0) = make_float4(yc[0], yc[1], yc[2], yc[3]);
        } else {
            for (int k = 0; k < m; ++k) {
                a.y[i0 + k] = y[k];
                if (a.y_clamped) a.y_clamped[i0 + k] = yc[k];
            }
        }
    }
}

}  // namespace

extern "C" int ige_dr_noise(const IgeNoiseArgs *a, void *stream) {
    if (!a) return ige_fail(IGE_ENULL, "args is NULL");
    if (a->count < 0 || a->count > ((int64_t)1 << 33)) return ige_fail(IGE_ESIZE, "ige_dr_noise: count=%lld out of range", (long long)a->count);
    if (a->operation != IGE_NOISE_ADDITIVE && a->operation != IGE_NOISE_SCALING) return ige_fail(IGE_EPARAM, "ige_dr_noise: operation invalid");
    if (a->distribution != IGE_NOISE_GAUSSIAN && a->distribution != IGE_NOISE_UNIFORM) return ige_fail(IGE_EPARAM, "ige_dr_noise: distribution invalid");
    if (a->rand_mode != IGE_RAND_INJECTED && a->rand_mode != IGE_RAND_PHILOX) return ige_fail(IGE_EPARAM, "ige_dr_noise: rand_mode invalid");
    if (!a->x || !a->y || !a->corr) return ige_fail(IGE_ENULL, "ige_dr_noise: x / y / corr must be non-NULL");
    if (a->rand_mode == IGE_RAND_INJECTED && !a->draw) return ige_fail(IGE_ENULL, "ige_dr_noise: draw is NULL (INJECTED mode)");
    if (a->rand_mode == IGE_RAND_INJECTED && a->init_corr) return ige_fail(IGE_EPARAM, "ige_dr_noise: init_corr needs PHILOX mode (INJECTED: the caller fills corr)");
    if (!(a->clip >= 0.0f)) return ige_fail(IGE_EPARAM, "ige_dr_noise: clip must be >= 0 (INFINITY = none)");
    if (a->count == 0) return IGE_OK;
    const int64_t quads = (a->count + 3) / 4;
    const int64_t want = (quads + kThreads - 1) / kThreads, cap = (int64_t)ige_current_device_sms() * 8;
    const int grid = (int)(want < cap ? want : cap);
    ige_launch(dr_noise_kernel, grid, kThreads, 0, (cudaStream_t)stream, *a);
    return ige_check_launch("ige_dr_noise");
}
