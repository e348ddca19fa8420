// cartpole.cu -- fused Cartpole.post_physics_step (reference: isaacgymenvs/tasks/cartpole.py:165-173).
//
// One thread per env; per env the whole state is ONE 16-byte dof_state row (cart pos, cart vel,
// pole pos, pole vel), so a warp reads 512 contiguous bytes with one LDG.128 per thread and writes
// obs the same way.  Algorithmic traffic 68 B/env: HBM-bound, no staging needed.
// The grid is capped at the resident capacity (SM count x 8 CTAs) and the warps stride over the
// 32-env tiles: statistics and the retirement ticket are then a few atomics per CTA, not per 128 envs
// (32 k CTAs bumping one counter serialised 4 M-env launches to 20 % of the HBM roofline).
#include "ige_common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kCtasPerSm = 8;   // 8 x 256 threads = the SM's 2048 thread slots

__global__ void __launch_bounds__(kThreads) cartpole_post_physics_kernel(IgeCartpoleParams p, IgeCartpoleBuffers b) {
    ige::pdl_prologue_early();   // PDL: let the next kernel in; wait for the previous one before touching memory
    const uint64_t step = ige::load_step(p.rand);
    float st_rew = 0.f, st_done = 0.f, st_to = 0.f, st_len = 0.f, st_n = 0.f;
    const int stride = gridDim.x * kThreads;
    const int n_round = (p.num_envs + 31) & ~31;  // whole warps iterate together (ballot below)
    for (int e = blockIdx.x * kThreads + threadIdx.x; e < n_round; e += stride) {
    bool was_reset = false;
    if (e < p.num_envs) {
        float4 ds = reinterpret_cast<const float4 *>(b.dof_state)[e];  // (x, xd, th, thd)
        long long progress = b.progress_buf[e] + 1;                   // cartpole.py:166
        long long reset = b.reset_buf[e];
        was_reset = reset != 0;
        if (reset != 0) {                                              // reset_idx cartpole.py:144-157
            ds.x = 0.2f * (ige::reset_uniform(p.rand, step, 0, e, 2, 0) - 0.5f);
            ds.z = 0.2f * (ige::reset_uniform(p.rand, step, 0, e, 2, 1) - 0.5f);
            ds.y = 0.5f * (ige::reset_uniform(p.rand, step, 1, e, 2, 0) - 0.5f);
            ds.w = 0.5f * (ige::reset_uniform(p.rand, step, 1, e, 2, 1) - 0.5f);
            reinterpret_cast<float4 *>(b.dof_state)[e] = ds;
            reset = 0;
            progress = 0;
        }
        // compute_observations cartpole.py:137-140: obs = (x, xd, theta, thetad) == the dof row
        reinterpret_cast<float4 *>(b.obs_buf)[e] = ds;
        const float cart_pos = ds.x, cart_vel = ds.y, pole_angle = ds.z, pole_vel = ds.w;
        // compute_cartpole_reward cartpole.py:180-196
        float reward = 1.0f - pole_angle * pole_angle - 0.01f * fabsf(cart_vel) - 0.005f * fabsf(pole_vel);
        const bool out_x = fabsf(cart_pos) > p.reset_dist;
        const bool out_a = fabsf(pole_angle) > ige::kHalfPi;
        if (out_x || out_a) reward = -2.0f;
        const bool late = (float)progress >= p.max_episode_length - 1.0f;
        if (out_x || out_a || late) reset = 1;
        b.rew_buf[e] = reward;
        b.reset_buf[e] = reset;
        b.progress_buf[e] = progress;
        const bool timeout = late && (reset != 0);                     // vec_task.py:396
        if (b.timeout_buf) b.timeout_buf[e] = timeout ? 1 : 0;
        if (b.obs_clamped) {                                           // vec_task.py:404
            float4 c;
            c.x = fminf(fmaxf(ds.x, -p.clip_obs), p.clip_obs);
            c.y = fminf(fmaxf(ds.y, -p.clip_obs), p.clip_obs);
            c.z = fminf(fmaxf(ds.z, -p.clip_obs), p.clip_obs);
            c.w = fminf(fmaxf(ds.w, -p.clip_obs), p.clip_obs);
            reinterpret_cast<float4 *>(b.obs_clamped)[e] = c;
        }
        st_n += 1.f; st_rew += reward; st_done += reset != 0 ? 1.f : 0.f; st_to += timeout ? 1.f : 0.f;
        st_len += reset != 0 ? (float)progress : 0.f;
    }
    if (b.list.scratch) {  // warp == one 32-env tile of the reset list
        const uint32_t mask = __ballot_sync(0xffffffffu, was_reset);
        if ((threadIdx.x & 31) == 0 && e < p.num_envs) ige::publish_tile_mask(b.list, p.num_envs, e >> 5, mask);
    }
    }
    if (b.stats) {
        const float v[5] = {st_n, st_rew, st_done, st_to, st_len};
        const int slot[5] = {IGE_STAT_STEPS, IGE_STAT_REWARD, IGE_STAT_DONE, IGE_STAT_TIMEOUT, IGE_STAT_DONE_LENGTH};
        ige::warp_stats_add<5>(b.stats, v, slot);
    }
    if (b.list.scratch) ige::finish_reset_list(b.list, p.num_envs, p.rand.step, (threadIdx.x & 31) == 0);
}

}  // namespace

extern "C" int ige_cartpole_post_physics(const IgeCartpoleParams *p, const IgeCartpoleBuffers *b, void *stream) {
    if (!p || !b) return ige_fail(IGE_ENULL, "params / buffers is NULL");
    if (p->num_envs < 0) return ige_fail(IGE_ESIZE, "num_envs < 0");
    IGE_REQUIRE_PTR(b->dof_state);
    IGE_REQUIRE_PTR(b->obs_buf);
    IGE_REQUIRE_PTR(b->rew_buf);
    IGE_REQUIRE_PTR(b->reset_buf);
    IGE_REQUIRE_PTR(b->progress_buf);
    IGE_OPTIONAL_PTR(b->obs_clamped);
    if (int rc = ige_check_list(b->list)) return rc;
    if (int rc = ige_check_rand(p->rand)) return rc;
    if (p->num_envs == 0) return IGE_OK;
    static int sms[IGE_MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= IGE_MAX_DEVICES) return ige_fail(IGE_EPARAM, "device ordinal %d not supported", dev);
    if (sms[dev] == 0 && (cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms[dev] < 1))
        sms[dev] = 1;
    const int grid = min((p->num_envs + kThreads - 1) / kThreads, sms[dev] * kCtasPerSm);
ige_launch(cartpole_post_physics_kernel, grid, kThreads, 0, (cudaStream_t)stream, *p, *b);
    return ige_check_launch("ige_cartpole_post_physics");
}
