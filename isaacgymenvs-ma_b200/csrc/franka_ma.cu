// franka_ma.cu -- fused post_physics_step of the fork's multi-agent Franka tasks for sm_100a.
//
// Replaces (reference, isaacgymenvs/tasks/): FrankaReachMA (franka_reach_MA.py:519-612, :616-704,
// :821-828, :875-889, :928-960), FrankaCollectMA (franka_collect_MA.py:549-734, :737-860, :1082-1174),
// FrankaCombineMA (franka_combine_MA.py:565-789, :793-908, :1080-1166), FrankaPPMA (franka_PPMA.py:
// 561-772, :776-895, :958-966, :1082-1167): progress++, agent flags -> AND-filtered env resets
// (cube / destination / stack-base root rows, arm DOF rows, controls), nearest-cube argmin, the
// per-agent FSM, the global FSM, the agent-axis observation rows (cyclic shift: every agent sees
// itself first), the FSM-gated reward, the time-out reset flag and the logged reward-component means.
//
// One thread per ENV (it owns the env's A agent rows: the per-env quantities -- distance matrix,
// nearest ids, FSM of every agent -- are shared by the rows, so nothing is exchanged between
// threads).  The gathers from the rigid-body / root tensors are 28-byte pieces of 52-byte rows
// (sector granular by nature); the (envs x A x num_obs) observation tile is assembled in shared
// memory and leaves with one TMA bulk store, so the dominant write stream is fully coalesced.
//
// Combine / PP: compute_global_FSM pairs flat row k with d_ecp[k] of a (2, N)-shaped stack that was
// flatten()ed (franka_PPMA.py:606), i.e. row k reads the gripper / cube of env k or k - N.  Those
// cube rows may be rewritten by that other env's reset, so these two variants run the reset as its
// own launch first (grid-wide ordering by the stream); Reach / Collect are a single launch.
#include <cstdlib>

#include "ige_common.cuh"

#ifndef IGE_STAGE_ARGS
#define IGE_STAGE_ARGS 0
#endif

namespace {

constexpr int TE = 64;   // envs per CTA (two 32-env list tiles)
constexpr int NT = 64;
constexpr int MAXA = IGE_MA_MAX_AGENTS;
constexpr int NC = IGE_MA_NUM_COMPONENTS;
enum Phase { PH_FUSED = 0, PH_RESET = 1, PH_COMPUTE = 2 };

__device__ __forceinline__ float clampf(float x, float c) { return x < -c ? -c : (x > c ? c : x); }

__device__ __forceinline__ float ma_uniform(const IgeMAResetRand &r, uint64_t step, int draw, int rank, int env, int cols,
                                            int j) {
    if (r.mode == IGE_RAND_PHILOX) {
        uint32_t c[4] = {(uint32_t)env, ((uint32_t)draw << 16) | ((uint32_t)j >> 2), (uint32_t)step, (uint32_t)(step >> 32)};
        ige::philox4x32_10(c, (uint32_t)r.seed, (uint32_t)(r.seed >> 32));
        return (float)(c[j & 3] >> 8) * (1.0f / 16777216.0f);
    }
    return __ldg(r.u[draw] + (size_t)rank * cols + j);
}
__device__ __forceinline__ const float *body_row(const IgeFrankaMAParams &p, const IgeFrankaMABuffers &b, int e, int start,
                                                 int step, int i) {
    return b.rigid_body_state + ((size_t)e * p.num_bodies + start + step * i) * 13;
}
__device__ __forceinline__ float *actor_row(const IgeFrankaMAParams &p, const IgeFrankaMABuffers &b, int e, int actor) {
    return b.root_state + ((size_t)e * p.num_actors + actor) * 13;
}

// reset_idx for one env whose agents are all flagged (franka_reach_MA.py:616-704 and siblings), done by a
// whole warp: one uniform (one Philox call) per lane, then the rows are written item-parallel.  A single
// thread doing this serially (~25 Philox calls) used to hold its whole CTA back for microseconds.
// us[]: per-warp shared scratch (>= 32 floats).  Draw 2 (the z noise of the destination / stack-base cube)
// is multiplied by 0.0 in the reference (franka_PPMA.py:880), so its value never matters and is not drawn.
__device__ void reset_env_coop(const IgeFrankaMAParams &p, const IgeFrankaMABuffers &b, uint64_t step, int e, int lane,
                               float *us) {
    const int A = p.num_agents, T = p.num_targets, V = p.variant;
    const IgeMAResetRand &r = p.rand;
    const int rk = r.env_rank ? __ldg(r.env_rank + e) : 0;
    const int naux = V == IGE_FRANKA_PP ? T : (V == IGE_FRANKA_COMBINE ? 1 : 0);
    // uniform table: [0, T) draw 0 | [T, 3T) draw 1 | [3T, 3T + 2 naux) draw 3 | then 9 of draw 4
    const int o1 = T, o3 = 3 * T, o4 = o3 + 2 * naux, nu = o4 + 9;
    if (lane < nu) {
        int draw, cols, j;
        if (lane < o1) { draw = 0; cols = T; j = lane; }
        else if (lane < o3) { draw = 1; cols = 2 * T; j = lane - o1; }
        else if (lane < o4) { draw = 3; cols = 2 * naux; j = lane - o3; }
        else { draw = 4; cols = 9; j = lane - o4; }
        float u = ma_uniform(r, step, draw, rk, e, cols, j);
        if (draw == 4) {  // :638-644: the arm's joint sample (the same one is tiled over the A arms)
            float v = p.franka_default_dof_pos[j] + p.dof_noise_mult * (u - 0.5f);
            v = fminf(v, p.franka_dof_upper[j]);
            v = fmaxf(v, p.franka_dof_lower[j]);
            u = j >= 7 ? p.franka_default_dof_pos[j] : v;
        }
        us[lane] = u;
    }
    __syncwarp();
    for (int it = lane; it < T * 13; it += 32) {  // _reset_init_cube_state
        const int c = it / 13, i = it - c * 13;
        float v = 0.0f;
        if (i == 0) v = p.cube_xy_center[0] + p.cube_xy_mult * (us[o1 + 2 * c] - 0.5f);
        else if (i == 1) v = p.cube_xy_center[1] + p.cube_xy_mult * (us[o1 + 2 * c + 1] - 0.5f);
        else if (i == 2) v = p.cube_z_base + us[c] * p.cube_z_range;
        else if (i == 6) v = 1.0f;
        b.init_cubeA_state[((size_t)e * T + c) * 13 + i] = v;
        actor_row(p, b, e, p.cube_actor_start + p.cube_actor_step * c)[i] = v;
    }
    for (int it = lane; it < naux * 13; it += 32) {  // _reset_dest_cube_state / _reset_stack_base_cube_state
        const int c = it / 13, i = it - c * 13;
        float v = 0.0f;
        if (i == 0) v = p.aux_xy_center_x[c] + p.aux_xy_mult * (us[o3 + 2 * c] - 0.5f);
        else if (i == 1) v = p.aux_xy_center_y + p.aux_xy_mult * (us[o3 + 2 * c + 1] - 0.5f);
        else if (i == 2) v = p.aux_z_base + 0.0f;
        else if (i == 6) v = 1.0f;
        const int actor = V == IGE_FRANKA_PP ? p.dest_actor_start + p.dest_actor_step * c : p.stack_base_actor;
        actor_row(p, b, e, actor)[i] = v;
    }
    for (int it = lane; it < A * 9; it += 32) {  // :646-656
        const int j = it % 9;
        const size_t k = (size_t)e * p.num_dofs + it;
        const float pos = us[o4 + j];
        reinterpret_cast<float2 *>(b.dof_state)[k] = make_float2(pos, 0.0f);
        b.pos_control[k] = pos;
        b.effort_control[k] = 0.0f;
    }
    if (lane < A) { b.progress_buf[(size_t)e * A + lane] = 0; b.reset_buf[(size_t)e * A + lane] = 0; }
    __syncwarp();
}

__device__ __forceinline__ float pow2_small(int s) { return s >= 0 ? (float)(1 << s) : 0.5f; }  // torch.pow(2.0, FSM)

// torch.sum(actions ** 2, dim=-1) for fewer than 8 columns: ATen's scalar path (4 interleaved
// accumulators, remainder into the first)
__device__ __forceinline__ float sum_sq_small(const float *act, int d) {
    float p4[4] = {0.f, 0.f, 0.f, 0.f};
    const int q = d / 4;
    for (int i = 0; i < q; ++i)
#pragma unroll
        for (int k = 0; k < 4; ++k) { const float a = act[4 * i + k]; p4[k] += a * a; }
    for (int i = 4 * q; i < d; ++i) { const float a = act[i]; p4[0] += a * a; }
    return ((p4[0] + p4[1]) + p4[2]) + p4[3];
}

// Common tail of both kernels: the CTA's observation tile leaves with one bulk store (or a coalesced loop for
// ragged tiles), then the reward-component sums / statistics / reset list / last-CTA means.
__device__ __forceinline__ void franka_tile_out_and_reduce(const IgeFrankaMAParams &p, const IgeFrankaMABuffers &b,
                                                           const float *obs_tile, int *s_last_p, float (&comp)[NC],
                                                           float st_rew, float st_done, float st_to, float st_len,
                                                           bool valid, int phase, int use_tma, int A, int num_obs_) {
    const int tid = threadIdx.x, lane = tid & 31, N = p.num_envs;
    const int e0 = blockIdx.x * TE;
    const int ne = min(TE, N - e0);
    int &s_last = *s_last_p;
    // ---- observation tile out: one bulk store (or a coalesced loop for ragged tiles) ----
    const int count = ne * A * num_obs_;
    const bool bulk = use_tma && (count % 4 == 0);
    if (bulk) ige::fence_proxy_async();
    __syncthreads();
    float *dst = b.obs_buf + (size_t)e0 * A * num_obs_;
    if (bulk) {
        if (tid == 0) { ige::bulk_s2g(dst, obs_tile, (uint32_t)count * 4); ige::bulk_commit(); }
    } else {
        for (int i = tid; i < count; i += NT) dst[i] = obs_tile[i];
    }
    if (b.obs_clamped) {  // vec_task.py:404
        float *dc = b.obs_clamped + (size_t)e0 * A * num_obs_;
        for (int i = tid; i < count; i += NT) dc[i] = clampf(obs_tile[i], p.clip_obs);
    }

    // ---- reward-component sums (the .mean()s of extras) + statistics ----
#pragma unroll
    for (int i = 0; i < NC; ++i) comp[i] = ige::warp_sum(comp[i]);
    if (lane < NC) {
        float mine = 0.f;
#pragma unroll
        for (int i = 0; i < NC; ++i) if (lane == i) mine = comp[i];
        atomicAdd(b.reduce_scratch + lane, mine);
        if (b.stats) atomicAdd(b.stats + (blockIdx.x % IGE_STATS_ROWS) * IGE_STATS_STRIDE + IGE_STAT_TASK0 + lane, mine);
    }
    if (b.stats) {
        const float v[5] = {valid ? (float)A : 0.f, st_rew, st_done, st_to, st_len};
        const int slot[5] = {IGE_STAT_STEPS, IGE_STAT_REWARD, IGE_STAT_DONE, IGE_STAT_TIMEOUT, IGE_STAT_DONE_LENGTH};
        ige::warp_stats_add<5>(b.stats, v, slot);
    }
    // ---- ONE retirement ticket per CTA: the CTA that arrives last expands the reset list (single-launch
    //      variants) and turns the component sums into the logged means, lane-parallel ----
    if (lane < NC) __threadfence();  // the lanes that published (ballot mask by lane 0, sums by lanes < NC)
    __syncthreads();
    if (tid == 0) {
        unsigned *cnt = reinterpret_cast<unsigned *>(b.reduce_scratch + NC);
        s_last = atomicAdd(cnt, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        if (phase == PH_FUSED && b.list.scratch) ige::expand_reset_list(b.list, N, nullptr);
        if (tid < NC) {
            const float sum = __ldcg(b.reduce_scratch + tid);
            if (b.reward_components) b.reward_components[tid] = sum / (float)((long long)N * A);
            b.reduce_scratch[tid] = 0.0f;
        }
        if (tid == 0) {
            b.reduce_scratch[NC] = 0.0f;
            if (b.list.scratch && p.rand.step) *p.rand.step += 1ull;
        }
    }
    if (bulk && tid == 0) ige::bulk_wait_read_all();  // the tile must outlive the bulk store's reads
}

__global__ void __launch_bounds__(NT) franka_ma_kernel(const __grid_constant__ IgeFrankaMAParams p,
                                                       const __grid_constant__ IgeFrankaMABuffers b, const int phase,
                                                       const int use_tma) {
    ige::pdl_prologue();   // PDL: let the next kernel in; wait for the previous one before touching memory
    extern __shared__ __align__(128) float obs_tile[];  // TE * A * num_obs
    __shared__ int s_last;
    __shared__ float s_u[NT / 32][32];
    const int tid = threadIdx.x, lane = tid & 31;
    const int N = p.num_envs, A = p.num_agents, T = p.num_targets, NA = p.num_actions, NO = p.num_obs, V = p.variant;
    const int e0 = blockIdx.x * TE;
    const int ne = min(TE, N - e0);
    const int e = e0 + tid;
    const bool valid = tid < ne;
    const uint64_t step = (p.rand.mode == IGE_RAND_PHILOX && p.rand.step)
                              ? *reinterpret_cast<const volatile uint64_t *>(p.rand.step) : 0ull;

    // ---- post_physics_step :821-826: progress++, agent flags -> reset_idx (AND filter) ----
    bool did_reset = false;
    if (phase != PH_COMPUTE && valid) {
        bool all = true;
        for (int a = 0; a < A; ++a) {
            const size_t row = (size_t)e * A + a;
            b.progress_buf[row] += 1;
            all &= b.reset_buf[row] != 0;
        }
        did_reset = all;
    }
    if (phase != PH_COMPUTE) {
        const uint32_t mask = __ballot_sync(0xffffffffu, did_reset);
        if (b.list.scratch && lane == 0 && e < N) ige::publish_tile_mask(b.list, N, e >> 5, mask);
        for (uint32_t m = mask; m; m &= m - 1)  // the warp resets its flagged envs one after the other, together
            reset_env_coop(p, b, step, e - lane + __ffs(m) - 1, lane, s_u[tid >> 5]);
    }
    if (phase == PH_RESET) {
        if (b.list.scratch) ige::finish_reset_list(b.list, N, nullptr, lane == 0);
        return;
    }

    // ---- _update_states / compute_FSM / compute_observations / compute_franka_reward ----
    float comp[NC];
#pragma unroll
    for (int i = 0; i < NC; ++i) comp[i] = 0.0f;
    float st_rew = 0.f, st_done = 0.f, st_to = 0.f, st_len = 0.f;
    if (valid) {
        float eef[MAXA][7], cube[MAXA][7];
        for (int a = 0; a < A; ++a) {
            const float *r = body_row(p, b, e, p.eef_body_start, p.eef_body_step, a);
#pragma unroll
            for (int i = 0; i < 7; ++i) eef[a][i] = __ldg(r + i);
        }
        for (int c = 0; c < T; ++c) {
            const float *r = actor_row(p, b, e, p.cube_actor_start + p.cube_actor_step * c);
#pragma unroll
            for (int i = 0; i < 7; ++i) cube[c][i] = r[i];
        }
        // _get_min_relative_pos (franka_reach_MA.py:521-531)
        float ncube[MAXA][3];   // nearest cube position per agent, kept by value
        unsigned touched = 0u;   // bit c set <=> cube c is some agent's nearest cube
        for (int a = 0; a < MAXA; ++a) { ncube[a][0] = ncube[a][1] = ncube[a][2] = 0.f; }
        for (int a = 0; a < A; ++a) {
            float best = 0.0f, bx = 0.f, by = 0.f, bz = 0.f;
            int bi = 0;
            for (int c = 0; c < T; ++c) {
                const float d = ige::norm3(cube[c][0] - eef[a][0], cube[c][1] - eef[a][1], cube[c][2] - eef[a][2]);
                if (c == 0 || d < best) { best = d; bi = c; bx = cube[c][0]; by = cube[c][1]; bz = cube[c][2]; }
            }
            ncube[a][0] = bx; ncube[a][1] = by; ncube[a][2] = bz;
            touched |= 1u << bi;
        }
        int fsm[MAXA] = {0, 0, 0, 0};
        float md[MAXA] = {0.f, 0.f, 0.f, 0.f};
        const float on_th = (float)((double)p.cubeA_size * 0.5 * 0.9);
        for (int a = 0; a < A; ++a) {
            const float alast = __ldg(b.actions + ((size_t)e * A + a) * NA + NA - 1);
            const bool closed = alast < 0.0f;
            int s = 0;
            if (V == IGE_FRANKA_REACH || V == IGE_FRANKA_COLLECT) {
                const float *nc = ncube[a];
                md[a] = ige::norm3(nc[0] - eef[a][0], nc[1] - eef[a][1], nc[2] - eef[a][2]);
                if (V == IGE_FRANKA_COLLECT) {  // franka_collect_MA.py:549-606
                    const bool close = md[a] <= on_th;
                    const bool holding = close && closed;
                    const bool high = (nc[2] - 1.05f) > (float)((double)p.wall_height + (double)p.cubeA_size / 2);
                    const bool in_area = (nc[1] > (float)((double)p.wall_pos_y + (double)p.cubeA_size * 1.)) && (fabsf(nc[0]) < 0.6f);
                    const bool low = (nc[2] - 1.05f) < (float)((double)p.wall_height / 2.);
                    if (close) s = 1;
                    if (holding) s = 2;
                    if (holding && high) s = 3;
                    if (holding && in_area) s = 4;
                    if (holding && in_area && low) s = 5;
                    if (holding && in_area && low && !closed) s = 6;
                }
            } else {  // Combine :565-611 / PP :561-595: own cube
                md[a] = ige::norm3(cube[a][0] - eef[a][0], cube[a][1] - eef[a][1], cube[a][2] - eef[a][2]);
                const bool on = md[a] <= on_th;
                if (on) s = 1;
                if (on && closed) s = 2;
                if (V == IGE_FRANKA_PP) {
                    const float *dest = actor_row(p, b, e, p.dest_actor_start + p.dest_actor_step * a);
                    const float dz = dest[2] - cube[a][2];
                    const bool lifted = fabsf(dz) >= (float)(((double)p.dest_size + (double)p.cubeA_size) / 2.);
                    const bool align = ige::norm2(dest[0] - cube[a][0], dest[1] - cube[a][1]) < (float)((0.05) * 0.5);
                    const bool stackable = fabsf(dz) <= (float)((double)p.cubeA_size * 0.8660 + (double)p.dest_size / 2.);
                    if (on && lifted) s = 3;
                    if (align) s = 4;
                    if (align && stackable) s = 5;
                    if (align && stackable && !closed) s = 6;
                }
            }
            fsm[a] = s;
        }
        for (int a = 0; a < A; ++a) {
            const size_t row = (size_t)e * A + a;
            const float *act = b.actions + row * NA;
            float *o = obs_tile + ((size_t)tid * A + a) * NO;
            int gfsm = 0;
            if (V == IGE_FRANKA_COMBINE || V == IGE_FRANKA_PP) {
                // compute_global_FSM franka_PPMA.py:597-611 (the flatten() of a (2, N) stack, see header)
                const bool grip = ige::norm3(eef[1][0] - eef[0][0], eef[1][1] - eef[0][1], eef[1][2] - eef[0][2]) <= 0.18f;
                const bool first_half = row < (size_t)N;
                const int j = first_half ? (int)row : (int)(row - N);
                const float *je = body_row(p, b, j, p.eef_body_start, p.eef_body_step, first_half ? 0 : 1);
                const float *jc = actor_row(p, b, j, p.cube_actor_start + p.cube_actor_step * (first_half ? 1 : 0));
                const bool cub = ige::norm3(__ldg(je) - jc[0], __ldg(je + 1) - jc[1], __ldg(je + 2) - jc[2]) <= 0.18f;
                gfsm = (grip || cub) ? -1 : 0;
            }
            int n = 0;
            if (V == IGE_FRANKA_REACH || V == IGE_FRANKA_COLLECT) {
                for (int c = 0; c < T; ++c) { o[n++] = cube[c][0]; o[n++] = cube[c][1]; o[n++] = cube[c][2]; }
#pragma unroll
                for (int i = 0; i < 4; ++i) o[n++] = eef[a][3 + i];
#pragma unroll
                for (int i = 0; i < 3; ++i) o[n++] = eef[a][i];
#pragma unroll
                for (int i = 0; i < 3; ++i) o[n++] = ncube[a][i] - eef[a][i];
                if (V == IGE_FRANKA_COLLECT) {
                    const float *bs = body_row(p, b, e, p.base_body_start, p.base_body_step, a);
#pragma unroll
                    for (int i = 0; i < 7; ++i) o[n++] = __ldg(bs + i);
                }
                for (int s = 1; s < A; ++s) {
                    const int q = (a + s) % A;
                    o[n++] = eef[q][0]; o[n++] = eef[q][1]; o[n++] = eef[q][2];
                }
                if (V == IGE_FRANKA_COLLECT) { o[n++] = (float)fsm[a]; o[n++] = (float)fsm[a]; }
            } else {
                for (int s = 0; s < T; ++s) { const int q = (a + s) % T;
#pragma unroll
                    for (int i = 0; i < 4; ++i) o[n++] = cube[q][3 + i]; }
                for (int s = 0; s < T; ++s) { const int q = (a + s) % T; o[n++] = cube[q][0]; o[n++] = cube[q][1]; o[n++] = cube[q][2]; }
                const float *aux;
                if (V == IGE_FRANKA_COMBINE) {
                    aux = actor_row(p, b, e, p.stack_base_actor);
                    o[n++] = aux[0]; o[n++] = aux[1]; o[n++] = aux[2];
                } else {
                    for (int s = 0; s < T; ++s) {
                        const float *d = actor_row(p, b, e, p.dest_actor_start + p.dest_actor_step * ((a + s) % T));
                        o[n++] = d[0]; o[n++] = d[1]; o[n++] = d[2];
                    }
                    aux = actor_row(p, b, e, p.dest_actor_start + p.dest_actor_step * a);
                }
#pragma unroll
                for (int i = 0; i < 3; ++i) o[n++] = cube[a][i] - eef[a][i];
#pragma unroll
                for (int i = 0; i < 3; ++i) o[n++] = aux[i] - cube[a][i];
                const float *bs = body_row(p, b, e, p.base_body_start, p.base_body_step, a);
#pragma unroll
                for (int i = 0; i < 7; ++i) o[n++] = __ldg(bs + i);
                for (int s = 0; s < A; ++s) { const int q = (a + s) % A; o[n++] = eef[q][0]; o[n++] = eef[q][1]; o[n++] = eef[q][2]; }
                for (int s = 0; s < A; ++s) { const int q = (a + s) % A;
#pragma unroll
                    for (int i = 0; i < 4; ++i) o[n++] = eef[q][3 + i]; }
                for (int s = 0; s < A; ++s) o[n++] = pow2_small(fsm[(a + s) % A]);
                o[n++] = pow2_small(gfsm);
                if (V == IGE_FRANKA_COMBINE) o[n++] = (a == A - 1) ? 1.0f : 0.0f;  // agent_idx :526-527
            }
            // ---- compute_franka_reward ----
            float rewards;
            float c10[NC];
#pragma unroll
            for (int i = 0; i < NC; ++i) c10[i] = 0.0f;
            const float alast = __ldg(act + NA - 1);
            if (V == IGE_FRANKA_REACH) {  // franka_reach_MA.py:928-960
                const float d = md[a];
                const float dist_reward = 1.0f / (0.5f + d * d);
                float av[8];
                for (int j = 0; j < NA; ++j) av[j] = __ldg(act + j);
                const float actions_cost = sum_sq_small(av, NA) * 0.01f;
                rewards = dist_reward - actions_cost;
                const int distinct = __popc(touched);  // mask.sum(dim=-1), franka_reach_MA.py:944-948
                const float all_touched = ((float)distinct / (float)A) * 1.0f;
                const float *cf = b.contact_forces + ((size_t)e * p.num_bodies + p.hand_body_start + p.hand_body_step * a) * 3;
                const float colliding = (ige::norm3(__ldg(cf), __ldg(cf + 1), __ldg(cf + 2)) >= 0.1f ? 1.0f : 0.0f) * -10.0f;
                rewards = rewards + (all_touched + colliding);
            } else {
                const float m = md[a];
                const int s = fsm[a];
                c10[0] = s == 0 ? expf(-5.0f * (m * m)) : 0.0f;
                c10[1] = s == 1 ? expf(-1.0f * alast) : 0.0f;
                if (V == IGE_FRANKA_COLLECT) {  // franka_collect_MA.py:1082-1174
                    const float *hc = ncube[a];
                    c10[2] = s == 2 ? (hc[2] - 1.05f) / (float)(0.3 + 0.025) : 0.0f;
                    const float dy = fabsf(hc[1] - (float)(0.3 + 0.05 * 2.));
                    c10[3] = s == 3 ? expf(-5.0f * (dy * dy)) : 0.0f;
                    const float dz = fabsf(hc[2] - (float)(0.3 / 2. + 1.05));
                    c10[4] = s == 4 ? expf(-5.0f * (dz * dz)) : 0.0f;
                    c10[5] = s == 5 ? expf(4.0f * alast) : 0.0f;
                    c10[6] = s == 6 ? 1.0f * 3.0f : 0.0f;
                    c10[7] = (float)s;
                    rewards = c10[0] + c10[1] + c10[2] + c10[3] + c10[4] + c10[5] + c10[6] + c10[7] + c10[8] + 0.0f;
                } else if (V == IGE_FRANKA_COMBINE) {  // franka_combine_MA.py:1080-1166
                    const float *sb = actor_row(p, b, e, p.stack_base_actor);
                    float lift = fabsf(sb[2] - cube[a][2]) / (float)((0.05 + 0.05) * 0.5);
                    lift = lift < 0.0f ? 0.0f : (lift > 1.0f ? 1.0f : lift);
                    c10[2] = s == 2 ? lift : 0.0f;
                    c10[7] = (float)s;
                    rewards = c10[0] + c10[1] + c10[7];
                } else {  // franka_PPMA.py:1082-1167
                    const float *dest = actor_row(p, b, e, p.dest_actor_start + p.dest_actor_step * a);
                    const float rx = dest[0] - cube[a][0], ry = dest[1] - cube[a][1], rz = dest[2] - cube[a][2];
                    float lift = fabsf(rz) / (float)((0.05 + 0.05) * 0.5);
                    lift = lift < 0.0f ? 0.0f : (lift > 1.0f ? 1.0f : lift);
                    c10[2] = s == 2 ? lift : 0.0f;
                    const float dt = ige::norm3(rx + 0.0f, ry + 0.0f, rz + 0.05f);
                    c10[3] = s == 3 ? expf(-3.0f * dt) : 0.0f;
                    c10[4] = s == 4 ? expf(-10.0f * dt) : 0.0f;
                    c10[5] = s == 5 ? expf(5.0f * alast) : 0.0f;
                    c10[6] = s == 6 ? tanhf(m) : 0.0f;
                    c10[7] = (float)s;
                    c10[8] = (float)gfsm * 0.5f;
                    rewards = c10[0] + c10[1] + c10[2] + c10[3] + c10[4] + c10[5] + c10[6] + c10[7] + c10[8] + c10[9];
                }
            }
            rewards = rewards < 0.0f ? 0.0f : rewards;  // torch.clip(rewards, 0., None)
#pragma unroll
            for (int i = 0; i < NC; ++i) comp[i] += c10[i];
            long long reset = b.reset_buf[row];
            const long long progress = b.progress_buf[row];
            const bool late = (float)progress >= p.max_episode_length - 1.0f;
            if (late) reset = 1;
            b.rew_buf[row] = rewards;
            b.reset_buf[row] = reset;
            const bool timeout = late && reset != 0;  // vec_task.py:396
            if (b.timeout_buf) b.timeout_buf[row] = timeout ? 1 : 0;
            st_rew += rewards; st_done += reset != 0 ? 1.f : 0.f; st_to += timeout ? 1.f : 0.f;
            st_len += reset != 0 ? (float)progress : 0.f;
        }
    }

    franka_tile_out_and_reduce(p, b, obs_tile, &s_last, comp, st_rew, st_done, st_to, st_len, valid, phase, use_tma, A, NO);
}


// ---------------------------------------------------------------------------------------------
// Rows kernel: the shipped configuration of all four tasks (numAgents = numTargets = 2,
// cfg/task/Franka*MA.yaml), ONE launch for every variant.
// ---------------------------------------------------------------------------------------------
// One thread per AGENT ROW (row = env * 2 + agent; the two lanes of an env sit next to each other and trade the
// partner's gripper / cube / FSM with shuffles), 32-env tiles, persistent CTAs over contiguous tile ranges:
// BASELINE.json's 8192 x 2 configuration is 512 warps on 256 CTAs instead of 256 warps on 128 (round 1: 3 % of
// the warp slots active), the per-thread dependent chain is half as long, and every load whose address depends
// only on the row index is issued before anything is consumed.
//
// Resets happen in registers: the two lanes of a flagged env draw the env's uniforms (<= 6 Philox calls per
// lane; injected mode: loads by reset rank), keep the new cube / destination rows for the arithmetic below and
// write them out with fire-and-forget stores -- no thread ever re-reads a row its own launch rewrote.
//
// compute_global_FSM pairs flat row k with d_ecp[k] of a flatten()ed (2, N) stack (franka_PPMA.py:606), i.e.
// row k needs the gripper and one cube of ANOTHER env j = k or k - N, whose cube row env j's own reset may be
// rewriting in this very launch.  Round 1 ran the reset as a separate launch for that reason.  Here the reader
// never uses a row that is being rewritten: post-reset cube positions are a pure function of (env, draw,
// step | reset rank), so it recomputes env j's when env j resets and uses memory when it does not.  "Does env j
// reset" is env j's ENTRY flag, which every CTA snapshots for all of its rows in its prologue (a ballot per tile in
// shared memory) BEFORE it publishes its reset-list slots ({launch tag : entry-flag ballot}, ige::list_publish_slot).
// The ascending reset id list is written by one extra scanner CTA from those slots (ige::list_scanner_cta); once it
// has seen every CTA's slots it stores the launch tag, and only then -- one tile of work later, so the poll never
// actually spins -- do the workers overwrite reset_buf.  No worker waits for a prefix or a ticket.  (Polling the
// partner's SLOT instead of snapshotting its flag was measured: 16 k threads spinning on a few L2 lines cost
// +1.4 us.)  The loads of a CTA's first tile are issued BEFORE the prologue: the flag round trip and the tile's
// DRAM round trip overlap instead of following each other.
constexpr int RT_E = 32;                // envs per tile
constexpr int RT_NT = 2 * RT_E;         // threads per CTA = rows per tile
constexpr int RT_MAX_TILES = 96;        // tiles per CTA the partner-flag snapshot of Combine / PP has room for

// four consecutive uniforms j = 4 blk .. 4 blk + 3 of draw `draw` for env `env` (reset rank `rank`)
__device__ __forceinline__ void ma_uniform4(const IgeMAResetRand &r, uint64_t step, int draw, int rank, int env, int cols,
                                            int blk, float (&u)[4]) {
    if (r.mode == IGE_RAND_PHILOX) {
        uint32_t c[4] = {(uint32_t)env, ((uint32_t)draw << 16) | (uint32_t)blk, (uint32_t)step, (uint32_t)(step >> 32)};
        ige::philox4x32_10(c, (uint32_t)r.seed, (uint32_t)(r.seed >> 32));
#pragma unroll
        for (int k = 0; k < 4; ++k) u[k] = (float)(c[k] >> 8) * (1.0f / 16777216.0f);
    } else {
        const float *src = r.u[draw] + (size_t)rank * cols + blk * 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) u[k] = blk * 4 + k < cols ? __ldg(src + k) : 0.0f;
    }
}

template <int V>
__global__ void __launch_bounds__(RT_NT) franka_ma_rows_kernel(const __grid_constant__ IgeFrankaMAParams p_in,
                                                               const __grid_constant__ IgeFrankaMABuffers b_in,
                                                               const int num_tiles, const int tiles_per_cta,
                                                               const int use_tma, const int workers) {
#if IGE_STAGE_ARGS
    // (experiment) kernel arguments -> shared memory once
    __shared__ __align__(16) IgeFrankaMAParams p;
    __shared__ __align__(16) IgeFrankaMABuffers b;
    {
        static_assert(sizeof(IgeFrankaMAParams) % 8 == 0 && sizeof(IgeFrankaMABuffers) % 8 == 0, "argument structs copy as 8-B words");
        constexpr int WP = sizeof(IgeFrankaMAParams) / 8, WB = sizeof(IgeFrankaMABuffers) / 8;
        const uint64_t *sp = reinterpret_cast<const uint64_t *>(&p_in), *sb = reinterpret_cast<const uint64_t *>(&b_in);
        for (int w = threadIdx.x; w < WP + WB; w += blockDim.x) {
            if (w < WB) reinterpret_cast<uint64_t *>(&b)[w] = sb[w];
            else reinterpret_cast<uint64_t *>(&p)[w - WB] = sp[w - WB];
        }
    }
    __syncthreads();
#else
    const IgeFrankaMAParams &p = p_in;
    const IgeFrankaMABuffers &b = b_in;
#endif
    constexpr int A = 2, T = 2;
    constexpr int NA = V == IGE_FRANKA_REACH ? 6 : 7;
    constexpr int NO = V == IGE_FRANKA_REACH ? 19 : V == IGE_FRANKA_COLLECT ? 28 : V == IGE_FRANKA_COMBINE ? 48 : 50;
    constexpr bool PAIRED = V == IGE_FRANKA_COMBINE || V == IGE_FRANKA_PP;
    constexpr bool MEANS = V != IGE_FRANKA_REACH;   // Reach logs no component means (franka_reach_MA.py:928-960)
    __shared__ __align__(128) float obs_tile[2][RT_NT * NO];
    __shared__ uint64_t s_step;
    __shared__ uint32_t s_tag;
    __shared__ uint32_t s_pmask[PAIRED ? RT_MAX_TILES : 1][2];   // bit = "the partner env of this row resets in this launch"
    __shared__ int s_last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int N = p.num_envs;
    const int first = blockIdx.x * tiles_per_cta;
    const int my_tiles = min(tiles_per_cta, num_tiles - first);
    const bool emit = b.list.scratch != nullptr;
    const int a = tid & 1;                     // my agent
    auto partner_env = [&](int row, bool &first_half) {
        first_half = row < N;
        return first_half ? row : row - N;
    };

    ige::pdl_wait();   // global memory from here on
    if ((int)blockIdx.x >= workers) {   // the scanner CTA: slots -> ascending id list, count, launch tag, Philox step
        ige::pdl_trigger();
        if (num_tiles <= 2 * 32 * ige::kListUnitSlots) ige::list_scanner_cta<RT_NT>(b_in.list, num_tiles, p_in.rand.step);
        else ige::list_scanner_cta<RT_NT, 32>(b_in.list, num_tiles, p_in.rand.step);   // large env counts: 1024-tile units
        return;
    }
    if (tid == 0) {
        s_step = (p.rand.mode == IGE_RAND_PHILOX && p.rand.step) ? *reinterpret_cast<const volatile uint64_t *>(p.rand.step) : 0ull;
        s_tag = emit ? ige::list_epoch_tag(b.list) : 0u;
    }

    // registers of one tile's row-indexed loads
    long long prog = 0, rst = 0;
    float eef[7], base[7], act[NA], cf[3], cube[7], dest[3], sb[3], je[3], jc[3];
    bool j_first = false;
    int j = 0;
    // ---- every load whose address depends only on the row (issued for tile 0 before the prologue below) ----
    auto load_tile = [&](int t) {
        const int e0 = (first + t) * RT_E;
        const int ne = min(RT_E, N - e0);
        const int ec = (tid >> 1) < ne ? e0 + (tid >> 1) : N - 1;   // clamped: invalid lanes load a real row and discard it
        const size_t row = (size_t)ec * A + a;
        prog = b.progress_buf[row];
        rst = b.reset_buf[row];
        {
            const float *r = body_row(p, b, ec, p.eef_body_start, p.eef_body_step, a);
#pragma unroll
            for (int i = 0; i < 7; ++i) eef[i] = __ldg(r + i);
        }
        if (V != IGE_FRANKA_REACH) {
            const float *r = body_row(p, b, ec, p.base_body_start, p.base_body_step, a);
#pragma unroll
            for (int i = 0; i < 7; ++i) base[i] = __ldg(r + i);
        }
        {
            const float *r = b.actions + row * NA;
#pragma unroll
            for (int jj = 0; jj < NA; ++jj) act[jj] = __ldg(r + jj);
        }
        if (V == IGE_FRANKA_REACH) {
            const float *r = b.contact_forces + ((size_t)ec * p.num_bodies + p.hand_body_start + p.hand_body_step * a) * 3;
#pragma unroll
            for (int i = 0; i < 3; ++i) cf[i] = __ldg(r + i);
        }
        {
            const float *r = actor_row(p, b, ec, p.cube_actor_start + p.cube_actor_step * a);   // cube `a` (pre-reset)
#pragma unroll
            for (int i = 0; i < 7; ++i) cube[i] = r[i];
        }
        if (V == IGE_FRANKA_PP) {
            const float *r = actor_row(p, b, ec, p.dest_actor_start + p.dest_actor_step * a);
#pragma unroll
            for (int i = 0; i < 3; ++i) dest[i] = r[i];
        }
        if (V == IGE_FRANKA_COMBINE) {
            const float *r = actor_row(p, b, ec, p.stack_base_actor);
#pragma unroll
            for (int i = 0; i < 3; ++i) sb[i] = r[i];
        }
        if (PAIRED) {
            j = partner_env((int)row, j_first);
            const float *r = body_row(p, b, j, p.eef_body_start, p.eef_body_step, j_first ? 0 : 1);
#pragma unroll
            for (int i = 0; i < 3; ++i) je[i] = __ldg(r + i);
            // env j's cube row: used only when env j does NOT reset in this launch (then nobody writes the row; a value
            // read while env j's owner rewrites it is discarded below)
            const volatile float *c = actor_row(p, b, j, p.cube_actor_start + p.cube_actor_step * (j_first ? 1 : 0));
#pragma unroll
            for (int i = 0; i < 3; ++i) jc[i] = c[i];
        }
    };
    if (my_tiles > 0) load_tile(0);

    // ---- prologue.  Combine / PP: snapshot the ENTRY flags of the partner envs of every row of every tile (warp w
    //      takes rows [32 w, 32 w + 32) of each tile).  Only then are this CTA's slots published, and reset_buf is
    //      not overwritten by anybody before the scanner has seen EVERY CTA's slots (the gate below) ----
    if (PAIRED) {
        for (int t = 0; t < my_tiles; ++t) {
            const int row = ((first + t) * RT_E) * A + tid;
            bool f = false;
            if (row < N * A) {
                bool fh;
                const int pj = partner_env(row, fh);
                const longlong2 r = *reinterpret_cast<const longlong2 *>(b.reset_buf + (size_t)pj * A);
                f = r.x != 0 && r.y != 0;
            }
            const uint32_t m = __ballot_sync(0xffffffffu, f);
            if (lane == 0) s_pmask[t][warp] = m;
        }
    }
    __syncthreads();   // step word, launch tag and every partner flag were read before this CTA's first publish
    const uint64_t step = s_step;
    const uint32_t ltag = s_tag;
    if (emit && warp == 0) {
        for (int t = 0; t < my_tiles; ++t) {
            const int e = (first + t) * RT_E + lane;
            bool f = false;
            if (e < N) {
                const longlong2 r = *reinterpret_cast<const longlong2 *>(b.reset_buf + (size_t)e * A);
                f = r.x != 0 && r.y != 0;
            }
            const uint32_t m = __ballot_sync(0xffffffffu, f);
            if (lane == 0) ige::list_publish_slot(b.list, ltag, first + t, m);
        }
    }

    float comp[NC];
#pragma unroll
    for (int i = 0; i < NC; ++i) comp[i] = 0.0f;
    float st_rew = 0.f, st_done = 0.f, st_to = 0.f, st_len = 0.f, st_n = 0.f;
    bool flags_open = !PAIRED;   // may reset_buf be overwritten yet?
    for (int t = 0; t < my_tiles; ++t) {
        const int e0 = (first + t) * RT_E;
        const int ne = min(RT_E, N - e0);
        const int e = e0 + (tid >> 1);
        const bool valid = (tid >> 1) < ne;
        const int ec = valid ? e : N - 1;
        const size_t row = (size_t)ec * A + a;
        float *otile = obs_tile[t & 1];
        const bool j_reset = PAIRED && ((s_pmask[PAIRED ? t : 0][warp] >> lane) & 1u);   // does the partner env reset in this launch?

        // ---- post_physics_step :821-826: progress++, agent flags -> reset_idx (AND filter) ----
        const long long rst_o = __shfl_xor_sync(0xffffffffu, rst, 1);
        const bool did_reset = valid && rst != 0 && rst_o != 0;
        if (did_reset) {
            const IgeMAResetRand &rr = p.rand;
            const int rk = rr.env_rank ? __ldg(rr.env_rank + e) : 0;
            float u0[4], u1[4], u3[4], u4[3][4];
            ma_uniform4(rr, step, 0, rk, e, T, 0, u0);
            ma_uniform4(rr, step, 1, rk, e, 2 * T, 0, u1);
            if (PAIRED) ma_uniform4(rr, step, 3, rk, e, V == IGE_FRANKA_PP ? 2 * T : 2, 0, u3);
#pragma unroll
            for (int k = 0; k < 3; ++k) ma_uniform4(rr, step, 4, rk, e, 9, k, u4[k]);
            // _reset_init_cube_state: cube `a` (franka_reach_MA.py:676-704)
            cube[0] = p.cube_xy_center[0] + p.cube_xy_mult * ((a ? u1[2] : u1[0]) - 0.5f);
            cube[1] = p.cube_xy_center[1] + p.cube_xy_mult * ((a ? u1[3] : u1[1]) - 0.5f);
            cube[2] = p.cube_z_base + (a ? u0[1] : u0[0]) * p.cube_z_range;
            cube[3] = cube[4] = cube[5] = 0.0f; cube[6] = 1.0f;
            float *ci = b.init_cubeA_state + ((size_t)e * T + a) * 13;
            float *cr = actor_row(p, b, e, p.cube_actor_start + p.cube_actor_step * a);
#pragma unroll
            for (int i = 0; i < 13; ++i) { const float v = i < 7 ? cube[i] : 0.0f; ci[i] = v; cr[i] = v; }
            if (V == IGE_FRANKA_PP) {  // _reset_dest_cube_state franka_PPMA.py:862-895
                dest[0] = (a ? p.aux_xy_center_x[1] : p.aux_xy_center_x[0]) + p.aux_xy_mult * ((a ? u3[2] : u3[0]) - 0.5f);
                dest[1] = p.aux_xy_center_y + p.aux_xy_mult * ((a ? u3[3] : u3[1]) - 0.5f);
                dest[2] = p.aux_z_base + 0.0f;
                float *dr = actor_row(p, b, e, p.dest_actor_start + p.dest_actor_step * a);
#pragma unroll
                for (int i = 0; i < 13; ++i) dr[i] = i < 3 ? dest[i] : (i == 6 ? 1.0f : 0.0f);
            }
            if (V == IGE_FRANKA_COMBINE) {  // _reset_stack_base_cube_state franka_combine_MA.py:885-908
                sb[0] = p.aux_xy_center_x[0] + p.aux_xy_mult * (u3[0] - 0.5f);
                sb[1] = p.aux_xy_center_y + p.aux_xy_mult * (u3[1] - 0.5f);
                sb[2] = p.aux_z_base + 0.0f;
                if (a == 0) {
                    float *dr = actor_row(p, b, e, p.stack_base_actor);
#pragma unroll
                    for (int i = 0; i < 13; ++i) dr[i] = i < 3 ? sb[i] : (i == 6 ? 1.0f : 0.0f);
                }
            }
            // arm `a`: the SAME joint sample is tiled over the arms (franka_reach_MA.py:638-656)
#pragma unroll
            for (int jj = 0; jj < 9; ++jj) {
                float v = p.franka_default_dof_pos[jj] + p.dof_noise_mult * (u4[jj >> 2][jj & 3] - 0.5f);
                v = fminf(v, p.franka_dof_upper[jj]);
                v = fmaxf(v, p.franka_dof_lower[jj]);
                const float pos = jj >= 7 ? p.franka_default_dof_pos[jj] : v;
                const size_t k = (size_t)e * p.num_dofs + a * 9 + jj;
                reinterpret_cast<float2 *>(b.dof_state)[k] = make_float2(pos, 0.0f);
                b.pos_control[k] = pos;
                b.effort_control[k] = 0.0f;
            }
            prog = 0; rst = 0;
        } else {
            prog += 1;
        }
        if (PAIRED && j_reset) {   // env j's cube (j_first ? 1 : 0) as env j's own reset writes it
            const IgeMAResetRand &rr = p.rand;
            const int rk = rr.env_rank ? __ldg(rr.env_rank + j) : 0;
            const int c = j_first ? 1 : 0;
            float u0[4], u1[4];
            ma_uniform4(rr, step, 0, rk, j, T, 0, u0);
            ma_uniform4(rr, step, 1, rk, j, 2 * T, 0, u1);
            jc[0] = p.cube_xy_center[0] + p.cube_xy_mult * ((c ? u1[2] : u1[0]) - 0.5f);
            jc[1] = p.cube_xy_center[1] + p.cube_xy_mult * ((c ? u1[3] : u1[1]) - 0.5f);
            jc[2] = p.cube_z_base + (c ? u0[1] : u0[0]) * p.cube_z_range;
        }

        // ---- the partner lane's rows ----
        float eef_o[7], cube_o[7], dest_o[3];
#pragma unroll
        for (int i = 0; i < 7; ++i) {
            eef_o[i] = __shfl_xor_sync(0xffffffffu, eef[i], 1);
            cube_o[i] = __shfl_xor_sync(0xffffffffu, cube[i], 1);
        }
        if (V == IGE_FRANKA_PP) {
#pragma unroll
            for (int i = 0; i < 3; ++i) dest_o[i] = __shfl_xor_sync(0xffffffffu, dest[i], 1);
        }
        // cube[0] / cube[1] in the reference's indexing
        const float *c0 = a == 0 ? cube : cube_o, *c1 = a == 0 ? cube_o : cube;

        // ---- _update_states / compute_FSM ----
        float ncube[3];
        int near;
        {   // _get_min_relative_pos (franka_reach_MA.py:521-531): first minimum wins
            const float d0 = ige::norm3(c0[0] - eef[0], c0[1] - eef[1], c0[2] - eef[2]);
            const float d1 = ige::norm3(c1[0] - eef[0], c1[1] - eef[1], c1[2] - eef[2]);
            near = d1 < d0 ? 1 : 0;
            const float *nc = near ? c1 : c0;
            ncube[0] = nc[0]; ncube[1] = nc[1]; ncube[2] = nc[2];
        }
        const int near_o = __shfl_xor_sync(0xffffffffu, near, 1);
        const unsigned touched = (1u << near) | (1u << near_o);
        int fsm = 0;
        float md;
        const float on_th = (float)((double)p.cubeA_size * 0.5 * 0.9);
        const bool closed = act[NA - 1] < 0.0f;
        if (V == IGE_FRANKA_REACH || V == IGE_FRANKA_COLLECT) {
            md = ige::norm3(ncube[0] - eef[0], ncube[1] - eef[1], ncube[2] - eef[2]);
            if (V == IGE_FRANKA_COLLECT) {  // franka_collect_MA.py:549-606
                const bool close = md <= on_th;
                const bool holding = close && closed;
                const bool high = (ncube[2] - 1.05f) > (float)((double)p.wall_height + (double)p.cubeA_size / 2);
                const bool in_area = (ncube[1] > (float)((double)p.wall_pos_y + (double)p.cubeA_size * 1.)) && (fabsf(ncube[0]) < 0.6f);
                const bool low = (ncube[2] - 1.05f) < (float)((double)p.wall_height / 2.);
                if (close) fsm = 1;
                if (holding) fsm = 2;
                if (holding && high) fsm = 3;
                if (holding && in_area) fsm = 4;
                if (holding && in_area && low) fsm = 5;
                if (holding && in_area && low && !closed) fsm = 6;
            }
        } else {  // Combine :565-611 / PP :561-595: own cube
            md = ige::norm3(cube[0] - eef[0], cube[1] - eef[1], cube[2] - eef[2]);
            const bool on = md <= on_th;
            if (on) fsm = 1;
            if (on && closed) fsm = 2;
            if (V == IGE_FRANKA_PP) {
                const float dz = dest[2] - cube[2];
                const bool lifted = fabsf(dz) >= (float)(((double)p.dest_size + (double)p.cubeA_size) / 2.);
                const bool align = ige::norm2(dest[0] - cube[0], dest[1] - cube[1]) < (float)((0.05) * 0.5);
                const bool stackable = fabsf(dz) <= (float)((double)p.cubeA_size * 0.8660 + (double)p.dest_size / 2.);
                if (on && lifted) fsm = 3;
                if (align) fsm = 4;
                if (align && stackable) fsm = 5;
                if (align && stackable && !closed) fsm = 6;
            }
        }
        const int fsm_o = __shfl_xor_sync(0xffffffffu, fsm, 1);
        int gfsm = 0;
        if (PAIRED) {  // compute_global_FSM franka_PPMA.py:597-611: eef[1] - eef[0] of my env, d_ecp of env j
            const float *e0p = a == 0 ? eef : eef_o, *e1p = a == 0 ? eef_o : eef;
            const bool grip = ige::norm3(e1p[0] - e0p[0], e1p[1] - e0p[1], e1p[2] - e0p[2]) <= 0.18f;
            const bool cub = ige::norm3(je[0] - jc[0], je[1] - jc[1], je[2] - jc[2]) <= 0.18f;
            gfsm = (grip || cub) ? -1 : 0;
        }

        // ---- compute_observations: my row of the tile ----
        float *o = otile + tid * NO;
        int n = 0;
        if (V == IGE_FRANKA_REACH || V == IGE_FRANKA_COLLECT) {
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = c0[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = c1[i];
#pragma unroll
            for (int i = 0; i < 4; ++i) o[n++] = eef[3 + i];
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = eef[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = ncube[i] - eef[i];
            if (V == IGE_FRANKA_COLLECT) {
#pragma unroll
                for (int i = 0; i < 7; ++i) o[n++] = base[i];
            }
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = eef_o[i];
            if (V == IGE_FRANKA_COLLECT) { o[n++] = (float)fsm; o[n++] = (float)fsm; }
        } else {
#pragma unroll
            for (int i = 0; i < 4; ++i) o[n++] = cube[3 + i];
#pragma unroll
            for (int i = 0; i < 4; ++i) o[n++] = cube_o[3 + i];
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = cube[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = cube_o[i];
            float aux[3];
            if (V == IGE_FRANKA_COMBINE) {
#pragma unroll
                for (int i = 0; i < 3; ++i) { aux[i] = sb[i]; o[n++] = sb[i]; }
            } else {
#pragma unroll
                for (int i = 0; i < 3; ++i) o[n++] = dest[i];
#pragma unroll
                for (int i = 0; i < 3; ++i) o[n++] = dest_o[i];
#pragma unroll
                for (int i = 0; i < 3; ++i) aux[i] = dest[i];
            }
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = cube[i] - eef[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = aux[i] - cube[i];
#pragma unroll
            for (int i = 0; i < 7; ++i) o[n++] = base[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = eef[i];
#pragma unroll
            for (int i = 0; i < 3; ++i) o[n++] = eef_o[i];
#pragma unroll
            for (int i = 0; i < 4; ++i) o[n++] = eef[3 + i];
#pragma unroll
            for (int i = 0; i < 4; ++i) o[n++] = eef_o[3 + i];
            o[n++] = pow2_small(fsm);
            o[n++] = pow2_small(fsm_o);
            o[n++] = pow2_small(gfsm);
            if (V == IGE_FRANKA_COMBINE) o[n++] = (a == A - 1) ? 1.0f : 0.0f;  // agent_idx :526-527
        }

        // ---- compute_franka_reward ----
        float rewards;
        float c10[NC];
#pragma unroll
        for (int i = 0; i < NC; ++i) c10[i] = 0.0f;
        const float alast = act[NA - 1];
        if (V == IGE_FRANKA_REACH) {  // franka_reach_MA.py:928-960
            const float dist_reward = 1.0f / (0.5f + md * md);
            const float actions_cost = sum_sq_small(act, NA) * 0.01f;
            rewards = dist_reward - actions_cost;
            const float all_touched = ((float)__popc(touched) / (float)A) * 1.0f;   // mask.sum(dim=-1) :944-948
            const float colliding = (ige::norm3(cf[0], cf[1], cf[2]) >= 0.1f ? 1.0f : 0.0f) * -10.0f;
            rewards = rewards + (all_touched + colliding);
        } else {
            c10[0] = fsm == 0 ? expf(-5.0f * (md * md)) : 0.0f;
            c10[1] = fsm == 1 ? expf(-1.0f * alast) : 0.0f;
            if (V == IGE_FRANKA_COLLECT) {  // franka_collect_MA.py:1082-1174
                c10[2] = fsm == 2 ? (ncube[2] - 1.05f) / (float)(0.3 + 0.025) : 0.0f;
                const float dy = fabsf(ncube[1] - (float)(0.3 + 0.05 * 2.));
                c10[3] = fsm == 3 ? expf(-5.0f * (dy * dy)) : 0.0f;
                const float dz = fabsf(ncube[2] - (float)(0.3 / 2. + 1.05));
                c10[4] = fsm == 4 ? expf(-5.0f * (dz * dz)) : 0.0f;
                c10[5] = fsm == 5 ? expf(4.0f * alast) : 0.0f;
                c10[6] = fsm == 6 ? 1.0f * 3.0f : 0.0f;
                c10[7] = (float)fsm;
                rewards = c10[0] + c10[1] + c10[2] + c10[3] + c10[4] + c10[5] + c10[6] + c10[7] + c10[8] + 0.0f;
            } else if (V == IGE_FRANKA_COMBINE) {  // franka_combine_MA.py:1080-1166
                float lift = fabsf(sb[2] - cube[2]) / (float)((0.05 + 0.05) * 0.5);
                lift = lift < 0.0f ? 0.0f : (lift > 1.0f ? 1.0f : lift);
                c10[2] = fsm == 2 ? lift : 0.0f;
                c10[7] = (float)fsm;
                rewards = c10[0] + c10[1] + c10[7];
            } else {  // franka_PPMA.py:1082-1167
                const float rx = dest[0] - cube[0], ry = dest[1] - cube[1], rz = dest[2] - cube[2];
                float lift = fabsf(rz) / (float)((0.05 + 0.05) * 0.5);
                lift = lift < 0.0f ? 0.0f : (lift > 1.0f ? 1.0f : lift);
                c10[2] = fsm == 2 ? lift : 0.0f;
                const float dt = ige::norm3(rx + 0.0f, ry + 0.0f, rz + 0.05f);
                c10[3] = fsm == 3 ? expf(-3.0f * dt) : 0.0f;
                c10[4] = fsm == 4 ? expf(-10.0f * dt) : 0.0f;
                c10[5] = fsm == 5 ? expf(5.0f * alast) : 0.0f;
                c10[6] = fsm == 6 ? tanhf(md) : 0.0f;
                c10[7] = (float)fsm;
                c10[8] = (float)gfsm * 0.5f;
                rewards = c10[0] + c10[1] + c10[2] + c10[3] + c10[4] + c10[5] + c10[6] + c10[7] + c10[8] + c10[9];
            }
        }
        rewards = rewards < 0.0f ? 0.0f : rewards;  // torch.clip(rewards, 0., None)
        const bool late = (float)prog >= p.max_episode_length - 1.0f;
        if (late) rst = 1;
        const bool timeout = late && rst != 0;  // vec_task.py:396
        if (valid) {
#pragma unroll
            for (int i = 0; i < NC; ++i) comp[i] += c10[i];
            st_n += 1.f; st_rew += rewards; st_done += rst != 0 ? 1.f : 0.f; st_to += timeout ? 1.f : 0.f;
            st_len += rst != 0 ? (float)prog : 0.f;
        }

        // ---- outputs.  The ENTRY flags of other CTAs' partner envs are only overwritten once every CTA has
        //      snapshot them, i.e. once the launch epoch has advanced (see the header); by now it has ----
        const int count = ne * A * NO;
        const bool bulk = use_tma && (count % 4 == 0);
        if (bulk) ige::fence_proxy_async();
#ifndef IGE_EXP_NOGATE   // (timing experiment only: without the gate the partner snapshot of another CTA may race)
        if (!flags_open && tid == 0) {
            // the gate: the scanner CTA stores the launch tag once it has seen the slots of EVERY CTA, i.e. once every
            // CTA is past its partner-flag snapshot.  By now (one tile of work later) it has long done so.
            const uint64_t *sc = reinterpret_cast<const uint64_t *>(b.list.scratch);
            while ((uint32_t)ige::ld_relaxed_gpu_u64(sc) != ltag) __nanosleep(100);
        }
#endif
        if (t >= 1 && tid == 0) ige::bulk_wait_read_all();  // tile t - 1 has left obs_tile[(t + 1) & 1]: free for tile t + 1
        __syncthreads();
        flags_open = true;
        if (valid) {
            b.rew_buf[row] = rewards;
            b.reset_buf[row] = rst;
            b.progress_buf[row] = prog;
            if (b.timeout_buf) b.timeout_buf[row] = timeout ? 1 : 0;
        }
        float *dst = b.obs_buf + (size_t)e0 * A * NO;
        if (bulk) {
            if (tid == 0) { ige::bulk_s2g(dst, otile, (uint32_t)count * 4); ige::bulk_commit(); }
        } else {
            for (int i = tid; i < count; i += RT_NT) dst[i] = otile[i];
        }
        if (b.obs_clamped) {  // vec_task.py:404
            float *dc = b.obs_clamped + (size_t)e0 * A * NO;
            for (int i = tid; i < count; i += RT_NT) dc[i] = clampf(otile[i], p.clip_obs);
        }
        if (t + 1 < my_tiles) load_tile(t + 1);   // (every register of tile t is dead by now)
    }
    ige::pdl_trigger();   // PDL: the next step's grid may be scheduled behind this CTA's tail (ige_common.cuh)

    // ---- reward-component sums (the .mean()s of extras) + statistics ----
#pragma unroll
    for (int i = 0; i < NC; ++i) comp[i] = ige::warp_sum(comp[i]);
    if (lane < NC) {
        float mine = 0.f;
#pragma unroll
        for (int i = 0; i < NC; ++i) if (lane == i) mine = comp[i];
        if (MEANS) atomicAdd(b.reduce_scratch + lane, mine);
        if (b.stats && MEANS) atomicAdd(b.stats + (blockIdx.x % IGE_STATS_ROWS) * IGE_STATS_STRIDE + IGE_STAT_TASK0 + lane, mine);
    }
    if (b.stats) {
        const float v[5] = {st_n, st_rew, st_done, st_to, st_len};
        const int slot[5] = {IGE_STAT_STEPS, IGE_STAT_REWARD, IGE_STAT_DONE, IGE_STAT_TIMEOUT, IGE_STAT_DONE_LENGTH};
        ige::warp_stats_add<5>(b.stats, v, slot);
    }
    if (MEANS) {   // the CTA that retires last turns the component sums into the logged means
        if (lane < NC) __threadfence();
        __syncthreads();
        if (tid == 0) {
            unsigned *c = reinterpret_cast<unsigned *>(b.reduce_scratch + NC);
            s_last = atomicAdd(c, 1u) == (unsigned)workers - 1u;
        }
        __syncthreads();
        if (s_last) {
            __threadfence();
            if (tid < NC) {
                const float sum = __ldcg(b.reduce_scratch + tid);
                if (b.reward_components) b.reward_components[tid] = sum / (float)((long long)N * A);
                b.reduce_scratch[tid] = 0.0f;
            }
            if (tid == 0) b.reduce_scratch[NC] = 0.0f;
        }
    } else if (b.reward_components && blockIdx.x == 0 && tid < NC) {
        b.reward_components[tid] = 0.0f;
    }
    if (tid == 0) ige::bulk_wait_read_all();  // the tiles must outlive the bulk stores' reads
}
}  // namespace

extern "C" int ige_franka_ma_post_physics(const IgeFrankaMAParams *p, const IgeFrankaMABuffers *b, void *stream) {
    if (!p || !b) return ige_fail(IGE_ENULL, "params / buffers is NULL");
    if (p->variant < 0 || p->variant > 3) return ige_fail(IGE_EPARAM, "variant=%d invalid", p->variant);
    if (p->num_envs < 0) return ige_fail(IGE_ESIZE, "num_envs < 0");
    if (p->num_agents < 1 || p->num_agents > MAXA || p->num_targets < 1 || p->num_targets > MAXA)
        return ige_fail(IGE_ESIZE, "num_agents=%d num_targets=%d outside 1..%d", p->num_agents, p->num_targets, MAXA);
    const bool pair_only = p->variant == IGE_FRANKA_COMBINE || p->variant == IGE_FRANKA_PP;
    if (pair_only && (p->num_agents != 2 || p->num_targets != 2))
        return ige_fail(IGE_EPARAM, "Combine / PP need exactly 2 agents and 2 targets");
    if (!pair_only && p->variant != IGE_FRANKA_REACH && p->num_targets < 1) return ige_fail(IGE_EPARAM, "bad targets");
    if (p->num_actions < 1 || p->num_actions > 7) return ige_fail(IGE_ESIZE, "num_actions=%d invalid", p->num_actions);
    if (p->num_dofs != 9 * p->num_agents) return ige_fail(IGE_ESIZE, "num_dofs=%d != 9 * num_agents", p->num_dofs);
    const int last_eef = p->eef_body_start + p->eef_body_step * (p->num_agents - 1);
    const int last_base = p->base_body_start + p->base_body_step * (p->num_agents - 1);
    const int last_hand = p->hand_body_start + p->hand_body_step * (p->num_agents - 1);
    if (last_eef >= p->num_bodies || last_base >= p->num_bodies || last_hand >= p->num_bodies || p->eef_body_start < 0 ||
        p->base_body_start < 0 || p->hand_body_start < 0)
        return ige_fail(IGE_ESIZE, "rigid-body slice outside [0, %d)", p->num_bodies);
    if (p->cube_actor_start + p->cube_actor_step * (p->num_targets - 1) >= p->num_actors)
        return ige_fail(IGE_ESIZE, "cube actor slice outside [0, %d)", p->num_actors);
    IGE_REQUIRE_PTR(b->rigid_body_state);
    IGE_REQUIRE_PTR(b->root_state);
    IGE_REQUIRE_PTR(b->dof_state);
    IGE_REQUIRE_PTR(b->actions);
    IGE_REQUIRE_PTR(b->pos_control);
    IGE_REQUIRE_PTR(b->effort_control);
    IGE_REQUIRE_PTR(b->init_cubeA_state);
    IGE_REQUIRE_PTR(b->obs_buf);
    IGE_REQUIRE_PTR(b->rew_buf);
    IGE_REQUIRE_PTR(b->reset_buf);
    IGE_REQUIRE_PTR(b->progress_buf);
    if (p->variant == IGE_FRANKA_REACH) IGE_REQUIRE_PTR(b->contact_forces);
    if (!b->reduce_scratch) return ige_fail(IGE_ENULL, "reduce_scratch is NULL");
    IGE_OPTIONAL_PTR(b->obs_clamped);
    if (int rc = ige_check_list(b->list)) return rc;
    if (p->rand.mode == IGE_RAND_INJECTED && !p->rand.env_rank) return ige_fail(IGE_ENULL, "rand.env_rank is NULL (INJECTED mode)");
    if (p->rand.mode != IGE_RAND_INJECTED && p->rand.mode != IGE_RAND_PHILOX) return ige_fail(IGE_EPARAM, "rand.mode invalid");
    if (p->num_envs == 0) return IGE_OK;
    cudaStream_t s = (cudaStream_t)stream;
    static const int kPairObs[4] = {19, 28, 48, 50}, kPairActions[4] = {6, 7, 7, 7};
    if (ige_fast_paths() && p->num_agents == 2 && p->num_targets == 2 && p->num_obs == kPairObs[p->variant] &&
        p->num_actions == kPairActions[p->variant] && (!pair_only || b->list.scratch != nullptr)) {
        // rows kernel: ONE launch for every variant (Combine / PP need the reset list's launch epoch, see its header)
        typedef void (*Rows)(const IgeFrankaMAParams, const IgeFrankaMABuffers, const int, const int, const int, const int);
        static const Rows rows[4] = {franka_ma_rows_kernel<IGE_FRANKA_REACH>, franka_ma_rows_kernel<IGE_FRANKA_COLLECT>,
                                     franka_ma_rows_kernel<IGE_FRANKA_COMBINE>, franka_ma_rows_kernel<IGE_FRANKA_PP>};
        int dev = 0;
        cudaGetDevice(&dev);
        static int max_ctas[4][IGE_MAX_DEVICES];   // resident CTAs per variant and device (0 = not queried yet)
        static int ctas_per_sm[4][IGE_MAX_DEVICES];
        if (dev < 0 || dev >= IGE_MAX_DEVICES) return ige_fail(IGE_EPARAM, "device ordinal %d not supported", dev);
        if (max_ctas[p->variant][dev] == 0) {
            int sms = 0, per_sm = 0;
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, rows[p->variant], RT_NT, 0);
            if (e != cudaSuccess || per_sm < 1 || sms < 1)
                return ige_fail(e != cudaSuccess ? (int)e : IGE_EPARAM, "ige_franka_ma_post_physics: occupancy query failed");
            int m = sms * per_sm;
            if (m > ige::kListMaxCtas) m = ige::kListMaxCtas;
            ctas_per_sm[p->variant][dev] = per_sm;
            max_ctas[p->variant][dev] = m;
        }
        const int tiles = (p->num_envs + RT_E - 1) / RT_E;
        const int scanner = b->list.scratch ? 1 : 0;
        // Combine / PP: workers wait for the scanner's gate and the scanner waits for every worker, so ALL CTAs must be
        // resident together.  The occupancy figure is that of an otherwise empty GPU; a kernel of another stream (the
        // statistics exchange polling its peers, an NCCL kernel) may hold a slot.  Keep one SM's worth of CTAs free.
        const int headroom = pair_only ? ctas_per_sm[p->variant][dev] : 0;
        const int avail = max_ctas[p->variant][dev] - headroom - scanner;
        const int cap = avail > 0 ? avail : 1;
        const int per_cta = (tiles + cap - 1) / cap;
        if (!pair_only || per_cta <= RT_MAX_TILES) {
            const int workers = (tiles + per_cta - 1) / per_cta;   // workers + scanner <= resident capacity: all CTAs co-resident
            // Combine / PP are launched stream-serialised: with a dependent grid resident behind them they measured
            // 12.4 -> 14.1 us at 8192 x 2 (profiles/r2_pdl_trigger.txt); Reach / Collect gain from PDL
            ige_launch_opt(!pair_only || getenv("IGE_PDL_PAIRED") != nullptr, rows[p->variant], workers + scanner, RT_NT, 0, s,
                           *p, *b, tiles, per_cta, ige_use_tma(), workers);
            return ige_check_launch("ige_franka_ma_post_physics");
        }
    }
    const size_t smem = (size_t)TE * p->num_agents * p->num_obs * sizeof(float);
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(franka_ma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return ige_fail((int)e, "cudaFuncSetAttribute(%zu): %s", smem, cudaGetErrorString(e));
    }
    const int grid = (p->num_envs + TE - 1) / TE;
    if (pair_only) {
        ige_launch(franka_ma_kernel, grid, NT, 0, s, *p, *b, (int)PH_RESET, ige_use_tma());
        if (int rc = ige_check_launch("ige_franka_ma_post_physics(reset)")) return rc;
        ige_launch(franka_ma_kernel, grid, NT, smem, s, *p, *b, (int)PH_COMPUTE, ige_use_tma());
    } else {
        ige_launch(franka_ma_kernel, grid, NT, smem, s, *p, *b, (int)PH_FUSED, ige_use_tma());
    }
    return ige_check_launch("ige_franka_ma_post_physics");
}
