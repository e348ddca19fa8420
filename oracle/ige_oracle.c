/*
 * ige_oracle.c -- CPU restatement of the reference's post-physics hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may call into this file; the product (isaacgymenvs-ma_b200/) never does.
 *
 * Parity status: PINNED.  The reference holds no golden vectors for this path (SURVEY.md
 * section 4), so the oracle is pinned against outputs of the reference itself: the fixtures in
 * tests/golden/*.npz are produced by importing the unmodified reference task classes in the
 * build container (tests/golden/gen_golden.py) and tests/test_oracle_golden.py checks this file
 * against them (flags / counters / reset ids bit-exact, floats <= 1e-5 relative; in practice
 * bit-exact except where libm and torch's vectorised atan2/asin/exp differ in the last ulp).
 *
 * Every function cites the reference file:line it restates (paths relative to
 * /root/reference/isaacgymenvs/).  Arithmetic is fp32 with the reference's (torch CPU) operation
 * order; compile with -ffp-contract=off so that only the fmaf() calls written here fuse.
 * The structs (include/ige_b200.h) are the product's C ABI, used here with HOST pointers.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/ige_b200.h"

#define TWO_PI_F 6.28318530717958647692f /* float(2*np.pi) */
#define HALF_PI_F 1.57079632679489661923f

/* ------------------------------------------------------------------------------------------
 * torch-CPU primitive semantics (measured against torch 2.11 CPU in the build container)
 * ------------------------------------------------------------------------------------------ */

/* torch.norm(p=2, dim=-1) over 3 elements == sqrt of an fma chain (ATen norm kernel). */
static inline float norm3(float x, float y, float z) {
    float s = x * x;
    s = fmaf(y, y, s);
    s = fmaf(z, z, s);
    return sqrtf(s);
}
static inline float norm2(float x, float y) {
    float s = x * x;
    s = fmaf(y, y, s);
    return sqrtf(s);
}
/* torch.bmm (1x3)(3x1): plain multiply-add chain, no fma. utils/torch_jit_utils.py:88-89 */
static inline float dot3(const float *a, const float *b) {
    return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2];
}
/* torch.sum(dim=-1) of a contiguous row: 8 vector lanes accumulated over full groups of 8,
 * then tail elements, then the lanes, all added sequentially (ATen SumKernel inner reduction). */
static float row_sum(const float *x, int d) {
    if (d < 8) { /* scalar path: 4 interleaved accumulators, remainder into the first, then a chain */
        float p[4] = {0, 0, 0, 0};
        int q = d / 4;
        for (int i = 0; i < q; ++i)
            for (int k = 0; k < 4; ++k) p[k] += x[4 * i + k];
        for (int i = 4 * q; i < d; ++i) p[0] += x[i];
        return ((p[0] + p[1]) + p[2]) + p[3];
    }
    float lanes[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int nv = d / 8;
    if (nv >= 4) { /* 4 interleaved accumulators per lane; not reached for d <= 31 */
        float p[4][8];
        memset(p, 0, sizeof(p));
        int q = nv / 4;
        for (int i = 0; i < q; ++i)
            for (int k = 0; k < 4; ++k)
                for (int l = 0; l < 8; ++l) p[k][l] += x[(4 * i + k) * 8 + l];
        for (int i = 4 * q; i < nv; ++i)
            for (int l = 0; l < 8; ++l) p[0][l] += x[i * 8 + l];
        for (int l = 0; l < 8; ++l) lanes[l] = ((p[0][l] + p[1][l]) + p[2][l]) + p[3][l];
    } else {
        for (int i = 0; i < nv; ++i)
            for (int l = 0; l < 8; ++l) lanes[l] += x[i * 8 + l];
    }
    float acc = 0.0f;
    for (int k = nv * 8; k < d; ++k) acc += x[k];
    if (nv > 0)
        for (int l = 0; l < 8; ++l) acc += lanes[l];
    return acc;
}
/* python-style remainder, torch.remainder(Tensor, Scalar): torch_jit_utils.py:195 */
static inline float py_remainder(float a, float b) {
    float m = fmodf(a, b);
    if (m != 0.0f && ((b < 0.0f) != (m < 0.0f))) m += b;
    return m;
}

/* utils/torch_jit_utils.py:41-62 quat_mul (xyzw), the 9-product form */
static void quat_mul(const float *a, const float *b, float *out) {
    float x1 = a[0], y1 = a[1], z1 = a[2], w1 = a[3];
    float x2 = b[0], y2 = b[1], z2 = b[2], w2 = b[3];
    float ww = (z1 + x1) * (x2 + y2);
    float yy = (w1 - y1) * (w2 + z2);
    float zz = (w1 + y1) * (w2 - z2);
    float xx = ww + yy + zz;
    float qq = 0.5f * (xx + (z1 - x1) * (x2 - y2));
    float w = qq - ww + (z1 - y1) * (y2 - z2);
    float x = qq - xx + (x1 + w1) * (x2 + w2);
    float y = qq - yy + (w1 - x1) * (y2 + z2);
    float z = qq - zz + (z1 + y1) * (w2 - x2);
    out[0] = x; out[1] = y; out[2] = z; out[3] = w;
}
/* utils/torch_jit_utils.py:106-110 */
static void quat_conjugate(const float *a, float *out) {
    out[0] = -a[0]; out[1] = -a[1]; out[2] = -a[2]; out[3] = a[3];
}
/* torch.cross(dim=-1) */
static void cross3(const float *a, const float *b, float *o) {
    o[0] = a[1] * b[2] - a[2] * b[1];
    o[1] = a[2] * b[0] - a[0] * b[2];
    o[2] = a[0] * b[1] - a[1] * b[0];
}
/* utils/torch_jit_utils.py:80-90 (sign=+1) and :93-103 (sign=-1, quat_rotate_inverse) */
static void quat_rotate_impl(const float *q, const float *v, int inverse, float *out) {
    float qw = q[3];
    float s = 2.0f * (qw * qw) - 1.0f;
    float cr[3];
    cross3(q, v, cr);
    float d = dot3(q, v);
    for (int i = 0; i < 3; ++i) {
        float a = v[i] * s;
        float b = cr[i] * qw * 2.0f;
        float c = q[i] * d * 2.0f;
        out[i] = inverse ? (a - b + c) : (a + b + c);
    }
}
/* utils/torch_jit_utils.py:65-67 normalize: x / clamp(norm, min=1e-9) */
static void normalize3(const float *x, float *out) {
    float n = norm3(x[0], x[1], x[2]);
    if (n < 1e-9f) n = 1e-9f;
    out[0] = x[0] / n; out[1] = x[1] / n; out[2] = x[2] / n;
}
/* utils/torch_jit_utils.py:175-195 get_euler_xyz; only roll and yaw are consumed by the tasks */
static void euler_roll_yaw(const float *q, float *roll, float *yaw) {
    float qx = q[0], qy = q[1], qz = q[2], qw = q[3];
    float sinr_cosp = 2.0f * (qw * qx + qy * qz);
    float cosr_cosp = qw * qw - qx * qx - qy * qy + qz * qz;
    float r = atan2f(sinr_cosp, cosr_cosp);
    float siny_cosp = 2.0f * (qw * qz + qx * qy);
    float cosy_cosp = qw * qw + qx * qx - qy * qy - qz * qz;
    float y = atan2f(siny_cosp, cosy_cosp);
    *roll = py_remainder(r, TWO_PI_F);
    *yaw = py_remainder(y, TWO_PI_F);
}
/* utils/torch_jit_utils.py:126-128 */
static inline float normalize_angle(float x) { return atan2f(sinf(x), cosf(x)); }

/* Philox4x32-10 (Salmon et al. 2011), the counter-based generator the production kernels use
 * when IgeResetRand.mode == IGE_RAND_PHILOX.  counter = (row, draw, step_lo, step_hi), key = seed. */
static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
        uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}
/* uniform in [0,1) with 24 bits, element `idx` of draw `draw` for row `row` */
static float philox_uniform(const IgeResetRand *r, uint32_t row, uint32_t draw, uint32_t idx) {
    uint64_t step = r->step ? *r->step : 0;
    uint32_t c[4] = {row, (draw << 16) | (idx >> 2), (uint32_t)step, (uint32_t)(step >> 32)};
    philox4x32_10(c, (uint32_t)r->seed, (uint32_t)(r->seed >> 32));
    return (float)(c[idx & 3] >> 8) * (1.0f / 16777216.0f);
}
static float reset_uniform(const IgeResetRand *r, int which, int row, int d, int j) {
    if (r->mode == IGE_RAND_PHILOX) return philox_uniform(r, (uint32_t)row, (uint32_t)which, (uint32_t)j);
    const float *u = which ? r->u1 : r->u0;
    return u[(size_t)r->reset_rank[row] * d + j];
}

static void add_stat(float *stats, int slot, double v) {
    if (stats) stats[slot] += (float)v;
}

/* ------------------------------------------------------------------------------------------
 * Minimal pthread parallel-for (no OpenMP runtime in this image).  Env ranges are independent.
 * ------------------------------------------------------------------------------------------ */
#include <pthread.h>
#define IGE_ORACLE_MAX_THREADS 256
#define IGE_ORACLE_NACC 8
static int g_threads = 1;
void ige_oracle_set_threads(int n) { g_threads = n < 1 ? 1 : (n > IGE_ORACLE_MAX_THREADS ? IGE_ORACLE_MAX_THREADS : n); }
int ige_oracle_get_threads(void) { return g_threads; }
typedef void (*range_fn)(const void *p, const void *b, int e0, int e1, double *acc);
typedef struct { range_fn fn; const void *p, *b; int e0, e1; double acc[IGE_ORACLE_NACC]; } range_job;
static void *range_thread(void *arg) {
    range_job *j = (range_job *)arg;
    j->fn(j->p, j->b, j->e0, j->e1, j->acc);
    return NULL;
}
static void parallel_ranges(range_fn fn, const void *p, const void *b, int n, double *acc) {
    int t = g_threads;
    if (t > n / 64) t = n / 64;
    if (t < 1) t = 1;
    range_job jobs[IGE_ORACLE_MAX_THREADS];
    pthread_t th[IGE_ORACLE_MAX_THREADS];
    int chunk = (n + t - 1) / t;
    for (int i = 0; i < t; ++i) {
        jobs[i].fn = fn; jobs[i].p = p; jobs[i].b = b;
        jobs[i].e0 = i * chunk; jobs[i].e1 = (i + 1) * chunk < n ? (i + 1) * chunk : n;
        memset(jobs[i].acc, 0, sizeof(jobs[i].acc));
        if (i > 0) pthread_create(&th[i], NULL, range_thread, &jobs[i]);
    }
    range_thread(&jobs[0]);
    for (int i = 1; i < t; ++i) pthread_join(th[i], NULL);
    for (int k = 0; k < IGE_ORACLE_NACC; ++k) {
        acc[k] = 0;
        for (int i = 0; i < t; ++i) acc[k] += jobs[i].acc[k];
    }
}

/* ------------------------------------------------------------------------------------------
 * Observation / action noise lambdas (vec_task.py:373-374, :399-400; closures :675-685 gaussian, :703-711 uniform)
 * + the clamp of step() (:376 / :404).  Restated op by op in float32, the order torch evaluates
 *   op(tensor, corr * s_c + m_c  +  draw * s + m):   t1 = corr*s_c; t1 += m_c; t2 = draw*s; n = t1 + t2; n += m.
 * INJECTED mode takes the reference's own randn_like / rand_like draws; PHILOX mode restates the kernel's generator
 * (uniforms as the reset noise, normals by Box-Muller) so the two can be compared.
 * ------------------------------------------------------------------------------------------ */
static void noise_u4(const IgeNoiseArgs *a, uint64_t step, uint32_t block4, uint32_t stream, float u[4]) {
    uint32_t c[4] = {block4, stream, (uint32_t)step, (uint32_t)(step >> 32)};
    philox4x32_10(c, (uint32_t)a->seed, (uint32_t)(a->seed >> 32));
    for (int k = 0; k < 4; ++k) u[k] = (float)(c[k] >> 8) * (1.0f / 16777216.0f);
}
static void noise_n4(const IgeNoiseArgs *a, uint64_t step, uint32_t block4, uint32_t stream, float z[4]) {
    float u[4];
    noise_u4(a, step, block4, stream, u);
    for (int k = 0; k < 2; ++k) {
        volatile float l = logf(1.0f - u[2 * k]);
        volatile float m2 = -2.0f * l;
        float r = sqrtf(m2);
        volatile float t = 6.28318530717958647692f * u[2 * k + 1];
        volatile float zc = r * cosf(t), zs = r * sinf(t);
        z[2 * k] = zc;
        z[2 * k + 1] = zs;
    }
}
int ige_oracle_dr_noise(const IgeNoiseArgs *a) {
    const uint64_t step = (a->rand_mode == IGE_RAND_PHILOX && a->step) ? *a->step : 0;
    for (int64_t q = 0; q * 4 < a->count; ++q) {
        float corr4[4], draw4[4];
        if (a->rand_mode == IGE_RAND_PHILOX) {
            if (a->init_corr) noise_n4(a, step, (uint32_t)q, 2u * a->stream + 1u, corr4);
            if (a->distribution == IGE_NOISE_GAUSSIAN) noise_n4(a, step, (uint32_t)q, 2u * a->stream, draw4);
            else noise_u4(a, step, (uint32_t)q, 2u * a->stream, draw4);
        }
        for (int k = 0; k < 4 && q * 4 + k < a->count; ++k) {
            const int64_t i = q * 4 + k;
            if (a->rand_mode == IGE_RAND_PHILOX && a->init_corr) a->corr[i] = corr4[k];
            const float corr = a->corr[i];
            const float draw = a->rand_mode == IGE_RAND_PHILOX ? draw4[k] : a->draw[i];
            volatile float t1 = corr * a->corr_scale;     /* volatile: no contraction into fma */
            volatile float c = t1 + a->corr_shift;
            volatile float t2 = draw * a->scale;
            volatile float n0 = c + t2;
            volatile float n = n0 + a->shift;
            const float x = a->x[i];
            volatile float y = a->operation == IGE_NOISE_SCALING ? x * n : x + n;
            a->y[i] = y;
            if (a->y_clamped) {
                float yc = y;
                if (yc == yc) { yc = yc < -a->clip ? -a->clip : yc; yc = yc > a->clip ? a->clip : yc; }
                a->y_clamped[i] = yc;
            }
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * reset_buf.nonzero() with the multi-agent AND filter (franka_reach_MA.py:875-884)
 * ------------------------------------------------------------------------------------------ */
int ige_oracle_compact_resets(const IgeCompactArgs *a) {
    int g = a->group < 1 ? 1 : a->group;
    int n = a->num_rows / g;
    int cnt = 0;
    for (int e = 0; e < n; ++e) {
        int all = 1;
        for (int k = 0; k < g; ++k) all &= (a->reset_buf[(size_t)e * g + k] != 0);
        if (a->reset_rank) a->reset_rank[e] = cnt;
        if (all) a->reset_ids[cnt++] = e;
    }
    *a->count = cnt;
    return 0;
}

/* IgeResetList: ids of the envs reset in this step == reset_buf.nonzero() at entry, ascending
 * (group > 1: envs whose agent rows are all flagged).  Must run BEFORE the step rewrites reset_buf. */
static void emit_reset_list(const IgeResetList *l, const int64_t *reset_buf, int num_envs, int group) {
    if (!l->scratch) return;
    int cnt = 0;
    for (int e = 0; e < num_envs; ++e) {
        int all = 1;
        for (int k = 0; k < group; ++k) all &= (reset_buf[(size_t)e * group + k] != 0);
        if (all) l->ids[cnt++] = e;
    }
    *l->count = cnt;
}
/* every step kernel that is given a reset list advances the Philox step word on exit */
static void advance_step(const IgeResetList *l, const IgeResetRand *r) {
    if (l->scratch && r->step) *r->step += 1;
}

/* ------------------------------------------------------------------------------------------
 * Cartpole: tasks/cartpole.py:165-173 post_physics_step
 * ------------------------------------------------------------------------------------------ */
static void cartpole_range(const void *vp, const void *vb, int e0, int e1, double *acc) {
    const IgeCartpoleParams *p = (const IgeCartpoleParams *)vp;
    const IgeCartpoleBuffers *b = (const IgeCartpoleBuffers *)vb;
    double s_rew = 0, s_done = 0, s_to = 0, s_len = 0;
    for (int e = e0; e < e1; ++e) {
        float *ds = b->dof_state + (size_t)e * 4; /* (cart pos, cart vel, pole pos, pole vel) */
        int64_t progress = b->progress_buf[e] + 1; /* :166 */
        int64_t reset = b->reset_buf[e];
        if (reset != 0) { /* reset_idx :144-157 */
            for (int j = 0; j < 2; ++j) {
                ds[2 * j + 0] = 0.2f * (reset_uniform(&p->rand, 0, e, 2, j) - 0.5f); /* :145 */
                ds[2 * j + 1] = 0.5f * (reset_uniform(&p->rand, 1, e, 2, j) - 0.5f); /* :146 */
            }
            reset = 0; progress = 0; /* :156-157 */
        }
        /* compute_observations :137-140 */
        float cart_pos = ds[0], cart_vel = ds[1], pole_angle = ds[2], pole_vel = ds[3];
        float *obs = b->obs_buf + (size_t)e * 4;
        obs[0] = cart_pos; obs[1] = cart_vel; obs[2] = pole_angle; obs[3] = pole_vel;
        /* compute_cartpole_reward :180-196 */
        float reward = 1.0f - pole_angle * pole_angle - 0.01f * fabsf(cart_vel) - 0.005f * fabsf(pole_vel);
        if (fabsf(cart_pos) > p->reset_dist) reward = -2.0f;
        if (fabsf(pole_angle) > HALF_PI_F) reward = -2.0f;
        if (fabsf(cart_pos) > p->reset_dist) reset = 1;
        if (fabsf(pole_angle) > HALF_PI_F) reset = 1;
        if ((float)progress >= p->max_episode_length - 1.0f) reset = 1;
        b->rew_buf[e] = reward;
        b->reset_buf[e] = reset;
        b->progress_buf[e] = progress;
        /* VecTask.step epilogue, tasks/base/vec_task.py:396,404 */
        int timeout = ((float)progress >= p->max_episode_length - 1.0f) && (reset != 0);
        if (b->timeout_buf) b->timeout_buf[e] = (uint8_t)timeout;
        if (b->obs_clamped)
            for (int j = 0; j < 4; ++j) {
                float v = obs[j];
                v = v < -p->clip_obs ? -p->clip_obs : v;
                v = v > p->clip_obs ? p->clip_obs : v;
                b->obs_clamped[(size_t)e * 4 + j] = v;
            }
        s_rew += reward; s_done += (reset != 0); s_to += timeout; s_len += (reset != 0) ? (double)progress : 0.0;
    }
    acc[0] = s_rew; acc[1] = s_done; acc[2] = s_to; acc[3] = s_len;
}
int ige_oracle_cartpole_post_physics(const IgeCartpoleParams *p, const IgeCartpoleBuffers *b) {
    double acc[IGE_ORACLE_NACC];
    emit_reset_list(&b->list, b->reset_buf, p->num_envs, 1);
    parallel_ranges(cartpole_range, p, b, p->num_envs, acc);
    advance_step(&b->list, &p->rand);
    add_stat(b->stats, IGE_STAT_STEPS, p->num_envs); add_stat(b->stats, IGE_STAT_REWARD, acc[0]);
    add_stat(b->stats, IGE_STAT_DONE, acc[1]); add_stat(b->stats, IGE_STAT_TIMEOUT, acc[2]);
    add_stat(b->stats, IGE_STAT_DONE_LENGTH, acc[3]);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Ant / Humanoid: tasks/ant.py:287-297, tasks/humanoid.py:287-296
 * ------------------------------------------------------------------------------------------ */
static inline float neg_norm_over_dt(const IgeLocomotionParams *p, float tx, float ty) {
    float nn = -norm3(tx, ty, 0.0f);
    return p->potentials_recip_mul ? nn * (1.0f / p->dt) : nn / p->dt;
}

static void locomotion_range(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, int humanoid,
                             int e0, int e1, double *acc) {
    const int d = p->num_dof, ns = p->num_sensor_floats;
    const int num_obs = 12 + 3 * d + ns + (humanoid ? d : 0);
    double s_rew = 0, s_done = 0, s_to = 0, s_len = 0, s_obj = 0;
    for (int e = e0; e < e1; ++e) {
        const float *root = b->root_states + (size_t)e * 13;
        float *dof = b->dof_state + (size_t)e * d * 2;
        const float *act = b->actions + (size_t)e * d;
        float *obs = b->obs_buf + (size_t)e * num_obs;
        int64_t progress = b->progress_buf[e] + 1; /* ant.py:288 */
        if (b->randomize_buf) b->randomize_buf[e] += 1; /* ant.py:289 */
        int64_t reset = b->reset_buf[e];
        float potentials = b->potentials[e];

        if (reset != 0) { /* reset_idx: ant.py:252-279, humanoid.py:253-279 */
            for (int j = 0; j < d; ++j) {
                /* torch_rand_float(-0.2, 0.2): (upper - lower) * rand + lower, torch_jit_utils.py:215-218 */
                float pos = 0.4f * reset_uniform(&p->rand, 0, e, d, j) + (-0.2f);
                float vel = 0.2f * reset_uniform(&p->rand, 1, e, d, j) + (-0.1f);
                float v = b->initial_dof_pos[(size_t)e * d + j] + pos;
                v = fminf(v, p->dof_limits_upper[j]); /* tensor_clamp, torch_jit_utils.py:228-230 */
                v = fmaxf(v, p->dof_limits_lower[j]);
                dof[2 * j + 0] = v;   /* ant.py:260 */
                dof[2 * j + 1] = vel; /* ant.py:261 */
            }
            const float *init = b->initial_root_states + (size_t)e * 13;
            float tx = p->targets[0] - init[0], ty = p->targets[1] - init[1]; /* ant.py:273-274 */
            potentials = neg_norm_over_dt(p, tx, ty); /* ant.py:275-276 (prev == potentials) */
            progress = 0; reset = 0; /* ant.py:278-279 */
        }

        /* compute_ant_observations ant.py:374-408 / compute_humanoid_observations humanoid.py:378-413 */
        float to_target[3] = {p->targets[0] - root[0], p->targets[1] - root[1], 0.0f};
        float prev_potentials = potentials;                             /* ant.py:390 */
        potentials = neg_norm_over_dt(p, to_target[0], to_target[1]);   /* ant.py:391 */

        /* compute_heading_and_up torch_jit_utils.py:247-262 */
        float target_dirs[3], torso_quat[4], up_vec[3], heading_vec[3];
        normalize3(to_target, target_dirs);
        quat_mul(root + 3, p->inv_start_rot, torso_quat);
        quat_rotate_impl(torso_quat, p->basis_vec1, 0, up_vec);
        quat_rotate_impl(torso_quat, p->basis_vec0, 0, heading_vec);
        float up_proj = up_vec[2];
        float heading_proj = dot3(heading_vec, target_dirs);

        /* compute_rot torch_jit_utils.py:265-276 */
        float vel_loc[3], angvel_loc[3], roll, yaw;
        quat_rotate_impl(torso_quat, root + 7, 1, vel_loc);
        quat_rotate_impl(torso_quat, root + 10, 1, angvel_loc);
        euler_roll_yaw(torso_quat, &roll, &yaw);
        float walk_target_angle = atan2f(p->targets[2] - root[2], p->targets[0] - root[0]);
        float angle_to_target = walk_target_angle - yaw;

        obs[0] = root[2];
        for (int i = 0; i < 3; ++i) obs[1 + i] = vel_loc[i];
        if (humanoid) {
            for (int i = 0; i < 3; ++i) obs[4 + i] = angvel_loc[i] * p->angular_velocity_scale;
            obs[7] = normalize_angle(yaw);              /* humanoid.py:402-404 */
            obs[8] = normalize_angle(roll);
            obs[9] = normalize_angle(angle_to_target);
        } else {
            for (int i = 0; i < 3; ++i) obs[4 + i] = angvel_loc[i];
            obs[7] = yaw; obs[8] = roll; obs[9] = angle_to_target; /* ant.py:403 */
        }
        obs[10] = up_proj;
        obs[11] = heading_proj;
        for (int j = 0; j < d; ++j) {
            float lo = p->dof_limits_lower[j], hi = p->dof_limits_upper[j];
            obs[12 + j] = (2.0f * dof[2 * j] - hi - lo) / (hi - lo); /* unscale torch_jit_utils.py:238-240 */
            obs[12 + d + j] = dof[2 * j + 1] * p->dof_vel_scale;
        }
        int o = 12 + 2 * d;
        if (humanoid) {
            for (int j = 0; j < d; ++j) obs[o + j] = b->dof_force_tensor[(size_t)e * d + j] * p->contact_force_scale;
            o += d;
        }
        for (int j = 0; j < ns; ++j) obs[o + j] = b->vec_sensor_tensor[(size_t)e * ns + j] * p->contact_force_scale;
        o += ns;
        for (int j = 0; j < d; ++j) obs[o + j] = act[j];

        /* compute_ant_reward ant.py:325-371 / compute_humanoid_reward humanoid.py:323-375 */
        float heading_reward = obs[11] > 0.8f ? p->heading_weight : p->heading_weight * obs[11] / 0.8f;
        float up_reward = obs[10] > 0.93f ? 0.0f + p->up_weight : 0.0f;
        float tmp[IGE_MAX_DOF];
        for (int j = 0; j < d; ++j) tmp[j] = act[j] * act[j];
        float actions_cost = row_sum(tmp, d);
        float total;
        float progress_reward = potentials - prev_potentials;
        if (humanoid) {
            for (int j = 0; j < d; ++j) {
                float a = fabsf(obs[12 + j]);
                float scaled = p->joints_at_limit_cost_scale * (a - 0.98f) / 0.02f; /* humanoid.py:356 */
                tmp[j] = (a > 0.98f ? 1.0f : 0.0f) * scaled * p->motor_effort_ratio[j];
            }
            float dof_at_limit_cost = row_sum(tmp, d);
            for (int j = 0; j < d; ++j) tmp[j] = fabsf(act[j] * obs[12 + d + j]) * p->motor_effort_ratio[j];
            float electricity_cost = row_sum(tmp, d);
            total = progress_reward + 2.0f + up_reward + heading_reward - p->actions_cost_scale * actions_cost -
                    p->energy_cost_scale * electricity_cost - dof_at_limit_cost;
        } else {
            for (int j = 0; j < d; ++j) tmp[j] = fabsf(act[j] * obs[12 + d + j]);
            float electricity_cost = row_sum(tmp, d);
            int64_t dof_at_limit = 0;
            for (int j = 0; j < d; ++j) dof_at_limit += obs[12 + j] > 0.99f;
            total = progress_reward + 0.5f + up_reward + heading_reward - p->actions_cost_scale * actions_cost -
                    p->energy_cost_scale * electricity_cost - (float)dof_at_limit * p->joints_at_limit_cost_scale;
        }
        if (obs[0] < p->termination_height) total = p->death_cost; /* ant.py:365 */
        if (obs[0] < p->termination_height) reset = 1;             /* ant.py:368 */
        if ((float)progress >= p->max_episode_length - 1.0f) reset = 1; /* ant.py:369 */

        b->rew_buf[e] = total;
        b->reset_buf[e] = reset;
        b->progress_buf[e] = progress;
        b->potentials[e] = potentials;
        b->prev_potentials[e] = prev_potentials;
        for (int i = 0; i < 3; ++i) {
            b->up_vec[(size_t)e * 3 + i] = up_vec[i];
            b->heading_vec[(size_t)e * 3 + i] = heading_vec[i];
        }
        int timeout = ((float)progress >= p->max_episode_length - 1.0f) && (reset != 0); /* vec_task.py:396 */
        if (b->timeout_buf) b->timeout_buf[e] = (uint8_t)timeout;
        if (b->obs_clamped)
            for (int j = 0; j < num_obs; ++j) {
                float v = obs[j];
                v = v < -p->clip_obs ? -p->clip_obs : v;
                v = v > p->clip_obs ? p->clip_obs : v;
                b->obs_clamped[(size_t)e * num_obs + j] = v;
            }
        s_rew += total; s_done += (reset != 0); s_to += timeout;
        s_len += (reset != 0) ? (double)progress : 0.0; s_obj += root[7];
    }
    acc[0] = s_rew; acc[1] = s_done; acc[2] = s_to; acc[3] = s_len; acc[4] = s_obj;
}
static void ant_range(const void *p, const void *b, int e0, int e1, double *acc) {
    locomotion_range((const IgeLocomotionParams *)p, (const IgeLocomotionBuffers *)b, 0, e0, e1, acc);
}
static void humanoid_range(const void *p, const void *b, int e0, int e1, double *acc) {
    locomotion_range((const IgeLocomotionParams *)p, (const IgeLocomotionBuffers *)b, 1, e0, e1, acc);
}
static int locomotion_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, int humanoid) {
    double acc[IGE_ORACLE_NACC];
    emit_reset_list(&b->list, b->reset_buf, p->num_envs, 1);
    parallel_ranges(humanoid ? humanoid_range : ant_range, p, b, p->num_envs, acc);
    advance_step(&b->list, &p->rand);
    add_stat(b->stats, IGE_STAT_STEPS, p->num_envs); add_stat(b->stats, IGE_STAT_REWARD, acc[0]);
    add_stat(b->stats, IGE_STAT_DONE, acc[1]); add_stat(b->stats, IGE_STAT_TIMEOUT, acc[2]);
    add_stat(b->stats, IGE_STAT_DONE_LENGTH, acc[3]); add_stat(b->stats, IGE_STAT_TASK0, acc[4]);
    return 0;
}

int ige_oracle_ant_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b) {
    return locomotion_post_physics(p, b, 0);
}
int ige_oracle_humanoid_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b) {
    return locomotion_post_physics(p, b, 1);
}

/* ------------------------------------------------------------------------------------------
 * ShadowHand: tasks/shadow_hand.py
 * ------------------------------------------------------------------------------------------ */
static float hand_uniform(const IgeHandResetRand *r, int draw, int row_rank, int env, int cols, int j) {
    /* draw: 0 = A (goal ids), 1 = B, 2 = C, 3 = D (env ids) */
    if (r->mode == IGE_RAND_PHILOX) {
        IgeResetRand q = {IGE_RAND_PHILOX, 0, r->seed, r->step, 0, 0, 0};
        return philox_uniform(&q, (uint32_t)env, (uint32_t)draw, (uint32_t)j);
    }
    const float *u = draw == 0 ? r->uA : draw == 1 ? r->uB : draw == 2 ? r->uC : r->uD;
    return u[(size_t)row_rank * cols + j];
}
/* torch_rand_float(-1, 1): (upper - lower) * rand + lower, torch_jit_utils.py:215-218 */
static inline float sym_uniform(float u) { return 2.0f * u + (-1.0f); }

/* randomize_rotation shadow_hand.py:803-806 with quat_from_angle_axis torch_jit_utils.py:118-123:
 * quat_mul(aa(rand0 * pi, x), aa(rand1 * pi, y)); each aa() result goes through quat_unit(). */
static void quat_from_angle_axis_unit(float angle, int axis, float *q) {
    float theta = angle / 2.0f;
    float s = sinf(theta), c = cosf(theta);
    float v[4] = {0, 0, 0, c};
    v[axis] = 1.0f * s; /* normalize(axis) * sin: the unit axis has norm exactly 1 */
    float n2 = v[0] * v[0];
    n2 = fmaf(v[1], v[1], n2); n2 = fmaf(v[2], v[2], n2); n2 = fmaf(v[3], v[3], n2);
    float n = sqrtf(n2);
    if (n < 1e-9f) n = 1e-9f;
    for (int i = 0; i < 4; ++i) q[i] = v[i] / n;
}
static void randomize_rotation(float rand0, float rand1, float *out) {
    const float PI_F = 3.14159265358979323846f;
    float qa[4], qb[4];
    quat_from_angle_axis_unit(rand0 * PI_F, 0, qa);
    quat_from_angle_axis_unit(rand1 * PI_F, 1, qb);
    quat_mul(qa, qb, out);
}

/* reset_target_pose shadow_hand.py:586-602 for one env */
static void hand_reset_target(const IgeShadowHandParams *p, const IgeShadowHandBuffers *b, int e, float r0, float r1) {
    float rot[4];
    randomize_rotation(r0, r1, rot);
    float *gs = b->goal_states + (size_t)e * 13;
    const float *gi = b->goal_init_state + (size_t)e * 13;
    float *root = b->root_state_tensor + (size_t)b->goal_object_indices[e] * 13;
    for (int i = 0; i < 3; ++i) { gs[i] = gi[i]; root[i] = gs[i] + p->goal_displacement[i]; }
    for (int i = 0; i < 4; ++i) { gs[3 + i] = rot[i]; root[3 + i] = rot[i]; }
    for (int i = 7; i < 13; ++i) root[i] = 0.0f;
    if (b->actor_dirty) b->actor_dirty[b->goal_object_indices[e]] = 1;
    b->reset_goal_buf[e] = 0;
}

/* the reset section of pre_physics_step, shadow_hand.py:670-682 */
int ige_oracle_shadow_hand_reset(const IgeShadowHandParams *p, const IgeShadowHandBuffers *b) {
    const int n = p->num_envs, d = p->num_hand_dofs, dpe = p->num_dofs_per_env;
    const IgeHandResetRand *r = &p->rand;
    emit_reset_list(&b->list, b->reset_buf, n, 1);
    for (int e = 0; e < n; ++e) {
        const int goal = b->reset_goal_buf[e] != 0, env = b->reset_buf[e] != 0;
        if (goal) { /* :675-679 reset_target_pose(goal_env_ids) */
            const int rk = r->goal_rank ? r->goal_rank[e] : 0;
            hand_reset_target(p, b, e, sym_uniform(hand_uniform(r, 0, rk, e, 4, 0)), sym_uniform(hand_uniform(r, 0, rk, e, 4, 1)));
        }
        if (!env) continue;
        /* reset_idx :604-668 */
        const int rk = r->env_rank ? r->env_rank[e] : 0;
        const int cols = 2 * d + 5;
        float rf[2 * IGE_MAX_DOF + 5];
        for (int j = 0; j < cols; ++j) rf[j] = sym_uniform(hand_uniform(r, 1, rk, e, cols, j));
        hand_reset_target(p, b, e, sym_uniform(hand_uniform(r, 2, rk, e, 4, 0)), sym_uniform(hand_uniform(r, 2, rk, e, 4, 1))); /* :613 */
        for (int i = 0; i < p->num_bodies * 3; ++i) b->rb_forces[(size_t)e * p->num_bodies * 3 + i] = 0.0f; /* :616 */
        float *obj = b->root_state_tensor + (size_t)b->object_indices[e] * 13;
        const float *oi = b->object_init_state + (size_t)e * 13;
        for (int i = 0; i < 13; ++i) obj[i] = oi[i];                               /* :619 */
        obj[0] = oi[0] + p->reset_position_noise * rf[0];                          /* :620-621 */
        obj[1] = oi[1] + p->reset_position_noise * rf[1];
        obj[2] = oi[2] + p->reset_position_noise * rf[2];                          /* :622-623 (up axis z) */
        float rot[4];
        randomize_rotation(rf[3], rf[4], rot);                                     /* :625 */
        for (int i = 0; i < 4; ++i) obj[3 + i] = rot[i];
        for (int i = 7; i < 13; ++i) obj[i] = 0.0f;
        if (b->actor_dirty) b->actor_dirty[b->object_indices[e]] = 1;
        /* :644-645 */
        float l0 = logf(p->force_prob_range[0]), l1 = logf(p->force_prob_range[1]);
        b->random_force_prob[e] = expf((l0 - l1) * hand_uniform(r, 3, rk, e, 1, 0) + l1);
        for (int j = 0; j < d; ++j) { /* :647-657 */
            float delta_max = p->dof_upper[j] - p->dof_default_pos[j];
            float delta_min = p->dof_lower[j] - p->dof_default_pos[j];
            float rand_delta = delta_min + (delta_max - delta_min) * 0.5f * (rf[5 + j] + 1.0f);
            float pos = p->dof_reset_base[j] + p->reset_dof_pos_noise * rand_delta;
            float vel = p->dof_default_vel[j] + p->reset_dof_vel_noise * rf[5 + d + j];
            b->dof_state[((size_t)e * dpe + j) * 2 + 0] = pos;
            b->dof_state[((size_t)e * dpe + j) * 2 + 1] = vel;
            b->prev_targets[(size_t)e * dpe + j] = pos;
            b->cur_targets[(size_t)e * dpe + j] = pos;
        }
        b->progress_buf[e] = 0; b->reset_buf[e] = 0; b->successes[e] = 0.0f; /* :666-668 */
    }
    return 0;
}

/* observation row layouts, shadow_hand.py:473-584 */
enum { SEG_DOF_POS, SEG_DOF_VEL, SEG_DOF_FORCE, SEG_OBJ_POSE, SEG_OBJ_POS, SEG_OBJ_LINVEL, SEG_OBJ_ANGVEL,
       SEG_GOAL_POSE, SEG_QUAT_DIFF, SEG_FT_STATE, SEG_FT_POS, SEG_SENSORS, SEG_ACTIONS, SEG_END };
static const int HAND_LAYOUT[7][12][2] = {
    /* openai 42 */ {{SEG_FT_POS, 0}, {SEG_OBJ_POS, 15}, {SEG_QUAT_DIFF, 18}, {SEG_ACTIONS, 22}, {SEG_END, 42}},
    /* full_no_vel 77 */ {{SEG_DOF_POS, 0}, {SEG_OBJ_POSE, 24}, {SEG_GOAL_POSE, 31}, {SEG_QUAT_DIFF, 38}, {SEG_FT_POS, 42},
                          {SEG_ACTIONS, 57}, {SEG_END, 77}},
    /* full 157 */ {{SEG_DOF_POS, 0}, {SEG_DOF_VEL, 24}, {SEG_OBJ_POSE, 48}, {SEG_OBJ_LINVEL, 55}, {SEG_OBJ_ANGVEL, 58},
                    {SEG_GOAL_POSE, 61}, {SEG_QUAT_DIFF, 68}, {SEG_FT_STATE, 72}, {SEG_ACTIONS, 137}, {SEG_END, 157}},
    /* full_state 211 */ {{SEG_DOF_POS, 0}, {SEG_DOF_VEL, 24}, {SEG_DOF_FORCE, 48}, {SEG_OBJ_POSE, 72}, {SEG_OBJ_LINVEL, 79},
                          {SEG_OBJ_ANGVEL, 82}, {SEG_GOAL_POSE, 85}, {SEG_QUAT_DIFF, 92}, {SEG_FT_STATE, 96}, {SEG_SENSORS, 161},
                          {SEG_ACTIONS, 191}, {SEG_END, 211}},
    /* AllegroHand full_no_vel 50 (allegro_hand.py:439-449) */
    {{SEG_DOF_POS, 0}, {SEG_OBJ_POSE, 16}, {SEG_GOAL_POSE, 23}, {SEG_QUAT_DIFF, 30}, {SEG_ACTIONS, 34}, {SEG_END, 50}},
    /* AllegroHand full 72 (:450-465) */
    {{SEG_DOF_POS, 0}, {SEG_DOF_VEL, 16}, {SEG_OBJ_POSE, 32}, {SEG_OBJ_LINVEL, 39}, {SEG_OBJ_ANGVEL, 42},
     {SEG_GOAL_POSE, 45}, {SEG_QUAT_DIFF, 52}, {SEG_ACTIONS, 56}, {SEG_END, 72}},
    /* AllegroHand full_state 88 (:470-489; also the asymmetric states_buf :452-469) */
    {{SEG_DOF_POS, 0}, {SEG_DOF_VEL, 16}, {SEG_DOF_FORCE, 32}, {SEG_OBJ_POSE, 48}, {SEG_OBJ_LINVEL, 55},
     {SEG_OBJ_ANGVEL, 58}, {SEG_GOAL_POSE, 61}, {SEG_QUAT_DIFF, 68}, {SEG_ACTIONS, 72}, {SEG_END, 88}}};

static void hand_fill_row(const IgeShadowHandParams *p, const IgeShadowHandBuffers *b, int e, int layout, float *row,
                          const float *quat_diff) {
    const int d = p->num_hand_dofs, dpe = p->num_dofs_per_env;
    const float *obj = b->root_state_tensor + (size_t)b->object_indices[e] * 13;
    const float *goal = b->goal_states + (size_t)e * 13;
    for (int s = 0; HAND_LAYOUT[layout][s][0] != SEG_END; ++s) {
        float *o = row + HAND_LAYOUT[layout][s][1];
        switch (HAND_LAYOUT[layout][s][0]) {
        case SEG_DOF_POS:
            for (int j = 0; j < d; ++j) {
                float lo = p->dof_lower[j], hi = p->dof_upper[j];
                o[j] = (2.0f * b->dof_state[((size_t)e * dpe + j) * 2] - hi - lo) / (hi - lo);
            }
            break;
        case SEG_DOF_VEL: for (int j = 0; j < d; ++j) o[j] = p->vel_obs_scale * b->dof_state[((size_t)e * dpe + j) * 2 + 1]; break;
        case SEG_DOF_FORCE: for (int j = 0; j < d; ++j) o[j] = p->force_torque_obs_scale * b->dof_force_tensor[(size_t)e * d + j]; break;
        case SEG_OBJ_POSE: for (int i = 0; i < 7; ++i) o[i] = obj[i]; break;
        case SEG_OBJ_POS: for (int i = 0; i < 3; ++i) o[i] = obj[i]; break;
        case SEG_OBJ_LINVEL: for (int i = 0; i < 3; ++i) o[i] = obj[7 + i]; break;
        case SEG_OBJ_ANGVEL: for (int i = 0; i < 3; ++i) o[i] = p->vel_obs_scale * obj[10 + i]; break;
        case SEG_GOAL_POSE: for (int i = 0; i < 7; ++i) o[i] = goal[i]; break;
        case SEG_QUAT_DIFF: for (int i = 0; i < 4; ++i) o[i] = quat_diff[i]; break;
        case SEG_FT_STATE:
            for (int f = 0; f < IGE_HAND_NUM_FINGERTIPS; ++f)
                for (int i = 0; i < 13; ++i)
                    o[f * 13 + i] = b->rigid_body_states[((size_t)e * p->num_bodies + p->fingertip_handles[f]) * 13 + i];
            break;
        case SEG_FT_POS:
            for (int f = 0; f < IGE_HAND_NUM_FINGERTIPS; ++f)
                for (int i = 0; i < 3; ++i)
                    o[f * 3 + i] = b->rigid_body_states[((size_t)e * p->num_bodies + p->fingertip_handles[f]) * 13 + i];
            break;
        case SEG_SENSORS: for (int i = 0; i < 30; ++i) o[i] = p->force_torque_obs_scale * b->vec_sensor_tensor[(size_t)e * 30 + i]; break;
        case SEG_ACTIONS: for (int i = 0; i < p->num_actions; ++i) o[i] = b->actions[(size_t)e * p->num_actions + i]; break;
        }
    }
}
static int hand_num_obs(int layout) {
    int s = 0;
    while (HAND_LAYOUT[layout][s][0] != SEG_END) ++s;
    return HAND_LAYOUT[layout][s][1];
}

/* post_physics_step shadow_hand.py:710-715 */
int ige_oracle_shadow_hand_post_physics(const IgeShadowHandParams *p, const IgeShadowHandBuffers *b) {
    const int n = p->num_envs, na = p->num_actions;
    const int num_obs = hand_num_obs(p->obs_type);
    /* the asymmetric states_buf: ShadowHand 211-wide full_state (:529-556), AllegroHand 88-wide (allegro_hand.py:452-469) */
    const int st_layout = p->obs_type >= IGE_HAND_OBS_ALLEGRO_FULL_NO_VEL ? IGE_HAND_OBS_ALLEGRO_FULL_STATE : IGE_HAND_OBS_FULL_STATE;
    const int num_st = hand_num_obs(st_layout);
    float tol = p->success_tolerance;
    if (p->ignore_z_rot) tol = 2.0f * tol; /* :758-759 */
    int64_t num_resets = 0;
    float finished = 0.0f;
    double s_rew = 0, s_done = 0, s_to = 0, s_len = 0;
    for (int e = 0; e < n; ++e) {
        int64_t progress = b->progress_buf[e] + 1; /* :711 */
        if (b->randomize_buf) b->randomize_buf[e] += 1;
        const float *obj = b->root_state_tensor + (size_t)b->object_indices[e] * 13;
        const float *goal = b->goal_states + (size_t)e * 13;
        const float *act = b->actions + (size_t)e * na;
        /* compute_observations :437-471 */
        float conj[4], quat_diff[4];
        quat_conjugate(goal + 3, conj);
        quat_mul(obj + 3, conj, quat_diff);
        float *obs = b->obs_buf + (size_t)e * num_obs;
        hand_fill_row(p, b, e, p->obs_type, obs, quat_diff);
        if (p->asymmetric_obs) hand_fill_row(p, b, e, st_layout, b->states_buf + (size_t)e * num_st, quat_diff);
        /* compute_hand_reward :746-800 */
        float goal_dist = norm3(obj[0] - goal[0], obj[1] - goal[1], obj[2] - goal[2]);
        float qn = norm3(quat_diff[0], quat_diff[1], quat_diff[2]);
        float rot_dist = 2.0f * asinf(qn > 1.0f ? 1.0f : qn);
        float dist_rew = goal_dist * p->dist_reward_scale;
        float rot_rew = 1.0f / (fabsf(rot_dist) + p->rot_eps) * p->rot_reward_scale;
        float sq[IGE_MAX_DOF];
        for (int j = 0; j < na; ++j) sq[j] = act[j] * act[j];
        float action_penalty = row_sum(sq, na);
        float reward = dist_rew + rot_rew + action_penalty * p->action_penalty_scale;
        int64_t goal_resets = fabsf(rot_dist) <= tol ? 1 : b->reset_goal_buf[e];
        float successes = b->successes[e] + (float)goal_resets;
        if (goal_resets == 1) reward = reward + p->reach_goal_bonus;
        if (goal_dist >= p->fall_dist) reward = reward + p->fall_penalty;
        int64_t resets = goal_dist >= p->fall_dist ? 1 : b->reset_buf[e];
        if (p->max_consecutive_successes > 0) {
            if (fabsf(rot_dist) <= tol) progress = 0;
            if (successes >= (float)p->max_consecutive_successes) resets = 1;
        }
        const int late = (float)progress >= p->max_episode_length - 1.0f;
        if (late) resets = 1;
        if (p->max_consecutive_successes > 0 && late) reward = reward + 0.5f * p->fall_penalty;
        num_resets += resets;
        finished += successes * (float)resets;
        b->rew_buf[e] = reward; b->reset_buf[e] = resets; b->reset_goal_buf[e] = goal_resets;
        b->progress_buf[e] = progress; b->successes[e] = successes;
        /* VecTask.step epilogue vec_task.py:396-408 */
        int timeout = late && resets != 0;
        if (b->timeout_buf) b->timeout_buf[e] = (uint8_t)timeout;
        if (b->obs_clamped)
            for (int j = 0; j < num_obs; ++j) {
                float v = obs[j];
                v = v < -p->clip_obs ? -p->clip_obs : v; v = v > p->clip_obs ? p->clip_obs : v;
                b->obs_clamped[(size_t)e * num_obs + j] = v;
            }
        if (b->states_clamped && p->asymmetric_obs)
            for (int j = 0; j < num_st; ++j) {
                float v = b->states_buf[(size_t)e * num_st + j];
                v = v < -p->clip_obs ? -p->clip_obs : v; v = v > p->clip_obs ? p->clip_obs : v;
                b->states_clamped[(size_t)e * num_st + j] = v;
            }
        s_rew += reward; s_done += resets != 0; s_to += timeout; s_len += resets != 0 ? (double)progress : 0.0;
    }
    /* :795-798 */
    if (num_resets > 0)
        b->consecutive_successes[0] = p->av_factor * finished / (float)num_resets +
                                      (float)(1.0 - (double)p->av_factor) * b->consecutive_successes[0];
    add_stat(b->stats, IGE_STAT_STEPS, n); add_stat(b->stats, IGE_STAT_REWARD, s_rew);
    add_stat(b->stats, IGE_STAT_DONE, s_done); add_stat(b->stats, IGE_STAT_TIMEOUT, s_to);
    add_stat(b->stats, IGE_STAT_DONE_LENGTH, s_len); add_stat(b->stats, IGE_STAT_TASK0, (double)num_resets);
    add_stat(b->stats, IGE_STAT_TASK0 + 1, finished);
    if (b->list.scratch && p->rand.step) *p->rand.step += 1;
    return 0;
}


/* ------------------------------------------------------------------------------------------
 * Franka multi-agent tasks: tasks/franka_{reach,collect,combine}_MA.py, tasks/franka_PPMA.py
 * ------------------------------------------------------------------------------------------ */
static float ma_uniform(const IgeMAResetRand *r, int draw, int rank, int env, int cols, int j) {
    if (r->mode == IGE_RAND_PHILOX) {
        IgeResetRand q = {IGE_RAND_PHILOX, 0, r->seed, r->step, 0, 0, 0};
        return philox_uniform(&q, (uint32_t)env, (uint32_t)draw, (uint32_t)j);
    }
    return r->u[draw][(size_t)rank * cols + j];
}
static const float *ma_body(const IgeFrankaMAParams *p, const IgeFrankaMABuffers *b, int e, int start, int step, int i) {
    return b->rigid_body_state + ((size_t)e * p->num_bodies + start + step * i) * 13;
}
static float *ma_actor(const IgeFrankaMAParams *p, const IgeFrankaMABuffers *b, int e, int actor) {
    return b->root_state + ((size_t)e * p->num_actors + actor) * 13;
}

/* reset_idx for one env whose agents are all flagged (franka_reach_MA.py:616-704 and siblings) */
static void ma_reset_env(const IgeFrankaMAParams *p, const IgeFrankaMABuffers *b, int e) {
    const int A = p->num_agents, T = p->num_targets;
    const IgeMAResetRand *r = &p->rand;
    const int rk = r->env_rank ? r->env_rank[e] : 0;
    for (int c = 0; c < T; ++c) { /* _reset_init_cube_state */
        float *init = b->init_cubeA_state + ((size_t)e * T + c) * 13;
        for (int i = 0; i < 13; ++i) init[i] = 0.0f;
        init[2] = p->cube_z_base + ma_uniform(r, 0, rk, e, T, c) * p->cube_z_range;
        init[0] = p->cube_xy_center[0] + p->cube_xy_mult * (ma_uniform(r, 1, rk, e, T * 2, c * 2 + 0) - 0.5f);
        init[1] = p->cube_xy_center[1] + p->cube_xy_mult * (ma_uniform(r, 1, rk, e, T * 2, c * 2 + 1) - 0.5f);
        init[6] = 1.0f;
        float *cube = ma_actor(p, b, e, p->cube_actor_start + p->cube_actor_step * c);
        for (int i = 0; i < 13; ++i) cube[i] = init[i]; /* _cubeA_state[env_ids] = _init_cubeA_state[env_ids] */
    }
    if (p->variant == IGE_FRANKA_PP) { /* _reset_dest_cube_state franka_PPMA.py:867-895 */
        for (int c = 0; c < T; ++c) {
            float *d = ma_actor(p, b, e, p->dest_actor_start + p->dest_actor_step * c);
            for (int i = 0; i < 13; ++i) d[i] = 0.0f;
            d[2] = p->aux_z_base + ma_uniform(r, 2, rk, e, T, c) * 0.0f;
            d[0] = p->aux_xy_center_x[c] + p->aux_xy_mult * (ma_uniform(r, 3, rk, e, T * 2, c * 2 + 0) - 0.5f);
            d[1] = p->aux_xy_center_y + p->aux_xy_mult * (ma_uniform(r, 3, rk, e, T * 2, c * 2 + 1) - 0.5f);
            d[6] = 1.0f;
        }
    }
    if (p->variant == IGE_FRANKA_COMBINE) { /* _reset_stack_base_cube_state franka_combine_MA.py:880-908 */
        float *d = ma_actor(p, b, e, p->stack_base_actor);
        for (int i = 0; i < 13; ++i) d[i] = 0.0f;
        d[2] = p->aux_z_base + ma_uniform(r, 2, rk, e, 1, 0) * 0.0f;
        d[0] = p->aux_xy_center_x[0] + p->aux_xy_mult * (ma_uniform(r, 3, rk, e, 2, 0) - 0.5f);
        d[1] = p->aux_xy_center_y + p->aux_xy_mult * (ma_uniform(r, 3, rk, e, 2, 1) - 0.5f);
        d[6] = 1.0f;
    }
    float pos[9];
    for (int j = 0; j < 9; ++j) { /* :638-644 */
        float v = p->franka_default_dof_pos[j] + p->dof_noise_mult * (ma_uniform(r, 4, rk, e, 9, j) - 0.5f);
        v = fminf(v, p->franka_dof_upper[j]);
        v = fmaxf(v, p->franka_dof_lower[j]);
        pos[j] = j >= 7 ? p->franka_default_dof_pos[j] : v;
    }
    for (int a = 0; a < A; ++a)
        for (int j = 0; j < 9; ++j) { /* :646-656: the same sample tiled over the A arms */
            size_t k = (size_t)e * p->num_dofs + a * 9 + j;
            b->dof_state[k * 2 + 0] = pos[j];
            b->dof_state[k * 2 + 1] = 0.0f;
            b->pos_control[k] = pos[j];
            b->effort_control[k] = 0.0f;
        }
    for (int a = 0; a < A; ++a) { b->progress_buf[(size_t)e * A + a] = 0; b->reset_buf[(size_t)e * A + a] = 0; }
}

static inline float pow2_small(int s) { return s >= 0 ? (float)(1 << s) : 0.5f; } /* torch.pow(2.0, FSM) */

int ige_oracle_franka_ma_post_physics(const IgeFrankaMAParams *p, const IgeFrankaMABuffers *b) {
    const int N = p->num_envs, A = p->num_agents, T = p->num_targets, NA = p->num_actions, NO = p->num_obs, V = p->variant;
    /* post_physics_step :821-826: progress++, agent flags -> reset_idx (AND filter) */
    for (size_t i = 0; i < (size_t)N * A; ++i) b->progress_buf[i] += 1;
    emit_reset_list(&b->list, b->reset_buf, N, A);
    for (int e = 0; e < N; ++e) {
        int all = 1;
        for (int a = 0; a < A; ++a) all &= b->reset_buf[(size_t)e * A + a] != 0;
        if (all) ma_reset_env(p, b, e);
    }
    double comp[IGE_MA_NUM_COMPONENTS] = {0};
    double s_rew = 0, s_done = 0, s_to = 0, s_len = 0;
    for (int e = 0; e < N; ++e) {
        const float *eef[IGE_MA_MAX_AGENTS], *base[IGE_MA_MAX_AGENTS], *cube[IGE_MA_MAX_AGENTS];
        for (int a = 0; a < A; ++a) {
            eef[a] = ma_body(p, b, e, p->eef_body_start, p->eef_body_step, a);
            base[a] = ma_body(p, b, e, p->base_body_start, p->base_body_step, a);
        }
        for (int c = 0; c < T; ++c) cube[c] = ma_actor(p, b, e, p->cube_actor_start + p->cube_actor_step * c);
        /* _get_min_relative_pos (franka_reach_MA.py:521-531) */
        int nearest[IGE_MA_MAX_AGENTS];
        float min_rel[IGE_MA_MAX_AGENTS][3];
        for (int a = 0; a < A; ++a) {
            float best = 0.0f;
            int bi = 0;
            for (int c = 0; c < T; ++c) {
                float d = norm3(cube[c][0] - eef[a][0], cube[c][1] - eef[a][1], cube[c][2] - eef[a][2]);
                if (c == 0 || d < best) { best = d; bi = c; }
            }
            nearest[a] = bi;
            for (int i = 0; i < 3; ++i) min_rel[a][i] = cube[bi][i] - eef[a][i];
        }
        /* per-agent FSM */
        int fsm[IGE_MA_MAX_AGENTS] = {0, 0, 0, 0};
        float md[IGE_MA_MAX_AGENTS];
        const float on_cube_th = (float)(0.05 * 0.5 * 0.9); /* cubeA_size * 0.5 * 0.9 with cubeA_size = 0.05 */
        const float on_th = (float)((double)p->cubeA_size * 0.5 * 0.9);
        (void)on_cube_th;
        for (int a = 0; a < A; ++a) {
            const float *act = b->actions + ((size_t)e * A + a) * NA;
            const int closed = act[NA - 1] < 0.0f;
            if (V == IGE_FRANKA_REACH) {
                md[a] = norm3(min_rel[a][0], min_rel[a][1], min_rel[a][2]);
            } else if (V == IGE_FRANKA_COLLECT) { /* franka_collect_MA.py:549-606 */
                md[a] = norm3(min_rel[a][0], min_rel[a][1], min_rel[a][2]);
                const float *hc = cube[nearest[a]];
                const int close = md[a] <= on_th;
                const int holding = close && closed;
                const int high = (hc[2] - 1.05f) > (float)((double)p->wall_height + (double)p->cubeA_size / 2);
                const int in_area = (hc[1] > (float)((double)p->wall_pos_y + (double)p->cubeA_size * 1.)) && (fabsf(hc[0]) < 0.6f);
                const int low = (hc[2] - 1.05f) < (float)((double)p->wall_height / 2.);
                int s = 0;
                if (close) s = 1;
                if (holding) s = 2;
                if (holding && high) s = 3;
                if (holding && in_area) s = 4;
                if (holding && in_area && low) s = 5;
                if (holding && in_area && low && !closed) s = 6;
                fsm[a] = s;
            } else { /* Combine :565-611 / PP :561-595: own cube */
                md[a] = norm3(cube[a][0] - eef[a][0], cube[a][1] - eef[a][1], cube[a][2] - eef[a][2]);
                const int on = md[a] <= on_th;
                int s = 0;
                if (on) s = 1;
                if (on && closed) s = 2;
                if (V == IGE_FRANKA_PP) {
                    const float *dest = ma_actor(p, b, e, p->dest_actor_start + p->dest_actor_step * a);
                    const float dz = dest[2] - cube[a][2];
                    const int lifted = fabsf(dz) >= (float)(((double)p->dest_size + (double)p->cubeA_size) / 2.);
                    const int align = norm2(dest[0] - cube[a][0], dest[1] - cube[a][1]) < (float)((0.05) * 0.5);
                    const int stackable = fabsf(dz) <= (float)((double)p->cubeA_size * 0.8660 + (double)p->dest_size / 2.);
                    if (on && lifted) s = 3;
                    if (align) s = 4;
                    if (align && stackable) s = 5;
                    if (align && stackable && !closed) s = 6;
                }
                fsm[a] = s;
            }
        }
        /* rows */
        for (int a = 0; a < A; ++a) {
            const size_t row = (size_t)e * A + a;
            const float *act = b->actions + row * NA;
            float *o = b->obs_buf + row * NO;
            int gfsm = 0;
            if (V == IGE_FRANKA_COMBINE || V == IGE_FRANKA_PP) {
                /* compute_global_FSM franka_PPMA.py:597-611: d_ecp is stack([..]).flatten(), i.e. laid out
                 * [all envs | all envs]; flat row k pairs d_ep[k // 2] with d_ecp[k] */
                const float *e0 = eef[0], *e1 = eef[1];
                const int too_close_grip = norm3(e1[0] - e0[0], e1[1] - e0[1], e1[2] - e0[2]) <= 0.18f;
                const size_t k = row;
                const int j = k < (size_t)N ? (int)k : (int)(k - N);
                const float *je = ma_body(p, b, j, p->eef_body_start, p->eef_body_step, k < (size_t)N ? 0 : 1);
                const float *jc = ma_actor(p, b, j, p->cube_actor_start + p->cube_actor_step * (k < (size_t)N ? 1 : 0));
                const int too_close_cube = norm3(je[0] - jc[0], je[1] - jc[1], je[2] - jc[2]) <= 0.18f;
                gfsm = (too_close_grip || too_close_cube) ? -1 : 0;
            }
            int n = 0;
            if (V == IGE_FRANKA_REACH || V == IGE_FRANKA_COLLECT) {
                for (int c = 0; c < T; ++c) for (int i = 0; i < 3; ++i) o[n++] = cube[c][i];
                for (int i = 0; i < 4; ++i) o[n++] = eef[a][3 + i];
                for (int i = 0; i < 3; ++i) o[n++] = eef[a][i];
                for (int i = 0; i < 3; ++i) o[n++] = min_rel[a][i];
                if (V == IGE_FRANKA_COLLECT) {
                    for (int i = 0; i < 3; ++i) o[n++] = base[a][i];
                    for (int i = 0; i < 4; ++i) o[n++] = base[a][3 + i];
                }
                for (int s = 1; s < A; ++s) for (int i = 0; i < 3; ++i) o[n++] = eef[(a + s) % A][i];
                if (V == IGE_FRANKA_COLLECT) { o[n++] = (float)fsm[a]; o[n++] = (float)fsm[a]; }
            } else {
                for (int s = 0; s < T; ++s) for (int i = 0; i < 4; ++i) o[n++] = cube[(a + s) % T][3 + i];
                for (int s = 0; s < T; ++s) for (int i = 0; i < 3; ++i) o[n++] = cube[(a + s) % T][i];
                const float *aux = NULL;
                if (V == IGE_FRANKA_COMBINE) {
                    aux = ma_actor(p, b, e, p->stack_base_actor);
                    for (int i = 0; i < 3; ++i) o[n++] = aux[i];
                } else {
                    for (int s = 0; s < T; ++s) {
                        const float *d = ma_actor(p, b, e, p->dest_actor_start + p->dest_actor_step * ((a + s) % T));
                        for (int i = 0; i < 3; ++i) o[n++] = d[i];
                    }
                    aux = ma_actor(p, b, e, p->dest_actor_start + p->dest_actor_step * a);
                }
                for (int i = 0; i < 3; ++i) o[n++] = cube[a][i] - eef[a][i];
                for (int i = 0; i < 3; ++i) o[n++] = aux[i] - cube[a][i];
                for (int i = 0; i < 3; ++i) o[n++] = base[a][i];
                for (int i = 0; i < 4; ++i) o[n++] = base[a][3 + i];
                for (int s = 0; s < A; ++s) for (int i = 0; i < 3; ++i) o[n++] = eef[(a + s) % A][i];
                for (int s = 0; s < A; ++s) for (int i = 0; i < 4; ++i) o[n++] = eef[(a + s) % A][3 + i];
                for (int s = 0; s < A; ++s) o[n++] = pow2_small(fsm[(a + s) % A]);
                o[n++] = pow2_small(gfsm);
                if (V == IGE_FRANKA_COMBINE) o[n++] = (a == A - 1) ? 1.0f : 0.0f; /* agent_idx :526-527 */
            }
            /* compute_franka_reward */
            float rewards;
            float c10[IGE_MA_NUM_COMPONENTS] = {0};
            const float alast = act[NA - 1];
            if (V == IGE_FRANKA_REACH) { /* franka_reach_MA.py:928-960 */
                const float d = md[a];
                const float dist_reward = 1.0f / (0.5f + d * d);
                float sq[8];
                for (int j = 0; j < NA; ++j) sq[j] = act[j] * act[j];
                const float actions_cost = row_sum(sq, NA) * 0.01f;
                rewards = dist_reward - actions_cost;
                int64_t distinct = 0;
                for (int c = 0; c < T; ++c) { int hit = 0; for (int q = 0; q < A; ++q) hit |= nearest[q] == c; distinct += hit; }
                const float all_touched = ((float)distinct / (float)A) * 1.0f;
                const float *cf = b->contact_forces + ((size_t)e * p->num_bodies + p->hand_body_start + p->hand_body_step * a) * 3;
                const float colliding = (norm3(cf[0], cf[1], cf[2]) >= 0.1f ? 1.0f : 0.0f) * -10.0f;
                rewards = rewards + (all_touched + colliding);
            } else {
                const float m = md[a];
                const int s = fsm[a];
                c10[0] = s == 0 ? expf(-5.0f * (m * m)) : 0.0f;
                c10[1] = s == 1 ? expf(-1.0f * alast) : 0.0f;
                if (V == IGE_FRANKA_COLLECT) { /* franka_collect_MA.py:1082-1174 */
                    const float *hc = cube[nearest[a]];
                    c10[2] = s == 2 ? (hc[2] - 1.05f) / (float)(0.3 + 0.025) : 0.0f;
                    const float dy = fabsf(hc[1] - (float)(0.3 + 0.05 * 2.));
                    c10[3] = s == 3 ? expf(-5.0f * (dy * dy)) : 0.0f;
                    const float dz = fabsf(hc[2] - (float)(0.3 / 2. + 1.05));
                    c10[4] = s == 4 ? expf(-5.0f * (dz * dz)) : 0.0f;
                    c10[5] = s == 5 ? expf(4.0f * alast) : 0.0f;
                    c10[6] = s == 6 ? 1.0f * 3.0f : 0.0f;
                    c10[7] = (float)s; /* BSR */
                    c10[8] = 0.0f;     /* gBSR */
                    rewards = c10[0] + c10[1] + c10[2] + c10[3] + c10[4] + c10[5] + c10[6] + c10[7] + c10[8] + 0.0f;
                } else if (V == IGE_FRANKA_COMBINE) { /* franka_combine_MA.py:1080-1166 */
                    const float *sb = ma_actor(p, b, e, p->stack_base_actor);
                    float lift = fabsf(sb[2] - cube[a][2]) / (float)((0.05 + 0.05) * 0.5);
                    lift = lift < 0.0f ? 0.0f : (lift > 1.0f ? 1.0f : lift);
                    c10[2] = s == 2 ? lift : 0.0f;
                    c10[7] = (float)s;
                    rewards = c10[0] + c10[1] + c10[7];
                } else { /* franka_PPMA.py:1082-1167 */
                    const float *dest = ma_actor(p, b, e, p->dest_actor_start + p->dest_actor_step * a);
                    const float rx = dest[0] - cube[a][0], ry = dest[1] - cube[a][1], rz = dest[2] - cube[a][2];
                    float lift = fabsf(rz) / (float)((0.05 + 0.05) * 0.5);
                    lift = lift < 0.0f ? 0.0f : (lift > 1.0f ? 1.0f : lift);
                    c10[2] = s == 2 ? lift : 0.0f;
                    const float dt = norm3(rx + 0.0f, ry + 0.0f, rz + 0.05f);
                    c10[3] = s == 3 ? expf(-3.0f * dt) : 0.0f;
                    c10[4] = s == 4 ? expf(-10.0f * dt) : 0.0f;
                    c10[5] = s == 5 ? expf(5.0f * alast) : 0.0f;
                    c10[6] = s == 6 ? tanhf(m) : 0.0f;
                    c10[7] = (float)s;
                    c10[8] = (float)gfsm * 0.5f;
                    c10[9] = 0.0f;
                    rewards = c10[0] + c10[1] + c10[2] + c10[3] + c10[4] + c10[5] + c10[6] + c10[7] + c10[8] + c10[9];
                }
            }
            rewards = rewards < 0.0f ? 0.0f : rewards; /* torch.clip(rewards, 0., None) */
            for (int i = 0; i < IGE_MA_NUM_COMPONENTS; ++i) comp[i] += c10[i];
            int64_t reset = b->reset_buf[row];
            const int64_t progress = b->progress_buf[row];
            const int late = (float)progress >= p->max_episode_length - 1.0f;
            if (late) reset = 1;
            b->rew_buf[row] = rewards;
            b->reset_buf[row] = reset;
            const int timeout = late && reset != 0;
            if (b->timeout_buf) b->timeout_buf[row] = (uint8_t)timeout;
            if (b->obs_clamped)
                for (int j = 0; j < NO; ++j) {
                    float v = o[j];
                    v = v < -p->clip_obs ? -p->clip_obs : v; v = v > p->clip_obs ? p->clip_obs : v;
                    b->obs_clamped[row * NO + j] = v;
                }
            s_rew += rewards; s_done += reset != 0; s_to += timeout; s_len += reset != 0 ? (double)progress : 0.0;
        }
    }
    if (b->reward_components)
        for (int i = 0; i < IGE_MA_NUM_COMPONENTS; ++i) b->reward_components[i] = (float)(comp[i] / ((double)N * A));
    add_stat(b->stats, IGE_STAT_STEPS, (double)N * A); add_stat(b->stats, IGE_STAT_REWARD, s_rew);
    add_stat(b->stats, IGE_STAT_DONE, s_done); add_stat(b->stats, IGE_STAT_TIMEOUT, s_to);
    add_stat(b->stats, IGE_STAT_DONE_LENGTH, s_len);
    for (int i = 0; i < IGE_MA_NUM_COMPONENTS; ++i) add_stat(b->stats, IGE_STAT_TASK0 + i, comp[i]);
    if (b->list.scratch && p->rand.step) *p->rand.step += 1;
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * pre_physics_step action scaling
 * ------------------------------------------------------------------------------------------ */
static float clamp_sym(float x, float c) { /* torch.clamp(x, -c, c), tasks/base/vec_task.py:373 */
    x = x < -c ? -c : x;
    return x > c ? c : x;
}

/* Ant.pre_physics_step tasks/ant.py:281-285 / Humanoid tasks/humanoid.py:281-285:
 * forces = self.actions * gears * self.power_scale (left to right), after the clamp of VecTask.step. */
int ige_oracle_locomotion_pre_physics(const IgeActionForceArgs *a) {
    const int D = a->num_dof;
    for (long long i = 0; i < (long long)a->num_envs * D; ++i) {
        const float x = clamp_sym(a->actions_in[i], a->clip_actions);
        a->actions[i] = x;
        const float t = x * a->gears[i % D];
        a->forces[i] = t * a->power_scale;
    }
    return 0;
}

/* ShadowHand.pre_physics_step tasks/shadow_hand.py:684-698 (targets part; utils/torch_jit_utils.py:229-235
 * scale / tensor_clamp). */
int ige_oracle_shadow_hand_pre_physics_targets(const IgeHandTargetArgs *a) {
    const int NA = a->num_actions, DPE = a->num_dofs_per_env;
    for (int e = 0; e < a->num_envs; ++e)
        for (int j = 0; j < NA; ++j) {
            const float x = clamp_sym(a->actions_in[(size_t)e * NA + j], a->clip_actions);
            a->actions[(size_t)e * NA + j] = x;
            const size_t t = (size_t)e * DPE + a->actuated[j];
            const float lo = a->lower[j], hi = a->upper[j], prev = a->prev_targets[t];
            float cur;
            if (a->use_relative_control) {
                const float d = a->speed_scale_dt * x; /* :686 */
                cur = prev + d;
            } else {
                const float s0 = 0.5f * (x + 1.0f);    /* scale(): 0.5 * (x + 1.0) * (upper - lower) + lower */
                const float s1 = s0 * (hi - lo);
                cur = s1 + lo;
                const float m0 = a->act_moving_average * cur, m1 = a->one_minus_mavg * prev; /* :692-693 */
                cur = m0 + m1;
            }
            cur = cur < hi ? cur : hi;                   /* torch.max(torch.min(t, upper), lower) */
            cur = cur > lo ? cur : lo;
            a->cur_targets[t] = cur;
            a->prev_targets[t] = cur;                    /* :697 */
        }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Agent-side bookkeeping of A2CAgentMA.play_steps, learning/A2CAgent_MA.py:35-56.
 * rl_games pieces (setup.py:23 "rl-games>=1.6.0", not vendored) restated from their published source:
 * DefaultRewardsShaper.__call__ = clamp((r + shift_value) * scale_value, min_val, max_val);
 * AverageMeter.update(values): new_mean = mean(values); size = clip(len, 0, max_size);
 * old_size = min(max_size - size, current_size); mean = (mean*old_size + new_mean*size) / (old_size+size).
 * ------------------------------------------------------------------------------------------ */
int ige_oracle_ma_rollout_step(const IgeRolloutStepArgs *a) {
    const int n = a->num_rows, A = a->num_agents;
    int rank = 0, kept = 0;
    double sum[3] = {0.0, 0.0, 0.0};
    for (int r = 0; r < n; ++r) {
        float s = a->rewards[r] + a->shaper_shift;          /* :34 self.rewards_shaper(rewards) */
        s = s * a->shaper_scale;
        s = s < a->shaper_min ? a->shaper_min : s;
        s = s > a->shaper_max ? a->shaper_max : s;
        if (a->value_bootstrap) {                            /* :35-36 */
            const float gv = a->gamma * a->values[r];
            s = s + gv * (float)a->time_outs[r];
        }
        a->mb_rewards[r] = s;                                /* :38 */
        const float cr = a->current_rewards[r] + a->rewards[r]; /* :40-42 */
        const float cs = a->current_shaped_rewards[r] + s;
        const float cl = a->current_lengths[r] + 1.0f;
        const int done = a->dones[r] != 0;
        if (done) {                                          /* :43-44 all_done_indices[::num_agents] */
            if (rank % A == 0) {
                a->done_rows[kept++] = r;
                sum[0] += cr, sum[1] += cs, sum[2] += cl;    /* :46-48 (values before the masking below) */
            }
            ++rank;
        }
        const float nd = 1.0f - (float)a->dones[r];          /* :51 */
        a->current_rewards[r] = cr * nd;                     /* :53-55 */
        a->current_shaped_rewards[r] = cs * nd;
        a->current_lengths[r] = cl * nd;
        if (a->dones_out) a->dones_out[r] = (uint8_t)a->dones[r];
    }
    *a->done_count = kept;
    if (kept > 0)
        for (int m = 0; m < 3; ++m) {                        /* AverageMeter.update */
            const float new_mean = (float)sum[m] / (float)kept;
            const int size = kept < a->meter_max_size ? kept : a->meter_max_size;
            int old_size = a->meter_max_size - size;
            if (a->meter_size[m] < old_size) old_size = a->meter_size[m];
            const int size_sum = old_size + size;
            a->meter_size[m] = size_sum;
            const float x = a->meter_mean[m] * (float)old_size, y = new_mean * (float)size;
            a->meter_mean[m] = (x + y) / (float)size_sum;
        }
    return 0;
}

/* rl_games A2CBase.discount_values as called from learning/A2CAgent_MA.py:60-66, and :67
 * mb_returns = mb_advs + mb_values. */
int ige_oracle_ma_discount_values(const IgeDiscountArgs *a) {
    const size_t n = (size_t)a->num_rows;
    for (size_t r = 0; r < n; ++r) {
        float lastgaelam = 0.0f, nextvalues = a->last_values[r], nnt = 1.0f - (float)a->dones[r];
        for (int t = a->horizon - 1; t >= 0; --t) {
            const size_t i = (size_t)t * n + r;
            const float v = a->mb_values[i];
            const float gv = a->gamma * nextvalues;
            const float t1 = gv * nnt;
            const float t2 = a->mb_rewards[i] + t1;
            const float delta = t2 - v;
            const float gl = a->gamma_tau * nnt;
            const float t3 = gl * lastgaelam;
            lastgaelam = delta + t3;
            if (a->mb_advs) a->mb_advs[i] = lastgaelam;
            a->mb_returns[i] = lastgaelam + v;
            nextvalues = v;
            nnt = 1.0f - (float)a->mb_dones[i];
        }
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * MA Franka pre_physics_step + _compute_osc_torques, tasks/franka_PPMA.py:897-956
 * (same text in franka_reach_MA.py:770-818, franka_collect_MA.py:892-951, franka_combine_MA.py:910-969).
 * torch.inverse is LAPACK LU; mm and J mm^-1 J^T are SPD, so Gauss-Jordan without pivoting is used here and
 * in the kernel (agreement with the reference: rounding x condition number, see tests/test_oracle_golden.py).
 * Products are re-associated where mathematically identical: j_eef_inv = m_eef @ (J @ mm_inv),
 * J^T @ (m_eef @ v) instead of (J^T @ m_eef) @ v.
 * ------------------------------------------------------------------------------------------ */
static void gj_inverse(float *A, int n) { /* in place, row-major n x n */
    for (int k = 0; k < n; ++k) {
        const float p = 1.0f / A[k * n + k];
        A[k * n + k] = 1.0f;
        for (int j = 0; j < n; ++j) A[k * n + j] = A[k * n + j] * p;
        for (int i = 0; i < n; ++i) {
            if (i == k) continue;
            const float f = A[i * n + k];
            A[i * n + k] = 0.0f;
            for (int j = 0; j < n; ++j) A[i * n + j] = fmaf(-f, A[k * n + j], A[i * n + j]);
        }
    }
}

int ige_oracle_franka_ma_pre_physics(const IgeFrankaOscArgs *a) {
    const int A = a->num_agents, NA = a->num_actions;
    const float PI = 3.14159265358979323846f, TWO_PI = 6.28318530717958647692f;
    for (int e = 0; e < a->num_envs; ++e)
        for (int g = 0; g < A; ++g) {
            const size_t r = (size_t)e * A + g;
            float act[8];
            for (int i = 0; i < NA; ++i) { /* vec_task.py:373 + :933 self.actions = actions.clone() */
                float x = a->actions_in[r * NA + i];
                x = x < -a->clip_actions ? -a->clip_actions : x;
                x = x > a->clip_actions ? a->clip_actions : x;
                act[i] = x;
                a->actions[r * NA + i] = x;
            }
            const float *ds = a->dof_state + r * 18;
            const float *ev = a->rigid_body_state + ((size_t)e * a->num_bodies + a->eef_body_start + (size_t)g * a->eef_body_step) * 13 + 7;
            float M[49], J[42], JM[42], E[36], v[6], un[7], w[7], t[6], y[6], z[6], u[7];
            for (int i = 0; i < 49; ++i) M[i] = a->mm[r * 49 + i];
            for (int i = 0; i < 42; ++i) J[i] = a->j_eef[r * 42 + i];
            for (int i = 0; i < 6; ++i) {          /* :939, :913 kp * dpose - kd * eef_vel */
                const float dpose = (act[i] * a->cmd_limit[i]) / a->action_scale;
                const float x0 = a->kp[i] * dpose, x1 = a->kd[i] * ev[i];
                v[i] = x0 - x1;
            }
            for (int k = 0; k < 7; ++k) {          /* :920 */
                const float q = ds[2 * k], qd = ds[2 * k + 1];
                const float d0 = (a->default_dof_pos[k] - q) + PI;
                const float wrap = py_remainder(d0, TWO_PI) - PI;
                const float x0 = a->kd_null[k] * (-qd), x1 = a->kp_null[k] * wrap;
                un[k] = x0 + x1;
            }
            for (int i = 0; i < 7; ++i) {          /* :922 u_null = mm @ u_null */
                float s = 0.0f;
                for (int k = 0; k < 7; ++k) s = fmaf(M[i * 7 + k], un[k], s);
                w[i] = s;
            }
            gj_inverse(M, 7);                      /* :907 */
            for (int i = 0; i < 6; ++i)            /* J @ mm_inv */
                for (int j = 0; j < 7; ++j) {
                    float s = 0.0f;
                    for (int k = 0; k < 7; ++k) s = fmaf(J[i * 7 + k], M[k * 7 + j], s);
                    JM[i * 7 + j] = s;
                }
            for (int i = 0; i < 6; ++i)            /* :908 m_eef_inv */
                for (int j = 0; j < 6; ++j) {
                    float s = 0.0f;
                    for (int k = 0; k < 7; ++k) s = fmaf(JM[i * 7 + k], J[j * 7 + k], s);
                    E[i * 6 + j] = s;
                }
            gj_inverse(E, 6);                      /* :909 m_eef */
            for (int i = 0; i < 6; ++i) {
                float s = 0.0f, s2 = 0.0f;
                for (int k = 0; k < 6; ++k) s = fmaf(E[i * 6 + k], v[k], s);
                t[i] = s;
                for (int k = 0; k < 7; ++k) s2 = fmaf(JM[i * 7 + k], w[k], s2);
                y[i] = s2;
            }
            for (int i = 0; i < 6; ++i) {
                float s = 0.0f;
                for (int k = 0; k < 6; ++k) s = fmaf(E[i * 6 + k], y[k], s);
                z[i] = s;
            }
            for (int k = 0; k < 7; ++k) {          /* :913 + :923, :926 clamp */
                float s = 0.0f, s2 = 0.0f;
                for (int i = 0; i < 6; ++i) s = fmaf(J[i * 7 + k], t[i], s);
                for (int i = 0; i < 6; ++i) s2 = fmaf(J[i * 7 + k], z[i], s2);
                float x = s + (w[k] - s2);
                const float lim = a->effort_limits[k];
                x = x < -lim ? -lim : x;
                u[k] = x > lim ? lim : x;
            }
            for (int k = 0; k < 6; ++k) a->effort_control[r * 9 + k] = u[k]; /* :944 _arm_control[:, :6] = u_arm[:, :6] */
            if (a->has_gripper) {                  /* :946-953 */
                const int open = act[6] >= 0.0f;
                a->pos_control[r * 9 + 7] = open ? a->gripper_upper[0] : a->gripper_lower[0];
                a->pos_control[r * 9 + 8] = open ? a->gripper_upper[1] : a->gripper_lower[1];
            }
        }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Anymal: tasks/anymal.py:231-240 post_physics_step, :277-304 reset_idx,
 * :311-351 compute_anymal_reward, :355-386 compute_anymal_observations; epilogue vec_task.py:396-404.
 * ------------------------------------------------------------------------------------------ */
static float anymal_uniform(const IgeAnymalParams *p, int draw, int rank, int env, int cols, int j) {
    if (p->rand_mode == IGE_RAND_PHILOX) {
        IgeResetRand r;
        memset(&r, 0, sizeof(r));
        r.seed = p->seed; r.step = p->step;
        return philox_uniform(&r, (uint32_t)env, (uint32_t)draw, (uint32_t)j);
    }
    return p->u[draw][(size_t)rank * cols + j];
}
/* torch_rand_float (utils/torch_jit_utils.py:216-218): (upper - lower) * rand + lower, scalars rounded to fp32 */
static inline float rand_float(double lower, double upper, float u) {
    const float t = (float)(upper - lower) * u;
    return t + (float)lower;
}

int ige_oracle_anymal_post_physics(const IgeAnymalParams *p, const IgeAnymalBuffers *b) {
    const int n = p->num_envs;
    int rank = 0;
    for (int e = 0; e < n; ++e) {
        float *ds = b->dof_state + (size_t)e * 24;
        float *cmd = b->commands + (size_t)e * 3;
        int64_t progress = b->progress_buf[e] + 1;             /* :232 */
        if (b->reset_buf[e] != 0) {                            /* :234-236 reset_idx */
            for (int j = 0; j < 12; ++j) {
                const float off = rand_float(0.5, 1.5, anymal_uniform(p, 0, rank, e, 12, j));   /* :282 */
                const float vel = rand_float(-0.1, 0.1, anymal_uniform(p, 1, rank, e, 12, j));  /* :283 */
                ds[2 * j] = p->default_dof_pos[j] * off;       /* :285 */
                ds[2 * j + 1] = vel;                           /* :286 */
            }
            cmd[0] = rand_float(p->command_x_range[0], p->command_x_range[1], anymal_uniform(p, 2, rank, e, 1, 0));
            cmd[1] = rand_float(p->command_y_range[0], p->command_y_range[1], anymal_uniform(p, 3, rank, e, 1, 0));
            cmd[2] = rand_float(p->command_yaw_range[0], p->command_yaw_range[1], anymal_uniform(p, 4, rank, e, 1, 0));
            progress = 0;                                      /* :303 */
            b->reset_ids[rank++] = e;
        }
        const float *root = b->root_states + (size_t)e * 13;
        const float *q = root + 3;
        const float grav[3] = {0.0f, 0.0f, -1.0f};
        float blv[3], bav[3], pg[3];
        quat_rotate_impl(q, root + 7, 1, blv);                 /* :369-370 / :328-329 */
        quat_rotate_impl(q, root + 10, 1, bav);
        quat_rotate_impl(q, grav, 0, pg);                      /* :371 */
        float *obs = b->obs_buf + (size_t)e * 48;
        const float *act = b->actions + (size_t)e * 12;
        for (int i = 0; i < 3; ++i) {
            obs[i] = blv[i] * p->lin_vel_scale;
            obs[3 + i] = bav[i] * p->ang_vel_scale;
            obs[6 + i] = pg[i];
            obs[9 + i] = cmd[i] * (i < 2 ? p->lin_vel_scale : p->ang_vel_scale);   /* :374 */
        }
        for (int j = 0; j < 12; ++j) {
            obs[12 + j] = (ds[2 * j] - p->default_dof_pos[j]) * p->dof_pos_scale;  /* :372 */
            obs[24 + j] = ds[2 * j + 1] * p->dof_vel_scale;
            obs[36 + j] = act[j];
        }
        /* compute_anymal_reward :331-351 */
        const float d0 = cmd[0] - blv[0], d1 = cmd[1] - blv[1], d2 = cmd[2] - bav[2];
        const float sq[2] = {d0 * d0, d1 * d1};
        const float lin_err = row_sum(sq, 2);
        const float ang_err = d2 * d2;
        const float r_lin = expf(-lin_err / 0.25f) * p->rew_lin_vel_xy;
        const float r_ang = expf(-ang_err / 0.25f) * p->rew_ang_vel_z;
        float tq2[12];
        for (int j = 0; j < 12; ++j) { const float t = b->torques[(size_t)e * 12 + j]; tq2[j] = t * t; }
        const float r_tq = row_sum(tq2, 12) * p->rew_torque;
        float total = (r_lin + r_ang) + r_tq;
        total = total < 0.0f ? 0.0f : total;                   /* torch.clip(total, 0., None) */
        const float *cf = b->contact_forces + (size_t)e * p->num_bodies * 3;
        int reset = norm3(cf[p->base_index * 3], cf[p->base_index * 3 + 1], cf[p->base_index * 3 + 2]) > 1.0f;
        for (int k = 0; k < p->num_knees; ++k) {
            const float *c = cf + p->knee_indices[k] * 3;
            reset |= norm3(c[0], c[1], c[2]) > 1.0f;
        }
        const int time_out = (float)progress >= p->max_episode_length - 1.0f;     /* :348 */
        reset |= time_out;
        b->rew_buf[e] = total;
        b->reset_buf[e] = reset;
        b->progress_buf[e] = progress;
        if (b->timeout_buf) b->timeout_buf[e] = (uint8_t)(time_out && reset);     /* vec_task.py:396 */
        if (b->obs_clamped)
            for (int j = 0; j < 48; ++j) {
                float v = obs[j];
                v = v < -p->clip_obs ? -p->clip_obs : v;
                v = v > p->clip_obs ? p->clip_obs : v;
                b->obs_clamped[(size_t)e * 48 + j] = v;
            }
    }
    *b->reset_count = rank;
    if (p->rand_mode == IGE_RAND_PHILOX && p->step) *p->step += 1;
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Remaining single-agent tasks: BallBalance, Quadcopter, Ingenuity.
 * These reset in PRE-physics position, so each has a pre_physics and a post_physics restatement.
 * ------------------------------------------------------------------------------------------ */
static float task_uniform(const IgeTaskRand *r, const int32_t *rank, int draw, int env, int cols, int j) {
    if (r->mode == IGE_RAND_PHILOX) {
        IgeResetRand rr;
        memset(&rr, 0, sizeof(rr));
        rr.seed = r->seed; rr.step = r->step;
        return philox_uniform(&rr, (uint32_t)env, (uint32_t)draw, (uint32_t)j);
    }
    return r->u[draw][(size_t)rank[env] * cols + j];
}
static void task_list(const IgeResetList *l, const IgeTaskRand *r, const int64_t *reset_buf, int n) {
    emit_reset_list(l, reset_buf, n, 1);
}
static void task_advance(const IgeResetList *l, const IgeTaskRand *r) {
    if (l->scratch && r->step) *r->step += 1;
}
static inline float tensor_clamp1(float t, float lo, float hi) { return fmaxf(fminf(t, hi), lo); } /* utils/torch_jit_utils.py:228-230 */
static inline float clamp_obs1(float x, float c) { return fminf(fmaxf(x, -c), c); }

/* tasks/ball_balance.py:410-430 pre_physics_step (+ reset_idx :360-408) */
int ige_oracle_ball_balance_pre_physics(const IgeBallBalanceArgs *a) {
    const int n = a->num_envs;
    task_list(&a->list, &a->rand, a->reset_buf, n);
    for (int e = 0; e < n; ++e) {
        const int flagged = a->reset_buf[e] != 0;
        float *tg = a->dof_position_targets + (size_t)e * 6;
        if (flagged) {
            float *root = a->root_states + (size_t)e * 26, *ball = root + 13;
            memcpy(root, a->initial_root_states + (size_t)e * 26, 26 * sizeof(float));              /* :364 */
            const float dists = rand_float(0.001, 0.5, task_uniform(&a->rand, a->rand.rank, 0, e, 1, 0));   /* :373 */
            const float angle = rand_float(-M_PI, M_PI, task_uniform(&a->rand, a->rand.rank, 1, e, 1, 0)); /* :374, torch_jit_utils.py:222-225 */
            const float dc = cosf(angle), ds = sinf(angle);
            const float hx = dists * dc, hy = dists * ds;                                            /* :375 */
            const float speedscales = (dists - 0.001f) / (float)(0.5 - 0.001);                       /* :377 */
            const float hspeeds = rand_float(0.0, 5.0, task_uniform(&a->rand, a->rand.rank, 2, e, 1, 0));   /* :378 */
            const float hs = -speedscales * hspeeds;                                                 /* :379 */
            const float vspeed = -rand_float(5.0, 5.0, task_uniform(&a->rand, a->rand.rank, 3, e, 1, 0));   /* :380 */
            const float height = rand_float(1.0, 2.0, task_uniform(&a->rand, a->rand.rank, 4, e, 1, 0));    /* :383 */
            ball[0] = hx; ball[1] = hy; ball[2] = height;
            ball[3] = 0.f; ball[4] = 0.f; ball[5] = 0.f; ball[6] = 1.f;                              /* :385-386 */
            ball[7] = hs * dc; ball[8] = hs * ds; ball[9] = vspeed;
            ball[10] = ball[11] = ball[12] = 0.f;                                                    /* :390 */
            memcpy(a->dof_states + (size_t)e * 12, a->initial_dof_states + (size_t)e * 12, 12 * sizeof(float)); /* :398 */
            a->reset_buf[e] = 0;
            a->progress_buf[e] = 0;
        }
        for (int k = 0; k < 3; ++k) tg[2 * k + 1] += a->target_step * a->actions[(size_t)e * 3 + k];  /* :421 */
        for (int j = 0; j < 6; ++j) {
            tg[j] = tensor_clamp1(tg[j], a->dof_lower[j], a->dof_upper[j]);                          /* :422 */
            if (flagged) tg[j] = 0.0f;                                                               /* :425 */
        }
    }
    task_advance(&a->list, &a->rand);
    return 0;
}

/* tasks/ball_balance.py:432-441 post_physics_step; compute_observations :336-349; compute_bbot_reward :460-476 */
int ige_oracle_ball_balance_post_physics(const IgeBallBalanceArgs *a) {
    for (int e = 0; e < a->num_envs; ++e) {
        const int64_t progress = a->progress_buf[e] + 1;
        const float *dof = a->dof_states + (size_t)e * 12, *ball = a->root_states + (size_t)e * 26 + 13;
        const float *sens = a->sensor_tensor + (size_t)e * 18;
        float *obs = a->obs_buf + (size_t)e * 24;
        for (int k = 0; k < 3; ++k) {
            obs[k] = dof[2 * (2 * k + 1)];
            obs[3 + k] = dof[2 * (2 * k + 1) + 1];
            obs[6 + k] = ball[k];
            obs[9 + k] = ball[7 + k];
            obs[12 + k] = sens[6 * k] / 20.0f;
            obs[15 + k] = sens[6 * k + 3] / 20.0f;
            obs[18 + k] = sens[6 * k + 4] / 20.0f;
            obs[21 + k] = sens[6 * k + 5] / 20.0f;
        }
        if (a->obs_clamped)
            for (int i = 0; i < 24; ++i) a->obs_clamped[(size_t)e * 24 + i] = clamp_obs1(obs[i], a->clip_obs);
        const float zz = ball[2] - 0.7f;
        const float ball_dist = sqrtf((ball[0] * ball[0] + zz * zz) + ball[1] * ball[1]);
        const float ball_speed = sqrtf((ball[7] * ball[7] + ball[8] * ball[8]) + ball[9] * ball[9]);
        const float pos_reward = 1.0f / (1.0f + ball_dist), speed_reward = 1.0f / (1.0f + ball_speed);
        int64_t reset = a->reset_buf[e];
        const int late = (float)progress >= a->max_episode_length - 1.0f;
        if (late) reset = 1;
        if (ball[2] < (float)((double)a->ball_radius * 1.5)) reset = 1;
        a->rew_buf[e] = pos_reward * speed_reward;
        a->reset_buf[e] = reset;
        a->progress_buf[e] = progress;
        if (a->timeout_buf) a->timeout_buf[e] = (late && reset != 0) ? 1 : 0;                       /* vec_task.py:396 */
    }
    return 0;
}

/* tasks/quadcopter.py:301-331 pre_physics_step (+ reset_idx :280-299) */
int ige_oracle_quadcopter_pre_physics(const IgeQuadcopterArgs *a) {
    const int n = a->num_envs, nb = a->num_bodies;
    task_list(&a->list, &a->rand, a->reset_buf, n);
    for (int e = 0; e < n; ++e) {
        const int flagged = a->reset_buf[e] != 0;
        float *tg = a->dof_position_targets + (size_t)e * 8, *th = a->thrusts + (size_t)e * 4;
        float *f = a->forces + (size_t)e * nb * 3, *dof = a->dof_states + (size_t)e * 16;
        const float *act = a->actions + (size_t)e * 12;
        if (flagged) {
            float *root = a->root_states + (size_t)e * 13;
            memcpy(root, a->initial_root_states + (size_t)e * 13, 13 * sizeof(float));                      /* :288 */
            root[0] += rand_float(-1.5, 1.5, task_uniform(&a->rand, a->rand.rank, 0, e, 1, 0));              /* :289 */
            root[1] += rand_float(-1.5, 1.5, task_uniform(&a->rand, a->rand.rank, 1, e, 1, 0));
            root[2] += rand_float(-0.2, 1.5, task_uniform(&a->rand, a->rand.rank, 2, e, 1, 0));
            for (int j = 0; j < 8; ++j) {
                dof[2 * j] = rand_float(-0.2, 0.2, task_uniform(&a->rand, a->rand.rank, 3, e, 8, j));        /* :294 */
                dof[2 * j + 1] = 0.0f;                                                                       /* :295 */
            }
            a->reset_buf[e] = 0;
            a->progress_buf[e] = 0;
        }
        for (int j = 0; j < 8; ++j) {
            tg[j] += a->dof_target_step * act[j];                                                            /* :310 */
            tg[j] = tensor_clamp1(tg[j], a->dof_lower[j], a->dof_upper[j]);                                  /* :311 */
        }
        for (int k = 0; k < 4; ++k) {
            th[k] += a->thrust_step * act[8 + k];                                                            /* :314 */
            th[k] = tensor_clamp1(th[k], a->thrust_lower[k], a->thrust_upper[k]);                            /* :315 */
            f[a->rotor_bodies[k] * 3 + 2] = th[k];                                                           /* :317-320 */
        }
        if (flagged) {                                                                                       /* :323-325 */
            for (int k = 0; k < 4; ++k) th[k] = 0.0f;
            for (int i = 0; i < nb * 3; ++i) f[i] = 0.0f;
            for (int j = 0; j < 8; ++j) tg[j] = dof[2 * j];
        }
    }
    task_advance(&a->list, &a->rand);
    return 0;
}

/* reward terms shared by compute_quadcopter_reward (quadcopter.py:391-425) and compute_ingenuity_reward
 * (ingenuity.py:420-455); quat_axis(q, 2) = quat_rotate(q, z) (utils/torch_jit_utils.py:280-284) */
static float copter_reward(const float *root, float target_dist, float up_scale) {
    const float pos_reward = 1.0f / (1.0f + target_dist * target_dist);
    const float z[3] = {0.0f, 0.0f, 1.0f};
    float ups[3];
    quat_rotate_impl(root + 3, z, 0, ups);
    const float tiltage = fabsf(1.0f - ups[2]);
    const float up_reward = up_scale / (1.0f + tiltage * tiltage);
    const float spinnage = fabsf(root[12]);
    const float spinnage_reward = 1.0f / (1.0f + spinnage * spinnage);
    return pos_reward + pos_reward * (up_reward + spinnage_reward);
}

/* tasks/quadcopter.py:333-341 post_physics_step; compute_observations :359-370 */
int ige_oracle_quadcopter_post_physics(const IgeQuadcopterArgs *a) {
    const float pi_f = 3.14159274101257324219f;
    for (int e = 0; e < a->num_envs; ++e) {
        const int64_t progress = a->progress_buf[e] + 1;
        const float *root = a->root_states + (size_t)e * 13, *dof = a->dof_states + (size_t)e * 16;
        float *obs = a->obs_buf + (size_t)e * 21;
        obs[0] = (0.0f - root[0]) / 3.0f;
        obs[1] = (0.0f - root[1]) / 3.0f;
        obs[2] = (1.0f - root[2]) / 3.0f;
        for (int i = 0; i < 4; ++i) obs[3 + i] = root[3 + i];
        for (int i = 0; i < 3; ++i) { obs[7 + i] = root[7 + i] / 2.0f; obs[10 + i] = root[10 + i] / pi_f; }
        for (int j = 0; j < 8; ++j) obs[13 + j] = dof[2 * j];
        if (a->obs_clamped)
            for (int i = 0; i < 21; ++i) a->obs_clamped[(size_t)e * 21 + i] = clamp_obs1(obs[i], a->clip_obs);
        const float dz = 1.0f - root[2];
        const float target_dist = sqrtf((root[0] * root[0] + root[1] * root[1]) + dz * dz);
        const int die = target_dist > 3.0f || root[2] < 0.3f;
        const int late = (float)progress >= a->max_episode_length - 1.0f;
        const int64_t reset = (late || die) ? 1 : 0;
        a->rew_buf[e] = copter_reward(root, target_dist, 1.0f);
        a->reset_buf[e] = reset;
        a->progress_buf[e] = progress;
        if (a->timeout_buf) a->timeout_buf[e] = (late && reset != 0) ? 1 : 0;
    }
    return 0;
}

/* tasks/ingenuity.py:322-364 pre_physics_step (set_targets :283-294, reset_idx :296-320) */
static void ingenuity_set_target(const IgeIngenuityArgs *a, int e, float ux, float uy, float uz) {
    float *t = a->target_root_positions + (size_t)e * 3, *marker = a->root_tensor + (size_t)e * 26 + 13;
    t[0] = ux * 10.0f - 5.0f;                                                                                /* :286 */
    t[1] = uy * 10.0f - 5.0f;
    t[2] = uz + 1.0f;                                                                                        /* :287 */
    marker[0] = t[0]; marker[1] = t[1]; marker[2] = t[2] + 0.4f;                                             /* :288-290 */
}
int ige_oracle_ingenuity_pre_physics(const IgeIngenuityArgs *a) {
    const int n = a->num_envs, nb = a->num_bodies;
    task_list(&a->list, &a->rand, a->reset_buf, n);
    int any = 0;
    for (int e = 0; e < n; ++e) any |= a->reset_buf[e] != 0;
    if (any)                                                                                                 /* :299-300 */
        for (int e = 0; e < n; ++e) {
            a->dof_states[(size_t)e * 8 + 3] = -50.0f;
            a->dof_states[(size_t)e * 8 + 7] = 50.0f;
        }
    for (int e = 0; e < n; ++e) {
        const int flagged = a->reset_buf[e] != 0;
        const int retarget = a->progress_buf[e] % 500 == 0;                                                  /* :325 */
        if (retarget)
            ingenuity_set_target(a, e, task_uniform(&a->rand, a->target_rank, 0, e, 2, 0),
                                 task_uniform(&a->rand, a->target_rank, 0, e, 2, 1),
                                 task_uniform(&a->rand, a->target_rank, 1, e, 1, 0));
        if (flagged) {
            float *root = a->root_tensor + (size_t)e * 26;
            ingenuity_set_target(a, e, task_uniform(&a->rand, a->rand.rank, 2, e, 2, 0),
                                 task_uniform(&a->rand, a->rand.rank, 2, e, 2, 1),
                                 task_uniform(&a->rand, a->rand.rank, 3, e, 1, 0));                          /* :304 */
            memcpy(root, a->initial_root_states + (size_t)e * 13, 13 * sizeof(float));                       /* :308 */
            root[0] += rand_float(-1.5, 1.5, task_uniform(&a->rand, a->rand.rank, 4, e, 1, 0));
            root[1] += rand_float(-1.5, 1.5, task_uniform(&a->rand, a->rand.rank, 5, e, 1, 0));
            root[2] += rand_float(-0.2, 1.5, task_uniform(&a->rand, a->rand.rank, 6, e, 1, 0));
            a->reset_buf[e] = 0;
            a->progress_buf[e] = 0;
        }
        if (a->actor_dirty) {                                                                                /* :334, :341 */
            a->actor_dirty[(size_t)e * 2] = flagged ? 1 : 0;
            a->actor_dirty[(size_t)e * 2 + 1] = (flagged || retarget) ? 1 : 0;
        }
        const float *act = a->actions + (size_t)e * 6;
        float *th = a->thrusts + (size_t)e * 6, *f = a->forces + (size_t)e * nb * 3;
        for (int r = 0; r < 2; ++r) {                                                                        /* :345-355 */
            const float vert = fminf(fmaxf(act[3 * r + 2] * 2000.0f, -a->thrust_upper_limit), a->thrust_upper_limit);
            th[3 * r + 2] = a->dt * vert;
            th[3 * r] = th[3 * r + 2] * fminf(fmaxf(act[3 * r], -a->thrust_lateral_component), a->thrust_lateral_component);
            th[3 * r + 1] = th[3 * r + 2] * fminf(fmaxf(act[3 * r + 1], -a->thrust_lateral_component), a->thrust_lateral_component);
        }
        for (int i = 0; i < 3; ++i) { f[3 + i] = th[i]; f[9 + i] = th[3 + i]; }                              /* :357-358 */
        if (flagged) {                                                                                       /* :361-362 */
            for (int i = 0; i < 6; ++i) th[i] = 0.0f;
            for (int i = 0; i < nb * 3; ++i) f[i] = 0.0f;
        }
    }
    task_advance(&a->list, &a->rand);
    return 0;
}

/* tasks/ingenuity.py:366-374 post_physics_step; compute_observations :396-401 */
int ige_oracle_ingenuity_post_physics(const IgeIngenuityArgs *a) {
    const float pi_f = 3.14159274101257324219f;
    for (int e = 0; e < a->num_envs; ++e) {
        const int64_t progress = a->progress_buf[e] + 1;
        const float *root = a->root_tensor + (size_t)e * 26, *tgt = a->target_root_positions + (size_t)e * 3;
        float *obs = a->obs_buf + (size_t)e * 13;
        for (int i = 0; i < 3; ++i) obs[i] = (tgt[i] - root[i]) / 3.0f;
        for (int i = 0; i < 4; ++i) obs[3 + i] = root[3 + i];
        for (int i = 0; i < 3; ++i) { obs[7 + i] = root[7 + i] / 2.0f; obs[10 + i] = root[10 + i] / pi_f; }
        if (a->obs_clamped)
            for (int i = 0; i < 13; ++i) a->obs_clamped[(size_t)e * 13 + i] = clamp_obs1(obs[i], a->clip_obs);
        const float dx = tgt[0] - root[0], dy = tgt[1] - root[1], dz = tgt[2] - root[2];
        const float target_dist = sqrtf((dx * dx + dy * dy) + dz * dz);                                      /* :424 sum(-1) */
        const int die = target_dist > 8.0f || root[2] < 0.5f;
        const int late = (float)progress >= a->max_episode_length - 1.0f;
        const int64_t reset = (late || die) ? 1 : 0;
        a->rew_buf[e] = copter_reward(root, target_dist, 5.0f);
        a->reset_buf[e] = reset;
        a->progress_buf[e] = progress;
        if (a->timeout_buf) a->timeout_buf[e] = (late && reset != 0) ? 1 : 0;
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Single-agent Franka cube tasks: FrankaReach, FrankaCubeStack2, FrankaCubeStack.
 * Reference: tasks/franka_reach.py, tasks/franka_cube_stack2.py, tasks/franka_cube_stack.py.
 * ------------------------------------------------------------------------------------------ */
static float sa_uniform(const IgeFrankaSAArgs *a, int which, int e, int cols, int j) {
    return task_uniform(&a->rand, a->rand.rank, a->draw_slot[which], e, cols, j);
}
/* axisangle2quat((0, 0, az)) followed by quat_mul(., (0,0,0,1)): franka_reach.py:18-49, :626-629 */
static void sa_yaw_quat(float az, float *q) {
    const float angle = norm3(0.0f, 0.0f, az);
    float aa[4] = {0.0f, 0.0f, 0.0f, 1.0f};
    if (angle > 1e-6f) {
        const float s = sinf(angle / 2.0f);
        aa[0] = 0.0f * s / angle; aa[1] = 0.0f * s / angle; aa[2] = az * s / angle;
        aa[3] = cosf(angle / 2.0f);
    }
    const float ident[4] = {0.0f, 0.0f, 0.0f, 1.0f};
    quat_mul(aa, ident, q);
}
/* _reset_init_cube_state for one env (franka_reach.py:538-632, franka_cube_stack2.py:557-647,
 * franka_cube_stack.py:516-599); which: 0 = cubeB, 1 = cubeA.  other_xy: the other cube's fresh init xy (Stack). */
static void sa_sample_cube(const IgeFrankaSAArgs *a, int e, int which, const float *other_xy, float *row) {
    const int base = which ? 3 : 0;
    memset(row, 0, 13 * sizeof(float));
    row[2] = which ? a->cubeA_z_base : a->cubeB_z_base;
    if (a->cube_z_range != 0.0f) row[2] = row[2] + sa_uniform(a, base + 0, e, 1, 0) * a->cube_z_range;
    row[6] = 1.0f;
    float ux = sa_uniform(a, base + 1, e, 2, 0), uy = sa_uniform(a, base + 1, e, 2, 1);
    row[0] = a->cube_xy_center[0] + a->cube_xy_mult * (ux - 0.5f);
    row[1] = a->cube_xy_center[1] + a->cube_xy_mult * (uy - 0.5f);
    if (which && a->variant == IGE_FRANKA_SA_STACK && a->rand.mode == IGE_RAND_PHILOX) {
        /* check_valid: re-sample while the cubes would touch (:567-585); INJECTED mode is handed the accepted sample */
        for (int t = 1; t < 100 && norm2(row[0] - other_xy[0], row[1] - other_xy[1]) < a->min_dist; ++t) {
            IgeTaskRand r = a->rand;
            ux = task_uniform(&r, r.rank, 8 + t, e, 2, 0); uy = task_uniform(&r, r.rank, 8 + t, e, 2, 1);
            row[0] = a->cube_xy_center[0] + a->cube_xy_mult * (ux - 0.5f);
            row[1] = a->cube_xy_center[1] + a->cube_xy_mult * (uy - 0.5f);
        }
    }
    if (a->cube_yaw_mult > 0.0f) sa_yaw_quat(a->cube_yaw_mult * (sa_uniform(a, base + 2, e, 1, 0) - 0.5f), row + 3);
}

int ige_oracle_franka_sa_post_physics(const IgeFrankaSAArgs *a) {
    const int n = a->num_envs, V = a->variant;
    const int nobs = V == IGE_FRANKA_SA_REACH ? 14 : (V == IGE_FRANKA_SA_STACK2 ? 21 : 19);
    double comp[IGE_FRANKA_SA_COMPONENTS];
    memset(comp, 0, sizeof(comp));
    task_list(&a->list, &a->rand, a->reset_buf, n);
    for (int e = 0; e < n; ++e) {
        int64_t progress = a->progress_buf[e] + 1;
        int64_t reset = a->reset_buf[e];
        float *cubeA = a->root_state + ((size_t)e * a->num_actors + a->cubeA_actor) * 13;
        float *cubeB = a->root_state + ((size_t)e * a->num_actors + a->cubeB_actor) * 13;
        float *dof = a->dof_state + (size_t)e * 18;
        if (reset != 0) { /* reset_idx */
            float row[13], bxy[2] = {0.0f, 0.0f};
            if (V != IGE_FRANKA_SA_REACH) { /* cube B first, then A */
                sa_sample_cube(a, e, 0, bxy, row);
                memcpy(a->init_cubeB_state + (size_t)e * 13, row, sizeof(row));
                memcpy(cubeB, row, sizeof(row));
                bxy[0] = row[0]; bxy[1] = row[1];
            }
            sa_sample_cube(a, e, 1, bxy, row);
            memcpy(a->init_cubeA_state + (size_t)e * 13, row, sizeof(row));
            memcpy(cubeA, row, sizeof(row));
            for (int j = 0; j < 9; ++j) {
                float v = a->default_dof_pos[j] + a->dof_noise_mult * (sa_uniform(a, 6, e, 9, j) - 0.5f);
                v = tensor_clamp1(v, a->dof_lower[j], a->dof_upper[j]);
                if (j >= 7) v = a->default_dof_pos[j];
                dof[2 * j] = v; dof[2 * j + 1] = 0.0f;
                a->pos_control[(size_t)e * 9 + j] = v;
                a->effort_control[(size_t)e * 9 + j] = 0.0f;
            }
            progress = 0; reset = 0;
        }
        /* _update_states */
        const float *rb = a->rigid_body_state + (size_t)e * a->num_bodies * 13;
        const float *eef = rb + (size_t)a->eef_body * 13;
        const float rel[3] = {cubeA[0] - eef[0], cubeA[1] - eef[1], cubeA[2] - eef[2]};
        const float a2b[3] = {cubeB[0] - cubeA[0], cubeB[1] - cubeA[1], cubeB[2] - cubeA[2]};
        const float height = cubeA[2] - a->table_height;
        const float qg[2] = {dof[14], dof[16]};
        const float agrip = a->actions[(size_t)e * 7 + 6];
        const float d = norm3(rel[0], rel[1], rel[2]);
        float *obs = a->obs_buf + (size_t)e * nobs;
        float reward;
        if (V == IGE_FRANKA_SA_STACK) {
            const float *lf = rb + (size_t)a->lf_body * 13, *rf = rb + (size_t)a->rf_body * 13;
            int k = 0;
            for (int i = 0; i < 4; ++i) obs[k++] = cubeA[3 + i];
            for (int i = 0; i < 3; ++i) obs[k++] = cubeA[i];
            for (int i = 0; i < 3; ++i) obs[k++] = a2b[i];
            for (int i = 0; i < 3; ++i) obs[k++] = eef[i];
            for (int i = 0; i < 4; ++i) obs[k++] = eef[3 + i];
            obs[k++] = qg[0]; obs[k++] = qg[1];
            /* compute_franka_reward franka_cube_stack.py:698-747 */
            const float target_height = a->cubeB_size + a->cubeA_size / 2.0f;
            const float d_lf = norm3(cubeA[0] - lf[0], cubeA[1] - lf[1], cubeA[2] - lf[2]);
            const float d_rf = norm3(cubeA[0] - rf[0], cubeA[1] - rf[1], cubeA[2] - rf[2]);
            float dist_reward = 1.0f - tanhf(10.0f * ((d + d_lf) + d_rf) / 3.0f);
            const int lifted = (height - a->cubeA_size) > 0.04f;
            const float d_ab = norm3(a2b[0] + 0.0f, a2b[1] + 0.0f, a2b[2] + (a->cubeA_size + a->cubeB_size) / 2.0f);
            const float align_reward = (1.0f - tanhf(10.0f * d_ab)) * (lifted ? 1.0f : 0.0f);
            dist_reward = fmaxf(dist_reward, align_reward);
            const int stack = (norm2(a2b[0], a2b[1]) < 0.02f) && (fabsf(height - target_height) < 0.02f) && (d > 0.04f);
            reward = stack ? a->r_stack_scale * 1.0f
                           : (a->r_dist_scale * dist_reward + a->r_lift_scale * (lifted ? 1.0f : 0.0f)) + a->r_align_scale * align_reward;
            if (stack) reset = 1;
        } else {
            /* FSM franka_reach.py:410-430 / franka_cube_stack2.py:408-439 */
            const int on = d <= (a->cubeA_size / 2.0f) * 0.9f;
            const int closed = agrip < 0.0f;
            int fsm = 0;
            if (on) fsm = 1;
            if (on && closed) fsm = 2;
            float tdist = 0.0f, tdist2 = 0.0f;
            if (V == IGE_FRANKA_SA_STACK2) {
                const int lifted = height >= (a->cubeB_size + a->cubeA_size * 0.5f);
                if (on && closed && lifted) fsm = 3;
                const int aligned = norm2(a2b[0], a2b[1]) <= a->cubeA_size * 0.5f;
                if (aligned) fsm = 4;
                const int super_close = norm3(a2b[0], a2b[1], a2b[2]) < (a->cubeA_size * 0.7071f + a->cubeB_size * 0.5f);
                if (aligned && super_close) fsm = 5;
                if (aligned && super_close && !closed) fsm = 6;
            }
            int k = 0;
            for (int i = 0; i < 4; ++i) obs[k++] = eef[3 + i];
            for (int i = 0; i < 3; ++i) obs[k++] = eef[i];
            for (int i = 0; i < 3; ++i) obs[k++] = rel[i];
            obs[k++] = height;
            if (V == IGE_FRANKA_SA_STACK2) {
                for (int i = 0; i < 4; ++i) obs[k++] = cubeA[3 + i];
                for (int i = 0; i < 3; ++i) obs[k++] = a2b[i];
            }
            obs[k++] = qg[0]; obs[k++] = qg[1];
            obs[k++] = pow2_small(fsm);
            const float dist_reward = 1.0f / (0.5f + d * d) * 0.5f;
            const float sf = (float)fsm;
            if (V == IGE_FRANKA_SA_REACH) { /* franka_reach.py:877-928 */
                const float s0 = (dist_reward + tanhf(agrip * 3.0f)) / 2.0f;
                const float s1 = (dist_reward + tanhf(-agrip * 3.0f)) / 2.0f;
                const float s2 = expf(2.5f * height) - 1.0f;
                reward = fsm == 0 ? s0 : 0.0f;
                reward += s1;
                reward += fsm == 1 ? s1 : 0.0f;
                reward += fsm == 2 ? s2 : 0.0f;
                reward += sf;
                comp[0] += s0; comp[1] += s1; comp[2] += s2; comp[7] += sf; comp[8] += height;
            } else { /* franka_cube_stack2.py:774-850 */
                const float s0 = dist_reward;
                const float s1 = (dist_reward + fmaxf(tanhf(-agrip * 3.0f), 0.0f)) / 2.0f;
                const float s2 = fsm == 2 ? fminf(height / 0.095f, 1.0f) : 0.0f;
                const float off = (a->cubeA_size + a->cubeB_size) / 2.0f;
                const float tz = a2b[2] + off;
                tdist = norm3(a2b[0], a2b[1], tz);
                tdist2 = norm3(a2b[0], a2b[1], tz + off); /* the debug block adds the offset once more (:752-756) */
                const float s3 = tanhf(5.0f * -tdist) + 1.0f;
                const float s4 = tanhf(6.0f * -tdist) + 1.0f;
                const float s5 = tanhf(agrip * 7.0f) + 1.0f;
                const float s6 = tanhf(7.0f * d);
                reward = fsm == 0 ? s0 : 0.0f;
                reward += fsm == 1 ? s1 : 0.0f;
                reward += fsm == 2 ? s2 : 0.0f;
                reward += fsm == 3 ? s3 : 0.0f;
                reward += fsm == 4 ? s4 : 0.0f;
                reward += fsm == 5 ? s5 : 0.0f;
                reward += fsm == 6 ? s6 + 10.0f : 0.0f;
                reward += sf;
                comp[0] += s0; comp[1] += s1; comp[2] += s2; comp[3] += s3; comp[4] += s4; comp[5] += s5; comp[6] += s6;
                comp[7] += sf; comp[8] += height; comp[9] += tdist2;
            }
            reward = fmaxf(reward, 0.0f);
        }
        const int late = (float)progress >= a->max_episode_length - 1.0f;
        if (late) reset = 1;
        if (a->obs_clamped)
            for (int i = 0; i < nobs; ++i) a->obs_clamped[(size_t)e * nobs + i] = clamp_obs1(obs[i], a->clip_obs);
        a->rew_buf[e] = reward;
        a->reset_buf[e] = reset;
        a->progress_buf[e] = progress;
        if (a->timeout_buf) a->timeout_buf[e] = (late && reset != 0) ? 1 : 0;
    }
    if (a->reward_components)
        for (int i = 0; i < IGE_FRANKA_SA_COMPONENTS; ++i) a->reward_components[i] = n ? (float)(comp[i] / n) : 0.0f;
    task_advance(&a->list, &a->rand);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * FrankaCabinet: tasks/franka_cabinet.py
 * ------------------------------------------------------------------------------------------ */
/* utils/torch_jit_utils.py:71-77 quat_apply */
static void quat_apply3(const float *q, const float *b, float *out) {
    float t[3], c2[3];
    cross3(q, b, t);
    for (int i = 0; i < 3; ++i) t[i] = t[i] * 2.0f;
    cross3(q, t, c2);
    for (int i = 0; i < 3; ++i) out[i] = (b[i] + q[3] * t[i]) + c2[i];
}
int ige_oracle_franka_cabinet_pre_physics(const IgeFrankaCabinetArgs *a) { /* :417-425 */
    for (int e = 0; e < a->num_envs; ++e)
        for (int j = 0; j < 9; ++j) {
            float *t = a->franka_dof_targets + (size_t)e * a->num_dofs + j;
            const float v = *t + a->dof_speed_scales[j] * a->target_step * a->actions[(size_t)e * 9 + j] * a->action_scale;
            *t = tensor_clamp1(v, a->dof_lower[j], a->dof_upper[j]);
        }
    return 0;
}
int ige_oracle_franka_cabinet_post_physics(const IgeFrankaCabinetArgs *a) { /* :427-436 */
    const int n = a->num_envs, D = a->num_dofs;
    task_list(&a->list, &a->rand, a->reset_buf, n);
    for (int e = 0; e < n; ++e) {
        int64_t progress = a->progress_buf[e] + 1;
        int64_t reset = a->reset_buf[e];
        float *dof = a->dof_state + (size_t)e * D * 2;
        if (reset != 0) { /* reset_idx :385-415 */
            for (int j = 0; j < 9; ++j) {
                float v = a->default_dof_pos[j] + 0.25f * (task_uniform(&a->rand, a->rand.rank, 0, e, 9, j) - 0.5f);
                v = tensor_clamp1(v, a->dof_lower[j], a->dof_upper[j]);
                dof[2 * j] = v; dof[2 * j + 1] = 0.0f;
                a->franka_dof_targets[(size_t)e * D + j] = v;
            }
            for (int j = 9; j < D; ++j) { dof[2 * j] = 0.0f; dof[2 * j + 1] = 0.0f; }
            if (a->num_props > 0)
                memcpy(a->root_state_tensor + ((size_t)e * a->num_actors + 2) * 13,
                       a->default_prop_states + (size_t)e * a->num_props * 13, (size_t)a->num_props * 13 * sizeof(float));
            progress = 0; reset = 0;
        }
        const float *rb = a->rigid_body_states + (size_t)e * a->num_bodies * 13;
        const float *hand = rb + (size_t)a->hand_body * 13, *drawer = rb + (size_t)a->drawer_body * 13;
        const float *lf = rb + (size_t)a->lfinger_body * 13, *rf = rb + (size_t)a->rfinger_body * 13;
        float fg_rot[4], fg_pos[3], dg_rot[4], dg_pos[3], t3[3];
        quat_mul(hand + 3, a->franka_local_grasp_rot, fg_rot); /* tf_combine :148-149 */
        quat_apply3(hand + 3, a->franka_local_grasp_pos, t3);
        for (int i = 0; i < 3; ++i) fg_pos[i] = t3[i] + hand[i];
        quat_mul(drawer + 3, a->drawer_local_grasp_rot, dg_rot);
        quat_apply3(drawer + 3, a->drawer_local_grasp_pos, t3);
        for (int i = 0; i < 3; ++i) dg_pos[i] = t3[i] + drawer[i];
        if (a->franka_grasp_pos) memcpy(a->franka_grasp_pos + (size_t)e * 3, fg_pos, sizeof(fg_pos));
        if (a->franka_grasp_rot) memcpy(a->franka_grasp_rot + (size_t)e * 4, fg_rot, sizeof(fg_rot));
        if (a->drawer_grasp_pos) memcpy(a->drawer_grasp_pos + (size_t)e * 3, dg_pos, sizeof(dg_pos));
        if (a->drawer_grasp_rot) memcpy(a->drawer_grasp_rot + (size_t)e * 4, dg_rot, sizeof(dg_rot));
        float *obs = a->obs_buf + (size_t)e * 23;
        for (int j = 0; j < 9; ++j) {
            obs[j] = 2.0f * (dof[2 * j] - a->dof_lower[j]) / (a->dof_upper[j] - a->dof_lower[j]) - 1.0f; /* :378-379 */
            obs[9 + j] = dof[2 * j + 1] * a->dof_vel_scale;
        }
        for (int i = 0; i < 3; ++i) obs[18 + i] = dg_pos[i] - fg_pos[i];
        const float cab_pos = dof[2 * 12], cab_vel = dof[2 * 12 + 1]; /* cabinet_dof_*[:, 3] */
        obs[21] = cab_pos; obs[22] = cab_vel;
        /* compute_franka_reward :475-551 */
        const float d = norm3(fg_pos[0] - dg_pos[0], fg_pos[1] - dg_pos[1], fg_pos[2] - dg_pos[2]);
        float dist_reward = 1.0f / (1.0f + d * d);
        dist_reward *= dist_reward;
        if (d <= 0.02f) dist_reward = dist_reward * 2.0f;
        float ax1[3], ax2[3], ax3[3], ax4[3];
        quat_apply3(fg_rot, a->gripper_forward_axis, ax1);
        quat_apply3(dg_rot, a->drawer_inward_axis, ax2);
        quat_apply3(fg_rot, a->gripper_up_axis, ax3);
        quat_apply3(dg_rot, a->drawer_up_axis, ax4);
        const float dot1 = dot3(ax1, ax2), dot2 = dot3(ax3, ax4);
        const float sg1 = (float)((dot1 > 0.0f) - (dot1 < 0.0f)), sg2 = (float)((dot2 > 0.0f) - (dot2 < 0.0f));
        const float rot_reward = 0.5f * (sg1 * (dot1 * dot1) + sg2 * (dot2 * dot2));
        const int around = lf[2] > dg_pos[2] && rf[2] < dg_pos[2];
        const float around_handle_reward = around ? 0.0f + 0.5f : 0.0f;
        const float lfd = fabsf(lf[2] - dg_pos[2]), rfd = fabsf(rf[2] - dg_pos[2]);
        const float finger_dist_reward = around ? (0.04f - lfd) + (0.04f - rfd) : 0.0f;
        float sq[9];
        for (int j = 0; j < 9; ++j) { const float x = a->actions[(size_t)e * 9 + j]; sq[j] = x * x; }
        const float action_penalty = row_sum(sq, 9);
        const float open_reward = cab_pos * around_handle_reward + cab_pos;
        float rew = ((((a->dist_reward_scale * dist_reward + a->rot_reward_scale * rot_reward) +
                       a->around_handle_reward_scale * around_handle_reward) + a->open_reward_scale * open_reward) +
                     a->finger_dist_reward_scale * finger_dist_reward) - a->action_penalty_scale * action_penalty;
        if (cab_pos > 0.01f) rew = rew + 0.5f;
        if (cab_pos > 0.2f) rew = rew + around_handle_reward;
        if (cab_pos > 0.39f) rew = rew + (2.0f * around_handle_reward);
        if (lf[0] < dg_pos[0] - a->distX_offset) rew = -1.0f;
        if (rf[0] < dg_pos[0] - a->distX_offset) rew = -1.0f;
        if (cab_pos > 0.39f) reset = 1;
        const int late = (float)progress >= a->max_episode_length - 1.0f;
        if (late) reset = 1;
        if (a->obs_clamped)
            for (int i = 0; i < 23; ++i) a->obs_clamped[(size_t)e * 23 + i] = clamp_obs1(obs[i], a->clip_obs);
        a->rew_buf[e] = rew;
        a->reset_buf[e] = reset;
        a->progress_buf[e] = progress;
        if (a->timeout_buf) a->timeout_buf[e] = (late && reset != 0) ? 1 : 0;
    }
    task_advance(&a->list, &a->rand);
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Helper probe: the helpers above on arbitrary inputs, same columns as the device probe.
 * ------------------------------------------------------------------------------------------ */
int ige_oracle_debug_helper_probe(const IgeHelperProbeArgs *a) {
    for (int i = 0; i < a->n; ++i) {
        const float *q = a->q + (size_t)i * 4, *q2 = a->q2 + (size_t)i * 4, *v = a->v + (size_t)i * 3;
        float *o = a->out + (size_t)i * IGE_PROBE_COLS;
        quat_mul(q, q2, o);
        quat_rotate_impl(q, v, 0, o + 4);
        quat_rotate_impl(q, v, 1, o + 7);
        normalize3(v, o + 10);
        euler_roll_yaw(q, &o[13], &o[14]);
        o[15] = normalize_angle(v[0]);
        quat_conjugate(q, o + 16);
        o[20] = norm3(v[0], v[1], v[2]);
        o[21] = dot3(q, v);
        o[22] = (2.0f * v[0] - v[2] - v[1]) / (v[2] - v[1]);
        o[23] = py_remainder(v[0], 6.28318530717958647692f);
    }
    return 0;
}
