/*
 * ige_b200.h -- C ABI of the B200-native post-physics step kernels for IsaacGymEnvs-MA.
 *
 * The reference has no FFI: its plug-in point for this path is the Python virtual method
 * VecTask.post_physics_step (isaacgymenvs/tasks/base/vec_task.py:358-360), called from
 * VecTask.step (vec_task.py:391).  Each entry point below replaces ONE task's implementation
 * of that method (plus the fused step() epilogue, vec_task.py:396-404) with a single fused
 * sm_100a CUDA kernel.  INTEGRATION.md shows the ctypes binding a maintainer adds.
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, no C++/torch types, never throws, never exits.
 *   - All pointers are DEVICE pointers borrowed for the duration of the call; the launch is
 *     asynchronous on `stream` (a cudaStream_t passed as void*); the callee never allocates,
 *     frees or synchronises.  The call is CUDA-graph capturable (ige_compact_resets_static: see its own
 *     comment instead).
 *   - Layouts are the gymtorch layouts, unchanged: root / rigid-body rows = 13 floats
 *     (pos3, quat xyzw 4, linvel3, angvel3), DOF rows = (pos, vel), force sensors = 6 floats,
 *     contact forces = 3 floats; reset/progress/randomize flags are int64 (vec_task.py:317-324).
 *   - Return value: 0 (IGE_OK), a negative IGE_E* code for bad arguments, or a positive
 *     cudaError_t from the launch.  ige_last_error_string() describes the last failure of the
 *     calling thread.
 *   - The SAME structs (with host pointers) drive the CPU oracle in oracle/ige_oracle.c
 *     (test infrastructure only).
 */
#ifndef IGE_B200_H
#define IGE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IGE_ABI_VERSION 1

#define IGE_OK 0
#define IGE_ENULL (-1)   /* a required pointer is NULL */
#define IGE_EALIGN (-2)  /* a pointer is not 16-byte aligned */
#define IGE_ESIZE (-3)   /* a size / count is out of range */
#define IGE_EPARAM (-4)  /* an inconsistent parameter combination */

#define IGE_MAX_DOF 32
#define IGE_NUM_STATS 16
/* The statistics accumulator is IGE_STATS_ROWS rows of IGE_STATS_STRIDE floats (first IGE_NUM_STATS
 * of each row used), one 128-byte line per row: a CTA adds to row (blockIdx % ROWS) so the L2
 * atomics of a launch spread over 32 lines instead of serialising on one.  The statistic is the
 * column sum over rows. */
#define IGE_STATS_ROWS 32
#define IGE_STATS_STRIDE 32

/* How reset_idx obtains its uniforms. */
#define IGE_RAND_INJECTED 0 /* parity mode: caller pre-draws torch.rand(k, D) like the reference */
#define IGE_RAND_PHILOX 1   /* production mode: Philox4x32-10 inside the kernel, no host sync */

typedef struct IgeResetRand {
    int32_t mode;              /* IGE_RAND_* */
    int32_t _pad;
    uint64_t seed;             /* PHILOX: key */
    uint64_t *step;            /* PHILOX: device word mixed into the counter; read at kernel entry and
                                  incremented by one at kernel exit by every step kernel that is given a
                                  reset list (IgeResetList.scratch != NULL); NULL = 0 */
    const int32_t *reset_rank; /* INJECTED: [N] exclusive prefix sum of (reset_buf != 0), i.e. the row
                                  of env n in u0/u1 == its position in reset_buf.nonzero() */
    const float *u0;           /* INJECTED: uniforms of the 1st torch.rand call of reset_idx, (k, D) */
    const float *u1;           /* INJECTED: uniforms of the 2nd torch.rand call, (k, D) */
} IgeResetRand;

/* Episode-statistics accumulator slots (device float[IGE_STATS_ROWS * IGE_STATS_STRIDE], atomically
 * added to by every step kernel; column sums all-reduced across GPUs once per horizon). */
#define IGE_STAT_STEPS 0        /* env (or agent) rows stepped */
#define IGE_STAT_REWARD 1       /* sum rew_buf */
#define IGE_STAT_DONE 2         /* sum (reset_buf != 0) after the step */
#define IGE_STAT_TIMEOUT 3      /* sum time_outs */
#define IGE_STAT_DONE_LENGTH 4  /* sum progress_buf over done rows */
#define IGE_STAT_TASK0 5        /* task specific, see each task */

/* Reset index list emitted by every step kernel: the ids of the envs that were reset IN this step
 * (== reset_buf.nonzero() at entry; for agent-axis tasks the envs whose agents were ALL flagged),
 * ascending int32 -- what gym.set_*_tensor_indexed consumes (ant.py:263-271).  Produced inside the
 * fused kernel (persistent kernels: contiguous tile ranges per CTA + a CTA-level look-back over the
 * per-CTA counts, fully overlapped with the step; small kernels: ballot masks expanded by the CTA that
 * retires last).  Optional as a whole (scratch == NULL: no list, rand.step is not advanced). */
typedef struct IgeResetList {
    int32_t *ids;      /* [num_envs] out; the first *count entries are valid */
    int32_t *count;    /* [1] out */
    uint32_t *scratch; /* [ige_reset_list_scratch_words(num_envs)] device words, 16-B aligned, zeroed ONCE by
                          the caller and then owned by the kernels (launch epoch, per-CTA look-back slots,
                          retirement counters, tile masks) */
} IgeResetList;

int ige_reset_list_scratch_words(int32_t num_envs);

/* ------------------------------------------------------------------------------------------
 * Reset compaction: reset_buf.nonzero() (ant.py:291, cartpole.py:168, franka_reach_MA.py:824)
 * ------------------------------------------------------------------------------------------ */
typedef struct IgeCompactArgs {
    int32_t num_rows;          /* N (or N*A for the agent axis) */
    int32_t group;             /* 1, or A: a row is emitted only if ALL `group` consecutive flags of
                                  its env are set (franka_reach_MA.py:879-882); output ids are then
                                  ENV ids */
    const int64_t *reset_buf;  /* [num_rows] */
    int32_t *reset_ids;        /* [num_rows/group] out: ascending ids, first *count entries valid */
    int32_t *reset_rank;       /* [num_rows/group] out: exclusive prefix of the per-env flag */
    int32_t *count;            /* [1] out */
    uint32_t *scratch;         /* [ige_compact_scratch_words(num_rows/group)] device words, zeroed ONCE by
                                  the caller; every launch leaves them zero again (graph re-playable) */
} IgeCompactArgs;

int ige_compact_scratch_words(int32_t num_envs);
int ige_compact_resets(const IgeCompactArgs *args, void *stream);

/* No counterpart in the reference (which lets torch allocate whatever nonzero() needs): the same compaction for a
 * caller that cannot provide `scratch`.  Meant for plug-ins living inside someone else's memory budget -- a simulator
 * that pre-sized its GPU heap, a trainer whose caching allocator owns the card up to a hard cap, a foreign-language
 * binding without a CUDA runtime -- which hold only the tensors they were given and cannot cudaMalloc the look-back
 * words.  The words live in an 8200-byte __device__ array inside the library (held by every GPU that loads it, until
 * process exit), so:
 *   - args->scratch is ignored; num_rows / group must be <= IGE_COMPACT_STATIC_MAX_ENVS, larger -> IGE_ESIZE;
 *   - the call allocates and frees nothing; it enqueues a small cudaMemsetAsync of the words it will use (the
 *     array keeps the previous call's contents) and one kernel, both on `stream`;
 *   - all calls on a device share the array: calls that may overlap on the GPU (two streams, two threads) race and
 *     give wrong ids -- keep them on one stream or otherwise ordered;
 *   - the first call on each device also resolves the array's device address (cudaGetSymbolAddress, with that
 *     device current).  Treat it as set-up: make it once per device before any stream capture begins.
 * Outputs are identical to ige_compact_resets. */
#define IGE_COMPACT_STATIC_MAX_ENVS 4194304
int ige_compact_resets_static(const IgeCompactArgs *args, void *stream);

/* ------------------------------------------------------------------------------------------
 * Observation / action noise of VecTask.step -- replaces dr_randomizations['actions']['noise_lambda'](actions)
 * (vec_task.py:373-374) and dr_randomizations['observations']['noise_lambda'](self.obs_buf) (:399-400), i.e. the
 * closures of apply_randomizations (vec_task.py:675-685 gaussian, :703-711 uniform), fused with the clamp that
 * follows in step() (:376 actions, :404 observations):
 *     noise = (corr * corr_scale + corr_shift) + draw * scale + shift;   y = x + noise | x * noise
 * gaussian: (corr_scale, corr_shift, scale, shift) = (var_corr, mu_corr, var, mu), draw = randn_like(x);
 * uniform:  (hi_corr - lo_corr, lo_corr, hi - lo, lo), draw = rand_like(x).  `corr` is the per-tensor N(0,1) sample
 * the reference draws on the first call and keeps in params['corr'].  The schedule (sched_scaling, :651-673) is
 * host arithmetic on those four numbers, done by the caller.
 * ------------------------------------------------------------------------------------------ */
#define IGE_NOISE_ADDITIVE 0
#define IGE_NOISE_SCALING 1
#define IGE_NOISE_GAUSSIAN 0
#define IGE_NOISE_UNIFORM 1
typedef struct IgeNoiseArgs {
    int64_t count;         /* elements of the tensor */
    int32_t operation;     /* IGE_NOISE_ADDITIVE / IGE_NOISE_SCALING */
    int32_t distribution;  /* IGE_NOISE_GAUSSIAN / IGE_NOISE_UNIFORM (PHILOX: which draw the kernel makes) */
    int32_t rand_mode;     /* IGE_RAND_INJECTED: `draw` (and `corr`) come from the caller; IGE_RAND_PHILOX: in-kernel */
    int32_t init_corr;     /* PHILOX only: != 0 -> this launch also draws `corr` (N(0,1)) and stores it (first call) */
    float corr_scale, corr_shift, scale, shift;
    float clip;            /* clamp(y, +-clip) -> y_clamped; INFINITY = none */
    uint32_t stream;       /* PHILOX: distinguishes tensors that share seed and step (0 = actions, 1 = observations) */
    uint64_t seed;         /* PHILOX key */
    const uint64_t *step;  /* PHILOX: device word mixed into the counter (NULL = 0); not advanced by this kernel */
    const float *x;        /* [count] the tensor (actions / obs_buf) */
    float *y;              /* [count] out, may alias x */
    float *y_clamped;      /* [count] out, optional */
    float *corr;           /* [count] in (out when init_corr) */
    const float *draw;     /* [count] INJECTED: randn_like / rand_like of the reference */
} IgeNoiseArgs;

int ige_dr_noise(const IgeNoiseArgs *args, void *stream);

/* ------------------------------------------------------------------------------------------
 * Cartpole -- replaces Cartpole.post_physics_step (tasks/cartpole.py:165-173) incl. reset_idx
 * (:144-157), compute_observations (:131-142), compute_cartpole_reward (:180-196).
 * ------------------------------------------------------------------------------------------ */
typedef struct IgeCartpoleParams {
    int32_t num_envs;
    float reset_dist;          /* cfg env.resetDist, cartpole.py:41 */
    float max_episode_length;  /* 500, cartpole.py:44 */
    float clip_obs;            /* cfg env.clipObservations (vec_task.py:404); INFINITY = none */
    IgeResetRand rand;         /* u0 = positions draw (k,2), u1 = velocities draw (k,2) */
} IgeCartpoleParams;

typedef struct IgeCartpoleBuffers {
    float *dof_state;          /* (N*2, 2) in/out: reset rows rewritten */
    float *obs_buf;            /* (N, 4) out */
    float *rew_buf;            /* (N) out */
    int64_t *reset_buf;        /* (N) in/out */
    int64_t *progress_buf;     /* (N) in/out */
    float *obs_clamped;        /* (N, 4) out, optional: clamp(obs_buf, +-clip_obs) (vec_task.py:404) */
    uint8_t *timeout_buf;      /* (N) out, optional: bool time_outs (vec_task.py:396) */
    float *stats;              /* [IGE_STATS_ROWS * IGE_STATS_STRIDE] in/out, optional */
    IgeResetList list;         /* optional */
} IgeCartpoleBuffers;

int ige_cartpole_post_physics(const IgeCartpoleParams *p, const IgeCartpoleBuffers *b, void *stream);

/* ------------------------------------------------------------------------------------------
 * Ant / Humanoid -- replace Ant.post_physics_step (tasks/ant.py:287-297) and
 * Humanoid.post_physics_step (tasks/humanoid.py:287-296) incl. reset_idx, compute_*_observations,
 * compute_*_reward.  One struct, two entry points (the tasks differ in obs layout and reward).
 * ------------------------------------------------------------------------------------------ */
typedef struct IgeLocomotionParams {
    int32_t num_envs;
    int32_t num_dof;               /* 8 (Ant) / 21 (Humanoid) */
    int32_t num_sensor_floats;     /* 24 (4 feet x 6) / 12 (2 feet x 6) */
    int32_t potentials_recip_mul;  /* 0: potentials = -|d| / dt (torch CPU semantics, parity default)
                                      1: -|d| * (1/dt)  (torch CUDA div-by-scalar semantics) */
    float dt;
    float dof_vel_scale, contact_force_scale, angular_velocity_scale;
    float up_weight, heading_weight;
    float actions_cost_scale, energy_cost_scale, joints_at_limit_cost_scale;
    float termination_height, death_cost;
    float max_episode_length;
    float clip_obs;
    /* per-env constant rows of the reference (ant.py:103-114), identical for every env */
    float targets[3], inv_start_rot[4], basis_vec0[3], basis_vec1[3];
    float dof_limits_lower[IGE_MAX_DOF], dof_limits_upper[IGE_MAX_DOF];
    float motor_effort_ratio[IGE_MAX_DOF]; /* Humanoid: motor_efforts / max_motor_effort (humanoid.py:355) */
    IgeResetRand rand;             /* u0 = positions draw (k,D) in [0,1), u1 = velocities draw */
} IgeLocomotionParams;

typedef struct IgeLocomotionBuffers {
    const float *root_states;         /* (N, 13) sim view */
    float *dof_state;                 /* (N*D, 2) sim view, reset rows rewritten */
    const float *vec_sensor_tensor;   /* (N, num_sensor_floats) sim view */
    const float *dof_force_tensor;    /* (N, D) sim view; Humanoid only */
    const float *actions;             /* (N, D) */
    const float *initial_root_states; /* (N, 13) */
    const float *initial_dof_pos;     /* (N, D) */
    float *potentials;                /* (N) in/out */
    float *prev_potentials;           /* (N) in/out */
    float *up_vec;                    /* (N, 3) out */
    float *heading_vec;               /* (N, 3) out */
    float *obs_buf;                   /* (N, 60 | 108) out */
    float *rew_buf;                   /* (N) out */
    int64_t *reset_buf;               /* (N) in/out */
    int64_t *progress_buf;            /* (N) in/out */
    int64_t *randomize_buf;           /* (N) in/out, optional */
    float *obs_clamped;               /* optional, see Cartpole */
    uint8_t *timeout_buf;             /* optional */
    float *stats;                     /* optional; TASK0 = sum root_states[:,7] (true_objective, ant.py:245-250) */
    IgeResetList list;                /* optional */
} IgeLocomotionBuffers;

int ige_ant_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, void *stream);
int ige_humanoid_post_physics(const IgeLocomotionParams *p, const IgeLocomotionBuffers *b, void *stream);

/* ------------------------------------------------------------------------------------------
 * ShadowHand -- replaces the reset section of ShadowHand.pre_physics_step (tasks/shadow_hand.py:670-682:
 * reset_target_pose :586-602, reset_idx :604-668, randomize_rotation :803-806) and
 * ShadowHand.post_physics_step (:710-715: compute_observations :437-584 for the four observation
 * types + the asymmetric states_buf, compute_hand_reward :746-800 incl. the global
 * consecutive_successes update).
 * ------------------------------------------------------------------------------------------ */
#define IGE_HAND_OBS_OPENAI 0       /* 42  (shadow_hand.py:473-485) */
#define IGE_HAND_OBS_FULL_NO_VEL 1  /* 77  (:499-510) */
#define IGE_HAND_OBS_FULL 2         /* 157 (:511-526) */
#define IGE_HAND_OBS_FULL_STATE 3   /* 211 (:557-584) */
/* AllegroHand (tasks/allegro_hand.py): same entry points, 16 DOFs / 16 actions, no fingertip or
 * force-sensor segments; reset_idx (:507-573), reset_target_pose (:491-505) and compute_hand_reward (:664-729) are
 * ShadowHand's.  The asymmetric states_buf is then the 88-wide full_state layout (:452-469). */
#define IGE_HAND_OBS_ALLEGRO_FULL_NO_VEL 4  /* 50  (allegro_hand.py:439-449) */
#define IGE_HAND_OBS_ALLEGRO_FULL 5         /* 72  (:450-465) */
#define IGE_HAND_OBS_ALLEGRO_FULL_STATE 6   /* 88  (:470-489) */
#define IGE_HAND_NUM_FINGERTIPS 5

/* Uniforms of the four torch.rand draws of the pre-physics reset, in the reference's order:
 *   A (kg, 4)      reset_target_pose(goal_env_ids)                    shadow_hand.py:587
 *   B (k, 2*24+5)  reset_idx(env_ids)                                 shadow_hand.py:610
 *   C (k, 4)       reset_target_pose(env_ids) nested in reset_idx     shadow_hand.py:613
 *   D (k)          random_force_prob                                  shadow_hand.py:644
 * kg = #(reset_goal_buf != 0), k = #(reset_buf != 0); rows follow nonzero() order. */
typedef struct IgeHandResetRand {
    int32_t mode;             /* IGE_RAND_INJECTED | IGE_RAND_PHILOX */
    int32_t _pad;
    uint64_t seed;
    uint64_t *step;           /* PHILOX: step word, advanced by ige_shadow_hand_post_physics (needs list.scratch) */
    const int32_t *goal_rank; /* INJECTED: [N] exclusive prefix of (reset_goal_buf != 0) */
    const int32_t *env_rank;  /* INJECTED: [N] exclusive prefix of (reset_buf != 0) */
    const float *uA, *uB, *uC, *uD;
} IgeHandResetRand;

typedef struct IgeShadowHandParams {
    int32_t num_envs;
    int32_t num_hand_dofs;     /* 24 */
    int32_t num_dofs_per_env;  /* rows of dof_state / columns of prev_targets per env (>= num_hand_dofs) */
    int32_t num_bodies;        /* rigid bodies per env in rigid_body_states / rb_forces */
    int32_t num_actions;       /* 20 */
    int32_t obs_type;          /* IGE_HAND_OBS_* */
    int32_t asymmetric_obs;    /* also write the 211-wide states_buf (shadow_hand.py:470-471) */
    int32_t ignore_z_rot;      /* object_type == "pen" doubles the success tolerance (:758-759) */
    int32_t max_consecutive_successes;
    int32_t fingertip_handles[IGE_HAND_NUM_FINGERTIPS];
    float max_episode_length;
    float dist_reward_scale, rot_reward_scale, rot_eps, action_penalty_scale;
    float success_tolerance, reach_goal_bonus, fall_dist, fall_penalty, av_factor;
    float vel_obs_scale, force_torque_obs_scale, clip_obs;
    float reset_position_noise, reset_dof_pos_noise, reset_dof_vel_noise;
    float goal_displacement[3];
    float force_prob_range[2];
    float dof_lower[IGE_MAX_DOF], dof_upper[IGE_MAX_DOF];
    float dof_default_pos[IGE_MAX_DOF];   /* shadow_hand_dof_default_pos (limits delta base, :648-649) */
    float dof_default_vel[IGE_MAX_DOF];   /* shadow_hand_dof_default_vel (:654) */
    float dof_reset_base[IGE_MAX_DOF];    /* shadow_hand_default_dof_pos (:652; zeros in the reference) */
    IgeHandResetRand rand;
} IgeShadowHandParams;

typedef struct IgeShadowHandBuffers {
    /* simulator views */
    float *dof_state;                 /* (N*num_dofs_per_env, 2) reset rows rewritten */
    float *root_state_tensor;         /* (num_actors, 13) object / goal rows rewritten by the reset */
    const float *rigid_body_states;   /* (N, num_bodies, 13) */
    const float *vec_sensor_tensor;   /* (N, 30); full_state / asymmetric only */
    const float *dof_force_tensor;    /* (N, 24); full_state / asymmetric only */
    /* task state */
    const float *actions;             /* (N, 20) */
    const int64_t *hand_indices, *object_indices, *goal_object_indices; /* (N) actor rows */
    float *goal_states;               /* (N, 13) */
    const float *goal_init_state;     /* (N, 13) */
    const float *object_init_state;   /* (N, 13) */
    float *prev_targets, *cur_targets;/* (N, num_dofs_per_env) */
    float *rb_forces;                 /* (N, num_bodies, 3) */
    float *random_force_prob;         /* (N) */
    float *successes;                 /* (N) */
    float *consecutive_successes;     /* (1) */
    float *obs_buf;                   /* (N, 42|77|157|211) */
    float *states_buf;                /* (N, 211) when asymmetric_obs */
    float *rew_buf;
    int64_t *reset_buf, *reset_goal_buf, *progress_buf, *randomize_buf;
    /* outputs for the simulator calls of the reset (optional) */
    int64_t *actor_dirty;             /* (num_actors) 0/1: root rows to hand to set_actor_root_state_tensor_indexed
                                         (== the sorted unique() of shadow_hand.py:634-636 once compacted) */
    /* fused step() epilogue (optional) */
    float *obs_clamped, *states_clamped;
    uint8_t *timeout_buf;
    float *stats;                     /* TASK0 = sum resets, TASK0+1 = sum successes*resets */
    uint32_t *reduce_scratch;         /* [4] zeroed once: global reduction of compute_hand_reward */
    IgeResetList list;                /* envs reset by ige_shadow_hand_reset (== reset_buf.nonzero()) */
} IgeShadowHandBuffers;

/* pre-physics position: resets for (reset_goal_buf != 0) and (reset_buf != 0) */
int ige_shadow_hand_reset(const IgeShadowHandParams *p, const IgeShadowHandBuffers *b, void *stream);
/* post-physics position: progress++, observations, reward, termination, consecutive_successes */
int ige_shadow_hand_post_physics(const IgeShadowHandParams *p, const IgeShadowHandBuffers *b, void *stream);

/* ------------------------------------------------------------------------------------------
 * Franka multi-agent tasks (the fork's agent axis: obs/rew/reset/progress rows = env * A + agent).
 * Replaces <Task>.post_physics_step incl. reset_idx (+ _agent_ids_to_env_ids AND filter,
 * _reset_init_cube_state, _reset_dest_cube_state / _reset_stack_base_cube_state), _update_states,
 * compute_FSM, compute_global_FSM, compute_observations and compute_franka_reward of
 *   FrankaReachMA   tasks/franka_reach_MA.py:519-612, :616-704, :821-828, :875-889, :928-960
 *   FrankaCollectMA tasks/franka_collect_MA.py:549-734, :737-860, :1082-1174
 *   FrankaCombineMA tasks/franka_combine_MA.py:565-789, :793-908, :1080-1166
 *   FrankaPPMA      tasks/franka_PPMA.py:561-772, :776-895, :958-966, :1082-1167
 * Reach / Collect accept 1..4 agents and targets; Combine / PP are written for exactly two
 * (franka_PPMA.py:599, :606 hard-code the pair).
 * ------------------------------------------------------------------------------------------ */
#define IGE_FRANKA_REACH 0
#define IGE_FRANKA_COLLECT 1
#define IGE_FRANKA_COMBINE 2
#define IGE_FRANKA_PP 3
#define IGE_MA_MAX_AGENTS 4
#define IGE_MA_NUM_COMPONENTS 10

/* Uniforms of reset_idx, in the reference's torch.rand order (k = #envs whose agents are ALL flagged):
 *   u[0] (k, T)     cube z         u[1] (k, T, 2)  cube xy      (_reset_init_cube_state)
 *   u[2] (k, T)|(k) dest / stack-base z   u[3] (k, T, 2)|(k, 2) dest / stack-base xy   (PP | Combine only)
 *   u[4] (k, 9)     franka DOF noise, shared by the A arms of the env (franka_reach_MA.py:638-647) */
typedef struct IgeMAResetRand {
    int32_t mode;
    int32_t _pad;
    uint64_t seed;
    uint64_t *step;
    const int32_t *env_rank;  /* INJECTED: [N] exclusive prefix of the per-env AND flag */
    const float *u[5];
} IgeMAResetRand;

typedef struct IgeFrankaMAParams {
    int32_t variant;          /* IGE_FRANKA_* */
    int32_t num_envs, num_agents, num_targets;
    int32_t num_actions;      /* 6 (Reach) / 7 */
    int32_t num_obs;          /* 19 / 28 / 48 / 50 for A = T = 2 */
    int32_t num_bodies, num_actors, num_dofs; /* per env: rigid bodies, actors, DOFs (= 9 A) */
    /* strided slices of the reference (franka_reach_MA.py:463-469, :509; franka_PPMA.py:489-517) */
    int32_t eef_body_start, eef_body_step;
    int32_t base_body_start, base_body_step;
    int32_t hand_body_start, hand_body_step;
    int32_t cube_actor_start, cube_actor_step;
    int32_t dest_actor_start, dest_actor_step; /* PP */
    int32_t stack_base_actor;                  /* Combine */
    float max_episode_length, clip_obs;
    float cubeA_size, dest_size, wall_pos_y, wall_height;
    /* reset constants, pre-folded by the host exactly as the reference's Python expressions do */
    float cube_z_base;        /* table z + cube size / 2 (fp32 ops in Reach/Collect, double in Combine/PP) */
    float cube_z_range;       /* rand_z_range: 0.5 (Reach) / 0.2 */
    float cube_xy_center[2];
    float cube_xy_mult;       /* 2 * startPositionNoise */
    float aux_z_base;         /* dest / stack-base z */
    float aux_xy_center_x[IGE_MA_MAX_AGENTS], aux_xy_center_y;
    float aux_xy_mult;        /* 2 * startPositionNoise (PP) / 2 * 0.15 (Combine) */
    float dof_noise_mult;     /* frankaDofNoise * 2 */
    float franka_default_dof_pos[9], franka_dof_lower[9], franka_dof_upper[9];
    IgeMAResetRand rand;
} IgeFrankaMAParams;

typedef struct IgeFrankaMABuffers {
    const float *rigid_body_state; /* (N, num_bodies, 13) */
    float *root_state;             /* (N, num_actors, 13) cube (+dest / stack-base) rows rewritten on reset */
    float *dof_state;              /* (N, num_dofs, 2) */
    const float *contact_forces;   /* (N, num_bodies, 3); Reach only */
    const float *actions;          /* (N*A, num_actions) */
    float *pos_control, *effort_control; /* (N, num_dofs) */
    float *init_cubeA_state;       /* (N, T, 13) */
    float *obs_buf;                /* (N*A, num_obs) */
    float *rew_buf;                /* (N*A) */
    int64_t *reset_buf, *progress_buf; /* (N*A) */
    float *reward_components;      /* [IGE_MA_NUM_COMPONENTS] out: the .mean()s logged in extras
                                      (state0..6, BSR, gBSR, side_rule; unused slots = 0) */
    float *obs_clamped;            /* optional */
    uint8_t *timeout_buf;          /* optional */
    float *stats;                  /* optional; TASK0.. = component SUMS (what the multi-GPU all-reduce needs) */
    float *reduce_scratch;         /* [IGE_MA_NUM_COMPONENTS + 2] zeroed once: component sums + retirement counter */
    IgeResetList list;             /* ENV ids that were reset (all agents flagged) */
} IgeFrankaMABuffers;

int ige_franka_ma_post_physics(const IgeFrankaMAParams *p, const IgeFrankaMABuffers *b, void *stream);

/* ------------------------------------------------------------------------------------------
 * pre_physics_step action scaling: the clamp of VecTask.step
 * (tasks/base/vec_task.py:373) fused with the task's action -> actuation map.  One launch each.
 * actions_in must not alias actions.
 * ------------------------------------------------------------------------------------------ */
typedef struct IgeActionForceArgs {        /* Ant tasks/ant.py:281-285, Humanoid tasks/humanoid.py:281-285 */
    int32_t num_envs, num_dof;
    float clip_actions;                    /* cfg env.clipActions; INFINITY = no clamp */
    float power_scale;
    float gears[IGE_MAX_DOF];              /* Ant joint_gears / Humanoid motor_efforts */
    const float *actions_in;               /* (N, D) policy output */
    float *actions;                        /* (N, D) self.actions = clamp(actions_in) */
    float *forces;                         /* (N, D) actions * gears * power_scale -> set_dof_actuation_force_tensor */
} IgeActionForceArgs;
int ige_locomotion_pre_physics(const IgeActionForceArgs *a, void *stream);

typedef struct IgeHandTargetArgs {         /* ShadowHand tasks/shadow_hand.py:684-698 */
    int32_t num_envs, num_actions, num_dofs_per_env;
    int32_t use_relative_control;
    float clip_actions;
    float speed_scale_dt;                  /* float32(shadow_hand_dof_speed_scale * dt), the Python product */
    float act_moving_average, one_minus_mavg; /* float32(a), float32(1.0 - a) */
    int32_t actuated[IGE_MAX_DOF];         /* actuated_dof_indices, by action column */
    float lower[IGE_MAX_DOF], upper[IGE_MAX_DOF]; /* DOF limits of the actuated DOFs, by action column */
    const float *actions_in;               /* (N, num_actions) */
    float *actions;                        /* (N, num_actions) */
    float *prev_targets, *cur_targets;     /* (N, num_dofs_per_env) */
} IgeHandTargetArgs;
int ige_shadow_hand_pre_physics_targets(const IgeHandTargetArgs *a, void *stream);

/* ------------------------------------------------------------------------------------------
 * Agent-side per-step bookkeeping of A2CAgentMA.play_steps: the immediate
 * consumer of rew_buf / reset_buf / time_outs.  ONE launch replaces learning/A2CAgent_MA.py:35-56
 * (rewards_shaper, time-out value bootstrap, experience_buffer 'rewards'[n] and next 'dones',
 * current_rewards / current_shaped_rewards / current_lengths accumulate + done masking,
 * env_done_indices = dones.nonzero()[::num_agents], the three AverageMeter updates) -- about 15 torch
 * launches and one host sync (len(env_done_indices)) in the reference.
 * rl_games (setup.py:23 "rl-games>=1.6.0", not vendored) supplies DefaultRewardsShaper
 * ((r + shift) * scale, clamp) and AverageMeter (windowed mean); their published algorithms are
 * restated here and in the oracle.
 * ------------------------------------------------------------------------------------------ */
#define IGE_METER_REWARDS 0        /* game_rewards         <- current_rewards[env_done_indices] */
#define IGE_METER_SHAPED_REWARDS 1 /* game_shaped_rewards  <- current_shaped_rewards[...] */
#define IGE_METER_LENGTHS 2        /* game_lengths         <- current_lengths[...] */
typedef struct IgeRolloutStepArgs {
    int32_t num_rows;              /* batch rows = num_envs * num_agents (row = env*A + agent) */
    int32_t num_agents;            /* A: every A-th entry of dones.nonzero() is kept (:45) */
    int32_t value_bootstrap;       /* config value_bootstrap and 'time_outs' in infos (:36) */
    int32_t meter_max_size;        /* games_to_track */
    float gamma;
    float shaper_shift, shaper_scale; /* DefaultRewardsShaper: (r + shift_value) * scale_value */
    float shaper_min, shaper_max;  /* then clamp; -INFINITY / INFINITY = none */
    const float *rewards;          /* (rows) rew_buf as returned by env.step */
    const int64_t *dones;          /* (rows) reset_buf as returned by env.step */
    const uint8_t *time_outs;      /* (rows) extras["time_outs"]; may be NULL when !value_bootstrap */
    const float *values;           /* (rows) res_dict['values']; may be NULL when !value_bootstrap */
    float *mb_rewards;             /* (rows) experience_buffer 'rewards'[n] out */
    uint8_t *dones_out;            /* (rows) optional out: dones as uint8 == what update_data('dones', n+1, self.dones)
                                      stores at the next step (:23) */
    float *current_rewards, *current_shaped_rewards, *current_lengths; /* (rows) in/out */
    int32_t *done_rows;            /* (rows) out: env_done_indices (ROW ids, ascending); first *done_count valid */
    int32_t *done_count;           /* [1] out */
    float *meter_mean;             /* [3] in/out: AverageMeter.mean of IGE_METER_* */
    int32_t *meter_size;           /* [3] in/out: AverageMeter.current_size */
    double *reduce_scratch;        /* [4] zeroed ONCE by the caller; left zero by every launch */
    uint32_t *scratch;             /* [ige_rollout_scratch_words(num_rows)] zeroed ONCE; left zero by every launch */
} IgeRolloutStepArgs;
int ige_rollout_scratch_words(int32_t num_rows);
int ige_ma_rollout_step(const IgeRolloutStepArgs *a, void *stream);

/* A2CBase.discount_values (rl_games GAE) as called from learning/A2CAgent_MA.py:60-66, plus
 * mb_returns = mb_advs + mb_values (:67).  Thread per row walks the horizon backwards. */
typedef struct IgeDiscountArgs {
    int32_t num_rows, horizon;
    float gamma, gamma_tau;        /* float32(gamma), float32(gamma * tau) (the Python double product) */
    const uint8_t *dones;          /* (rows) self.dones after the last step */
    const float *last_values;      /* (rows) */
    const uint8_t *mb_dones;       /* (horizon, rows) */
    const float *mb_values;        /* (horizon, rows) */
    const float *mb_rewards;       /* (horizon, rows) */
    float *mb_advs;                /* (horizon, rows) out, optional */
    float *mb_returns;             /* (horizon, rows) out */
} IgeDiscountArgs;
int ige_ma_discount_values(const IgeDiscountArgs *a, void *stream);

/* ------------------------------------------------------------------------------------------
 * MA Franka pre_physics_step with the operational-space torque controller:
 * Franka*MA.pre_physics_step + _compute_osc_torques (tasks/franka_reach_MA.py:770-818,
 * franka_collect_MA.py:892-951, franka_combine_MA.py:910-969, franka_PPMA.py:897-956) fused with the
 * action clamp of VecTask.step (vec_task.py:373).  One launch, one thread per agent row
 * (row = env*A + agent): mm^-1 (7x7) and (J mm^-1 J^T)^-1 (6x6) are inverted in registers.  Both matrices
 * are symmetric positive definite for a physical mass matrix and a full-row-rank Jacobian, so the inverses
 * use Gauss-Jordan without pivoting (torch.inverse = LAPACK LU with partial pivoting: results agree to
 * rounding x condition number, not bit-exact).  mm / j_eef are the tensors _get_mm / _get_j_eef return
 * (franka_PPMA.py:1041-1061), already stacked over agents.  controlType "osc" only ("joint_tor" cannot run
 * in the reference: (rows,6) * (1,7) does not broadcast, franka_PPMA.py:235,939).
 * ------------------------------------------------------------------------------------------ */
typedef struct IgeFrankaOscArgs {
    int32_t num_envs, num_agents;
    int32_t num_actions;           /* 6 (Reach: arm only) or 7 (7th column = gripper command) */
    int32_t has_gripper;           /* Collect / Combine / PP: franka_PPMA.py:946-953 */
    int32_t num_bodies;            /* rigid bodies per env */
    int32_t eef_body_start, eef_body_step; /* agent a's end-effector body = start + a*step (its [7:13] = eef_vel) */
    float clip_actions;            /* cfg env.clipActions; INFINITY = none */
    float action_scale;            /* u_arm = actions[:, :6] * cmd_limit / action_scale */
    float cmd_limit[6];
    float kp[6], kd[6], kp_null[7], kd_null[7];
    float default_dof_pos[7];      /* franka_default_dof_pos[:7] */
    float effort_limits[7];        /* _franka_effort_limits[:7] */
    float gripper_lower[2], gripper_upper[2]; /* franka_dof_{lower,upper}_limits[-2:] */
    const float *actions_in;       /* (N*A, num_actions) policy output */
    float *actions;                /* (N*A, num_actions) out: self.actions (clamped copy); may alias actions_in */
    const float *mm;               /* (N*A, 7, 7) */
    const float *j_eef;            /* (N*A, 6, 7) */
    const float *dof_state;        /* (N, 9*A, 2): q / qd = [..., 0] / [..., 1], 9 DOFs per agent */
    const float *rigid_body_state; /* (N, num_bodies, 13) */
    float *effort_control;         /* (N, 9*A) in/out: columns 0..5 of each agent written (:944 assigns [:, :6]) */
    float *pos_control;            /* (N, 9*A) in/out: columns 7..8 of each agent written when has_gripper */
} IgeFrankaOscArgs;
int ige_franka_ma_pre_physics(const IgeFrankaOscArgs *a, void *stream);

/* ------------------------------------------------------------------------------------------
 * Anymal (first of the remaining single-agent tasks) -- replaces
 * Anymal.post_physics_step (tasks/anymal.py:231-240) incl. reset_idx (:277-304),
 * compute_anymal_observations (:355-386) and compute_anymal_reward (:311-351), plus the step() epilogue
 * (vec_task.py:396-404).  One launch; the reset list (ascending env ids == reset_buf.nonzero()) and each
 * reset env's row in the injected uniforms come from an in-kernel chained scan.
 * ------------------------------------------------------------------------------------------ */
#define IGE_ANYMAL_DOF 12
#define IGE_ANYMAL_MAX_KNEES 8
typedef struct IgeAnymalParams {
    int32_t num_envs, num_bodies;  /* rigid bodies per env (contact_forces rows) */
    int32_t base_index, num_knees;
    int32_t knee_indices[IGE_ANYMAL_MAX_KNEES];
    float lin_vel_scale, ang_vel_scale, dof_pos_scale, dof_vel_scale;
    float rew_lin_vel_xy, rew_ang_vel_z, rew_torque; /* rew_scales[...] already multiplied by dt (anymal.py:57-59) */
    float max_episode_length;      /* int in the reference; compared as progress >= max - 1 */
    float clip_obs;                /* cfg env.clipObservations; INFINITY = none */
    float command_x_range[2], command_y_range[2], command_yaw_range[2];
    float default_dof_pos[IGE_ANYMAL_DOF]; /* every row of self.default_dof_pos (anymal.py:132-138) */
    int32_t rand_mode;             /* IGE_RAND_* */
    int32_t _pad;
    uint64_t seed;                 /* PHILOX */
    uint64_t *step;                /* PHILOX: device counter, advanced by one per launch; NULL = 0 */
    const float *u[5];             /* INJECTED, rows by reset rank: positions_offset (k,12), velocities (k,12),
                                      commands x (k,1), y (k,1), yaw (k,1) -- the torch.rand calls of :282-302 */
} IgeAnymalParams;

typedef struct IgeAnymalBuffers {
    const float *root_states;      /* (N, 13) */
    float *dof_state;              /* (N, 12, 2) in/out: reset rows rewritten */
    const float *contact_forces;   /* (N, num_bodies, 3) */
    const float *torques;          /* (N, 12) */
    const float *actions;          /* (N, 12) */
    float *commands;               /* (N, 3) in/out: re-drawn on reset */
    float *obs_buf;                /* (N, 48) out */
    float *rew_buf;                /* (N) out */
    int64_t *reset_buf, *progress_buf; /* (N) in/out */
    float *obs_clamped;            /* optional (N, 48) */
    uint8_t *timeout_buf;          /* optional (N) */
    int32_t *reset_ids;            /* (N) out: ids of the envs reset in this step, ascending */
    int32_t *reset_count;          /* [1] out */
    uint32_t *scratch;             /* [ige_anymal_scratch_words(num_envs)] zeroed ONCE; left zero by every launch */
} IgeAnymalBuffers;
int ige_anymal_scratch_words(int32_t num_envs);
int ige_anymal_post_physics(const IgeAnymalParams *p, const IgeAnymalBuffers *b, void *stream);

/* ------------------------------------------------------------------------------------------
 * Remaining single-agent tasks.  One Args struct per task (scalars first, then the
 * borrowed device pointers), two entry points for the tasks whose reset_idx runs in pre_physics_step.
 * Every kernel that resets envs emits the ascending reset id list itself (IgeResetList) and advances
 * rand.step once per launch.
 * ------------------------------------------------------------------------------------------ */
typedef struct IgeTaskRand {
    int32_t mode;          /* IGE_RAND_INJECTED | IGE_RAND_PHILOX */
    int32_t _pad;
    uint64_t seed;         /* PHILOX: key */
    uint64_t *step;        /* PHILOX: device word mixed into the counter, advanced by one per resetting launch */
    const int32_t *rank;   /* INJECTED: [N] exclusive prefix of the reset flag == row of env n in u[] */
    const float *u[8];     /* INJECTED: the torch.rand draws of the reset path, in call order (shapes per task) */
} IgeTaskRand;

/* BallBalance -- tasks/ball_balance.py: pre_physics_step :410-430 (reset_idx :360-408 + position targets),
 * post_physics_step :432-441 (compute_observations :330-345, compute_bbot_reward :460-476) and the step() epilogue.
 * Reset draws (rows by reset rank, each (k,1)): u[0] dists, u[1] direction angle, u[2] horizontal speed,
 * u[3] vertical speed (its value is multiplied by 5.0 - 5.0), u[4] height. */
typedef struct IgeBallBalanceArgs {
    int32_t num_envs;
    float max_episode_length;        /* cfg env.maxEpisodeLength (500) */
    float ball_radius;               /* 0.1 (:99 of _create_envs) */
    float target_step;               /* float32(dt * actionSpeedScale): the Python double product of :421 */
    float clip_obs;                  /* INFINITY = none */
    float dof_lower[6], dof_upper[6];/* bbot_dof_lower / upper_limits */
    IgeTaskRand rand;
    float *root_states;              /* (N, 2, 13) actor 0 = balance bot, actor 1 = ball; reset rows rewritten */
    float *dof_states;               /* (N, 6, 2) reset rows rewritten */
    const float *sensor_tensor;      /* (N, 3, 6) force sensors (force xyz, torque xyz) */
    const float *initial_root_states;/* (N, 2, 13) */
    const float *initial_dof_states; /* (N, 6, 2) */
    const float *actions;            /* (N, 3) as passed to pre_physics_step (already clamped by VecTask.step) */
    float *dof_position_targets;     /* (N, 6) in/out */
    float *obs_buf;                  /* (N, 24) */
    float *rew_buf;
    int64_t *reset_buf, *progress_buf;
    float *obs_clamped;              /* optional */
    uint8_t *timeout_buf;            /* optional */
    IgeResetList list;               /* pre_physics: envs reset in this call */
} IgeBallBalanceArgs;
int ige_ball_balance_pre_physics(const IgeBallBalanceArgs *a, void *stream);
int ige_ball_balance_post_physics(const IgeBallBalanceArgs *a, void *stream);

/* Quadcopter -- tasks/quadcopter.py: pre_physics_step :301-331 (reset_idx :280-299 + DOF targets, thrusts, rotor
 * forces), post_physics_step :333-341 (compute_observations :362-373, compute_quadcopter_reward :391-425).
 * Reset draws: u[0], u[1], u[2] root x / y / z offsets (k,1); u[3] DOF positions (k,8). */
typedef struct IgeQuadcopterArgs {
    int32_t num_envs;
    int32_t num_bodies;              /* rigid bodies per env in `forces` (9) */
    float max_episode_length;
    float dof_target_step;           /* float32(dt * 8 pi)   (:309-310) */
    float thrust_step;               /* float32(dt * 200)    (:313-314) */
    float clip_obs;
    float dof_lower[8], dof_upper[8];
    float thrust_lower[4], thrust_upper[4];
    int32_t rotor_bodies[4];         /* forces[:, body, 2] = thrust (2, 4, 6, 8; :317-320) */
    IgeTaskRand rand;
    float *root_states;              /* (N, 13) reset rows rewritten */
    float *dof_states;               /* (N, 8, 2) reset rows rewritten */
    const float *initial_root_states, *initial_dof_states;
    const float *actions;            /* (N, 12) */
    float *dof_position_targets;     /* (N, 8) in/out */
    float *thrusts;                  /* (N, 4) in/out */
    float *forces;                   /* (N, num_bodies, 3) in/out */
    float *obs_buf;                  /* (N, 21) */
    float *rew_buf;
    int64_t *reset_buf, *progress_buf;
    float *obs_clamped;
    uint8_t *timeout_buf;
    IgeResetList list;
} IgeQuadcopterArgs;
int ige_quadcopter_pre_physics(const IgeQuadcopterArgs *a, void *stream);
int ige_quadcopter_post_physics(const IgeQuadcopterArgs *a, void *stream);

/* Ingenuity -- tasks/ingenuity.py: pre_physics_step :322-364 (set_targets :283-294 for progress % 500 == 0,
 * reset_idx :296-320, rotor thrusts), post_physics_step :366-374 (compute_observations :396-401,
 * compute_ingenuity_reward :420-455).
 * Draws: u[0] (kt, 2) / u[1] (kt) set_targets of the envs with progress % 500 == 0 (rows by target_rank);
 *        u[2] (k, 2) / u[3] (k) set_targets nested in reset_idx, u[4], u[5], u[6] (k,1) root x / y / z offsets
 *        (rows by rand.rank). */
typedef struct IgeIngenuityArgs {
    int32_t num_envs;
    int32_t num_bodies;              /* rigid bodies per env in `forces` (6) */
    float max_episode_length;
    float dt;
    float thrust_upper_limit;        /* 2000 */
    float thrust_lateral_component;  /* 0.2 */
    float clip_obs;
    int32_t _pad;
    IgeTaskRand rand;
    const int32_t *target_rank;      /* INJECTED: [N] exclusive prefix of (progress_buf % 500 == 0) */
    float *root_tensor;              /* (N, 2, 13) actor 0 = copter, actor 1 = target marker */
    float *dof_states;               /* (N, 4, 2): rotor velocities of ALL envs are re-set whenever any env resets (:299-300) */
    const float *initial_root_states;/* (N, 13) copter rows */
    const float *actions;            /* (N, 6) */
    float *target_root_positions;    /* (N, 3) in/out */
    float *thrusts;                  /* (N, 2, 3) */
    float *forces;                   /* (N, num_bodies, 3) */
    float *obs_buf;                  /* (N, 13) */
    float *rew_buf;
    int64_t *reset_buf, *progress_buf;
    float *obs_clamped;
    uint8_t *timeout_buf;
    int64_t *actor_dirty;            /* optional (N * 2): 0/1 per actor row to hand to set_actor_root_state_tensor_indexed
                                        (== the sorted unique() of :341 once compacted) */
    IgeResetList list;               /* envs reset in this call (== reset_buf.nonzero()) */
} IgeIngenuityArgs;
int ige_ingenuity_pre_physics(const IgeIngenuityArgs *a, void *stream);
int ige_ingenuity_post_physics(const IgeIngenuityArgs *a, void *stream);

/* ------------------------------------------------------------------------------------------
 * Single-agent Franka cube tasks: FrankaReach (the fork's FSM lifting task,
 * tasks/franka_reach.py), FrankaCubeStack2 (the fork's 7-state stacking FSM, tasks/franka_cube_stack2.py) and
 * FrankaCubeStack (NVIDIA's original, tasks/franka_cube_stack.py).  One entry point replaces post_physics_step
 * (franka_reach.py:695-704, franka_cube_stack2.py:707-716, franka_cube_stack.py:657-665): progress++, reset_idx
 * (:479-535 / :501-558 / :458-514 with _reset_init_cube_state), _update_states + FSM (:410-458 / :408-466 / :416-433),
 * compute_observations, compute_franka_reward (:877-928 / :774-850 / :698-747) and the step() epilogue.
 *
 * torch.rand draws of reset_idx, rows by reset rank (slot = index into rand.u[], -1 = the variant does not draw it):
 *   draw_slot[0] cubeB z (k)   [1] cubeB xy (k,2)   [2] cubeB yaw (k)   [3] cubeA z (k)   [4] cubeA xy (k,2)
 *   [5] cubeA yaw (k)   [6] franka joints (k,9).
 * FrankaCubeStack re-samples cubeA's xy until it clears cubeB (check_valid, up to 100 tries, :567-585): in INJECTED
 * mode u[draw_slot[4]] holds the ACCEPTED sample of each env; in PHILOX mode the kernel runs the retry loop itself.
 * ------------------------------------------------------------------------------------------ */
#define IGE_FRANKA_SA_REACH 0
#define IGE_FRANKA_SA_STACK2 1
#define IGE_FRANKA_SA_STACK 2
#define IGE_FRANKA_SA_COMPONENTS 12
/* reward_components[]: Reach     0 r/state0, 1 r/state1, 2 r/state2, 7 r/state, 8 mean_cube_height
 *                      Stack2    0..6 r/state0, r/state1, r/state2, r/stat3, r/stat4, r/stat5, r/state6, 7 r/state,
 *                                8 mean_cube_height, 9 target_dist          (means over the envs of this launch) */
typedef struct IgeFrankaSAArgs {
    int32_t variant;                 /* IGE_FRANKA_SA_* */
    int32_t num_envs;
    int32_t num_bodies, num_actors;  /* rows per env of rigid_body_state / root_state */
    int32_t eef_body, lf_body, rf_body;  /* panda_grip_site / leftfinger_tip / rightfinger_tip handles (:355-362) */
    int32_t cubeA_actor, cubeB_actor;
    int32_t draw_slot[7];
    float max_episode_length, clip_obs;
    float cubeA_size, cubeB_size, table_height;
    float cube_xy_center[2], cube_xy_mult;   /* float32(table xy), float32(2 * startPositionNoise) */
    float cubeA_z_base, cubeB_z_base, cube_z_range;  /* z = base + rand * range (range 0: no draw) */
    float cube_yaw_mult;             /* float32(2 * startRotationNoise); 0 = no rotation draw */
    float dof_noise_mult;            /* float32(frankaDofNoise * 2) */
    float min_dist;                  /* Stack: collision-free distance of the two cubes (:550-553) */
    float r_dist_scale, r_lift_scale, r_align_scale, r_stack_scale;   /* Stack reward_settings */
    float default_dof_pos[9], dof_lower[9], dof_upper[9];
    IgeTaskRand rand;
    const float *rigid_body_state;   /* (N, num_bodies, 13) */
    float *root_state;               /* (N, num_actors, 13) cube rows rewritten by the reset */
    float *dof_state;                /* (N, 9, 2) reset rows rewritten */
    const float *actions;            /* (N, 7) */
    float *pos_control, *effort_control;         /* (N, 9) reset rows rewritten */
    float *init_cubeA_state, *init_cubeB_state;  /* (N, 13); cubeB unused by Reach */
    float *obs_buf;                  /* (N, 14 | 21 | 19) */
    float *rew_buf;
    int64_t *reset_buf, *progress_buf;
    float *obs_clamped;              /* optional */
    uint8_t *timeout_buf;            /* optional */
    float *reward_components;        /* optional [IGE_FRANKA_SA_COMPONENTS] */
    float *reduce_scratch;           /* [IGE_FRANKA_SA_COMPONENTS + 2], zeroed once */
    IgeResetList list;
} IgeFrankaSAArgs;
int ige_franka_sa_post_physics(const IgeFrankaSAArgs *a, void *stream);

/* FrankaCabinet -- tasks/franka_cabinet.py: pre_physics_step :417-425 (DOF position targets) and post_physics_step
 * :427-436 (reset_idx :385-415, compute_observations :352-383 with compute_grasp_transforms :554-567,
 * compute_franka_reward :475-551) + the step() epilogue.  Reset draw: u[0] (k, 9) franka joints. */
typedef struct IgeFrankaCabinetArgs {
    int32_t num_envs;
    int32_t num_bodies, num_actors, num_dofs;   /* per env: rigid bodies, actors (2 + num_props), DOFs (9 + 4) */
    int32_t num_props;
    int32_t hand_body, drawer_body, lfinger_body, rfinger_body;
    float max_episode_length, clip_obs;
    float dof_vel_scale, target_step;  /* target_step = float32(dt) applied as speed_scale * dt * action * action_scale */
    float action_scale;
    float dist_reward_scale, rot_reward_scale, around_handle_reward_scale, open_reward_scale;
    float finger_dist_reward_scale, action_penalty_scale, distX_offset;
    float franka_local_grasp_pos[3], franka_local_grasp_rot[4];
    float drawer_local_grasp_pos[3], drawer_local_grasp_rot[4];
    float gripper_forward_axis[3], drawer_inward_axis[3], gripper_up_axis[3], drawer_up_axis[3];
    float default_dof_pos[9], dof_lower[9], dof_upper[9], dof_speed_scales[9];
    IgeTaskRand rand;
    const float *rigid_body_states;  /* (N, num_bodies, 13) */
    float *root_state_tensor;        /* (N, num_actors, 13) prop rows rewritten by the reset */
    const float *default_prop_states;/* (N, num_props, 13) */
    float *dof_state;                /* (N, num_dofs, 2): franka 0..8, cabinet 9..12; reset rows rewritten */
    const float *actions;            /* (N, 9) */
    float *franka_dof_targets;       /* (N, num_dofs) */
    float *franka_grasp_pos, *franka_grasp_rot, *drawer_grasp_pos, *drawer_grasp_rot;   /* optional (N,3)/(N,4) outputs */
    float *obs_buf;                  /* (N, 23) */
    float *rew_buf;
    int64_t *reset_buf, *progress_buf;
    float *obs_clamped;
    uint8_t *timeout_buf;
    IgeResetList list;
} IgeFrankaCabinetArgs;
int ige_franka_cabinet_pre_physics(const IgeFrankaCabinetArgs *a, void *stream);
int ige_franka_cabinet_post_physics(const IgeFrankaCabinetArgs *a, void *stream);

/* ------------------------------------------------------------------------------------------
 * Episode statistics: fold the striped accumulator a step kernel adds into (Buffers.stats) into the
 * <= IGE_NUM_STATS-float vector that is all-reduced across GPUs once per horizon (the
 * reference keeps per-rank statistics, utils/rlgames_utils.py:89-107).  One tiny launch, off the step's
 * critical path (the host double-buffers the accumulator, see distributed.StatsReducer).
 * ------------------------------------------------------------------------------------------ */
typedef struct IgeStatsSnapshotArgs {
    float *stats;          /* [IGE_STATS_ROWS * IGE_STATS_STRIDE] accumulator rows */
    float *out;            /* [IGE_NUM_STATS] column sums over the rows */
    int32_t zero_stats;    /* != 0: leave the rows zeroed (a horizon's numbers never mix with the next one's) */
    int32_t accumulate;    /* != 0: out += sums instead of out = sums */
} IgeStatsSnapshotArgs;
int ige_stats_snapshot(const IgeStatsSnapshotArgs *a, void *stream);

/* ------------------------------------------------------------------------------------------
 * Statistics all-reduce fused with the horizon fold, over NVLink / NVSwitch peer memory.
 *
 * ONE single-CTA kernel per horizon: fold the accumulator rows into the 16-float vector (as ige_stats_snapshot),
 * STORE the vector into the mailbox of every rank (peer-mapped device memory: the stores travel over NVLink), wait
 * until every rank's vector of this epoch has landed in the LOCAL mailbox, sum the slots in rank order (identical
 * bits on every rank) into `out`.  Every mailbox word is 64 bits, {epoch : float}, written with one store -- a word
 * whose epoch matches is complete, so no fence and no separate flag travel over the link (the LL idea of NCCL).
 * The mailbox is double-buffered by epoch parity; a rank can only be one exchange ahead of any other, because it
 * needs that rank's vector of the previous epoch first.  The kernel polls local memory only, takes no SM resources
 * to speak of (IGE_NUM_STATS * world threads) and is CUDA-graph capturable: as a forked branch of the horizon graph
 * it costs the step stream nothing.  Replaces torch.distributed.all_reduce of the reference's statistics
 * (utils/rlgames_utils.py:89-107 keeps them per rank; rl_games sums them with NCCL) -- NCCL stays the bootstrap
 * (handle exchange, barriers) and the selectable fallback (IGE_STATS_REDUCE=nccl).
 *
 * Mailbox memory comes from ige_mailbox_create (cudaMalloc + zero fill + cudaIpcGetMemHandle); the 64-byte handles are
 * exchanged by the host (one all-gather at start-up) and opened with ige_mailbox_open.
 * ------------------------------------------------------------------------------------------ */
#define IGE_MAX_RANKS 16
#define IGE_IPC_HANDLE_BYTES 64
typedef struct IgeStatsExchangeArgs {
    float *stats;                       /* [IGE_STATS_ROWS * IGE_STATS_STRIDE] accumulator rows of the finished horizon */
    float *out;                         /* [IGE_NUM_STATS] sum over all ranks */
    float *local;                       /* optional [IGE_NUM_STATS]: this rank's own folded vector */
    uint64_t *mailbox[IGE_MAX_RANKS];   /* mailbox[r]: rank r's mailbox as mapped into THIS process (own: the local one) */
    uint64_t *epoch;                    /* device word, starts at 0; advanced by the kernel (graph replays count themselves) */
    int32_t *error;                     /* device word: set to 1 + the first missing rank when the wait timed out */
    int32_t rank, world;
    int32_t zero_stats;                 /* != 0: leave the rows zeroed */
    uint32_t timeout_ms;                /* give up waiting for a peer after this long (0 = 2000 ms) */
} IgeStatsExchangeArgs;
int ige_stats_exchange(const IgeStatsExchangeArgs *a, void *stream);
/* bytes of one rank's mailbox for `world` ranks */
uint64_t ige_mailbox_bytes(int32_t world);
/* allocate + zero this rank's mailbox on the current device; `handle` receives the IPC handle to send to the peers */
int ige_mailbox_create(int32_t world, void **mailbox, unsigned char handle[IGE_IPC_HANDLE_BYTES]);
/* map a peer's mailbox (enables peer access to its device on first use) */
int ige_mailbox_open(const unsigned char handle[IGE_IPC_HANDLE_BYTES], void **mailbox);
int ige_mailbox_close(void *peer_mailbox);      /* cudaIpcCloseMemHandle */
int ige_mailbox_destroy(void *own_mailbox);     /* cudaFree */

/* ------------------------------------------------------------------------------------------
 * Helper probe (test hook for SURVEY.md 8a row a12): evaluates the device versions of the utils/torch_jit_utils.py helpers
 * the kernels are built from on arbitrary inputs, one row per thread.  out row (IGE_PROBE_COLS floats):
 *   0..3  quat_mul(q, q2) :41-62        4..6  quat_rotate(q, v) :80-90      7..9  quat_rotate_inverse(q, v) :93-103
 *   10..12 normalize(v) :65-67          13,14 get_euler_xyz roll, yaw :175-195   15 normalize_angle(v.x) :126-128
 *   16..19 quat_conjugate(q) :106-110   20 norm(v) (torch.norm dim=-1)      21 dot(q.xyz, v) (bmm 1x3 3x1 :88-89)
 *   22 unscale(v.x, v.y, v.z) :238-240  23 remainder(v.x, 2 pi) (the `%` of :195)
 * ------------------------------------------------------------------------------------------ */
#define IGE_PROBE_COLS 24
typedef struct IgeHelperProbeArgs {
    int32_t n;
    int32_t _pad;
    const float *q, *q2;   /* (n, 4) xyzw */
    const float *v;        /* (n, 3) */
    float *out;            /* (n, IGE_PROBE_COLS) */
} IgeHelperProbeArgs;
int ige_debug_helper_probe(const IgeHelperProbeArgs *a, void *stream);

/* ------------------------------------------------------------------------------------------
 * Library
 * ------------------------------------------------------------------------------------------ */
int ige_version(void);
const char *ige_last_error_string(void);
/* Kernel launches issued by this library since load (bench.py's gpu_launches claim). */
uint64_t ige_launch_count(void);
/* Debug knob: 0 = stage tiles with plain coalesced loads/stores instead of TMA bulk copies
 * (also settable with the environment variable IGE_NO_TMA=1).  Results are identical. */
void ige_set_use_tma(int on);
/* Debug knob: 0 = take the generic any-shape kernels instead of the shape-specialised fast paths (ShadowHand with
 * 24 DOFs / 20 actions, Franka MA with 2 agents / 2 targets); also IGE_GENERIC=1.  Results are identical. */
void ige_set_fast_paths(int on);
/* Debug knob: 0 = launch the kernels fully stream-serialised instead of with programmatic dependent launch (each
 * kernel waits for its predecessor's completion before its first memory access either way); also IGE_PDL=0. */
void ige_set_use_pdl(int on);
/* PCI bus id ("0000:1b:00.0") of CUDA device `device` (host NUMA placement of pinned arenas, one rank per GPU). */
int ige_device_pci_bus_id(int device, char *out, int len);

#ifdef __cplusplus
}
#endif
#endif /* IGE_B200_H */
