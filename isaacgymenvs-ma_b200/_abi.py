"""ctypes mirror of include/ige_b200.h (the C ABI).  Field order and types must match the header
exactly; tests/test_abi.py cross-checks sizeof()/offsetof() against a compiled probe."""
import ctypes as C

IGE_MAX_DOF = 32
IGE_NUM_STATS = 16
IGE_STATS_ROWS = 32
IGE_STATS_STRIDE = 32
IGE_RAND_INJECTED = 0
IGE_RAND_PHILOX = 1
IGE_FRANKA_SA_REACH, IGE_FRANKA_SA_STACK2, IGE_FRANKA_SA_STACK = 0, 1, 2
IGE_FRANKA_SA_COMPONENTS = 12

IGE_OK, IGE_ENULL, IGE_EALIGN, IGE_ESIZE, IGE_EPARAM = 0, -1, -2, -3, -4

STAT_STEPS, STAT_REWARD, STAT_DONE, STAT_TIMEOUT, STAT_DONE_LENGTH, STAT_TASK0 = 0, 1, 2, 3, 4, 5

_p = C.c_void_p
_f = C.c_float
_i = C.c_int32


class IgeResetRand(C.Structure):
    _fields_ = [("mode", _i), ("_pad", _i), ("seed", C.c_uint64), ("step", _p),
                ("reset_rank", _p), ("u0", _p), ("u1", _p)]


class IgeResetList(C.Structure):
    _fields_ = [("ids", _p), ("count", _p), ("scratch", _p)]


class IgeCompactArgs(C.Structure):
    _fields_ = [("num_rows", _i), ("group", _i), ("reset_buf", _p), ("reset_ids", _p),
                ("reset_rank", _p), ("count", _p), ("scratch", _p)]


IGE_NOISE_ADDITIVE, IGE_NOISE_SCALING = 0, 1
IGE_NOISE_GAUSSIAN, IGE_NOISE_UNIFORM = 0, 1


class IgeNoiseArgs(C.Structure):
    _fields_ = [("count", C.c_int64), ("operation", _i), ("distribution", _i), ("rand_mode", _i), ("init_corr", _i),
                ("corr_scale", _f), ("corr_shift", _f), ("scale", _f), ("shift", _f), ("clip", _f),
                ("stream", C.c_uint32), ("seed", C.c_uint64), ("step", _p), ("x", _p), ("y", _p),
                ("y_clamped", _p), ("corr", _p), ("draw", _p)]


class IgeCartpoleParams(C.Structure):
    _fields_ = [("num_envs", _i), ("reset_dist", _f), ("max_episode_length", _f), ("clip_obs", _f),
                ("rand", IgeResetRand)]


class IgeCartpoleBuffers(C.Structure):
    _fields_ = [("dof_state", _p), ("obs_buf", _p), ("rew_buf", _p), ("reset_buf", _p),
                ("progress_buf", _p), ("obs_clamped", _p), ("timeout_buf", _p), ("stats", _p),
                ("list", IgeResetList)]


class IgeLocomotionParams(C.Structure):
    _fields_ = [("num_envs", _i), ("num_dof", _i), ("num_sensor_floats", _i),
                ("potentials_recip_mul", _i),
                ("dt", _f), ("dof_vel_scale", _f), ("contact_force_scale", _f),
                ("angular_velocity_scale", _f), ("up_weight", _f), ("heading_weight", _f),
                ("actions_cost_scale", _f), ("energy_cost_scale", _f),
                ("joints_at_limit_cost_scale", _f), ("termination_height", _f), ("death_cost", _f),
                ("max_episode_length", _f), ("clip_obs", _f),
                ("targets", _f * 3), ("inv_start_rot", _f * 4), ("basis_vec0", _f * 3),
                ("basis_vec1", _f * 3),
                ("dof_limits_lower", _f * IGE_MAX_DOF), ("dof_limits_upper", _f * IGE_MAX_DOF),
                ("motor_effort_ratio", _f * IGE_MAX_DOF),
                ("rand", IgeResetRand)]


class IgeLocomotionBuffers(C.Structure):
    _fields_ = [("root_states", _p), ("dof_state", _p), ("vec_sensor_tensor", _p),
                ("dof_force_tensor", _p), ("actions", _p), ("initial_root_states", _p),
                ("initial_dof_pos", _p), ("potentials", _p), ("prev_potentials", _p),
                ("up_vec", _p), ("heading_vec", _p), ("obs_buf", _p), ("rew_buf", _p),
                ("reset_buf", _p), ("progress_buf", _p), ("randomize_buf", _p),
                ("obs_clamped", _p), ("timeout_buf", _p), ("stats", _p), ("list", IgeResetList)]


IGE_HAND_OBS = {"openai": 0, "full_no_vel": 1, "full": 2, "full_state": 3}
IGE_HAND_OBS_ALLEGRO = {"full_no_vel": 4, "full": 5, "full_state": 6}


class IgeHandResetRand(C.Structure):
    _fields_ = [("mode", _i), ("_pad", _i), ("seed", C.c_uint64), ("step", _p), ("goal_rank", _p),
                ("env_rank", _p), ("uA", _p), ("uB", _p), ("uC", _p), ("uD", _p)]


class IgeShadowHandParams(C.Structure):
    _fields_ = [("num_envs", _i), ("num_hand_dofs", _i), ("num_dofs_per_env", _i), ("num_bodies", _i),
                ("num_actions", _i), ("obs_type", _i), ("asymmetric_obs", _i), ("ignore_z_rot", _i),
                ("max_consecutive_successes", _i), ("fingertip_handles", _i * 5),
                ("max_episode_length", _f), ("dist_reward_scale", _f), ("rot_reward_scale", _f), ("rot_eps", _f),
                ("action_penalty_scale", _f), ("success_tolerance", _f), ("reach_goal_bonus", _f),
                ("fall_dist", _f), ("fall_penalty", _f), ("av_factor", _f), ("vel_obs_scale", _f),
                ("force_torque_obs_scale", _f), ("clip_obs", _f), ("reset_position_noise", _f),
                ("reset_dof_pos_noise", _f), ("reset_dof_vel_noise", _f), ("goal_displacement", _f * 3),
                ("force_prob_range", _f * 2), ("dof_lower", _f * IGE_MAX_DOF), ("dof_upper", _f * IGE_MAX_DOF),
                ("dof_default_pos", _f * IGE_MAX_DOF), ("dof_default_vel", _f * IGE_MAX_DOF),
                ("dof_reset_base", _f * IGE_MAX_DOF), ("rand", IgeHandResetRand)]


class IgeShadowHandBuffers(C.Structure):
    _fields_ = [("dof_state", _p), ("root_state_tensor", _p), ("rigid_body_states", _p),
                ("vec_sensor_tensor", _p), ("dof_force_tensor", _p), ("actions", _p), ("hand_indices", _p),
                ("object_indices", _p), ("goal_object_indices", _p), ("goal_states", _p),
                ("goal_init_state", _p), ("object_init_state", _p), ("prev_targets", _p), ("cur_targets", _p),
                ("rb_forces", _p), ("random_force_prob", _p), ("successes", _p),
                ("consecutive_successes", _p), ("obs_buf", _p), ("states_buf", _p), ("rew_buf", _p),
                ("reset_buf", _p), ("reset_goal_buf", _p), ("progress_buf", _p), ("randomize_buf", _p),
                ("actor_dirty", _p), ("obs_clamped", _p), ("states_clamped", _p), ("timeout_buf", _p),
                ("stats", _p), ("reduce_scratch", _p), ("list", IgeResetList)]


IGE_FRANKA_VARIANT = {"reach": 0, "collect": 1, "combine": 2, "pp": 3}
IGE_MA_MAX_AGENTS = 4
IGE_MA_NUM_COMPONENTS = 10


class IgeMAResetRand(C.Structure):
    _fields_ = [("mode", _i), ("_pad", _i), ("seed", C.c_uint64), ("step", _p), ("env_rank", _p), ("u", _p * 5)]


class IgeFrankaMAParams(C.Structure):
    _fields_ = [("variant", _i), ("num_envs", _i), ("num_agents", _i), ("num_targets", _i), ("num_actions", _i),
                ("num_obs", _i), ("num_bodies", _i), ("num_actors", _i), ("num_dofs", _i),
                ("eef_body_start", _i), ("eef_body_step", _i), ("base_body_start", _i), ("base_body_step", _i),
                ("hand_body_start", _i), ("hand_body_step", _i), ("cube_actor_start", _i), ("cube_actor_step", _i),
                ("dest_actor_start", _i), ("dest_actor_step", _i), ("stack_base_actor", _i),
                ("max_episode_length", _f), ("clip_obs", _f), ("cubeA_size", _f), ("dest_size", _f),
                ("wall_pos_y", _f), ("wall_height", _f), ("cube_z_base", _f), ("cube_z_range", _f),
                ("cube_xy_center", _f * 2), ("cube_xy_mult", _f), ("aux_z_base", _f),
                ("aux_xy_center_x", _f * IGE_MA_MAX_AGENTS), ("aux_xy_center_y", _f), ("aux_xy_mult", _f),
                ("dof_noise_mult", _f), ("franka_default_dof_pos", _f * 9), ("franka_dof_lower", _f * 9),
                ("franka_dof_upper", _f * 9), ("rand", IgeMAResetRand)]


class IgeFrankaMABuffers(C.Structure):
    _fields_ = [("rigid_body_state", _p), ("root_state", _p), ("dof_state", _p), ("contact_forces", _p),
                ("actions", _p), ("pos_control", _p), ("effort_control", _p), ("init_cubeA_state", _p),
                ("obs_buf", _p), ("rew_buf", _p), ("reset_buf", _p), ("progress_buf", _p),
                ("reward_components", _p), ("obs_clamped", _p), ("timeout_buf", _p), ("stats", _p),
                ("reduce_scratch", _p), ("list", IgeResetList)]


class IgeActionForceArgs(C.Structure):
    _fields_ = [("num_envs", _i), ("num_dof", _i), ("clip_actions", _f), ("power_scale", _f),
                ("gears", _f * IGE_MAX_DOF), ("actions_in", _p), ("actions", _p), ("forces", _p)]


class IgeHandTargetArgs(C.Structure):
    _fields_ = [("num_envs", _i), ("num_actions", _i), ("num_dofs_per_env", _i), ("use_relative_control", _i),
                ("clip_actions", _f), ("speed_scale_dt", _f), ("act_moving_average", _f), ("one_minus_mavg", _f),
                ("actuated", _i * IGE_MAX_DOF), ("lower", _f * IGE_MAX_DOF), ("upper", _f * IGE_MAX_DOF),
                ("actions_in", _p), ("actions", _p), ("prev_targets", _p), ("cur_targets", _p)]


class IgeRolloutStepArgs(C.Structure):
    _fields_ = [("num_rows", _i), ("num_agents", _i), ("value_bootstrap", _i), ("meter_max_size", _i),
                ("gamma", _f), ("shaper_shift", _f), ("shaper_scale", _f), ("shaper_min", _f), ("shaper_max", _f),
                ("rewards", _p), ("dones", _p), ("time_outs", _p), ("values", _p), ("mb_rewards", _p),
                ("dones_out", _p), ("current_rewards", _p), ("current_shaped_rewards", _p),
                ("current_lengths", _p), ("done_rows", _p), ("done_count", _p), ("meter_mean", _p),
                ("meter_size", _p), ("reduce_scratch", _p), ("scratch", _p)]


class IgeDiscountArgs(C.Structure):
    _fields_ = [("num_rows", _i), ("horizon", _i), ("gamma", _f), ("gamma_tau", _f), ("dones", _p),
                ("last_values", _p), ("mb_dones", _p), ("mb_values", _p), ("mb_rewards", _p), ("mb_advs", _p),
                ("mb_returns", _p)]


class IgeFrankaOscArgs(C.Structure):
    _fields_ = [("num_envs", _i), ("num_agents", _i), ("num_actions", _i), ("has_gripper", _i), ("num_bodies", _i),
                ("eef_body_start", _i), ("eef_body_step", _i), ("clip_actions", _f), ("action_scale", _f),
                ("cmd_limit", _f * 6), ("kp", _f * 6), ("kd", _f * 6), ("kp_null", _f * 7), ("kd_null", _f * 7),
                ("default_dof_pos", _f * 7), ("effort_limits", _f * 7), ("gripper_lower", _f * 2),
                ("gripper_upper", _f * 2), ("actions_in", _p), ("actions", _p), ("mm", _p), ("j_eef", _p),
                ("dof_state", _p), ("rigid_body_state", _p), ("effort_control", _p), ("pos_control", _p)]


class IgeAnymalParams(C.Structure):
    _fields_ = [("num_envs", _i), ("num_bodies", _i), ("base_index", _i), ("num_knees", _i),
                ("knee_indices", _i * 8), ("lin_vel_scale", _f), ("ang_vel_scale", _f), ("dof_pos_scale", _f),
                ("dof_vel_scale", _f), ("rew_lin_vel_xy", _f), ("rew_ang_vel_z", _f), ("rew_torque", _f),
                ("max_episode_length", _f), ("clip_obs", _f), ("command_x_range", _f * 2),
                ("command_y_range", _f * 2), ("command_yaw_range", _f * 2), ("default_dof_pos", _f * 12),
                ("rand_mode", _i), ("_pad", _i), ("seed", C.c_uint64), ("step", _p), ("u", _p * 5)]


class IgeAnymalBuffers(C.Structure):
    _fields_ = [("root_states", _p), ("dof_state", _p), ("contact_forces", _p), ("torques", _p), ("actions", _p),
                ("commands", _p), ("obs_buf", _p), ("rew_buf", _p), ("reset_buf", _p), ("progress_buf", _p),
                ("obs_clamped", _p), ("timeout_buf", _p), ("reset_ids", _p), ("reset_count", _p), ("scratch", _p)]


class IgeTaskRand(C.Structure):
    _fields_ = [("mode", _i), ("_pad", _i), ("seed", C.c_uint64), ("step", _p), ("rank", _p), ("u", _p * 8)]


class IgeBallBalanceArgs(C.Structure):
    _fields_ = [("num_envs", _i), ("max_episode_length", _f), ("ball_radius", _f), ("target_step", _f),
                ("clip_obs", _f), ("dof_lower", _f * 6), ("dof_upper", _f * 6), ("rand", IgeTaskRand),
                ("root_states", _p), ("dof_states", _p), ("sensor_tensor", _p), ("initial_root_states", _p),
                ("initial_dof_states", _p), ("actions", _p), ("dof_position_targets", _p), ("obs_buf", _p),
                ("rew_buf", _p), ("reset_buf", _p), ("progress_buf", _p), ("obs_clamped", _p), ("timeout_buf", _p),
                ("list", IgeResetList)]


class IgeQuadcopterArgs(C.Structure):
    _fields_ = [("num_envs", _i), ("num_bodies", _i), ("max_episode_length", _f), ("dof_target_step", _f),
                ("thrust_step", _f), ("clip_obs", _f), ("dof_lower", _f * 8), ("dof_upper", _f * 8),
                ("thrust_lower", _f * 4), ("thrust_upper", _f * 4), ("rotor_bodies", _i * 4), ("rand", IgeTaskRand),
                ("root_states", _p), ("dof_states", _p), ("initial_root_states", _p), ("initial_dof_states", _p),
                ("actions", _p), ("dof_position_targets", _p), ("thrusts", _p), ("forces", _p), ("obs_buf", _p),
                ("rew_buf", _p), ("reset_buf", _p), ("progress_buf", _p), ("obs_clamped", _p), ("timeout_buf", _p),
                ("list", IgeResetList)]


class IgeIngenuityArgs(C.Structure):
    _fields_ = [("num_envs", _i), ("num_bodies", _i), ("max_episode_length", _f), ("dt", _f),
                ("thrust_upper_limit", _f), ("thrust_lateral_component", _f), ("clip_obs", _f), ("_pad", _i),
                ("rand", IgeTaskRand), ("target_rank", _p), ("root_tensor", _p), ("dof_states", _p),
                ("initial_root_states", _p), ("actions", _p), ("target_root_positions", _p), ("thrusts", _p),
                ("forces", _p), ("obs_buf", _p), ("rew_buf", _p), ("reset_buf", _p), ("progress_buf", _p),
                ("obs_clamped", _p), ("timeout_buf", _p), ("actor_dirty", _p), ("list", IgeResetList)]


class IgeFrankaSAArgs(C.Structure):
    _fields_ = [("variant", _i), ("num_envs", _i), ("num_bodies", _i), ("num_actors", _i), ("eef_body", _i),
                ("lf_body", _i), ("rf_body", _i), ("cubeA_actor", _i), ("cubeB_actor", _i), ("draw_slot", _i * 7),
                ("max_episode_length", _f), ("clip_obs", _f), ("cubeA_size", _f), ("cubeB_size", _f),
                ("table_height", _f), ("cube_xy_center", _f * 2), ("cube_xy_mult", _f), ("cubeA_z_base", _f),
                ("cubeB_z_base", _f), ("cube_z_range", _f), ("cube_yaw_mult", _f), ("dof_noise_mult", _f),
                ("min_dist", _f), ("r_dist_scale", _f), ("r_lift_scale", _f), ("r_align_scale", _f),
                ("r_stack_scale", _f), ("default_dof_pos", _f * 9), ("dof_lower", _f * 9), ("dof_upper", _f * 9),
                ("rand", IgeTaskRand), ("rigid_body_state", _p), ("root_state", _p), ("dof_state", _p),
                ("actions", _p), ("pos_control", _p), ("effort_control", _p), ("init_cubeA_state", _p),
                ("init_cubeB_state", _p), ("obs_buf", _p), ("rew_buf", _p), ("reset_buf", _p), ("progress_buf", _p),
                ("obs_clamped", _p), ("timeout_buf", _p), ("reward_components", _p), ("reduce_scratch", _p),
                ("list", IgeResetList)]


class IgeFrankaCabinetArgs(C.Structure):
    _fields_ = [("num_envs", _i), ("num_bodies", _i), ("num_actors", _i), ("num_dofs", _i), ("num_props", _i),
                ("hand_body", _i), ("drawer_body", _i), ("lfinger_body", _i), ("rfinger_body", _i),
                ("max_episode_length", _f), ("clip_obs", _f), ("dof_vel_scale", _f), ("target_step", _f),
                ("action_scale", _f), ("dist_reward_scale", _f), ("rot_reward_scale", _f),
                ("around_handle_reward_scale", _f), ("open_reward_scale", _f), ("finger_dist_reward_scale", _f),
                ("action_penalty_scale", _f), ("distX_offset", _f), ("franka_local_grasp_pos", _f * 3),
                ("franka_local_grasp_rot", _f * 4), ("drawer_local_grasp_pos", _f * 3),
                ("drawer_local_grasp_rot", _f * 4), ("gripper_forward_axis", _f * 3), ("drawer_inward_axis", _f * 3),
                ("gripper_up_axis", _f * 3), ("drawer_up_axis", _f * 3), ("default_dof_pos", _f * 9),
                ("dof_lower", _f * 9), ("dof_upper", _f * 9), ("dof_speed_scales", _f * 9), ("rand", IgeTaskRand),
                ("rigid_body_states", _p), ("root_state_tensor", _p), ("default_prop_states", _p), ("dof_state", _p),
                ("actions", _p), ("franka_dof_targets", _p), ("franka_grasp_pos", _p), ("franka_grasp_rot", _p),
                ("drawer_grasp_pos", _p), ("drawer_grasp_rot", _p), ("obs_buf", _p), ("rew_buf", _p),
                ("reset_buf", _p), ("progress_buf", _p), ("obs_clamped", _p), ("timeout_buf", _p),
                ("list", IgeResetList)]


IGE_MAX_RANKS = 16
IGE_IPC_HANDLE_BYTES = 64


class IgeStatsExchangeArgs(C.Structure):
    _fields_ = [("stats", _p), ("out", _p), ("local", _p), ("mailbox", _p * IGE_MAX_RANKS), ("epoch", _p),
                ("error", _p), ("rank", _i), ("world", _i), ("zero_stats", _i), ("timeout_ms", C.c_uint32)]


IGE_PROBE_COLS = 24


class IgeHelperProbeArgs(C.Structure):
    _fields_ = [("n", _i), ("_pad", _i), ("q", _p), ("q2", _p), ("v", _p), ("out", _p)]


class IgeStatsSnapshotArgs(C.Structure):
    _fields_ = [("stats", _p), ("out", _p), ("zero_stats", _i), ("accumulate", _i)]


def fill_array(carr, values):
    for i, v in enumerate(values):
        carr[i] = type(carr[0])(v) if isinstance(carr[0], int) else float(v)


def reset_list_scratch_words(num_envs):
    """Words of IgeResetList.scratch for `num_envs` envs: asks the library (ige_reset_list_scratch_words)."""
    from . import _lib
    return int(_lib.lib().ige_reset_list_scratch_words(int(num_envs)))
