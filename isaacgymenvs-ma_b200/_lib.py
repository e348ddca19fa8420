"""ctypes binding of libige_b200.so (the C ABI in include/ige_b200.h).

There is NO fallback: if the library is missing or a call fails, this raises.  The product path
never routes through the CPU oracle or eager torch.
"""
import ctypes as C
import os

from . import _abi
from .build import SO_PATH

_lib = None

_SIGS = {
    "ige_compact_resets": (_abi.IgeCompactArgs,),
    "ige_compact_resets_static": (_abi.IgeCompactArgs,),
    "ige_dr_noise": (_abi.IgeNoiseArgs,),
    "ige_cartpole_post_physics": (_abi.IgeCartpoleParams, _abi.IgeCartpoleBuffers),
    "ige_ant_post_physics": (_abi.IgeLocomotionParams, _abi.IgeLocomotionBuffers),
    "ige_humanoid_post_physics": (_abi.IgeLocomotionParams, _abi.IgeLocomotionBuffers),
    "ige_shadow_hand_reset": (_abi.IgeShadowHandParams, _abi.IgeShadowHandBuffers),
    "ige_shadow_hand_post_physics": (_abi.IgeShadowHandParams, _abi.IgeShadowHandBuffers),
    "ige_franka_ma_post_physics": (_abi.IgeFrankaMAParams, _abi.IgeFrankaMABuffers),
    "ige_locomotion_pre_physics": (_abi.IgeActionForceArgs,),
    "ige_shadow_hand_pre_physics_targets": (_abi.IgeHandTargetArgs,),
    "ige_franka_ma_pre_physics": (_abi.IgeFrankaOscArgs,),
    "ige_anymal_post_physics": (_abi.IgeAnymalParams, _abi.IgeAnymalBuffers),
    "ige_ma_rollout_step": (_abi.IgeRolloutStepArgs,),
    "ige_ma_discount_values": (_abi.IgeDiscountArgs,),
    "ige_ball_balance_pre_physics": (_abi.IgeBallBalanceArgs,),
    "ige_ball_balance_post_physics": (_abi.IgeBallBalanceArgs,),
    "ige_quadcopter_pre_physics": (_abi.IgeQuadcopterArgs,),
    "ige_quadcopter_post_physics": (_abi.IgeQuadcopterArgs,),
    "ige_ingenuity_pre_physics": (_abi.IgeIngenuityArgs,),
    "ige_ingenuity_post_physics": (_abi.IgeIngenuityArgs,),
    "ige_franka_sa_post_physics": (_abi.IgeFrankaSAArgs,),
    "ige_franka_cabinet_pre_physics": (_abi.IgeFrankaCabinetArgs,),
    "ige_franka_cabinet_post_physics": (_abi.IgeFrankaCabinetArgs,),
    "ige_stats_exchange": (_abi.IgeStatsExchangeArgs,),
    "ige_debug_helper_probe": (_abi.IgeHelperProbeArgs,),
    "ige_stats_snapshot": (_abi.IgeStatsSnapshotArgs,),
}


class IgeError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise IgeError(
                f"{SO_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  There is no CPU fallback for the post-physics kernels.")
        l = C.CDLL(SO_PATH)
        l.ige_version.restype = C.c_int
        l.ige_last_error_string.restype = C.c_char_p
        l.ige_launch_count.restype = C.c_uint64
        l.ige_compact_scratch_words.restype = C.c_int
        l.ige_compact_scratch_words.argtypes = [C.c_int32]
        l.ige_reset_list_scratch_words.restype = C.c_int
        l.ige_reset_list_scratch_words.argtypes = [C.c_int32]
        l.ige_rollout_scratch_words.restype = C.c_int
        l.ige_rollout_scratch_words.argtypes = [C.c_int32]
        l.ige_anymal_scratch_words.restype = C.c_int
        l.ige_anymal_scratch_words.argtypes = [C.c_int32]
        l.ige_mailbox_bytes.restype = C.c_uint64
        l.ige_mailbox_bytes.argtypes = [C.c_int32]
        l.ige_mailbox_create.restype = C.c_int
        l.ige_mailbox_create.argtypes = [C.c_int32, C.POINTER(C.c_void_p), C.c_char_p]
        l.ige_mailbox_open.restype = C.c_int
        l.ige_mailbox_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
        l.ige_mailbox_close.restype = C.c_int
        l.ige_mailbox_close.argtypes = [C.c_void_p]
        l.ige_mailbox_destroy.restype = C.c_int
        l.ige_mailbox_destroy.argtypes = [C.c_void_p]
        l.ige_set_use_tma.restype = None
        l.ige_set_use_tma.argtypes = [C.c_int]
        l.ige_device_pci_bus_id.restype = C.c_int
        l.ige_device_pci_bus_id.argtypes = [C.c_int, C.c_char_p, C.c_int]
        l.ige_set_use_pdl.restype = None
        l.ige_set_use_pdl.argtypes = [C.c_int]
        l.ige_set_fast_paths.restype = None
        l.ige_set_fast_paths.argtypes = [C.c_int]
        for name, structs in _SIGS.items():
            fn = getattr(l, name)
            fn.restype = C.c_int
            fn.argtypes = [C.POINTER(s) for s in structs] + [C.c_void_p]
        if l.ige_version() != 1:
            raise IgeError(f"ABI version mismatch: library {l.ige_version()}, binding 1")
        _lib = l
    return _lib


def exported_symbols():
    return ["ige_version", "ige_last_error_string", "ige_launch_count", "ige_compact_scratch_words",
            "ige_reset_list_scratch_words", "ige_rollout_scratch_words", "ige_anymal_scratch_words", "ige_set_use_tma", "ige_set_fast_paths", "ige_set_use_pdl", "ige_device_pci_bus_id",
            "ige_mailbox_bytes", "ige_mailbox_create", "ige_mailbox_open", "ige_mailbox_close", "ige_mailbox_destroy", *_SIGS.keys()]


def call(name, *structs, stream=0):
    """Invoke an entry point; raises IgeError with ige_last_error_string() on failure."""
    l = lib()
    rc = getattr(l, name)(*[C.byref(s) for s in structs], C.c_void_p(stream))
    if rc != 0:
        msg = l.ige_last_error_string().decode()
        raise IgeError(f"{name} failed (rc={rc}): {msg}")


def launch_count():
    return int(lib().ige_launch_count())
