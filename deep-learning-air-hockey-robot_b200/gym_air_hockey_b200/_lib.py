"""ctypes binding of lib/libahk.so (the C ABI in include/ahk.h).

There is deliberately no fallback: if the CUDA library is missing or fails to load, importing the engine raises.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(HERE), "lib", "libahk.so")

ACT_NONE = 255
EP_NORMAL, EP_MISSING, EP_RANDOM, EP_SAMPLE = 0, 1, 2, 255
USE_ARM, USE_PUCK, USE_TOP_MALLET, USE_ALL = 1, 2, 4, 7
OBS_U8, OBS_RGB_F32, OBS_GRAY_DIFF_F64 = 0, 1, 2
STATE_WORDS, STATS_WORDS, FRAME_BYTES, DIM = 16, 32, 3 * 128 * 128, 128


class Config(C.Structure):  # ahk_config
    _fields_ = [("n_envs", C.c_int32), ("device", C.c_int32), ("seed", C.c_uint64), ("env_id_base", C.c_uint32),
                ("auto_reset", C.c_int32), ("max_episode_steps", C.c_int32), ("ring_depth", C.c_int32), ("reserved", C.c_int32 * 4)]


class Buffers(C.Structure):  # ahk_buffers
    _fields_ = [(name, C.c_void_p) for name in
                ("state", "frames", "reward", "done", "robot_label", "human_label", "info", "stats", "drawlist", "warm")]


class Stack(C.Structure):  # ahk_stack
    _fields_ = [("frames", C.c_void_p), ("warm", C.c_void_p), ("meta", C.c_void_p), ("n", C.c_int32), ("depth", C.c_int32), ("head", C.c_int32)]


class StepLog(C.Structure):  # ahk_step_log
    _fields_ = [("robot_label", C.c_void_p), ("human_label", C.c_void_p), ("done", C.c_void_p), ("reward", C.c_void_p)]


class PolicyWeights(C.Structure):  # ahk_policy_weights
    NAMES = ("conv1_w", "conv1_b", "bn1_gamma", "bn1_beta", "bn1_mean", "bn1_var",
             "conv2_w", "conv2_b", "bn2_gamma", "bn2_beta", "bn2_mean", "bn2_var",
             "conv3_w", "conv3_b", "bn3_gamma", "bn3_beta", "bn3_mean", "bn3_var",
             "dense1_w", "dense1_b", "dense2_w", "dense2_b", "out_w", "out_b")
    _fields_ = [(name, C.c_void_p) for name in NAMES]


class PolicyActs(C.Structure):  # ahk_policy_acts
    _fields_ = [("act1", C.c_void_p), ("act2", C.c_void_p), ("act3", C.c_void_p), ("h1", C.c_void_p), ("h2", C.c_void_p), ("max_batch", C.c_int32)]


class ParamSegment(C.Structure):  # ahk_param_segment
    _fields_ = [("offset", C.c_int32), ("count", C.c_int32), ("flags", C.c_int32), ("stat_offset", C.c_int32)]


SEG_TRAINABLE, SEG_L2, SEG_BN_MEAN, SEG_BN_VAR, BN_STATS = 1, 2, 4, 8, 104


class GrayPolicyWeights(C.Structure):  # ahk_policy_gray_weights
    NAMES = tuple("%s%d_%s" % (k, i, f) if k == "bn" else "conv%d_%s" % (i, f)
                  for i in (1, 2, 3, 4) for k, f in (("conv", "w"), ("conv", "b"), ("bn", "gamma"), ("bn", "beta"), ("bn", "mean"), ("bn", "var"))) + \
        ("dense1_w", "dense1_b", "out_w", "out_b")
    _fields_ = [(name, C.c_void_p) for name in NAMES]


# every symbol include/ahk.h declares: name -> (restype, argtypes)
_vp, _i = C.c_void_p, C.c_int
SYMBOLS = {
    "ahk_create": (_i, [C.POINTER(Config), C.POINTER(_vp)]),
    "ahk_destroy": (None, [_vp]),
    "ahk_set_table_image": (_i, [_vp, _vp, _i, _i]),
    "ahk_bind_buffers": (_i, [_vp, C.POINTER(Buffers)]),
    "ahk_reset": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "ahk_step": (_i, [_vp, _vp, _vp, _vp]),
    "ahk_step_many": (_i, [_vp, _i, C.POINTER(StepLog), _vp]),
    "ahk_step_profiled": (_i, [_vp, _vp, _vp, _vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "ahk_step_host": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ahk_step_host_begin": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ahk_step_host_wait": (_i, [_vp, _i]),
    "ahk_step_host_results": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ahk_host_sync": (_i, [_vp]),
    "ahk_observe": (_i, [_vp, _i, _vp, _vp]),
    "ahk_head": (_i, [_vp]),
    "ahk_ring_depth": (_i, [_vp]),
    "ahk_stack_view": (_i, [_vp, C.POINTER(Stack)]),
    "ahk_policy_forward_stack": (_i, [_vp, C.POINTER(Stack), _vp, _vp, _vp]),
    "ahk_policy_forward_stack_profiled": (_i, [_vp, C.POINTER(Stack), _vp, _vp, _vp, C.POINTER(C.c_float)]),
    "ahk_set_head": (_i, [_vp, _i]),
    "ahk_set_pending": (_i, [_vp, _vp, _vp]),
    "ahk_launch_count": (_i, [_vp, C.POINTER(C.c_uint64)]),
    "ahk_set_trace": (_i, [_vp, _vp]),
    "ahk_trace_shape": (_i, [_vp, C.POINTER(_i), C.POINTER(_i)]),
    "ahk_get_partition": (_i, [_vp, C.POINTER(C.c_int32), _i]),
    "ahk_last_error": (C.c_char_p, []),
    "ahk_version": (C.c_char_p, []),
    "ahk_policy_create": (_i, [_i, _i, C.POINTER(PolicyWeights), C.POINTER(_vp)]),
    "ahk_policy_gray_create": (_i, [_i, _i, C.POINTER(GrayPolicyWeights), C.POINTER(_vp)]),
    "ahk_policy_set_weights": (_i, [_vp, C.POINTER(PolicyWeights)]),
    "ahk_policy_set_weights_dev": (_i, [_vp, C.POINTER(PolicyWeights), _vp]),
    "ahk_replay_gather": (_i, [_vp, _i, _i, _vp, _vp, _i, _vp, _vp]),
    "ahk_replay_record": (_i, [_vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ahk_rl_epsilon_greedy": (_i, [_vp, _i, C.c_float, C.c_uint64, C.c_uint32, C.c_uint32, _vp, _vp, _vp]),
    "ahk_rl_ddqn_targets": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, C.c_double, _vp, _vp, _vp, _vp, _vp]),
    "ahk_rl_pg_returns": (_i, [_vp, _vp, _i, _i, C.c_double, _vp, _vp]),
    "ahk_rl_pg_targets": (_i, [_vp, _vp, _vp, _i, C.c_float, _vp, _vp, _vp, _vp, _vp]),
    "ahk_policy_forward_train": (_i, [_vp, C.POINTER(PolicyWeights), _vp, _i, _i, _vp, _vp, _vp]),
    "ahk_policy_get_activations": (_i, [_vp, C.POINTER(PolicyActs)]),
    "ahk_policy_copy_activation": (_i, [_vp, _i, _vp, C.c_size_t, _vp]),
    "ahk_trainer_create": (_i, [_i, _i, C.POINTER(_vp)]),
    "ahk_trainer_destroy": (None, [_vp]),
    "ahk_train_backward": (_i, [_vp, _vp, C.POINTER(PolicyWeights), _vp, _i, _i, _vp, _vp, C.POINTER(PolicyWeights), _vp]),
    "ahk_train_sgd_step": (_i, [_vp, _vp, _vp, _vp, C.POINTER(ParamSegment), _i, C.c_float, C.c_float, C.c_float, C.c_float, _vp, C.c_float, _vp, _vp]),
    "ahk_train_adam_step": (_i, [_vp, _vp, _vp, _vp, _vp, C.POINTER(ParamSegment), _i, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, _vp, C.c_float, _vp, _vp]),
    "ahk_trainer_launches": (C.c_uint64, [_vp]),
    "ahk_trainer_copy_buffer": (_i, [_vp, _i, _vp, C.c_size_t, _vp]),
    "ahk_policy_destroy": (None, [_vp]),
    "ahk_policy_forward": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "ahk_policy_forward_profiled": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp, C.POINTER(C.c_float)]),
    "ahk_policy_forward_static": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
}

_lib = None


def load():
    """Load libahk.so and declare its signatures.  Raises if the library is absent: no fallback exists."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s is missing: build it with `python __graft_entry__.py` (or make -C deep-learning-air-hockey-robot_b200/csrc). "
            "The engine has no CPU or PyTorch fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (restype, argtypes) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the ABI and the binding disagree
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(rc):
    """Translate an ahk_status into the exception the reference's callers would expect."""
    if rc == 0:
        return
    msg = load().ahk_last_error().decode()
    if rc == -1:
        raise ValueError(msg)  # e.g. 'Processing mode is not supported' (processor.py:66)
    raise RuntimeError("libahk: %s (status %d)" % (msg, rc))
