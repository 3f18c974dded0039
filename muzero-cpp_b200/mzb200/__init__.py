"""mzb200 — Python host mirror of the reference interface for the self-play search hot path, over the C ABI of
libmzb200.so (include/mzb200.h).  ctypes only: no torch on the product path.

Names follow the reference: MuZeroConfig / MuZeroNetworkConfig fields (muzero-cpp/config.h:15-117), MCTSReturn
(types.h:70-76), VPRNetModel.InitialInference / RecurrentInference (vprnet.h:71-100), MCTS.run (mcts.h:133-136) —
batched: every call takes T trees / B samples at once.  The library must be present: there is no CPU fallback."""
import ctypes as C
import os
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MZB200_LIB") or os.path.join(os.path.dirname(_HERE), "libmzb200.so")   # MZB200_LIB: another build of the same library (kernel experiments)

MAX_ACTIONS = 32
MAX_HEAD_LAYERS = 8
FP32_CHECK, BF16 = 0, 1
EVAL_NETWORK, EVAL_SYNTHETIC, EVAL_HOST, EVAL_SYNTHETIC_OBS = 0, 1, 2, 3
INF_D = float(np.finfo(np.float64).max)    # types::INF_D  (types.h:78)
NINF_D = -float(np.finfo(np.float64).max)  # types::NINF_D (types.h:77)


class MzbError(RuntimeError):
    pass


class _CConfig(C.Structure):
    _fields_ = [
        ("observation_c", C.c_int32), ("observation_h", C.c_int32), ("observation_w", C.c_int32),
        ("num_actions", C.c_int32), ("action_channels", C.c_int32), ("num_players", C.c_int32),
        ("stacked_observations", C.c_int32),
        ("value_upperbound", C.c_double), ("value_lowerbound", C.c_double),
        ("min_reward", C.c_double), ("max_reward", C.c_double), ("min_value", C.c_double), ("max_value", C.c_double),
        ("use_contractive_mapping", C.c_int32),
        ("num_simulations", C.c_int32),
        ("discount", C.c_double), ("dirichlet_alpha", C.c_double), ("dirichlet_epsilon", C.c_double),
        ("pb_c_base", C.c_double), ("pb_c_init", C.c_double),
        ("seed", C.c_int32),
        ("downsample", C.c_int32), ("normalize_hidden_states", C.c_int32),
        ("resnet_channels", C.c_int32), ("representation_blocks", C.c_int32), ("dynamics_blocks", C.c_int32),
        ("prediction_blocks", C.c_int32),
        ("reward_reduced_channels", C.c_int32), ("policy_reduced_channels", C.c_int32),
        ("value_reduced_channels", C.c_int32),
        ("n_reward_head_layers", C.c_int32), ("n_policy_head_layers", C.c_int32), ("n_value_head_layers", C.c_int32),
        ("reward_head_layers", C.c_int32 * MAX_HEAD_LAYERS), ("policy_head_layers", C.c_int32 * MAX_HEAD_LAYERS),
        ("value_head_layers", C.c_int32 * MAX_HEAD_LAYERS),
        ("action_planes", C.POINTER(C.c_float)),
        ("max_trees", C.c_int32), ("precision", C.c_int32), ("evaluator", C.c_int32),
        ("se_value_lo", C.c_double), ("se_value_hi", C.c_double), ("se_reward_lo", C.c_double),
        ("se_reward_hi", C.c_double),
    ]


class _CTensorDesc(C.Structure):
    _fields_ = [("name", C.c_char_p), ("data", C.POINTER(C.c_float)), ("ndim", C.c_int32), ("dims", C.c_int64 * 4)]


class _CSearchResult(C.Structure):
    _fields_ = [("root_value", C.POINTER(C.c_double)), ("max_tree_depth", C.POINTER(C.c_int32)),
                ("n_children", C.POINTER(C.c_int32)), ("child_actions", C.POINTER(C.c_int32)),
                ("child_visits", C.POINTER(C.c_int32)), ("children_relative_visit", C.POINTER(C.c_double)),
                ("emperical_policy", C.POINTER(C.c_double))]


class _CBatch(C.Structure):   # mzb_batch = types::Batch (types.h:81-90)
    _fields_ = [("num_samples", C.c_int32), ("actions", C.POINTER(C.c_int32)), ("stacked_observations", C.POINTER(C.c_float)),
                ("target_values", C.POINTER(C.c_double)), ("target_rewards", C.POINTER(C.c_double)),
                ("target_policies", C.POINTER(C.c_double)), ("gradient_scale", C.POINTER(C.c_double)),
                ("priorities", C.POINTER(C.c_double))]


class _CLossInfo(C.Structure):   # mzb_loss_info = VPRNetModel::LossInfo (vprnet.h:33-50)
    _fields_ = [("total_loss", C.c_double), ("value", C.c_double), ("policy", C.c_double), ("reward", C.c_double),
                ("errors", C.POINTER(C.c_double))]


_RNG_FN = C.CFUNCTYPE(C.c_uint32, C.c_void_p)   # mzb_rng_fn: the caller's generator, 32 bits per draw (std::mt19937::operator())


class _CReplayConfig(C.Structure):   # mzb_replay_config
    _fields_ = [("capacity", C.c_int32), ("min_sample_size", C.c_int32), ("td_steps", C.c_int32), ("num_unroll_steps", C.c_int32),
                ("per_alpha", C.c_double), ("per_beta", C.c_double), ("per_epsilon", C.c_double), ("per_beta_increment", C.c_double),
                ("action_planes_initial", C.POINTER(C.c_float))]


class _CGameHistory(C.Structure):    # mzb_game_history = the fields of types::GameHistory the buffer reads (types.h:43-78)
    _fields_ = [("length", C.c_int32), ("observations", C.POINTER(C.c_float)), ("actions", C.POINTER(C.c_int32)),
                ("rewards", C.POINTER(C.c_double)), ("to_play", C.POINTER(C.c_int32)), ("root_values", C.POINTER(C.c_double)),
                ("child_visits", C.POINTER(C.c_double))]


class _CBatchOut(C.Structure):       # mzb_batch_out = types::Batch (types.h:81-104)
    _fields_ = [("num_samples", C.c_int32), ("priorities", C.POINTER(C.c_double)), ("indices", C.POINTER(C.c_int32)),
                ("actions", C.POINTER(C.c_int32)), ("stacked_observations", C.POINTER(C.c_float)),
                ("target_rewards", C.POINTER(C.c_double)), ("target_values", C.POINTER(C.c_double)),
                ("target_policies", C.POINTER(C.c_double)), ("gradient_scale", C.POINTER(C.c_double))]


class _CSearchTiming(C.Structure):
    _fields_ = [("total_ms", C.c_float), ("root_ms", C.c_float), ("sim_loop_ms", C.c_float),
                ("kernel_launches", C.c_int64), ("simulations", C.c_int64)]


class _CShapes(C.Structure):
    _fields_ = [("initial_channels", C.c_int32), ("initial_flat_size", C.c_int32), ("enc_c", C.c_int32),
                ("enc_h", C.c_int32), ("enc_w", C.c_int32), ("recurrent_flat_size", C.c_int32),
                ("value_support_size", C.c_int32), ("reward_support_size", C.c_int32),
                ("value_step", C.c_double), ("value_min", C.c_double), ("reward_step", C.c_double),
                ("reward_min", C.c_double), ("weight_blob_bytes", C.c_int64), ("node_pool_bytes", C.c_int64)]


@dataclass
class MuZeroNetworkConfig:  # config.h:15-31
    downsample: bool = False
    normalize_hidden_states: bool = True
    resnet_channels: int = 128
    representation_blocks: int = 4
    dynamics_blocks: int = 4
    prediction_blocks: int = 4
    reward_reduced_channels: int = 16
    policy_reduced_channels: int = 16
    value_reduced_channels: int = 16
    reward_head_layers: List[int] = field(default_factory=lambda: [64])
    policy_head_layers: List[int] = field(default_factory=lambda: [64])
    value_head_layers: List[int] = field(default_factory=lambda: [64])


@dataclass
class MuZeroConfig:  # config.h:33-117 (hot-path fields) + engine knobs
    observation_shape: tuple = (3, 6, 7)
    num_actions: int = 7                 # action_space = range(num_actions)
    action_channels: int = 1
    num_players: int = 2
    stacked_observations: int = 0
    value_upperbound: float = NINF_D
    value_lowerbound: float = INF_D
    min_reward: float = -300.0
    max_reward: float = 300.0
    min_value: float = -300.0
    max_value: float = 300.0
    use_contractive_mapping: bool = True
    num_simulations: int = 50
    discount: float = 0.997
    dirichlet_alpha: float = 0.3
    dirichlet_epsilon: float = 0.25
    pb_c_base: float = 19652.0
    pb_c_init: float = 1.25
    seed: int = 0
    network_config: MuZeroNetworkConfig = field(default_factory=MuZeroNetworkConfig)
    action_planes: Optional[np.ndarray] = None   # action_representation_recurrent pre-evaluated [A][ach*h*w]
    # engine
    max_trees: int = 1024
    precision: int = FP32_CHECK
    evaluator: int = EVAL_NETWORK
    se_ranges: tuple = (-1.0, 1.0, -0.5, 0.5)


@dataclass
class MCTSReturn:  # T x types::MCTSReturn (types.h:70-76), SoA
    root_value: np.ndarray
    max_tree_depth: np.ndarray
    n_children: np.ndarray
    child_actions: np.ndarray
    child_visits: np.ndarray
    children_relative_visit: np.ndarray
    emperical_policy: np.ndarray


@dataclass
class InferenceOutputs:  # B x VPRNetModel::InferenceOutputs (vprnet.h:64-69), SoA
    value: np.ndarray
    reward: np.ndarray
    policy: np.ndarray
    encoded_state: np.ndarray


# mzb_host_evaluator (MZB_EVAL_HOST): the two pure virtuals of the reference's Evaluator as batched callbacks
_HOST_INITIAL = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_float), C.c_int,
                            C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double))
_HOST_RECURRENT = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_int,
                              C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double))


class _CHostEvaluator(C.Structure):
    _fields_ = [("user", C.c_void_p), ("initial", _HOST_INITIAL), ("recurrent", _HOST_RECURRENT)]


_lib = None

EXPORTS = ["mzb_last_error", "mzb_version", "mzb_engine_create", "mzb_engine_destroy", "mzb_load_weights",
           "mzb_nccl_unique_id", "mzb_nccl_init", "mzb_broadcast_weights", "mzb_broadcast_weights_begin", "mzb_weights_commit",
           "mzb_weights_checksum",
           "mzb_initial_inference", "mzb_recurrent_inference", "mzb_bench_recurrent", "mzb_search_seed", "mzb_search_batch_slots",
           "mzb_search_batch", "mzb_search_batch_device_obs", "mzb_last_search_timing", "mzb_search_set_recording",
           "mzb_search_get_trace", "mzb_search_get_records", "mzb_get_shapes", "mzb_value_encode", "mzb_value_decode",
           "mzb_value_decode_static",
           "mzb_obs_ring_create", "mzb_obs_ring_reset", "mzb_obs_ring_push", "mzb_obs_ring_stacked", "mzb_search_batch_ring",
           "mzb_actor_seed", "mzb_select_actions", "mzb_actor_uniform_int", "mzb_set_host_evaluator",
           "mzb_checkpoint_open", "mzb_checkpoint_count", "mzb_checkpoint_tensor", "mzb_checkpoint_close", "mzb_load_checkpoint",
           "mzb_checkpoint_long_count", "mzb_checkpoint_long", "mzb_checkpoint_write", "mzb_save_checkpoint",
           "mzb_learn_forward", "mzb_learn_get_tensor", "mzb_learn_step", "mzb_learn_get_gradient", "mzb_learn_commit",
           "mzb_replay_create", "mzb_replay_destroy", "mzb_replay_save_game_history", "mzb_replay_size", "mzb_replay_can_sample",
           "mzb_replay_num_histories", "mzb_replay_total_priority", "mzb_replay_oldest_history", "mzb_replay_sample", "mzb_replay_update_priorities",
           "mzb_replay_sample_game", "mzb_replay_update_game_history",
           "mzb_debug_se_state", "mzb_debug_tree_write", "mzb_debug_tree_read", "mzb_debug_select_child", "mzb_debug_backpropagate"]


def load_library(path=None):
    """dlopen libmzb200.so.  Fails loudly when the CUDA extension is missing: there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise MzbError("libmzb200.so not found at %s — build it (python __graft_entry__.py / muzero-cpp_b200/build.py)" % path)
    lib = C.CDLL(path)
    lib.mzb_last_error.restype = C.c_char_p
    lib.mzb_version.restype = C.c_char_p
    lib.mzb_engine_create.argtypes = [C.POINTER(_CConfig), C.c_int, C.POINTER(C.c_void_p)]
    lib.mzb_engine_destroy.argtypes = [C.c_void_p]
    lib.mzb_engine_destroy.restype = None
    lib.mzb_load_weights.argtypes = [C.c_void_p, C.POINTER(_CTensorDesc), C.c_int]
    lib.mzb_nccl_unique_id.argtypes = [C.c_char_p]
    lib.mzb_nccl_init.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int]
    lib.mzb_broadcast_weights.argtypes = [C.c_void_p, C.c_int]
    lib.mzb_broadcast_weights_begin.argtypes = [C.c_void_p, C.c_int, C.c_int]
    lib.mzb_weights_commit.argtypes = [C.c_void_p]
    lib.mzb_weights_checksum.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    dp, fp, ip = C.POINTER(C.c_double), C.POINTER(C.c_float), C.POINTER(C.c_int32)
    lib.mzb_initial_inference.argtypes = [C.c_void_p, fp, C.c_int, dp, dp, dp, fp]
    lib.mzb_recurrent_inference.argtypes = [C.c_void_p, fp, ip, C.c_int, dp, dp, dp, fp]
    lib.mzb_bench_recurrent.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_int)]
    lib.mzb_search_seed.argtypes = [C.c_void_p, C.c_int, C.c_int, ip]
    lib.mzb_search_batch.argtypes = [C.c_void_p, fp, C.POINTER(C.c_uint32), ip, C.POINTER(C.c_uint8), C.c_int,
                                     C.POINTER(_CSearchResult)]
    lib.mzb_search_batch_slots.argtypes = [C.c_void_p, fp, ip, C.POINTER(C.c_uint32), ip, C.POINTER(C.c_uint8), C.c_int,
                                           C.POINTER(_CSearchResult)]
    lib.mzb_search_batch_device_obs.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint32), ip,
                                                C.POINTER(C.c_uint8), C.c_int, C.POINTER(_CSearchResult)]
    lib.mzb_last_search_timing.argtypes = [C.c_void_p, C.POINTER(_CSearchTiming)]
    lib.mzb_search_set_recording.argtypes = [C.c_void_p, C.c_int]
    lib.mzb_search_get_trace.argtypes = [C.c_void_p, C.c_int, ip, ip]
    lib.mzb_search_get_records.argtypes = [C.c_void_p, C.c_int, dp, dp, dp]
    lib.mzb_get_shapes.argtypes = [C.c_void_p, C.POINTER(_CShapes)]
    u32p, u8p = C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)
    lib.mzb_obs_ring_create.argtypes = [C.c_void_p, fp]
    lib.mzb_obs_ring_reset.argtypes = [C.c_void_p, ip, C.c_int]
    lib.mzb_obs_ring_push.argtypes = [C.c_void_p, ip, fp, ip, C.c_int]
    lib.mzb_obs_ring_stacked.argtypes = [C.c_void_p, ip, C.c_int, fp]
    lib.mzb_search_batch_ring.argtypes = [C.c_void_p, ip, u32p, ip, u8p, C.c_int, C.POINTER(_CSearchResult)]
    lib.mzb_actor_seed.argtypes = [C.c_void_p, ip, ip, C.c_int]
    lib.mzb_select_actions.argtypes = [C.c_void_p, ip, C.POINTER(_CSearchResult), C.c_int, dp, ip]
    lib.mzb_checkpoint_open.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    lib.mzb_checkpoint_count.argtypes = [C.c_void_p]
    lib.mzb_checkpoint_tensor.argtypes = [C.c_void_p, C.c_int, C.POINTER(_CTensorDesc)]
    lib.mzb_checkpoint_close.argtypes = [C.c_void_p]
    lib.mzb_checkpoint_close.restype = None
    lib.mzb_load_checkpoint.argtypes = [C.c_void_p, C.c_char_p]
    lib.mzb_save_checkpoint.argtypes = [C.c_void_p, C.c_char_p]
    lib.mzb_learn_forward.argtypes = [C.c_void_p, C.POINTER(_CBatch), C.c_int, C.c_double, C.POINTER(_CLossInfo)]
    lib.mzb_learn_get_tensor.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_float), C.c_int64]
    lib.mzb_learn_get_gradient.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_float), C.c_int64]
    lib.mzb_learn_step.argtypes = [C.c_void_p, C.POINTER(_CBatch), C.c_int, C.c_double, C.c_double, C.c_double, C.POINTER(_CLossInfo)]
    lib.mzb_learn_commit.argtypes = [C.c_void_p]
    lib.mzb_replay_create.argtypes = [C.c_void_p, C.POINTER(_CReplayConfig), C.POINTER(C.c_void_p)]
    lib.mzb_replay_destroy.argtypes = [C.c_void_p]
    lib.mzb_replay_destroy.restype = None
    lib.mzb_replay_save_game_history.argtypes = [C.c_void_p, C.POINTER(_CGameHistory)]
    for f in ("mzb_replay_size", "mzb_replay_can_sample", "mzb_replay_num_histories", "mzb_replay_oldest_history"):
        getattr(lib, f).argtypes = [C.c_void_p]
    lib.mzb_replay_total_priority.argtypes = [C.c_void_p]
    lib.mzb_replay_total_priority.restype = C.c_double
    lib.mzb_replay_sample.argtypes = [C.c_void_p, _RNG_FN, C.c_void_p, C.c_int, C.POINTER(_CBatchOut)]
    lib.mzb_replay_update_priorities.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_double), C.c_int]
    lib.mzb_replay_sample_game.argtypes = [C.c_void_p, _RNG_FN, C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.mzb_replay_update_game_history.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int]
    lib.mzb_checkpoint_write.argtypes = [C.c_char_p, C.POINTER(_CTensorDesc), C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int64), C.c_int]
    lib.mzb_checkpoint_long_count.argtypes = [C.c_void_p]
    lib.mzb_checkpoint_long.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int64)]
    lib.mzb_set_host_evaluator.argtypes = [C.c_void_p, C.POINTER(_CHostEvaluator)]
    lib.mzb_value_encode.argtypes = [C.c_void_p, C.c_int, fp, C.c_int, fp]
    lib.mzb_value_decode.argtypes = [C.c_void_p, C.c_int, fp, C.c_int, C.c_int, fp]
    lib.mzb_value_decode_static.argtypes = [C.c_void_p, C.c_int, fp, C.c_int, C.c_int, fp]
    _lib = lib
    return lib


def _ptr(a, ct):
    return a.ctypes.data_as(C.POINTER(ct))


def read_checkpoint(path):
    """The float32 tensors of a reference checkpoint file (torch::save module archive, vprnet.cpp:58-63), parsed natively by
    libmzb200 (no torch): {dotted name: ndarray}.  Needs no GPU."""
    lib = load_library()
    h = C.c_void_p()
    if lib.mzb_checkpoint_open(os.fsencode(path), C.byref(h)) != 0:
        raise MzbError(lib.mzb_last_error().decode())
    try:
        out = {}
        for i in range(lib.mzb_checkpoint_count(h)):
            d = _CTensorDesc()
            if lib.mzb_checkpoint_tensor(h, i, C.byref(d)) != 0:
                raise MzbError(lib.mzb_last_error().decode())
            shape = tuple(int(d.dims[k]) for k in range(d.ndim))
            n = int(np.prod(shape)) if shape else 1
            out[d.name.decode(errors="replace")] = np.ctypeslib.as_array(d.data, (n,)).reshape(shape).copy()
        return out
    finally:
        lib.mzb_checkpoint_close(h)


def write_checkpoint(path, tensors: dict, longs: dict = None):
    """VPRNetModel::SaveCheckpoint's write half (vprnet.cpp:58-63) without an engine: named float32 tensors (+ optional int64
    scalars such as '...num_batches_tracked') -> the module archive torch::load reads (mzb_checkpoint_write)."""
    lib = load_library()
    keep = []
    descs = (_CTensorDesc * len(tensors))()
    for i, (name, arr) in enumerate(tensors.items()):
        a = np.ascontiguousarray(arr, dtype=np.float32)
        keep.append(a)
        descs[i].name = name.encode()
        descs[i].data = _ptr(a, C.c_float)
        descs[i].ndim = a.ndim
        for k in range(a.ndim):
            descs[i].dims[k] = a.shape[k]
    longs = longs or {}
    names = (C.c_char_p * max(1, len(longs)))(*[k.encode() for k in longs])
    values = (C.c_int64 * max(1, len(longs)))(*[int(v) for v in longs.values()])
    rc = lib.mzb_checkpoint_write(os.fsencode(path), descs, len(tensors), names, values, len(longs))
    if rc != 0:
        raise MzbError("mzb200 error %d: %s" % (rc, lib.mzb_last_error().decode()))


def read_checkpoint_longs(path):
    """the int64 scalars of a checkpoint archive (BatchNorm num_batches_tracked) as {name: value}"""
    lib = load_library()
    h = C.c_void_p()
    rc = lib.mzb_checkpoint_open(os.fsencode(path), C.byref(h))
    if rc != 0:
        raise MzbError("mzb200 error %d: %s" % (rc, lib.mzb_last_error().decode()))
    try:
        out = {}
        for i in range(lib.mzb_checkpoint_long_count(h)):
            name = C.c_char_p(); val = C.c_int64()
            lib.mzb_checkpoint_long(h, i, C.byref(name), C.byref(val))
            out[name.value.decode()] = val.value
        return out
    finally:
        lib.mzb_checkpoint_close(h)


class Engine:
    """One engine per GPU: T tree slots (each with its own std::mt19937 stream, seed + slot), the node pool and the
    network.  Stands where the reference has N x MCTS + VPRNetEvaluator + DeviceManager + VPRNetModel."""

    def __init__(self, config: MuZeroConfig, device: int = 0):
        self.lib = load_library()
        self.config = config
        nc = config.network_config
        cc = _CConfig()
        cc.observation_c, cc.observation_h, cc.observation_w = config.observation_shape
        cc.num_actions = config.num_actions
        cc.action_channels = config.action_channels
        cc.num_players = config.num_players
        cc.stacked_observations = config.stacked_observations
        cc.value_upperbound, cc.value_lowerbound = config.value_upperbound, config.value_lowerbound
        cc.min_reward, cc.max_reward = config.min_reward, config.max_reward
        cc.min_value, cc.max_value = config.min_value, config.max_value
        cc.use_contractive_mapping = int(config.use_contractive_mapping)
        cc.num_simulations = config.num_simulations
        cc.discount, cc.dirichlet_alpha, cc.dirichlet_epsilon = config.discount, config.dirichlet_alpha, config.dirichlet_epsilon
        cc.pb_c_base, cc.pb_c_init = config.pb_c_base, config.pb_c_init
        cc.seed = config.seed
        cc.downsample = int(nc.downsample)
        cc.normalize_hidden_states = int(nc.normalize_hidden_states)
        cc.resnet_channels = nc.resnet_channels
        cc.representation_blocks, cc.dynamics_blocks, cc.prediction_blocks = nc.representation_blocks, nc.dynamics_blocks, nc.prediction_blocks
        cc.reward_reduced_channels, cc.policy_reduced_channels, cc.value_reduced_channels = nc.reward_reduced_channels, nc.policy_reduced_channels, nc.value_reduced_channels
        for name in ("reward", "policy", "value"):
            layers = list(getattr(nc, name + "_head_layers"))
            if len(layers) > MAX_HEAD_LAYERS:
                raise MzbError("too many head layers")
            setattr(cc, "n_%s_head_layers" % name, len(layers))
            arr = getattr(cc, name + "_head_layers")
            for i, v in enumerate(layers):
                arr[i] = v
        self._planes = None
        if config.action_planes is not None:
            self._planes = np.ascontiguousarray(config.action_planes, dtype=np.float32)
            cc.action_planes = _ptr(self._planes, C.c_float)
        cc.max_trees, cc.precision, cc.evaluator = config.max_trees, config.precision, config.evaluator
        cc.se_value_lo, cc.se_value_hi, cc.se_reward_lo, cc.se_reward_hi = config.se_ranges
        self._h = C.c_void_p()
        self._check(self.lib.mzb_engine_create(C.byref(cc), device, C.byref(self._h)))
        sh = _CShapes()
        self._check(self.lib.mzb_get_shapes(self._h, C.byref(sh)))
        self.shapes = sh

    def _check(self, rc):
        if rc != 0:
            raise MzbError("mzb200 error %d: %s" % (rc, self.lib.mzb_last_error().decode()))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self.lib.mzb_engine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def get_shapes(self):
        sh = _CShapes()
        self._check(self.lib.mzb_get_shapes(self._h, C.byref(sh)))
        return sh

    # ---- the callers' side: device-resident observation history (types.cpp:16-42) and action selection (self_play.cpp:21-39)
    def ring_create(self, action_planes_initial=None):
        if action_planes_initial is not None:
            self._planes_initial = np.ascontiguousarray(action_planes_initial, dtype=np.float32)
            self._check(self.lib.mzb_obs_ring_create(self._h, _ptr(self._planes_initial, C.c_float)))
        else:
            self._check(self.lib.mzb_obs_ring_create(self._h, None))

    def ring_reset(self, slots):
        sl = np.ascontiguousarray(slots, dtype=np.int32).reshape(-1)
        self._check(self.lib.mzb_obs_ring_reset(self._h, _ptr(sl, C.c_int32), sl.shape[0]))

    def ring_push(self, slots, frames, actions):
        sl = np.ascontiguousarray(slots, dtype=np.int32).reshape(-1)
        fr = np.ascontiguousarray(frames, dtype=np.float32).reshape(sl.shape[0], -1)
        ac = np.ascontiguousarray(actions, dtype=np.int32).reshape(-1)
        c = self.config
        if fr.shape[1] != c.observation_shape[0] * c.observation_shape[1] * c.observation_shape[2] or ac.shape[0] != sl.shape[0]:
            raise MzbError("frames / actions disagree with the slot list or the observation shape")
        self._check(self.lib.mzb_obs_ring_push(self._h, _ptr(sl, C.c_int32), _ptr(fr, C.c_float), _ptr(ac, C.c_int32), sl.shape[0]))

    def ring_stacked(self, slots):
        sl = np.ascontiguousarray(slots, dtype=np.int32).reshape(-1)
        out = np.empty((sl.shape[0], self.shapes.initial_flat_size), np.float32)
        self._check(self.lib.mzb_obs_ring_stacked(self._h, _ptr(sl, C.c_int32), sl.shape[0], _ptr(out, C.c_float)))
        return out

    def _result_arrays(self, T):
        A = self.config.num_actions
        res = MCTSReturn(np.empty(T, np.float64), np.empty(T, np.int32), np.empty(T, np.int32),
                         np.empty((T, A), np.int32), np.empty((T, A), np.int32), np.empty((T, A), np.float64),
                         np.empty((T, A), np.float64))
        return res, self._cresult(res)

    @staticmethod
    def _cresult(res):
        return _CSearchResult(_ptr(res.root_value, C.c_double), _ptr(res.max_tree_depth, C.c_int32),
                              _ptr(res.n_children, C.c_int32), _ptr(res.child_actions, C.c_int32),
                              _ptr(res.child_visits, C.c_int32), _ptr(res.children_relative_visit, C.c_double),
                              _ptr(res.emperical_policy, C.c_double))

    def run_ring(self, slots, legal_mask, to_play, add_exploration_noise) -> MCTSReturn:
        sl = np.ascontiguousarray(slots, dtype=np.int32).reshape(-1)
        legal = np.ascontiguousarray(legal_mask, dtype=np.uint32).reshape(-1)
        tp = np.ascontiguousarray(to_play, dtype=np.int32).reshape(-1)
        noise = np.ascontiguousarray(add_exploration_noise, dtype=np.uint8).reshape(-1)
        T = sl.shape[0]
        if legal.shape[0] != T or tp.shape[0] != T or noise.shape[0] != T:
            raise MzbError("per-tree inputs disagree on the number of trees")
        res, cr = self._result_arrays(T)
        self._check(self.lib.mzb_search_batch_ring(self._h, _ptr(sl, C.c_int32), _ptr(legal, C.c_uint32), _ptr(tp, C.c_int32),
                                                   _ptr(noise, C.c_uint8), T, C.byref(cr)))
        return res

    def actor_seed(self, slots, seeds):
        sl = np.ascontiguousarray(slots, dtype=np.int32).reshape(-1)
        sd = np.ascontiguousarray(seeds, dtype=np.int32).reshape(-1)
        self._check(self.lib.mzb_actor_seed(self._h, _ptr(sl, C.c_int32), _ptr(sd, C.c_int32), sl.shape[0]))

    def select_actions(self, slots, stats: MCTSReturn, temperature):
        sl = np.ascontiguousarray(slots, dtype=np.int32).reshape(-1)
        tmp = np.ascontiguousarray(np.broadcast_to(np.asarray(temperature, np.float64), sl.shape), dtype=np.float64)
        out = np.empty(sl.shape[0], np.int32)
        cr = self._cresult(stats)
        self._check(self.lib.mzb_select_actions(self._h, _ptr(sl, C.c_int32), C.byref(cr), sl.shape[0], _ptr(tmp, C.c_double), _ptr(out, C.c_int32)))
        return out

    # ValueEncoder::encode / decode (util.cpp:59-87); which = 0 value encoder, 1 reward encoder
    def value_encode(self, x, which=0):
        x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
        S = self.shapes.value_support_size if which == 0 else self.shapes.reward_support_size
        out = np.empty((x.shape[0], S), np.float32)
        self._check(self.lib.mzb_value_encode(self._h, which, _ptr(x, C.c_float), x.shape[0], _ptr(out, C.c_float)))
        return out

    def value_decode(self, logits, which=0, take_softmax=True):
        S = self.shapes.value_support_size if which == 0 else self.shapes.reward_support_size
        lg = np.ascontiguousarray(logits, dtype=np.float32).reshape(-1, S)
        out = np.empty(lg.shape[0], np.float32)
        self._check(self.lib.mzb_value_decode(self._h, which, _ptr(lg, C.c_float), lg.shape[0], int(bool(take_softmax)), _ptr(out, C.c_float)))
        return out

    def value_decode_static(self, logits, which=0, take_softmax=True):
        """value_decode without any device allocation (static scratch, at most VALUE_DECODE_STATIC_MAX_FLOATS = n * (S + 1) floats)"""
        S = self.shapes.value_support_size if which == 0 else self.shapes.reward_support_size
        lg = np.ascontiguousarray(logits, dtype=np.float32).reshape(-1, S)
        out = np.empty(lg.shape[0], np.float32)
        self._check(self.lib.mzb_value_decode_static(self._h, which, _ptr(lg, C.c_float), lg.shape[0], int(bool(take_softmax)), _ptr(out, C.c_float)))
        return out

    # VPRNetModel::LoadCheckpoint (vprnet.cpp:66-74): named fp32 tensors of the module archive
    def load_weights(self, tensors: dict):
        keep = []
        descs = (_CTensorDesc * len(tensors))()
        for i, (name, arr) in enumerate(tensors.items()):
            a = np.ascontiguousarray(arr, dtype=np.float32)
            keep.append(a)
            descs[i].name = name.encode()
            descs[i].data = _ptr(a, C.c_float)
            descs[i].ndim = a.ndim
            for k in range(a.ndim):
                descs[i].dims[k] = a.shape[k]
        self._check(self.lib.mzb_load_weights(self._h, descs, len(tensors)))

    def load_checkpoint(self, path):
        """VPRNetModel::LoadCheckpoint's read half (vprnet.cpp:71-74): the reference's own .pt file, parsed without libtorch"""
        self._check(self.lib.mzb_load_checkpoint(self._h, os.fsencode(path)))

    def save_checkpoint(self, path):
        """VPRNetModel::SaveCheckpoint's write half (vprnet.cpp:58-63): the tensors this engine was last loaded with, as the
        reference's own .pt archive"""
        self._check(self.lib.mzb_save_checkpoint(self._h, os.fsencode(path)))

    def learn_forward(self, stacked_observations, actions, target_values, target_rewards, target_policies, priorities,
                      gradient_scale=None, value_loss_weight=0.25):
        """The forward half of VPRNetModel::Learn (vprnet.cpp:192-311): K + 1 unrolled train-mode network calls + the loss.
        Returns dict(total_loss, value, policy, reward, errors[B]) = VPRNetModel::LossInfo.  K + 1 = actions.shape[1]."""
        act = np.ascontiguousarray(actions, dtype=np.int32)
        B, K1 = act.shape
        obs = np.ascontiguousarray(stacked_observations, dtype=np.float32).reshape(B, self.shapes.initial_flat_size)
        tv = np.ascontiguousarray(target_values, dtype=np.float64).reshape(B, K1)
        tr = np.ascontiguousarray(target_rewards, dtype=np.float64).reshape(B, K1)
        tp = np.ascontiguousarray(target_policies, dtype=np.float64).reshape(B, K1, self.config.num_actions)
        pr = np.ascontiguousarray(priorities, dtype=np.float64).reshape(B)
        gs = np.ascontiguousarray(gradient_scale if gradient_scale is not None else np.full(B, 1.0 / max(1, K1 - 1)), dtype=np.float64).reshape(B)
        errors = np.empty(B, np.float64)
        b = _CBatch(B, _ptr(act, C.c_int32), _ptr(obs, C.c_float), _ptr(tv, C.c_double), _ptr(tr, C.c_double), _ptr(tp, C.c_double),
                    _ptr(gs, C.c_double), _ptr(pr, C.c_double))
        li = _CLossInfo(0.0, 0.0, 0.0, 0.0, _ptr(errors, C.c_double))
        self._check(self.lib.mzb_learn_forward(self._h, C.byref(b), K1 - 1, float(value_loss_weight), C.byref(li)))
        return dict(total_loss=li.total_loss, value=li.value, policy=li.policy, reward=li.reward, errors=errors)

    def learn_step(self, stacked_observations, actions, target_values, target_rewards, target_policies, priorities,
                   gradient_scale=None, value_loss_weight=0.25, learning_rate=3e-4, l2_weight_decay=1e-4):
        """VPRNetModel::Learn (vprnet.cpp:192-311): forward, loss, backward with Learn's gradient hooks, one Adam step.  Same
        arguments and return value as learn_forward."""
        act = np.ascontiguousarray(actions, dtype=np.int32)
        B, K1 = act.shape
        obs = np.ascontiguousarray(stacked_observations, dtype=np.float32).reshape(B, self.shapes.initial_flat_size)
        tv = np.ascontiguousarray(target_values, dtype=np.float64).reshape(B, K1)
        tr = np.ascontiguousarray(target_rewards, dtype=np.float64).reshape(B, K1)
        tp = np.ascontiguousarray(target_policies, dtype=np.float64).reshape(B, K1, self.config.num_actions)
        pr = np.ascontiguousarray(priorities, dtype=np.float64).reshape(B)
        gs = np.ascontiguousarray(gradient_scale if gradient_scale is not None else np.full(B, 1.0 / max(1, K1 - 1)), dtype=np.float64).reshape(B)
        errors = np.empty(B, np.float64)
        b = _CBatch(B, _ptr(act, C.c_int32), _ptr(obs, C.c_float), _ptr(tv, C.c_double), _ptr(tr, C.c_double), _ptr(tp, C.c_double),
                    _ptr(gs, C.c_double), _ptr(pr, C.c_double))
        li = _CLossInfo(0.0, 0.0, 0.0, 0.0, _ptr(errors, C.c_double))
        self._check(self.lib.mzb_learn_step(self._h, C.byref(b), K1 - 1, float(value_loss_weight), float(learning_rate), float(l2_weight_decay), C.byref(li)))
        return dict(total_loss=li.total_loss, value=li.value, policy=li.policy, reward=li.reward, errors=errors)

    def learn_gradient(self, name, shape):
        """the gradient the last learn_step left on a parameter"""
        out = np.empty(shape, np.float32)
        self._check(self.lib.mzb_learn_get_gradient(self._h, name.encode(), _ptr(out, C.c_float), int(out.size)))
        return out

    def learn_commit(self):
        """the trained tensors become the engine's search weights (and what save_checkpoint writes)"""
        self._check(self.lib.mzb_learn_commit(self._h))

    def learn_tensor(self, name, shape):
        """a tensor of the training state (BatchNorm running statistics as moved by learn_forward calls)"""
        out = np.empty(shape, np.float32)
        self._check(self.lib.mzb_learn_get_tensor(self._h, name.encode(), _ptr(out, C.c_float), int(out.size)))
        return out

    def weights_checksum(self):
        out = C.c_uint64()
        self._check(self.lib.mzb_weights_checksum(self._h, C.byref(out)))
        return out.value

    # learner.cpp:93-101 weight sync
    @staticmethod
    def nccl_unique_id():
        lib = load_library()
        buf = C.create_string_buffer(128)
        rc = lib.mzb_nccl_unique_id(buf)
        if rc != 0:
            raise MzbError(lib.mzb_last_error().decode())
        return buf.raw

    def nccl_init(self, uid: bytes, rank: int, world_size: int):
        self._check(self.lib.mzb_nccl_init(self._h, uid, rank, world_size))

    def broadcast_weights(self, root: int = 0):
        self._check(self.lib.mzb_broadcast_weights(self._h, root))

    def broadcast_weights_begin(self, root: int, rank: int):
        """phase 1 of the overlapped hand-over: receive into the shadow buffer on its own stream (searches keep running)"""
        self._check(self.lib.mzb_broadcast_weights_begin(self._h, root, rank))

    def weights_commit(self):
        """phase 2: install the received weights between two searches (enqueued on the search stream)"""
        self._check(self.lib.mzb_weights_commit(self._h))

    # VPRNetModel::InitialInference (vprnet.cpp:77-129)
    def initial_inference(self, stacked_observations) -> InferenceOutputs:
        obs = np.ascontiguousarray(stacked_observations, dtype=np.float32).reshape(-1, self.shapes.initial_flat_size)
        B, A = obs.shape[0], self.config.num_actions
        v = np.empty(B, np.float64); r = np.empty(B, np.float64); p = np.empty((B, A), np.float64)
        h = np.empty((B, self.shapes.recurrent_flat_size), np.float32)
        self._check(self.lib.mzb_initial_inference(self._h, _ptr(obs, C.c_float), B, _ptr(v, C.c_double),
                                                   _ptr(r, C.c_double), _ptr(p, C.c_double), _ptr(h, C.c_float)))
        return InferenceOutputs(v, r, p, h)

    # VPRNetModel::RecurrentInference (vprnet.cpp:131-189)
    def recurrent_inference(self, encoded_state, action) -> InferenceOutputs:
        hid = np.ascontiguousarray(encoded_state, dtype=np.float32).reshape(-1, self.shapes.recurrent_flat_size)
        act = np.ascontiguousarray(action, dtype=np.int32).reshape(-1)
        B, A = hid.shape[0], self.config.num_actions
        if act.shape[0] != B:
            raise MzbError("action batch differs from hidden batch")
        v = np.empty(B, np.float64); r = np.empty(B, np.float64); p = np.empty((B, A), np.float64)
        h = np.empty((B, self.shapes.recurrent_flat_size), np.float32)
        self._check(self.lib.mzb_recurrent_inference(self._h, _ptr(hid, C.c_float), _ptr(act, C.c_int32), B,
                                                     _ptr(v, C.c_double), _ptr(r, C.c_double), _ptr(p, C.c_double),
                                                     _ptr(h, C.c_float)))
        return InferenceOutputs(v, r, p, h)

    def bench_recurrent(self, batch, iters):
        ms = C.c_float(); nl = C.c_int()
        self._check(self.lib.mzb_bench_recurrent(self._h, batch, iters, C.byref(ms), C.byref(nl)))
        return ms.value, nl.value

    # MCTS::MCTS rng_(seed) (mcts.cpp:26)
    def set_host_evaluator(self, initial, recurrent):
        """EVAL_HOST engines: `initial(slots, obs[n][stride]) -> (value[n], reward[n], policy[n][A])` and
        `recurrent(slots, parent_node[n], action[n], child_node) -> (value, reward, policy)` stand behind the reference's
        Evaluator::InitialInference / RecurrentInference (vprnet_evaluator.h:35-41); nodes are named by expansion order."""
        A = self.config.num_actions

        def _store(out, n, value, reward, policy):
            v, r, p = out
            np.ctypeslib.as_array(value, (n,))[:] = np.asarray(v, np.float64)
            np.ctypeslib.as_array(reward, (n,))[:] = np.asarray(r, np.float64)
            np.ctypeslib.as_array(policy, (n, A))[:] = np.asarray(p, np.float64).reshape(n, A)

        def _initial(_user, n, slot, obs, stride, value, reward, policy):
            try:
                _store(initial(np.ctypeslib.as_array(slot, (n,)).copy(), np.ctypeslib.as_array(obs, (n, stride)).copy()), n, value, reward, policy)
                return 0
            except Exception as e:  # surfaces as an MzbError from the search call
                self._host_error = e
                return 1

        def _recurrent(_user, n, slot, parent, action, child, value, reward, policy):
            try:
                _store(recurrent(np.ctypeslib.as_array(slot, (n,)).copy(), np.ctypeslib.as_array(parent, (n,)).copy(),
                                 np.ctypeslib.as_array(action, (n,)).copy(), child), n, value, reward, policy)
                return 0
            except Exception as e:
                self._host_error = e
                return 1

        self._host_cb = _CHostEvaluator(None, _HOST_INITIAL(_initial), _HOST_RECURRENT(_recurrent))   # keep the thunks alive
        self._check(self.lib.mzb_set_host_evaluator(self._h, C.byref(self._host_cb)))

    def se_state(self, tree, node):
        """EVAL_SYNTHETIC: the hidden state the stub holds for node `node` of tree row `tree` (checks the state mover)"""
        out = np.empty(self.shapes.recurrent_flat_size, np.float32)
        self._check(self.lib.mzb_debug_se_state(self._h, int(tree), int(node), _ptr(out, C.c_float)))
        return out

    def seed_slots(self, seeds, first_slot=0):
        s = np.ascontiguousarray(seeds, dtype=np.int32)
        self._check(self.lib.mzb_search_seed(self._h, first_slot, s.shape[0], _ptr(s, C.c_int32)))

    # T x MCTS::run (mcts.cpp:34-87)
    def run(self, stacked_observations, legal_mask, to_play, add_exploration_noise, device_obs_ptr=None, slots=None) -> MCTSReturn:
        legal = np.ascontiguousarray(legal_mask, dtype=np.uint32).reshape(-1)
        T, A = legal.shape[0], self.config.num_actions
        tp = np.ascontiguousarray(to_play, dtype=np.int32).reshape(-1)
        noise = np.ascontiguousarray(add_exploration_noise, dtype=np.uint8).reshape(-1)
        if tp.shape[0] != T or noise.shape[0] != T:
            raise MzbError("per-tree inputs disagree on the number of trees")
        res = MCTSReturn(np.empty(T, np.float64), np.empty(T, np.int32), np.empty(T, np.int32),
                         np.empty((T, A), np.int32), np.empty((T, A), np.int32), np.empty((T, A), np.float64),
                         np.empty((T, A), np.float64))
        cr = _CSearchResult(_ptr(res.root_value, C.c_double), _ptr(res.max_tree_depth, C.c_int32),
                            _ptr(res.n_children, C.c_int32), _ptr(res.child_actions, C.c_int32),
                            _ptr(res.child_visits, C.c_int32), _ptr(res.children_relative_visit, C.c_double),
                            _ptr(res.emperical_policy, C.c_double))
        if slots is not None:
            sl = np.ascontiguousarray(slots, dtype=np.int32).reshape(-1)
            if sl.shape[0] != T:
                raise MzbError("slot list disagrees on the number of trees")
            width = 1 if self.config.evaluator == EVAL_SYNTHETIC else self.shapes.initial_flat_size
            obs = np.ascontiguousarray(stacked_observations, dtype=np.float32).reshape(T, width)
            self._check(self.lib.mzb_search_batch_slots(self._h, _ptr(obs, C.c_float), _ptr(sl, C.c_int32), _ptr(legal, C.c_uint32),
                                                        _ptr(tp, C.c_int32), _ptr(noise, C.c_uint8), T, C.byref(cr)))
        elif device_obs_ptr is not None:
            self._check(self.lib.mzb_search_batch_device_obs(self._h, C.c_void_p(device_obs_ptr), _ptr(legal, C.c_uint32),
                                                             _ptr(tp, C.c_int32), _ptr(noise, C.c_uint8), T, C.byref(cr)))
        else:
            width = 1 if self.config.evaluator == EVAL_SYNTHETIC else self.shapes.initial_flat_size
            obs = np.ascontiguousarray(stacked_observations, dtype=np.float32).reshape(T, width)
            self._check(self.lib.mzb_search_batch(self._h, _ptr(obs, C.c_float), _ptr(legal, C.c_uint32),
                                                  _ptr(tp, C.c_int32), _ptr(noise, C.c_uint8), T, C.byref(cr)))
        return res

    def last_search_timing(self):
        t = _CSearchTiming()
        self._check(self.lib.mzb_last_search_timing(self._h, C.byref(t)))
        return dict(total_ms=t.total_ms, root_ms=t.root_ms, sim_loop_ms=t.sim_loop_ms,
                    kernel_launches=t.kernel_launches, simulations=t.simulations)

    def set_recording(self, enabled=True):
        self._check(self.lib.mzb_search_set_recording(self._h, int(enabled)))

    def get_trace(self, n_trees):
        n = self.config.num_simulations
        parent = np.empty((n_trees, n), np.int32); action = np.empty((n_trees, n), np.int32)
        self._check(self.lib.mzb_search_get_trace(self._h, n_trees, _ptr(parent, C.c_int32), _ptr(action, C.c_int32)))
        return parent, action

    def get_records(self, n_trees):
        n1, A = self.config.num_simulations + 1, self.config.num_actions
        v = np.empty((n_trees, n1), np.float64); r = np.empty((n_trees, n1), np.float64)
        p = np.empty((n_trees, n1, A), np.float64)
        self._check(self.lib.mzb_search_get_records(self._h, n_trees, _ptr(v, C.c_double), _ptr(r, C.c_double), _ptr(p, C.c_double)))
        return v, r, p


def legal_mask_from_actions(legal_actions, num_actions):
    """legal action lists (AbstractGame::legal_actions, abstract_game.h) -> bit masks.  Children are created in
    ascending action order, which is the order every shipped game reports them in."""
    out = np.zeros(len(legal_actions), np.uint32)
    for i, acts in enumerate(legal_actions):
        m = 0
        for a in acts:
            if a < 0 or a >= num_actions:
                raise MzbError("action outside the action space")
            m |= 1 << a
        out[i] = m
    return out


# game presets = the shapes of BASELINE.json's configs
VALUE_DECODE_STATIC_MAX_FLOATS = 1 << 20   # MZB_VALUE_DECODE_STATIC_MAX_FLOATS


class Mt19937:
    """std::mt19937(seed) as a mzb_rng_fn source: numpy's MT19937 with the legacy (init_genrand) seeding is the same generator."""

    def __init__(self, seed):
        self._bg = np.random.RandomState(int(seed))._bit_generator
        self.draws = 0

    def __call__(self):
        self.draws += 1
        return int(self._bg.random_raw())


class ReplayBuffer:
    """PrioritizedReplayBuffer (replay_buffer.h:212-300) on one engine's GPU: the sum tree on the host, the histories in HBM, one
    kernel launch per sampled batch (csrc/replay.cu).  `rng` arguments are callables returning std::mt19937's 32-bit draws."""

    def __init__(self, engine, capacity, min_sample_size=1, td_steps=10, num_unroll_steps=5, per_alpha=1.0, per_beta=1.0,
                 per_epsilon=0.01, per_beta_increment=0.001, action_planes_initial=None):
        self.engine, self.lib = engine, engine.lib
        self.K = int(num_unroll_steps)
        planes = None
        if action_planes_initial is not None:
            planes = np.ascontiguousarray(action_planes_initial, dtype=np.float32)
        rc = _CReplayConfig(int(capacity), int(min_sample_size), int(td_steps), self.K, float(per_alpha), float(per_beta),
                            float(per_epsilon), float(per_beta_increment), _ptr(planes, C.c_float) if planes is not None else None)
        h = C.c_void_p()
        engine._check(self.lib.mzb_replay_create(engine._h, C.byref(rc), C.byref(h)))
        self._h = h

    def close(self):
        if self._h:
            self.lib.mzb_replay_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def size(self):
        return self.lib.mzb_replay_size(self._h)

    def can_sample(self):
        return bool(self.lib.mzb_replay_can_sample(self._h))

    def num_histories(self):
        return self.lib.mzb_replay_num_histories(self._h)

    def total_priority(self):
        return self.lib.mzb_replay_total_priority(self._h)

    def oldest_history(self):
        return self.lib.mzb_replay_oldest_history(self._h)

    def save_game_history(self, observations, actions, rewards, to_play, root_values, child_visits):
        """observations [L+1][c*h*w] (single frames), actions / rewards / to_play [L+1], root_values [L], child_visits [L][A]"""
        rv = np.ascontiguousarray(root_values, dtype=np.float64)
        L = rv.shape[0]
        cfg = self.engine.config
        c, h, w = cfg.observation_shape
        obs = np.ascontiguousarray(observations, dtype=np.float32).reshape(L + 1, c * h * w)
        act = np.ascontiguousarray(actions, dtype=np.int32).reshape(L + 1)
        rew = np.ascontiguousarray(rewards, dtype=np.float64).reshape(L + 1)
        tp = np.ascontiguousarray(to_play, dtype=np.int32).reshape(L + 1)
        cv = np.ascontiguousarray(child_visits, dtype=np.float64).reshape(L, cfg.num_actions)
        g = _CGameHistory(L, _ptr(obs, C.c_float), _ptr(act, C.c_int32), _ptr(rew, C.c_double), _ptr(tp, C.c_int32),
                          _ptr(rv, C.c_double), _ptr(cv, C.c_double))
        self.engine._check(self.lib.mzb_replay_save_game_history(self._h, C.byref(g)))

    def sample(self, rng, batch_size):
        """types::Batch as a dict of arrays (same field names)"""
        B, K1, A = int(batch_size), self.K + 1, self.engine.config.num_actions
        out = dict(priorities=np.empty(B, np.float64), indices=np.empty(B, np.int32), actions=np.empty((B, K1), np.int32),
                   stacked_observations=np.empty((B, self.engine.shapes.initial_flat_size), np.float32),
                   target_rewards=np.empty((B, K1), np.float64), target_values=np.empty((B, K1), np.float64),
                   target_policies=np.empty((B, K1, A), np.float64), gradient_scale=np.empty(B, np.float64))
        b = _CBatchOut(0, _ptr(out["priorities"], C.c_double), _ptr(out["indices"], C.c_int32), _ptr(out["actions"], C.c_int32),
                       _ptr(out["stacked_observations"], C.c_float), _ptr(out["target_rewards"], C.c_double),
                       _ptr(out["target_values"], C.c_double), _ptr(out["target_policies"], C.c_double),
                       _ptr(out["gradient_scale"], C.c_double))
        fn = _RNG_FN(lambda _u: rng())
        self.engine._check(self.lib.mzb_replay_sample(self._h, fn, None, B, C.byref(b)))
        out["num_samples"] = b.num_samples
        return out

    def update_history_priorities(self, indices, errors):
        idx = np.ascontiguousarray(indices, dtype=np.int32)
        err = np.ascontiguousarray(errors, dtype=np.float64)
        self.engine._check(self.lib.mzb_replay_update_priorities(self._h, _ptr(idx, C.c_int32), _ptr(err, C.c_double), int(idx.size)))

    def sample_game(self, rng):
        """(history id, stored length or 0 if the buffer has lost it)"""
        hid, ln = C.c_int32(), C.c_int32()
        fn = _RNG_FN(lambda _u: rng())
        self.engine._check(self.lib.mzb_replay_sample_game(self._h, fn, None, C.byref(hid), C.byref(ln)))
        return hid.value, ln.value

    def update_game_history(self, history_id, root_values, child_visits):
        rv = np.ascontiguousarray(root_values, dtype=np.float64)
        cv = np.ascontiguousarray(child_visits, dtype=np.float64)
        self.engine._check(self.lib.mzb_replay_update_game_history(self._h, int(history_id), _ptr(rv, C.c_double), _ptr(cv, C.c_double), int(rv.shape[0])))


def preset(name, **over):
    nc = MuZeroNetworkConfig()
    kw = {}
    if name in ("c4d", "c4r", "ttt"):
        kw = dict(observation_shape=(3, 6, 7) if name != "ttt" else (3, 3, 3), num_actions=7 if name != "ttt" else 9,
                  num_players=2, value_upperbound=10.0, value_lowerbound=-10.0, min_reward=-10.0, max_reward=10.0,
                  min_value=-10.0, max_value=10.0)
        if name == "c4r":
            nc = MuZeroNetworkConfig(resnet_channels=64, representation_blocks=3, dynamics_blocks=3, prediction_blocks=3)
        if name == "ttt":
            kw["num_simulations"] = 25
    elif name == "ale":
        nc = MuZeroNetworkConfig(downsample=True, reward_reduced_channels=64, policy_reduced_channels=64,
                                 value_reduced_channels=64, reward_head_layers=[64, 64], policy_head_layers=[64, 64],
                                 value_head_layers=[64, 64])
        kw = dict(observation_shape=(1, 96, 96), num_actions=18, num_players=1, stacked_observations=6,
                  min_value=-21.0, max_value=21.0, min_reward=-1.0, max_reward=1.0)
    elif name == "tiny":
        nc = MuZeroNetworkConfig(resnet_channels=16, representation_blocks=1, dynamics_blocks=1, prediction_blocks=2,
                                 reward_reduced_channels=4, policy_reduced_channels=2, value_reduced_channels=3,
                                 reward_head_layers=[8], policy_head_layers=[8, 6], value_head_layers=[])
        kw = dict(observation_shape=(2, 3, 4), num_actions=5, num_players=2, stacked_observations=1,
                  value_upperbound=5.0, value_lowerbound=-5.0, min_reward=-3.0, max_reward=3.0, min_value=-5.0, max_value=5.0)
    elif name == "tinyds":
        nc = MuZeroNetworkConfig(downsample=True, resnet_channels=16, representation_blocks=1, dynamics_blocks=1,
                                 prediction_blocks=1, reward_reduced_channels=4, policy_reduced_channels=4,
                                 value_reduced_channels=4, reward_head_layers=[8], policy_head_layers=[8], value_head_layers=[8])
        kw = dict(observation_shape=(1, 32, 24), num_actions=6, num_players=1, stacked_observations=2,
                  min_value=-21.0, max_value=21.0, min_reward=-1.0, max_reward=1.0)
    else:
        raise MzbError("unknown preset " + name)
    kw["network_config"] = nc
    kw.update(over)
    return MuZeroConfig(**kw)
