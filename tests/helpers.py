"""Shared test helpers: oracle configs matching engine configs, random weights with the reference's tensor names."""
import json
import os

import numpy as np

from oracle import oracle_py as O
from oracle import vprnet_oracle as V

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def oracle_config_from(kw):
    kw = dict(kw)
    for k in ("lower", "upper"):
        if kw.get(k) == "INF":
            kw[k] = O.INF_D
        elif kw.get(k) == "NINF":
            kw[k] = O.NINF_D
    return O.make_config(**kw)


def load_mcts_golden():
    with open(os.path.join(GOLDEN, "mcts_se.json")) as f:
        cases = json.load(f)
    for c in cases:
        for r in c["results"]:
            r["root_value"] = float.fromhex(r["root_value"])
            r["relvisits"] = [float.fromhex(x) for x in r["relvisits"]]
            r["policy"] = [float.fromhex(x) for x in r["policy"]]
            r["trace"] = [tuple(x) for x in r["trace"]]
    return cases


def results_equal(a, b):
    """bit-exact comparison of two search results (dicts from oracle_py)"""
    return (a["root_value"] == b["root_value"] and a["depth"] == b["depth"] and a["actions"] == b["actions"]
            and a["relvisits"] == b["relvisits"] and a["policy"] == b["policy"]
            and [tuple(x) for x in a["trace"]] == [tuple(x) for x in b["trace"]])


def se_keys(T, run=0):
    """position keys `ref_driver mcts` / mzo_run_se use for tree t, run r"""
    return np.array([((t * 7919 + run * 104729 + 1) & 0xFFFFFF) for t in range(T)], dtype=np.uint32)


def se_legal_masks(keys, num_actions, mode):
    out = np.zeros(len(keys), np.uint32)
    full = 0xFFFFFFFF if num_actions >= 32 else (1 << num_actions) - 1
    for i, k in enumerate(keys):
        if mode == 0:
            out[i] = full
        else:
            m = _mix(int(k) ^ 0x5bd1e995) & full
            if m == 0:
                m = 1 << (int(k) % num_actions)
            out[i] = m
    return out


def _mix(x):
    x &= 0xFFFFFFFF
    x ^= x >> 16
    x = (x * 0x7feb352d) & 0xFFFFFFFF
    x ^= x >> 15
    x = (x * 0x846ca68b) & 0xFFFFFFFF
    x ^= x >> 16
    return x


def random_weights(cfg: "V.NetConfig", seed=0, randomize_bn=True):
    """Random module tensors with the reference's names/shapes; delegates to the product's generator."""
    import mzb200
    from mzb200.weights import random_weights as rw
    mc = mzb200.MuZeroConfig(
        observation_shape=cfg.obs_shape, num_actions=cfg.num_actions, action_channels=cfg.action_channels,
        stacked_observations=cfg.stacked_observations, min_value=cfg.min_value, max_value=cfg.max_value,
        min_reward=cfg.min_reward, max_reward=cfg.max_reward, use_contractive_mapping=cfg.contractive,
        network_config=mzb200.MuZeroNetworkConfig(
            downsample=cfg.downsample, normalize_hidden_states=cfg.normalize, resnet_channels=cfg.channels,
            representation_blocks=cfg.rep_blocks, dynamics_blocks=cfg.dyn_blocks, prediction_blocks=cfg.pred_blocks,
            reward_reduced_channels=cfg.reward_reduced, policy_reduced_channels=cfg.policy_reduced,
            value_reduced_channels=cfg.value_reduced, reward_head_layers=list(cfg.reward_layers),
            policy_head_layers=list(cfg.policy_layers), value_head_layers=list(cfg.value_layers)))
    return rw(mc, seed, randomize_bn)
