"""Random-init module tensors with exactly the reference's names and shapes (register_module calls in
muzero-cpp/models.cpp:31-209 and model_layers.cpp:46-135), for synthetic benchmarks and tests.  Plain numpy."""
import math

import numpy as np


def _contractive_f32(x):
    x = np.float32(x)
    return float(np.sign(x) * (np.sqrt(np.abs(x) + np.float32(1)) - np.float32(1)) + np.float32(0.001) * x)


def support_size(min_value, max_value, contractive=True):  # ValueEncoder::get_support_size (util.cpp:91-102)
    if contractive:
        max_value, min_value = _contractive_f32(max_value), _contractive_f32(min_value)
    return int(math.ceil(max_value) - math.floor(min_value) + 1)


def random_weights(config, seed=0, randomize_bn=True):
    """config: mzb200.MuZeroConfig.  Kaiming-uniform-like convs/linears; BN statistics randomised so that the
    eval-mode folding is exercised (a fresh torch module would have mean 0 / var 1)."""
    nc = config.network_config
    rng = np.random.default_rng(seed)
    W = {}
    C = nc.resnet_channels
    c0, h0, w0 = config.observation_shape
    cin = c0 * (config.stacked_observations + 1) + config.stacked_observations * config.action_channels
    eh, ew = ((h0 + 15) // 16, (w0 + 15) // 16) if nc.downsample else (h0, w0)
    hw = eh * ew

    def conv(name, cout, cin_, k=3, bias=False):
        bound = 1.0 / math.sqrt(cin_ * k * k)
        W[name + ".weight"] = rng.uniform(-bound, bound, (cout, cin_, k, k)).astype(np.float32)
        if bias:
            W[name + ".bias"] = rng.uniform(-bound, bound, (cout,)).astype(np.float32)

    def bn(name, c):
        if randomize_bn:
            W[name + ".weight"] = rng.uniform(0.5, 1.5, c).astype(np.float32)
            W[name + ".bias"] = (rng.standard_normal(c) * 0.2).astype(np.float32)
            W[name + ".running_mean"] = (rng.standard_normal(c) * 0.3).astype(np.float32)
            W[name + ".running_var"] = rng.uniform(0.5, 1.5, c).astype(np.float32)
        else:
            W[name + ".weight"] = np.ones(c, np.float32)
            W[name + ".bias"] = np.zeros(c, np.float32)
            W[name + ".running_mean"] = np.zeros(c, np.float32)
            W[name + ".running_var"] = np.ones(c, np.float32)

    def block(prefix, i, c):
        p = "%s.resnet_%d_" % (prefix, i)
        conv(p + "conv1", c, c); bn(p + "bn1", c); conv(p + "conv2", c, c); bn(p + "bn2", c)

    def mlp(prefix, sizes):
        for l in range(len(sizes) - 1):
            bound = 1.0 / math.sqrt(sizes[l])
            W["%s.linear_%d.weight" % (prefix, l)] = rng.uniform(-bound, bound, (sizes[l + 1], sizes[l])).astype(np.float32)
            W["%s.linear_%d.bias" % (prefix, l)] = rng.uniform(-bound, bound, (sizes[l + 1],)).astype(np.float32)

    rh = "representation_network.representation_head.0"
    if nc.downsample:
        conv(rh + ".resnet_head_downsample_conv1", C // 2, cin)
        conv(rh + ".resnet_head_downsample_conv2", C, C // 2)
        for i in range(2):
            block(rh + ".resnet_head_downsample_resblock1.%d" % i, i, C // 2)
        for i in range(2, 5):
            block(rh + ".resnet_head_downsample_resblock2.%d" % (i - 2), i, C)
        for i in range(5, 8):
            block(rh + ".resnet_head_downsample_resblock3.%d" % (i - 5), i, C)
    else:
        conv(rh + ".representation_resnet_head_conv", C, cin)
        bn(rh + ".representation_resnet_head_bn", C)
    for i in range(nc.representation_blocks):
        block("representation_network.representation_layers.%d" % i, i, C)
    conv("dynamics_network.dynamics_resnet_head.resnet_head_conv", C, C + config.action_channels)
    bn("dynamics_network.dynamics_resnet_head.resnet_head_bn", C)
    for i in range(nc.dynamics_blocks):
        block("dynamics_network.dynamics_resnet_body.%d" % i, i, C)
    for i in range(nc.prediction_blocks):
        block("prediction_network.prediction_resnet_body.%d" % i, i, C)
    sv = support_size(config.min_value, config.max_value, config.use_contractive_mapping)
    sr = support_size(config.min_reward, config.max_reward, config.use_contractive_mapping)
    conv("dynamics_network.dynamics_reward_conv", nc.reward_reduced_channels, C, 1, True)
    mlp("dynamics_network.dynamics_reward_mlp.dynamics_reward_mlp", [nc.reward_reduced_channels * hw] + list(nc.reward_head_layers) + [sr])
    conv("prediction_network.prediction_policy_conv", nc.policy_reduced_channels, C, 1, True)
    mlp("prediction_network.prediction_policy_mlp.prediction_policy_mlp", [nc.policy_reduced_channels * hw] + list(nc.policy_head_layers) + [config.num_actions])
    conv("prediction_network.prediction_value_conv", nc.value_reduced_channels, C, 1, True)
    mlp("prediction_network.prediction_value_mlp.prediction_value_mlp", [nc.value_reduced_channels * hw] + list(nc.value_head_layers) + [sv])
    return W


def recurrent_flops(config):
    """algorithmic FLOPs of one recurrent_inference state (SURVEY.md §8d formula)"""
    nc = config.network_config
    C = nc.resnet_channels
    c0, h0, w0 = config.observation_shape
    eh, ew = ((h0 + 15) // 16, (w0 + 15) // 16) if nc.downsample else (h0, w0)
    hw = eh * ew
    conv = lambda ci, co, k: 2.0 * ci * co * k * k * hw
    sv = support_size(config.min_value, config.max_value, config.use_contractive_mapping)
    sr = support_size(config.min_reward, config.max_reward, config.use_contractive_mapping)

    def mlp(red, layers, out):
        sizes = [red * hw] + list(layers) + [out]
        return sum(2.0 * sizes[i] * sizes[i + 1] for i in range(len(sizes) - 1))
    f = conv(C + config.action_channels, C, 3) + 2 * (nc.dynamics_blocks + nc.prediction_blocks) * conv(C, C, 3)
    f += conv(C, nc.reward_reduced_channels, 1) + conv(C, nc.policy_reduced_channels, 1) + conv(C, nc.value_reduced_channels, 1)
    f += mlp(nc.reward_reduced_channels, nc.reward_head_layers, sr) + mlp(nc.policy_reduced_channels, nc.policy_head_layers, config.num_actions)
    f += mlp(nc.value_reduced_channels, nc.value_head_layers, sv)
    return f
