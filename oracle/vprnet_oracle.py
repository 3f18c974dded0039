"""vprnet_oracle — TEST INFRASTRUCTURE.  CPU fp32 restatement (torch.nn.functional on CPU tensors) of the
reference network path: VPRNetModel::{InitialInference,RecurrentInference} (muzero-cpp/vprnet.cpp:77-189) over
MuzeroNetworkImpl (models.cpp:171-265), the layers (model_layers.cpp:46-162) and ValueEncoder::decode
(util.cpp:14-24,45-54,80-87).  The arithmetic itself lives in libtorch (un-vendored dependency, torch 2.11.0 here),
so the reference's own tests pin only shapes and the encoder; parity is PINNED by running the unmodified
reference (oracle/_ref/ref_driver net) on identical weights: golden fixtures under tests/golden/ (script:
tests/golden/make_golden.py).  Only tests/, smoke() and bench.py's cpu_baseline leg import this."""
import math

import numpy as np
import torch
import torch.nn.functional as F

EPS_BN = 1e-4  # model_layers.cpp:36-42


class NetConfig:
    """The MuZeroConfig fields the network reads (config.h:15-31,52-75)."""

    def __init__(self, obs_shape, num_actions, action_channels=1, stacked_observations=0, channels=128,
                 rep_blocks=4, dyn_blocks=4, pred_blocks=4, reward_reduced=16, policy_reduced=16, value_reduced=16,
                 reward_layers=(64,), policy_layers=(64,), value_layers=(64,), downsample=False, normalize=True,
                 min_value=-300.0, max_value=300.0, min_reward=-300.0, max_reward=300.0, contractive=True):
        self.obs_shape = tuple(obs_shape)
        self.num_actions = num_actions
        self.action_channels = action_channels
        self.stacked_observations = stacked_observations
        self.channels = channels
        self.rep_blocks, self.dyn_blocks, self.pred_blocks = rep_blocks, dyn_blocks, pred_blocks
        self.reward_reduced, self.policy_reduced, self.value_reduced = reward_reduced, policy_reduced, value_reduced
        self.reward_layers, self.policy_layers, self.value_layers = tuple(reward_layers), tuple(policy_layers), tuple(value_layers)
        self.downsample = downsample
        self.normalize = normalize
        self.min_value, self.max_value, self.min_reward, self.max_reward = min_value, max_value, min_reward, max_reward
        self.contractive = contractive

    @property
    def in_channels(self):  # models.cpp:183-187
        c = self.obs_shape[0]
        return c * (self.stacked_observations + 1) + self.stacked_observations * self.action_channels

    @property
    def enc_hw(self):  # model_layers.cpp:107-109,157-162
        h, w = self.obs_shape[1], self.obs_shape[2]
        if self.downsample:
            return (h + 15) // 16, (w + 15) // 16
        return h, w


PRESETS = {
    "c4d": dict(obs_shape=(3, 6, 7), num_actions=7, min_value=-10, max_value=10, min_reward=-10, max_reward=10),
    "c4r": dict(obs_shape=(3, 6, 7), num_actions=7, min_value=-10, max_value=10, min_reward=-10, max_reward=10,
                channels=64, rep_blocks=3, dyn_blocks=3, pred_blocks=3),
    "ttt": dict(obs_shape=(3, 3, 3), num_actions=9, min_value=-10, max_value=10, min_reward=-10, max_reward=10),
    "ale": dict(obs_shape=(1, 96, 96), num_actions=18, stacked_observations=6, downsample=True, reward_reduced=64,
                policy_reduced=64, value_reduced=64, reward_layers=(64, 64), policy_layers=(64, 64),
                value_layers=(64, 64), min_value=-21, max_value=21, min_reward=-1, max_reward=1),
    "tiny": dict(obs_shape=(2, 3, 4), num_actions=5, stacked_observations=1, channels=16, rep_blocks=1, dyn_blocks=1,
                 pred_blocks=2, reward_reduced=4, policy_reduced=2, value_reduced=3, reward_layers=(8,),
                 policy_layers=(8, 6), value_layers=(), min_value=-5, max_value=5, min_reward=-3, max_reward=3),
    "tinyds": dict(obs_shape=(1, 32, 24), num_actions=6, stacked_observations=2, downsample=True, channels=16,
                   rep_blocks=1, dyn_blocks=1, pred_blocks=1, reward_reduced=4, policy_reduced=4, value_reduced=4,
                   reward_layers=(8,), policy_layers=(8,), value_layers=(8,), min_value=-21, max_value=21,
                   min_reward=-1, max_reward=1),
}


def preset(name, **over):
    d = dict(PRESETS[name])
    d.update(over)
    return NetConfig(**d)


# ------------------------------------------------------------------ ValueEncoder (util.cpp)
def _contractive_f32(x):  # util.cpp:14-16 on a float32 tensor (torch::tensor({double}) is kFloat)
    eps = torch.tensor(0.001, dtype=torch.float32)
    return torch.sign(x) * (torch.sqrt(torch.abs(x) + 1) - 1) + eps * x


def _inverse_contractive(x):  # util.cpp:19-24
    eps = 0.001
    return torch.sign(x) * (torch.pow((torch.sqrt(1 + 4 * eps * (torch.abs(x) + 1 + eps)) - 1) / (2 * eps), 2) - 1)


class ValueEncoder:
    def __init__(self, min_value, max_value, contractive=True):  # util.cpp:45-54
        self.contractive = contractive
        if contractive:
            max_value = float(_contractive_f32(torch.tensor([max_value], dtype=torch.float32)).item())
            min_value = float(_contractive_f32(torch.tensor([min_value], dtype=torch.float32)).item())
        self.min_value, self.max_value = min_value, max_value
        self.support_size = int(math.ceil(max_value) - math.floor(min_value) + 1)
        self.step_size = (max_value - min_value) / (self.support_size - 1)

    def decode(self, logits, take_softmax=True):  # util.cpp:80-87
        p = torch.softmax(logits, 1) if take_softmax else logits
        support = torch.arange(0, self.support_size).expand(p.shape).to(torch.float32)
        x = (torch.sum(support * p, 1, True) * self.step_size) + self.min_value
        return _inverse_contractive(x) if self.contractive else x

    def encode(self, x):  # util.cpp:59-76 (training only; here for the reference's own round-trip tests)
        red = _contractive_f32(x) if self.contractive else x
        red = red.clamp(self.min_value, self.max_value).unsqueeze(-1)
        num_steps = (red - self.min_value) / self.step_size
        lower = num_steps.floor()
        upper_mod = num_steps - lower
        lower_mod = 1.0 - upper_mod
        lower = lower.to(torch.long)
        rng = torch.arange(0, self.support_size)
        return (lower == rng).float() * lower_mod + ((lower + 1) == rng).float() * upper_mod


# ------------------------------------------------------------------ layers (model_layers.cpp)
TRAIN = [False]   # model_->train() (vprnet.cpp:246): batch statistics, running statistics moved in place by momentum 0.01


def _bn(W, pre, x):
    return F.batch_norm(x, W[pre + ".running_mean"], W[pre + ".running_var"], W[pre + ".weight"], W[pre + ".bias"],
                        TRAIN[0], 0.01, EPS_BN)


def _resblock(W, pre, i, x):  # model_layers.cpp:81-88
    p = "%s.resnet_%d_" % (pre, i)
    out = F.relu(_bn(W, p + "bn1", F.conv2d(x, W[p + "conv1.weight"], None, 1, 1)))
    out = _bn(W, p + "bn2", F.conv2d(out, W[p + "conv2.weight"], None, 1, 1))
    return F.relu(out + x)


def _mlp(W, pre, n_layers, x):  # model_layers.cpp:46-65
    for i in range(n_layers):
        x = F.linear(x, W["%s.linear_%d.weight" % (pre, i)], W["%s.linear_%d.bias" % (pre, i)])
        if i < n_layers - 1:
            x = F.relu(x)
    return x


def normalize_hidden_state(x):  # models.cpp:16-26
    B, Cc, H, Wd = x.shape
    v = x.view(-1, Cc, H * Wd)
    mn = v.min(2, True)[0].unsqueeze(-1)
    mx = v.max(2, True)[0].unsqueeze(-1)
    scale = mx - mn
    scale = torch.where(scale < 1e-5, torch.full_like(scale, 1e-5), scale)
    return (x - mn) / scale


class VPRNetOracle:
    """weights: dict name -> np.ndarray / tensor, names exactly as in the reference's module archive."""

    def __init__(self, cfg, weights):
        self.cfg = cfg
        self.W = {k: torch.as_tensor(np.asarray(v), dtype=torch.float32) for k, v in weights.items()}
        self.value_encoder = ValueEncoder(cfg.min_value, cfg.max_value, cfg.contractive)
        self.reward_encoder = ValueEncoder(cfg.min_reward, cfg.max_reward, cfg.contractive)

    # representation (models.cpp:52-65,222-228; model_layers.cpp:101-104,140-155)
    def representation(self, obs):
        W, c = self.W, self.cfg
        head = "representation_network.representation_head.0"
        if c.downsample:
            x = F.conv2d(obs, W[head + ".resnet_head_downsample_conv1.weight"], None, 2, 1)
            for i in range(2):
                x = _resblock(W, head + ".resnet_head_downsample_resblock1.%d" % i, i, x)
            x = F.conv2d(x, W[head + ".resnet_head_downsample_conv2.weight"], None, 2, 1)
            for i in range(2, 5):
                x = _resblock(W, head + ".resnet_head_downsample_resblock2.%d" % (i - 2), i, x)
            x = F.avg_pool2d(x, 3, 2, 1)
            for i in range(5, 8):
                x = _resblock(W, head + ".resnet_head_downsample_resblock3.%d" % (i - 5), i, x)
            x = F.avg_pool2d(x, 3, 2, 1)
        else:
            x = F.relu(_bn(W, head + ".representation_resnet_head_bn",
                           F.conv2d(obs, W[head + ".representation_resnet_head_conv.weight"], None, 1, 1)))
        for i in range(c.rep_blocks):
            x = _resblock(W, "representation_network.representation_layers.%d" % i, i, x)
        self._raw = x
        return normalize_hidden_state(x) if c.normalize else x

    def dynamics(self, h, act):  # models.cpp:105-117,231-240
        W, c = self.W, self.cfg
        x = torch.cat([h, act], 1)
        p = "dynamics_network.dynamics_resnet_head"
        x = F.relu(_bn(W, p + ".resnet_head_bn", F.conv2d(x, W[p + ".resnet_head_conv.weight"], None, 1, 1)))
        for i in range(c.dyn_blocks):
            x = _resblock(W, "dynamics_network.dynamics_resnet_body.%d" % i, i, x)
        r = F.conv2d(x, W["dynamics_network.dynamics_reward_conv.weight"], W["dynamics_network.dynamics_reward_conv.bias"])
        r = r.reshape(r.shape[0], -1)
        r = _mlp(W, "dynamics_network.dynamics_reward_mlp.dynamics_reward_mlp", len(c.reward_layers) + 1, r)
        self._raw = x
        return (normalize_hidden_state(x) if c.normalize else x), r

    def prediction(self, x):  # models.cpp:153-166
        W, c = self.W, self.cfg
        for i in range(c.pred_blocks):
            x = _resblock(W, "prediction_network.prediction_resnet_body.%d" % i, i, x)
        p = F.conv2d(x, W["prediction_network.prediction_policy_conv.weight"], W["prediction_network.prediction_policy_conv.bias"])
        v = F.conv2d(x, W["prediction_network.prediction_value_conv.weight"], W["prediction_network.prediction_value_conv.bias"])
        p = _mlp(W, "prediction_network.prediction_policy_mlp.prediction_policy_mlp", len(c.policy_layers) + 1, p.reshape(p.shape[0], -1))
        v = _mlp(W, "prediction_network.prediction_value_mlp.prediction_value_mlp", len(c.value_layers) + 1, v.reshape(v.shape[0], -1))
        return p, v

    def _finish(self, value_logits, reward_logits, policy_logits, hidden):  # vprnet.cpp:103-127
        value = self.value_encoder.decode(value_logits).to(torch.float64)[:, 0]
        reward = self.reward_encoder.decode(reward_logits).to(torch.float64)[:, 0]
        policy = torch.softmax(policy_logits, -1).to(torch.float64)
        return dict(value=value.numpy(), reward=reward.numpy(), policy=policy.numpy(),
                    hidden=hidden.reshape(hidden.shape[0], -1).numpy(), value_logits=value_logits.numpy(),
                    reward_logits=reward_logits.numpy(), policy_logits=policy_logits.numpy(),
                    hidden_raw=self._raw.reshape(hidden.shape[0], -1).numpy())

    @torch.no_grad()
    def initial_inference(self, obs_flat):  # vprnet.cpp:77-129, models.cpp:248-258
        c = self.cfg
        obs = torch.as_tensor(np.asarray(obs_flat), dtype=torch.float32).reshape(-1, c.in_channels, c.obs_shape[1], c.obs_shape[2])
        h = self.representation(obs)
        p, v = self.prediction(h)
        S = self.reward_encoder.support_size
        onehot = torch.zeros(1, S)
        onehot[0, S // 2] = 1
        r = torch.log(onehot).repeat(obs.shape[0], 1)
        return self._finish(v, r, p, h)

    @torch.no_grad()
    def recurrent_inference(self, hidden_flat, actions, action_plane_fn=None):  # vprnet.cpp:131-189, models.cpp:261-265
        c = self.cfg
        eh, ew = c.enc_hw
        h = torch.as_tensor(np.asarray(hidden_flat), dtype=torch.float32).reshape(-1, c.channels, eh, ew)
        B = h.shape[0]
        act = torch.empty(B, c.action_channels, eh, ew)
        for b in range(B):
            if action_plane_fn is None:  # the examples' constant plane action/num_actions, computed in double then narrowed
                act[b].fill_(float(np.float32(float(int(actions[b])) / c.num_actions)))
            else:
                act[b] = torch.as_tensor(action_plane_fn(int(actions[b])), dtype=torch.float32).reshape(c.action_channels, eh, ew)
        nh, r = self.dynamics(h, act)
        p, v = self.prediction(nh)
        return self._finish(v, r, p, nh)

    @torch.no_grad()
    def learn_forward(self, obs_flat, actions, target_values, target_rewards, target_policies, priorities, value_loss_weight=0.25):
        """The forward half of VPRNetModel::Learn (vprnet.cpp:192-311): K + 1 unrolled TRAIN-mode calls, targets through
        ValueEncoder::encode, MuzeroNetworkImpl::loss (models.cpp:268-276), the reference's [B,1] x [B] broadcast of the
        importance weights.  Moves this oracle's BatchNorm running statistics like the reference's.  Returns the LossInfo fields."""
        c = self.cfg
        eh, ew = c.enc_hw
        actions = np.asarray(actions)
        B, K1 = actions.shape
        obs = torch.as_tensor(np.asarray(obs_flat), dtype=torch.float32).reshape(B, c.in_channels, c.obs_shape[1], c.obs_shape[2])
        tv = torch.as_tensor(np.asarray(target_values, np.float64)).to(torch.float32).reshape(B, K1)
        tr = torch.as_tensor(np.asarray(target_rewards, np.float64)).to(torch.float32).reshape(B, K1)
        tp = torch.as_tensor(np.asarray(target_policies, np.float64)).to(torch.float32).reshape(B, K1, c.num_actions)
        pri = torch.as_tensor(np.asarray(priorities, np.float64)).to(torch.float32).reshape(B, 1)
        tve, tre = self.value_encoder.encode(tv), self.reward_encoder.encode(tr)

        def loss(logits, target):
            return torch.sum(-target * torch.log_softmax(logits, 1), -1)
        TRAIN[0] = True
        try:
            h = self.representation(obs)
            p, v = self.prediction(h)
            value_loss, policy_loss = loss(v, tve[:, 0]), loss(p, tp[:, 0])
            reward_loss = torch.zeros_like(value_loss)
            errors = (self.value_encoder.decode(v) - tv[:, 0:1].to(torch.float64))[:, 0]
            for step in range(1, K1):
                act = torch.empty(B, c.action_channels, eh, ew)
                for b in range(B):
                    act[b].fill_(float(np.float32(float(int(actions[b, step])) / c.num_actions)))
                h, r = self.dynamics(h, act)
                p, v = self.prediction(h)
                value_loss = value_loss + loss(v, tve[:, step])
                reward_loss = reward_loss + loss(r, tre[:, step])
                policy_loss = policy_loss + loss(p, tp[:, step])
        finally:
            TRAIN[0] = False
        total = (pri * (value_loss * value_loss_weight + reward_loss + policy_loss)).mean()
        return dict(total_loss=float(total), value=float(value_loss.mean()), policy=float(policy_loss.mean()),
                    reward=float(reward_loss.mean()), errors=errors.numpy().astype(np.float64))

    def learn_step(self, obs_flat, actions, target_values, target_rewards, target_policies, priorities, gradient_scale,
                   value_loss_weight=0.25, learning_rate=3e-4, l2_weight_decay=1e-4):
        """The whole of VPRNetModel::Learn (vprnet.cpp:192-311) on torch autograd: the forward of learn_forward, the gradient hooks
        (x0.5 on every recurrent hidden state :272, gradient_scale on the step losses :278-286), backward, and
        torch::optim::Adam(lr).weight_decay(l2) (vprnet.cpp:35-36).  The optimizer persists across calls.  Returns the LossInfo
        fields; self.grads holds what backward() left on every parameter."""
        c = self.cfg
        eh, ew = c.enc_hw
        actions = np.asarray(actions)
        B, K1 = actions.shape
        names = [k for k in self.W if not (k.endswith("running_mean") or k.endswith("running_var"))]
        if not hasattr(self, "_opt"):
            for k in names:
                self.W[k] = self.W[k].clone().requires_grad_(True)
            self._opt = torch.optim.Adam([self.W[k] for k in names], lr=learning_rate, weight_decay=l2_weight_decay)
        obs = torch.as_tensor(np.asarray(obs_flat), dtype=torch.float32).reshape(B, c.in_channels, c.obs_shape[1], c.obs_shape[2])
        tv = torch.as_tensor(np.asarray(target_values, np.float64)).to(torch.float32).reshape(B, K1)
        tr = torch.as_tensor(np.asarray(target_rewards, np.float64)).to(torch.float32).reshape(B, K1)
        tp = torch.as_tensor(np.asarray(target_policies, np.float64)).to(torch.float32).reshape(B, K1, c.num_actions)
        pri = torch.as_tensor(np.asarray(priorities, np.float64)).to(torch.float32).reshape(B, 1)
        gs = torch.as_tensor(np.asarray(gradient_scale, np.float64)).to(torch.float32).reshape(B)
        tve, tre = self.value_encoder.encode(tv), self.reward_encoder.encode(tr)

        def loss(logits, target):
            return torch.sum(-target * torch.log_softmax(logits, 1), -1)
        TRAIN[0] = True
        try:
            with torch.enable_grad():
                h = self.representation(obs)
                p, v = self.prediction(h)
                value_loss, policy_loss = loss(v, tve[:, 0]), loss(p, tp[:, 0])
                reward_loss = torch.zeros_like(value_loss)
                errors = (self.value_encoder.decode(v.detach()) - tv[:, 0:1].to(torch.float64))[:, 0]
                for step in range(1, K1):
                    act = torch.empty(B, c.action_channels, eh, ew)
                    for b in range(B):
                        act[b].fill_(float(np.float32(float(int(actions[b, step])) / c.num_actions)))
                    h, r = self.dynamics(h, act)
                    p, v = self.prediction(h)
                    h.register_hook(lambda g: g * 0.5)
                    lv, lr_, lp = loss(v, tve[:, step]), loss(r, tre[:, step]), loss(p, tp[:, step])
                    for t in (lv, lr_, lp):
                        t.register_hook(lambda g: g * gs)
                    value_loss, reward_loss, policy_loss = value_loss + lv, reward_loss + lr_, policy_loss + lp
                total = (pri * (value_loss * value_loss_weight + reward_loss + policy_loss)).mean()
                self._opt.zero_grad()
                total.backward()
                self.grads = {k: self.W[k].grad.detach().clone() for k in names if self.W[k].grad is not None}
                self._opt.step()
        finally:
            TRAIN[0] = False
        return dict(total_loss=float(total.detach()), value=float(value_loss.detach().mean()), policy=float(policy_loss.detach().mean()),
                    reward=float(reward_loss.detach().mean()), errors=errors.numpy().astype(np.float64))
