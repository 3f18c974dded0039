// net_plan.h — host-side description of the VPRNet towers after eval-mode BN folding.
// Built once from the reference's named module tensors (mzb_load_weights); consumed by the fp32 check
// backend (net_fp32.cu) and the bf16 tcgen05 backend (net_bf16.cu).
//
// Reference structure restated here (not its code): MuzeroNetworkImpl (muzero-cpp/models.cpp:171-265),
// Representation/Dynamics/Prediction networks (models.cpp:31-166), ResidualBlock / ResidualHead /
// ResidualHeadDownsample / MLP (muzero-cpp/model_layers.cpp:46-162), bn() eps=1e-4 (model_layers.cpp:36-42).
#pragma once
#include <map>
#include <string>
#include <vector>

#include "common.h"

namespace mzb {

// 3x3 convolution, padding 1, stride 1|2, eval-mode BatchNorm folded in:  y = act(conv(x, w) + bias [+ table] [+ res])
struct Conv3Host {
    int cin = 0, cout = 0, stride = 1;
    int in_h = 0, in_w = 0, out_h = 0, out_w = 0;
    bool relu = false;
    std::vector<float> w;          // [cout][9][cin]   (tap = ky*3+kx), BN scale folded
    std::vector<float> bias;       // [cout]           BN shift (0 when the conv has no BN)
    // dynamics head only: contribution of the action planes, conv(plane_a)[pix][cout] * bn_scale, per action
    std::vector<float> act_table;  // [A][out_h*out_w][cout] or empty
};

// conv1x1(C -> red, bias) -> flatten channel-major (r*HW + p) -> MLP (Linear+ReLU ... Linear)
struct HeadHost {
    int C = 0, red = 0, hw = 0;
    std::vector<float> w1, b1;                 // [red][C], [red]
    std::vector<int> sizes;                    // {red*hw, hidden..., out}
    std::vector<std::vector<float>> W, b;      // per layer [out][in], [out]
};

struct TowerStep {
    enum Kind { CONV, BLOCK, POOL } kind;
    int conv0 = -1, conv1 = -1;  // indices into Tower::convs (BLOCK uses both)
};

struct Tower {
    std::vector<Conv3Host> convs;
    std::vector<TowerStep> steps;
    int in_c = 0, in_h = 0, in_w = 0, out_c = 0, out_h = 0, out_w = 0;
    int max_act_elems = 0;  // largest activation (per sample) any step produces or consumes
};

struct NetPlan {
    NetDims dims;
    bool normalize = true;
    Tower representation, dynamics, prediction;
    HeadHost reward_head, policy_head, value_head;
    EncoderConsts value_enc, reward_enc;
    double flops_initial = 0, flops_recurrent = 0;  // algorithmic FLOPs per sample (SURVEY.md §8d formula)
};

typedef std::map<std::string, std::pair<const float*, std::vector<int64_t>>> TensorMap;

// Derive shapes from the config alone (no weights needed).
int derive_dims(const mzb_config& cfg, NetDims* out, EncoderConsts* venc, EncoderConsts* renc);
// Build the folded plan.  Returns MZB_OK or sets the error string.
int build_plan(const mzb_config& cfg, const TensorMap& tensors, NetPlan* plan);

}  // namespace mzb
