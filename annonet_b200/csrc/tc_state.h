// State of the bf16 tensor-core path of one Engine (AB_PRECISION_BF16): padded flat NHWC workspaces,
// bf16 filter packs and the prepared launch descriptors (tensor maps) of every conv group.
#pragma once
#include "engine.h"
#include "tc.h"

#include <vector>

namespace ab {

struct TcGroup {
    // forward filters: one pack per launch (1, or the 4 parity classes of a stride-2 cont_)
    size_t fwd_pack_off[4] = {0, 0, 0, 0};
    int fwd_ntaps[4] = {0, 0, 0, 0};
    int fwd_tapmap[4][9] = {};
    int fwd_tapdy[4][9] = {}, fwd_tapdx[4][9] = {};  // tap offset = dy*P + dx of the M grid
    int nlaunch = 1;
    int npad = 0;
    std::vector<tc::ConvLaunch> infer;   // prepared for the current inference geometry
    // training
    std::vector<tc::ConvLaunch> fwd;     // conv -> z (or destination)
    size_t bwd_pack_off[4] = {0, 0, 0, 0};
    std::vector<tc::ConvLaunch> dgrad;
};

struct TcWorkspace {
    int n = 0, h = 0, w = 0;
    tc::Geo geo[8];
    DevBuf level[8];           // bf16 [positions][level_channels]
    DevBuf img;                // bf16 [positions][8]
    DevBuf pool;               // pooled inputs of the transition_down groups, dense
    std::vector<size_t> pool_off;  // per group, in elements
};

struct TcState {
    std::vector<TcGroup> groups;
    DevBuf fwd_packs, bwd_packs;
    bool fwd_packs_valid = false, bwd_packs_valid = false;
    TcWorkspace iw;            // inference
    // training
    TcWorkspace tw, tg;        // activations / gradients
    DevBuf z, dz, dpool, glogits16, stats_sums;
    std::vector<size_t> z_off;
};

}  // namespace ab
