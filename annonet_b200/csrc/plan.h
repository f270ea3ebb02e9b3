// Static execution plan of net_type / anet_type for the engine: the reference's 231-layer template
// graph (NetStructure.h:172-299, executor semantics dnn/core.h:932-948,998-1009) is folded into 49
// "conv groups" (conv [+ max-pool in front] [+ mish / bn|affine behind]) that read and write channel
// slices of ONE activation buffer per resolution level, so that no concat / tag layer ever copies
// (SURVEY §2.3 "concat: eliminated").
//
// Physical channel order inside a level buffer = reverse creation order of the tensors living at
// that resolution ("newest first").  With that order every conv group's input -- whatever
// concat_prev_/concat_ chain produced it (layers.h:2425-2455,3652-3667) -- is one contiguous channel
// range; where dlib's logical channel order differs from the physical one (concat2/concat3 list
// their tags oldest first) the group's filters are permuted along Cin when they are loaded.
#pragma once
#include "net_model.h"

namespace ab {

struct ConvGroup {
    int dlib_conv = -1;   // layer<i> of the con_/cont_
    int dlib_bn = -1;     // layer<i> of its bn_/affine_ (-1: none, the class head and the `up` conts)
    bool has_mish = false;
    bool bn_first = false;  // conv -> bn -> mish (level0, level0t) instead of conv -> mish -> bn (dense / down)
    bool is_cont = false;
    int ksize = 3, stride = 1, pad = 1;
    bool pool_input = false;  // max_pool 3x3/2 in front (transition_down)
    bool dead = false;        // output never consumed
    int in_level = 0;         // resolution level of the input slice; -1: the image tensor
    int in_c0 = 0, cin = 0;
    std::vector<int> perm;    // perm[logical ci] = physical ci (relative to in_c0)
    int out_level = 0;        // -2: logits
    int out_c0 = 0, cout = 0; // for dead groups: offset into the dead-scratch buffer (always 0)
    // offsets (floats) into the flat parameter arena
    size_t w_off = 0, b_off = 0, bn_off = 0;
    size_t nweights() const { return (size_t)cout * cin * ksize * ksize; }
};

struct Plan {
    int levels = 6;
    int class_count = 0;
    std::vector<ConvGroup> groups;        // forward order
    std::vector<int> level_channels;      // channels of the activation buffer of each level (0..levels)
    int max_dead_channels = 0;
    size_t num_params = 0;                // floats in the arena
    // per dlib layer index: arena offset and size of its params tensor (con/cont/bn/affine), else size 0
    std::vector<size_t> layer_off, layer_size;
};

Plan build_plan(const NetModel& net);

// spatial size of each level for an h x w input: level 0 = input size, level n = 1+(level n-1 - 3)/2
// (max_pool_<3,3,2,2>, layers.h:994-1012).  The `up` path comes back with 2*in+1 (cont_<3,3,2,2,0,0>, layers.h:521-535):
// up_h/up_w[n] is the size of the up-sampled tensor that joins level n.  For "canonical" sizes 64k+63 (what
// RuntimeNet::GetRecommendedInputDimension hands the CLIs) it equals the level's size; an even-sized level comes back one
// pixel short and concat_prev_ resizes it bilinearly (layers.h:2438,2494-2521): resized[n].
// first_resized: the finest level with such a join (-1: none).  In the reference the gradient does NOT pass a resized join
// (resize_if_needed hands the backward pass a temporary, layers.h:2452): every layer whose gradient arrives through it gets
// exactly zero -- restated by the engines (grad_blocked) and checked against the compiled reference.
struct LevelDims {
    int h[8], w[8];
    int up_h[8], up_w[8];
    bool resized[8];
    int first_resized;
};
LevelDims level_dims(int h, int w, int levels);
// does group g's parameter gradient vanish because its only gradient path crosses a resized join?
inline bool grad_blocked(const LevelDims& d, const ConvGroup& g)
{
    if (d.first_resized < 0) return false;
    const int ol = g.out_level == -2 ? 0 : g.out_level;
    return ol > d.first_resized || (g.is_cont && g.stride == 2 && ol == d.first_resized);
}

// source rows / columns and fractions of dlib's bilinear resize_image along one axis (image_transforms/interpolation.h:
// 722-806), bit for bit: rows accumulate y in double; columns run in groups of four SSE floats (x_j = float(-4 s + j s),
// advanced by float(4 s), left = trunc(x)) until a group's last `right` would leave the row, the rest in double
struct ResizeAxis { std::vector<int> lo, hi; std::vector<float> frac, ifrac; };   // ifrac = 1 - frac as dlib rounds it
ResizeAxis resize_axis(int n_in, int n_out, bool simd_columns);

// dlib layout <-> arena layout for one group's filters (permutation along Cin)
void pack_filters(const ConvGroup& g, const float* dlib_params, float* arena_w, float* arena_b);
void unpack_filters(const ConvGroup& g, const float* arena_w, const float* arena_b, float* dlib_params);

}  // namespace ab
