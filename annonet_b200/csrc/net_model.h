// Host-side model of the reference's network (net_type / anet_type) and its serialized form.
//
//  * static layer list     dlib-dnn-pimpl-wrapper/NetStructure.h:25-299  (loss first, input last;
//                          index == dlib's layer<i>(net))
//  * stream format         dlib/dlib/serialize.h:297-345,598-632,1209-1221; cuda/tensor.h:444-466,657-665;
//                          dnn/core.h:1056-1068,1425-1438,1662-1667,2675-2681; layer records
//                          dnn/layers.h:234-252,559-577,1017-1026,1428-1448,2231-2273,3344-3347,2463-2466,3675-3680;
//                          dnn/input.h:119-125
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

namespace ab {

struct serialization_error : std::runtime_error { using std::runtime_error::runtime_error; };

enum LayerType { L_LOSS, L_CON, L_CONT, L_BN, L_AFFINE, L_MISH, L_MAXPOOL, L_CONCAT_PREV, L_CONCAT, L_TAG, L_INPUT };

struct HostTensor {
    long long d[4] = {0, 0, 0, 0};
    std::vector<float> v;
    size_t size() const { return v.size(); }
};

struct Layer {
    LayerType type = L_TAG;
    int tag_id = 0;              // L_TAG: its id; L_CONCAT_PREV: the tag it concatenates
    std::vector<int> tags;       // L_CONCAT
    bool dead = false;           // never consumed
    // con / cont
    long num_filters = 0, nr = 0, nc = 0;
    int stride = 1, pad = 0;
    HostTensor params;           // con/cont: filters then biases; bn/affine: gamma then beta (dlib layout)
    long long falias[4] = {0, 0, 0, 0}, balias[4] = {0, 0, 0, 0};
    double mult[4] = {1, 1, 1, 0};  // lr, wd, bias lr, bias wd multipliers
    bool use_bias = true;
    // bn
    HostTensor means, invstds, running_means, running_variances;
    unsigned long long num_updates = 0, window = 100;
    double eps = 1e-4;
    int affine_mode = 0;
    int pool_geom[6] = {3, 3, 2, 2, 0, 0};
    // add_layer trailer
    bool setup_called = true, grad_stale = true, getters_disabled = false;
};

struct NetModel {
    std::vector<Layer> layers;
    int levels = 6;
    float avg[3] = {122.782f, 117.001f, 104.298f};
    bool is_training = true;  // bn_ records (net_type) vs affine_ (anet_type)

    int class_count() const { return (int)layers[1].num_filters; }
    long long num_params() const;
    bool params_allocated() const;
};

// NetStructure.h:172-299.  features: default filter counts {16,16,24,32,40,48,56}, deepest 64.
std::vector<Layer> static_layers(int levels, int class_count);

// SetNetWidthVisitor (NetPimpl.cpp:73-130) then SetClassCount (NetPimpl.cpp:37-41), the CLI's call
// order (annonet_train_main.cpp:402,405); then allocate parameters like con_::setup / bn_::setup
// (layers.h:170-193,1340-1365): Glorot-uniform filters, zero biases, gamma 1, beta 0, running stats 0/1... .
NetModel make_random_net(int class_count, double scaler, int min_filters, uint64_t seed, int levels = 6);

// consumed == nullptr: the net must fill the buffer; else *consumed = bytes read (dlib streams are not length-prefixed)
NetModel parse_net(const void* bytes, size_t n, int levels = 6, size_t* consumed = nullptr);
std::vector<char> serialize_net(const NetModel& net);
// RuntimeNet::operator=(TrainingNet): bn_ -> affine_ (layers.h:2144-2163)
NetModel to_runtime(const NetModel& net);

// output channel count of every layer, following dlib's executor (concat_prev: sub first, then tag)
std::vector<int> layer_channels(const std::vector<Layer>& layers);
int find_tag(const std::vector<Layer>& layers, int from, int tag_id);

// RuntimeNet::GetRecommendedInputDimension (NetPimpl.cpp:271-285); throws std::runtime_error when too large
int recommended_input_dimension(int minimum, int levels = 6);

}  // namespace ab
