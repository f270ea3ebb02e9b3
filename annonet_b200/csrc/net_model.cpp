// See net_model.h.  Host-only code (no CUDA).
#include "net_model.h"

#include <cmath>
#include <cstring>
#include <random>

namespace ab {

namespace {

// ---------------------------------------------------------------------------------------------
// static layer list (NetStructure.h)

const int FEATURES[7] = {16, 16, 24, 32, 40, 48, 56};   // NetStructure.h:118-124 (level N feature counts)
const int DEEPEST[7] = {16, 24, 32, 40, 48, 56, 64};
const int DENSE_KIND[7] = {0, 1, 2, 2, 3, 3, 4};        // NetStructure.h:142-147
enum { DINTAG = 2000, DOTAG0, DCTAG0, DOTAG1, DCTAG1, DOTAG2 };
inline int UTAG(int n) { return 2100 + n; }

Layer mk(LayerType t) { Layer l; l.type = t; return l; }
Layer con(int n, int k, bool dead = false) {
    Layer l = mk(L_CON); l.num_filters = n; l.nr = l.nc = k; l.stride = 1; l.pad = k / 2; l.dead = dead;
    return l;
}
Layer cont(int n, int stride) {
    Layer l = mk(L_CONT); l.num_filters = n; l.nr = l.nc = 3; l.stride = stride; l.pad = stride != 1 ? 0 : 1;
    return l;
}
Layer tag(int id) { Layer l = mk(L_TAG); l.tag_id = id; return l; }
Layer cprev(int id) { Layer l = mk(L_CONCAT_PREV); l.tag_id = id; return l; }
Layer concat(std::vector<int> ids) { Layer l = mk(L_CONCAT); l.tags = std::move(ids); return l; }

void dense_layer(std::vector<Layer>& L, int n, bool dead = false) {  // NetStructure.h:25: BN<mish<con 3x3>>
    L.push_back(mk(L_BN)); L.push_back(mk(L_MISH)); L.push_back(con(n, 3, dead));
}

void dense(std::vector<Layer>& L, int kind, int n) {  // dense_block{1..4}, NetStructure.h:40-66 (outermost first)
    if (kind == 1) {
        L.push_back(cprev(DINTAG)); dense_layer(L, n); L.push_back(tag(DINTAG));
    } else if (kind == 2) {
        L.push_back(cprev(DOTAG0)); dense_layer(L, n);
        L.push_back(cprev(DINTAG)); L.push_back(tag(DOTAG0)); dense_layer(L, n); L.push_back(tag(DINTAG));
    } else if (kind == 3) {
        L.push_back(concat({DOTAG0, DOTAG1})); dense_layer(L, n, true);
        L.push_back(cprev(DCTAG0)); L.push_back(tag(DOTAG1)); dense_layer(L, n);
        L.push_back(tag(DCTAG0)); L.push_back(cprev(DINTAG)); L.push_back(tag(DOTAG0)); dense_layer(L, n);
        L.push_back(tag(DINTAG));
    } else {
        L.push_back(concat({DOTAG0, DOTAG1, DOTAG2})); dense_layer(L, n, true);
        L.push_back(cprev(DCTAG1)); L.push_back(tag(DOTAG2)); dense_layer(L, n);
        L.push_back(tag(DCTAG1)); L.push_back(cprev(DCTAG0)); L.push_back(tag(DOTAG1)); dense_layer(L, n);
        L.push_back(tag(DCTAG0)); L.push_back(cprev(DINTAG)); L.push_back(tag(DOTAG0)); dense_layer(L, n);
        L.push_back(tag(DINTAG));
    }
}

}  // namespace

std::vector<Layer> static_layers(int levels, int class_count)
{
    if (levels != 6) throw std::invalid_argument("only DLIB_DNN_PIMPL_WRAPPER_LEVEL_COUNT = 6 is built");
    std::vector<Layer> L;
    L.push_back(mk(L_LOSS));
    L.push_back(con(class_count, 1));
    L.push_back(mk(L_MISH)); L.push_back(mk(L_BN)); L.push_back(cont(FEATURES[0], 1));  // level0t
    for (int n = 1; n <= levels; ++n) {                                                  // level1t ... level6t
        dense(L, DENSE_KIND[n], FEATURES[n]);
        L.push_back(cprev(UTAG(n)));
        L.push_back(cont(FEATURES[n], 2));
    }
    dense(L, 4, DEEPEST[levels]);
    for (int n = levels; n >= 1; --n) {                                                  // level6 ... level1
        L.push_back(mk(L_BN)); L.push_back(mk(L_MISH)); L.push_back(con(FEATURES[n], 1)); L.push_back(mk(L_MAXPOOL));
        L.push_back(tag(UTAG(n)));
        dense(L, DENSE_KIND[n], FEATURES[n]);
    }
    L.push_back(mk(L_MISH)); L.push_back(mk(L_BN)); L.push_back(con(FEATURES[0], 3));    // level0
    L.push_back(mk(L_INPUT));
    return L;
}

int find_tag(const std::vector<Layer>& layers, int from, int tag_id)
{
    for (size_t j = from + 1; j < layers.size(); ++j)
        if (layers[j].type == L_TAG && layers[j].tag_id == tag_id) return (int)j;
    throw std::logic_error("tag not found below layer " + std::to_string(from));
}

std::vector<int> layer_channels(const std::vector<Layer>& layers)
{
    const int n = (int)layers.size();
    std::vector<int> k(n, 0);
    k[n - 1] = 3;
    for (int i = n - 2; i >= 1; --i) {
        const Layer& l = layers[i];
        if (l.type == L_CON || l.type == L_CONT) k[i] = (int)l.num_filters;
        else if (l.type == L_CONCAT_PREV) k[i] = k[i + 1] + k[find_tag(layers, i, l.tag_id)];
        else if (l.type == L_CONCAT) { for (int t : l.tags) k[i] += k[find_tag(layers, i, t)]; }
        else k[i] = k[i + 1];
    }
    return k;
}

long long NetModel::num_params() const
{
    long long s = 0;
    for (const Layer& l : layers) s += (long long)l.params.size();
    return s;
}

bool NetModel::params_allocated() const
{
    for (const Layer& l : layers)
        if ((l.type == L_CON || l.type == L_CONT || l.type == L_BN || l.type == L_AFFINE) && l.params.size() == 0) return false;
    return true;
}

int recommended_input_dimension(int minimum, int levels)
{
    for (int w = 1; w < 500; ++w) {  // NetPimpl.cpp:246-285: table of Inputs<levels, w> for w = 1..499
        const int d = (w + 1) * (1 << levels) - 1;
        if (d >= minimum) return d;
    }
    throw std::runtime_error("Requested minimum input dimension " + std::to_string(minimum) + " is too large");
}

static void allocate_params(NetModel& net, std::mt19937_64* rng)
{
    std::vector<Layer>& L = net.layers;
    const std::vector<int> k = layer_channels(L);
    const int n = (int)L.size();
    for (int i = n - 2; i >= 1; --i) {
        Layer& l = L[i];
        if (l.type == L_CON || l.type == L_CONT) {
            const long cin = k[i + 1], cout = l.num_filters, ks = l.nr;
            const long num_inputs = ks * ks * cin;
            l.params.d[0] = num_inputs * cout + cout; l.params.d[1] = l.params.d[2] = l.params.d[3] = 1;
            l.params.v.assign((size_t)l.params.d[0], 0.f);
            if (rng) {
                // randomize_parameters(params, num_inputs+num_outputs, rnd), dnn/utilities.h:33-47
                std::uniform_real_distribution<float> u(0.f, 1.f);
                const double lim = std::sqrt(6.0 / (double)(num_inputs + cout));
                for (long j = 0; j < num_inputs * cout; ++j) {
                    float v = 2 * u(*rng) - 1;
                    v *= (float)lim;
                    l.params.v[j] = v;
                }
            }
            if (l.type == L_CON) { l.falias[0] = cout; l.falias[1] = cin; } else { l.falias[0] = cin; l.falias[1] = cout; }
            l.falias[2] = l.falias[3] = ks;
            l.balias[0] = 1; l.balias[1] = cout; l.balias[2] = l.balias[3] = 1;
            l.mult[0] = 1; l.mult[1] = 1; l.mult[2] = 1; l.mult[3] = 0;   // layers.h:56-59
        } else if (l.type == L_BN) {
            const long c = k[i + 1];
            l.params.d[0] = 2 * c; l.params.d[1] = l.params.d[2] = l.params.d[3] = 1;
            l.params.v.assign((size_t)2 * c, 0.f);
            for (long j = 0; j < c; ++j) l.params.v[j] = 1.f;
            l.falias[0] = 1; l.falias[1] = c; l.falias[2] = l.falias[3] = 1;
            std::memcpy(l.balias, l.falias, sizeof(l.falias));
            l.running_means.d[0] = 1; l.running_means.d[1] = c; l.running_means.d[2] = l.running_means.d[3] = 1;
            l.running_means.v.assign((size_t)c, 0.f);
            l.running_variances = l.running_means;
            l.running_variances.v.assign((size_t)c, 1.f);
            l.num_updates = 0;
            l.mult[0] = 1; l.mult[1] = 0; l.mult[2] = 1; l.mult[3] = 1;   // layers.h:1325-1328
        }
    }
}

NetModel make_random_net(int class_count, double scaler, int min_filters, uint64_t seed, int levels)
{
    if (class_count < 1 || class_count >= 65535) throw std::invalid_argument("class count out of range");
    NetModel net;
    net.levels = levels;
    net.is_training = true;
    net.layers = static_layers(levels, class_count);
    for (size_t i = 1; i < net.layers.size(); ++i) {
        Layer& l = net.layers[i];
        if (l.type == L_CON || l.type == L_CONT)
            l.num_filters = std::max<long>(min_filters, (long)std::lround(scaler * (double)l.num_filters));
    }
    net.layers[1].num_filters = class_count;
    std::mt19937_64 rng(seed);
    allocate_params(net, &rng);
    return net;
}

}  // namespace ab
