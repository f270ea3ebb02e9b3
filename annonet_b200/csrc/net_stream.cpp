// parse_net / serialize_net / to_runtime: the reference's serialized-net format (see net_model.h for
// the citations).  Stream order: versions on the way down (loss -> input), payloads on the way
// back up (input-side layer first, head last)  -- dnn/core.h:1056-1068.
#include "net_model.h"
#include "stream_primitives.h"

#include <cmath>
#include <cstring>

namespace ab {

static const char* LOSS_NAME = "loss_multiclass_log_per_pixel_weighted_";

static void read_details(Reader& r, Layer& l, bool& saw_bn, bool& saw_affine)
{
    const std::string name = r.string();
    switch (l.type) {
    case L_CON:
    case L_CONT: {
        const char* want = l.type == L_CON ? "con_5" : "cont_2";
        if (name != want) throw serialization_error(std::string("expected ") + want + ", found '" + name + "'");
        r.tensor(l.params);
        l.num_filters = (long)r.integer();
        const long nr = (long)r.integer(), nc = (long)r.integer();
        const int sy = (int)r.integer(), sx = (int)r.integer(), py = (int)r.integer(), px = (int)r.integer();
        if (nr != l.nr || nc != l.nc || sy != l.stride || sx != l.stride || py != l.pad || px != l.pad)
            throw serialization_error(name + ": geometry differs from NetStructure.h");
        r.alias(l.falias);
        r.alias(l.balias);
        for (int i = 0; i < 4; ++i) l.mult[i] = r.real();
        l.use_bias = r.boolean();
        if (!l.use_bias) throw serialization_error(name + ": use_bias = false is not part of net_type");
        break;
    }
    case L_BN:
    case L_AFFINE:
        if (name == "bn_con2") {
            l.type = L_BN;
            saw_bn = true;
            r.tensor(l.params);
            r.alias(l.falias);
            r.alias(l.balias);
            r.tensor(l.means);
            r.tensor(l.invstds);
            r.tensor(l.running_means);
            r.tensor(l.running_variances);
            l.num_updates = (unsigned long long)r.integer();
            l.window = (unsigned long long)r.integer();
            for (int i = 0; i < 4; ++i) l.mult[i] = r.real();
            l.eps = r.real();
        } else if (name == "affine_") {
            l.type = L_AFFINE;
            saw_affine = true;
            r.tensor(l.params);
            r.alias(l.falias);
            r.alias(l.balias);
            l.affine_mode = (int)r.integer();
        } else
            throw serialization_error("expected bn_con2 or affine_, found '" + name + "'");
        break;
    case L_MISH:
        if (name != "mish_") throw serialization_error("expected mish_, found '" + name + "'");
        break;
    case L_MAXPOOL:
        if (name != "max_pool_2") throw serialization_error("expected max_pool_2, found '" + name + "'");
        for (int i = 0; i < 6; ++i) l.pool_geom[i] = (int)r.integer();
        if (l.pool_geom[0] != 3 || l.pool_geom[1] != 3 || l.pool_geom[2] != 2 || l.pool_geom[3] != 2 || l.pool_geom[4] != 0 || l.pool_geom[5] != 0)
            throw serialization_error("max_pool_2: geometry differs from NetStructure.h");
        break;
    case L_CONCAT_PREV:
        if (name != "concat_prev_") throw serialization_error("expected concat_prev_, found '" + name + "'");
        break;
    case L_CONCAT:
        if (name != "concat_") throw serialization_error("expected concat_, found '" + name + "'");
        if ((size_t)r.integer() != l.tags.size()) throw serialization_error("concat_: tag count mismatch");
        break;
    default:
        throw serialization_error("unexpected layer type in stream");
    }
}

static void write_details(Writer& w, const Layer& l)
{
    switch (l.type) {
    case L_CON:
    case L_CONT:
        w.string(l.type == L_CON ? "con_5" : "cont_2");
        w.tensor(l.params);
        w.integer(l.num_filters);
        w.integer(l.nr); w.integer(l.nc);
        w.integer(l.stride); w.integer(l.stride);
        w.integer(l.pad); w.integer(l.pad);
        w.alias(l.falias);
        w.alias(l.balias);
        for (int i = 0; i < 4; ++i) w.real<double>(l.mult[i]);
        w.boolean(l.use_bias);
        break;
    case L_BN:
        w.string("bn_con2");
        w.tensor(l.params);
        w.alias(l.falias);
        w.alias(l.balias);
        w.tensor(l.means);
        w.tensor(l.invstds);
        w.tensor(l.running_means);
        w.tensor(l.running_variances);
        w.integer((long long)l.num_updates);
        w.integer((long long)l.window);
        for (int i = 0; i < 4; ++i) w.real<double>(l.mult[i]);
        w.real<double>(l.eps);
        break;
    case L_AFFINE:
        w.string("affine_");
        w.tensor(l.params);
        w.alias(l.falias);
        w.alias(l.balias);
        w.integer(l.affine_mode);
        break;
    case L_MISH: w.string("mish_"); break;
    case L_MAXPOOL:
        w.string("max_pool_2");
        for (int i = 0; i < 6; ++i) w.integer(l.pool_geom[i]);
        break;
    case L_CONCAT_PREV: w.string("concat_prev_"); break;
    case L_CONCAT:
        w.string("concat_");
        w.integer((long long)l.tags.size());
        break;
    default: break;
    }
}

NetModel parse_net(const void* bytes, size_t nbytes, int levels, size_t* consumed)
{
    NetModel net;
    net.levels = levels;
    net.layers = static_layers(levels, 2);
    std::vector<Layer>& L = net.layers;
    const int n = (int)L.size();
    Reader r(bytes, nbytes);
    if (r.integer() != 1) throw serialization_error("unexpected add_loss_layer version");
    if (r.string() != LOSS_NAME) throw serialization_error("unexpected loss layer; this engine serves loss_multiclass_log_per_pixel_weighted_");
    for (int i = 1; i < n - 1; ++i) {
        const long long v = r.integer();
        const long long want = L[i].type == L_TAG ? 1 : (i == n - 2 ? 3 : 2);
        if (v != want) throw serialization_error("layer<" + std::to_string(i) + ">: unexpected version (is this a LEVEL_COUNT=6 RGB net?)");
    }
    if (r.string() != "input_rgb_image") throw serialization_error("unexpected input layer (grayscale nets are not built)");
    for (int i = 0; i < 3; ++i) net.avg[i] = (float)r.real();
    bool saw_bn = false, saw_affine = false;
    HostTensor scratch;
    for (int i = n - 2; i >= 1; --i) {
        Layer& l = L[i];
        if (l.type == L_TAG) continue;
        read_details(r, l, saw_bn, saw_affine);
        l.setup_called = r.boolean();
        l.grad_stale = r.boolean();
        l.getters_disabled = r.boolean();
        r.tensor(scratch);  // x_grad
        r.tensor(scratch);  // cached_output
        r.tensor(scratch);  // params_grad, or grad_final on the input-side layer
        if (i == n - 2) r.integer();  // sample_expansion_factor
    }
    if (consumed) *consumed = r.pos;
    else if (r.pos != nbytes) throw serialization_error(std::to_string(nbytes - r.pos) + " trailing bytes after the net");
    if (saw_bn && saw_affine) throw serialization_error("stream mixes bn_con2 and affine_ records");
    net.is_training = !saw_affine;
    return net;
}

std::vector<char> serialize_net(const NetModel& net)
{
    Writer w;
    const std::vector<Layer>& L = net.layers;
    const int n = (int)L.size();
    w.integer(1);
    w.string(LOSS_NAME);
    for (int i = 1; i < n - 1; ++i) w.integer(L[i].type == L_TAG ? 1 : (i == n - 2 ? 3 : 2));
    w.string("input_rgb_image");
    for (int i = 0; i < 3; ++i) w.real<float>(net.avg[i]);
    for (int i = n - 2; i >= 1; --i) {
        const Layer& l = L[i];
        if (l.type == L_TAG) continue;
        write_details(w, l);
        w.boolean(l.setup_called);
        w.boolean(l.grad_stale);
        w.boolean(l.getters_disabled);
        w.empty_tensor();
        w.empty_tensor();
        w.empty_tensor();
        if (i == n - 2) w.integer(1);
    }
    return std::move(w.out);
}

NetModel to_runtime(const NetModel& net)
{
    NetModel out = net;
    out.is_training = false;
    for (Layer& l : out.layers) {
        if (l.type != L_BN) continue;
        const size_t c = l.params.size() / 2;
        // affine_(const bn_&), layers.h:2144-2163: gamma' = gamma/sqrt(running_var + eps), beta' = beta - gamma'*running_mean
        for (size_t j = 0; j < c; ++j) {
            const float g = l.params.v[j] / std::sqrt(l.running_variances.v[j] + (float)l.eps);
            const float b = l.params.v[c + j] - g * l.running_means.v[j];
            l.params.v[j] = g;
            l.params.v[c + j] = b;
        }
        l.type = L_AFFINE;
        l.affine_mode = 0;  // CONV_MODE
        l.means = l.invstds = l.running_means = l.running_variances = HostTensor();
    }
    return out;
}

}  // namespace ab
