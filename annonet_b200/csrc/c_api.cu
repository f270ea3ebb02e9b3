// extern "C" surface declared in include/annonet_b200.h: translates C++ exceptions into status codes.
#include "../../include/annonet_b200.h"

#include "engine.h"
#include "prof.h"
#include "kcommon.cuh"
#include "tiling.h"

#include <cstdlib>
#include <cstring>

namespace ab { void nccl_unique_id(void* id128); }
namespace ab { namespace noalloc {   // noalloc.cu
int max_classes();
void confusion_per_pixel(const uint16_t* truth_dev, const uint16_t* pred_dev, long long n, int K, unsigned long long* out_dev, cudaStream_t st);
} }

struct ab_net {
    std::unique_ptr<ab::Engine> eng;
};

namespace {

thread_local std::string g_error;

template <typename F>
ab_status guarded(F&& f)
{
    try {
        f();
        g_error.clear();
        return AB_OK;
    } catch (const ab::serialization_error& e) { g_error = e.what(); return AB_ERR_SERIALIZATION;
    } catch (const ab::cuda_error& e) { g_error = e.what(); return AB_ERR_CUDA;
    } catch (const ab::state_error& e) { g_error = e.what(); return AB_ERR_STATE;
    } catch (const std::invalid_argument& e) { g_error = e.what(); return AB_ERR_ARGUMENT;
    } catch (const std::bad_alloc&) { g_error = "out of host memory"; return AB_ERR_CUDA;
    } catch (const std::runtime_error& e) { g_error = e.what(); return AB_ERR_TOO_LARGE;
    } catch (const std::exception& e) { g_error = e.what(); return AB_ERR_STATE; }
}

ab::Engine& E(ab_net* n)
{
    if (!n || !n->eng) throw std::invalid_argument("null net");
    return *n->eng;
}

}  // namespace

extern "C" {

const char* ab_version(void) { return "annonet_b200 0.1 (sm_100a)"; }
const char* ab_last_error(void) { return g_error.c_str(); }
void ab_free(void* p) { std::free(p); }

ab_status ab_net_create(int device, int precision, ab_net** out)
{
    return guarded([&] {
        if (!out) throw std::invalid_argument("null out");
        if (precision != AB_PRECISION_FP32 && precision != AB_PRECISION_F16) throw std::invalid_argument("unknown precision");
        auto n = std::make_unique<ab_net>();
        n->eng = std::make_unique<ab::Engine>(device, precision);
        *out = n.release();
    });
}

void ab_net_destroy(ab_net* net) { delete net; }

ab_status ab_net_clone(const ab_net* net, ab_net** out)
{
    return guarded([&] {
        ab::Engine& src = E(const_cast<ab_net*>(net));
        auto n = std::make_unique<ab_net>();
        n->eng = std::make_unique<ab::Engine>(src.device(), src.precision());
        if (src.loaded()) n->eng->load(src.snapshot(false));
        n->eng->hyper = src.hyper;
        *out = n.release();
    });
}

ab_status ab_net_info(const ab_net* net, ab_net_info_t* info)
{
    return guarded([&] {
        ab::Engine& e = E(const_cast<ab_net*>(net));
        if (!e.loaded()) throw ab::state_error("no net loaded");
        info->levels = e.plan().levels;
        info->class_count = e.plan().class_count;
        info->is_training_net = e.is_training();
        info->precision = e.precision();
        info->num_params = (int64_t)e.plan().num_params;
        info->num_layers = (int32_t)e.model().layers.size();
        info->device = e.device();
    });
}

ab_status ab_net_init_random(ab_net* net, int class_count, double width_scaler, int min_filter_count, uint64_t seed)
{
    return guarded([&] { E(net).load(ab::make_random_net(class_count, width_scaler, min_filter_count, seed)); });
}

ab_status ab_net_load_dlib(ab_net* net, const void* bytes, size_t n)
{
    return guarded([&] { E(net).load(ab::parse_net(bytes, n)); });
}

ab_status ab_net_load_dlib_prefix(ab_net* net, const void* bytes, size_t n, size_t* consumed)
{
    return guarded([&] {
        if (!consumed) throw std::invalid_argument("consumed must not be null");
        E(net).load(ab::parse_net(bytes, n, 6, consumed));
    });
}

ab_status ab_net_save_dlib(ab_net* net, int as_runtime, void** bytes, size_t* n)
{
    return guarded([&] {
        const std::vector<char> s = ab::serialize_net(E(net).snapshot(as_runtime != 0));
        void* p = std::malloc(s.size());
        if (!p) throw std::bad_alloc();
        std::memcpy(p, s.data(), s.size());
        *bytes = p;
        *n = s.size();
    });
}

ab_status ab_dlib_transcode(const void* bytes, size_t n, int as_runtime, void** out_bytes, size_t* out_n)
{
    return guarded([&] {
        ab::NetModel m = ab::parse_net(bytes, n);
        if (as_runtime && m.is_training) m = ab::to_runtime(m);
        const std::vector<char> s = ab::serialize_net(m);
        void* p = std::malloc(s.size());
        if (!p) throw std::bad_alloc();
        std::memcpy(p, s.data(), s.size());
        *out_bytes = p;
        *out_n = s.size();
    });
}

int ab_required_input_dimension(void) { return (1 << 7) - 1; }  // Inputs<6, 1>: 127 (NetPimpl.cpp:137-146)

ab_status ab_recommended_input_dimension(int minimum, int* out)
{
    return guarded([&] { *out = ab::recommended_input_dimension(minimum); });
}

ab_status ab_forward(ab_net* net, const uint8_t* image_hwc, int h, int w, float* logits_host)
{
    return guarded([&] {
        ab::Engine& e = E(net);
        const float* d = e.forward_host(image_hwc, h, w);
        if (logits_host)
            AB_CUDA(cudaMemcpyAsync(logits_host, d, sizeof(float) * e.plan().class_count * h * w, cudaMemcpyDeviceToHost, e.stream()));
        e.synchronize();
    });
}

ab_status ab_process(ab_net* net, const uint8_t* image_hwc, int h, int w, const double* gains, int ngains, uint16_t* labels_host)
{
    return guarded([&] { E(net).process_host(image_hwc, h, w, gains, ngains, labels_host); });
}

ab_status ab_get_tiles(int width, int height, int max_tile_w, int max_tile_h, int overlap_x, int overlap_y, ab_tile* out, int cap,
                       int* count)
{
    return guarded([&] {
        const std::vector<ab::Tile> t = ab::get_tiles(width, height, max_tile_w, max_tile_h, overlap_x, overlap_y);
        static_assert(sizeof(ab::Tile) == sizeof(ab_tile), "tile layouts must agree");
        if (count) *count = (int)t.size();
        if (out) std::memcpy(out, t.data(), sizeof(ab_tile) * std::min<size_t>(t.size(), (size_t)std::max(cap, 0)));
    });
}

ab_status ab_plan_bands(int width, int height, int max_tile_w, int max_tile_h, int levels, int world, int* out)
{
    return guarded([&] {
        if (world < 1 || levels < 1 || levels > 12 || !out) throw std::invalid_argument("ab_plan_bands: bad arguments");
        const int ov = (1 << (levels + 1)) - 1;
        const std::vector<ab::Tile> t = ab::get_tiles(width, height, max_tile_w, max_tile_h, ov, ov);
        const std::vector<ab::Band> b = ab::plan_bands(t, world, levels);
        for (int r = 0; r < world; ++r) {
            const int v[6] = {b[r].t0, b[r].t1, b[r].own0, b[r].own1, b[r].f0, b[r].f1};
            std::memcpy(out + 6 * r, v, sizeof v);
        }
    });
}

ab_status ab_infer_tiled(ab_net* net, const uint8_t* image_hwc, int h, int w, int max_tile_w, int max_tile_h, const double* gains,
                         int ngains, uint16_t* labels_host, float* blended_host)
{
    return guarded([&] { E(net).infer_tiled_host(image_hwc, h, w, max_tile_w, max_tile_h, gains, ngains, labels_host, blended_host); });
}

ab_status ab_evaluate_labels(ab_net* net, const uint16_t* truth, const uint16_t* predicted, int h, int w, int class_count, uint64_t* per_pixel,
                             uint64_t* per_region)
{
    return guarded([&] {
        if (!truth || !predicted || !per_pixel || h < 1 || w < 1) throw std::invalid_argument("bad arguments");
        static_assert(sizeof(uint64_t) == sizeof(unsigned long long), "64-bit counters");
        E(net).evaluate_labels(truth, predicted, h, w, class_count, (unsigned long long*)per_pixel, (unsigned long long*)per_region);
    });
}

ab_status ab_confusion_per_pixel_noalloc(const uint16_t* truth_dev, const uint16_t* predicted_dev, long long n, int class_count,
                                         uint64_t* per_pixel_dev, void* stream)
{
    return guarded([&] {
        static_assert(AB_NOALLOC_MAX_CLASSES == 64, "keep the header's limit and noalloc.cu's array in step");
        if (ab::noalloc::max_classes() != AB_NOALLOC_MAX_CLASSES) throw ab::state_error("noalloc scratch size and header limit disagree");
        ab::noalloc::confusion_per_pixel(truth_dev, predicted_dev, n, class_count, (unsigned long long*)per_pixel_dev, (cudaStream_t)stream);
    });
}

ab_status ab_render_labels(ab_net* net, const uint16_t* labels, int h, int w, int out_h, int out_w, const uint8_t* palette_rgba, int class_count,
                           uint8_t* rgba_out)
{
    return guarded([&] {
        if (!labels || !palette_rgba || !rgba_out) throw std::invalid_argument("bad arguments");
        E(net).render_labels(labels, h, w, out_h, out_w, palette_rgba, class_count, rgba_out);
    });
}

ab_status ab_set_detection_levels(ab_net* net, const double* levels, int n)
{
    return guarded([&] {
        if (n < 0 || (n > 0 && !levels)) throw std::invalid_argument("bad detection levels");
        for (int i = 0; i < n; ++i) if (!(levels[i] >= 0.0)) throw std::invalid_argument("detection levels must be >= 0 (annonet_infer.cpp:189)");
        E(net).set_detection_levels(levels, n);
    });
}

ab_status ab_infer_tiled_device(ab_net* net, const void* image_dev, int h, int w, int max_tile_w, int max_tile_h, const double* gains,
                                int ngains, void* labels_dev)
{
    return guarded([&] {
        E(net).infer_tiled_device((const uint8_t*)image_dev, h, w, max_tile_w, max_tile_h, gains, ngains, (uint16_t*)labels_dev);
    });
}

ab_status ab_set_hyper(ab_net* net, int id, double v)
{
    return guarded([&] {
        ab::Engine& e = E(net);
        switch (id) {
        case AB_HYPER_LEARNING_RATE: e.hyper.learning_rate = v; break;
        case AB_HYPER_WEIGHT_DECAY: e.hyper.weight_decay = v; break;
        case AB_HYPER_MOMENTUM: e.hyper.momentum = v; break;
        case AB_HYPER_BN_STATS_WINDOW: e.set_bn_window((unsigned long long)v); break;
        case AB_HYPER_LR_SHRINK_FACTOR: e.hyper.lr_shrink_factor = v; break;
        case AB_HYPER_ITERS_WITHOUT_PROGRESS: e.hyper.iters_without_progress = v; break;
        case AB_HYPER_SOLVER: e.hyper.solver = (int)v; break;
        case AB_HYPER_ADAM_BETA1: e.hyper.adam_beta1 = v; break;
        case AB_HYPER_ADAM_BETA2: e.hyper.adam_beta2 = v; break;
        case AB_HYPER_SKIP_DEAD_LAYERS: e.hyper.skip_dead_layers = (int)v; break;
        case AB_HYPER_LOSS_SCALE: e.hyper.loss_scale = v; break;
        default: throw std::invalid_argument("unknown hyper-parameter id");
        }
    });
}

ab_status ab_get_hyper(ab_net* net, int id, double* v)
{
    return guarded([&] {
        ab::Engine& e = E(net);
        switch (id) {
        case AB_HYPER_LEARNING_RATE: *v = e.hyper.learning_rate; break;
        case AB_HYPER_WEIGHT_DECAY: *v = e.hyper.weight_decay; break;
        case AB_HYPER_MOMENTUM: *v = e.hyper.momentum; break;
        case AB_HYPER_LR_SHRINK_FACTOR: *v = e.hyper.lr_shrink_factor; break;
        case AB_HYPER_ITERS_WITHOUT_PROGRESS: *v = e.hyper.iters_without_progress; break;
        case AB_HYPER_SOLVER: *v = e.hyper.solver; break;
        case AB_HYPER_ADAM_BETA1: *v = e.hyper.adam_beta1; break;
        case AB_HYPER_ADAM_BETA2: *v = e.hyper.adam_beta2; break;
        case AB_HYPER_SKIP_DEAD_LAYERS: *v = e.hyper.skip_dead_layers; break;
        case AB_HYPER_LOSS_SCALE: *v = e.hyper.loss_scale; break;
        default: throw std::invalid_argument("unknown hyper-parameter id");
        }
    });
}

ab_status ab_train_step_async(ab_net* net, const uint8_t* images, const ab_wlabel* labels, int n, int h, int w)
{
    return guarded([&] { E(net).train_step_host(images, labels, n, h, w, true); });
}

ab_status ab_train_step_device(ab_net* net, const void* images_dev, const void* labels_dev, int n, int h, int w)
{
    return guarded([&] { E(net).train_step_device((const uint8_t*)images_dev, labels_dev, n, h, w, true); });
}

ab_status ab_train_sync(ab_net* net, double* loss)
{
    return guarded([&] {
        const double l = E(net).train_sync();
        if (loss) *loss = l;
    });
}

ab_status ab_compute_gradients(ab_net* net, const uint8_t* images, const ab_wlabel* labels, int n, int h, int w, double* loss)
{
    return guarded([&] {
        E(net).train_step_host(images, labels, n, h, w, false);
        const double l = E(net).train_sync();
        if (loss) *loss = l;
    });
}

ab_status ab_get_logits(ab_net* net, float* out, size_t capacity_floats)
{
    return guarded([&] {
        ab::Engine& e = E(net);
        const size_t n = (size_t)e.last_n() * e.plan().class_count * e.last_h() * e.last_w();
        if (n == 0) throw ab::state_error("no training step has run");
        if (capacity_floats < n) throw std::invalid_argument("logits buffer too small");
        e.synchronize();
        AB_CUDA(cudaMemcpy(out, e.train_logits(), n * sizeof(float), cudaMemcpyDeviceToHost));
    });
}

ab_status ab_get_level_tensor(ab_net* net, int level, int gradient, float* out, size_t capacity_floats, int32_t dims[4])
{
    return guarded([&] {
        std::vector<float> v;
        int d[4];
        E(net).level_values(level, gradient != 0, v, d);
        for (int i = 0; i < 4; ++i) dims[i] = d[i];
        if (out) {
            if (capacity_floats < v.size()) throw std::invalid_argument("level buffer too small");
            std::memcpy(out, v.data(), v.size() * sizeof(float));
        }
    });
}

static ab_status layer_values(ab_net* net, int idx, int which, float* out, size_t cap, size_t* count)
{
    return guarded([&] {
        std::vector<float> v;
        E(net).layer_values(idx, which, v);
        if (count) *count = v.size();
        if (out) {
            if (cap < v.size()) throw std::invalid_argument("buffer too small");
            std::memcpy(out, v.data(), v.size() * sizeof(float));
        }
    });
}

ab_status ab_get_layer_gradient(ab_net* net, int idx, float* out, size_t cap, size_t* count) { return layer_values(net, idx, 1, out, cap, count); }
ab_status ab_get_layer_params(ab_net* net, int idx, float* out, size_t cap, size_t* count) { return layer_values(net, idx, 0, out, cap, count); }
ab_status ab_get_layer_momentum(ab_net* net, int idx, float* out, size_t cap, size_t* count) { return layer_values(net, idx, 2, out, cap, count); }
ab_status ab_set_layer_momentum(ab_net* net, int idx, const float* v, size_t count) { return guarded([&] { E(net).set_layer_momentum(idx, v, count); }); }

ab_status ab_get_layer_types(ab_net* net, int32_t* kinds, int cap, int* count)
{
    return guarded([&] {
        const ab::NetModel& m = E(net).model();
        if (count) *count = (int)m.layers.size();
        if (kinds) {
            if (cap < (int)m.layers.size()) throw std::invalid_argument("buffer too small");
            for (size_t i = 0; i < m.layers.size(); ++i) kinds[i] = (int32_t)m.layers[i].type;
        }
    });
}

ab_status ab_gradient_arena(ab_net* net, void** grads_dev, size_t* count_floats)
{
    return guarded([&] {
        ab::Engine& e = E(net);
        if (!e.loaded()) throw ab::state_error("no net loaded");
        *grads_dev = e.grads();
        *count_floats = e.plan().num_params;
    });
}

ab_status ab_dataset_clear(ab_net* net) { return guarded([&] { E(net).dataset_clear(); }); }
ab_status ab_dataset_add(ab_net* net, const uint8_t* rgb, const uint16_t* labels, int h, int w, int* index)
{
    return guarded([&] { const int i = E(net).dataset_add(rgb, labels, h, w); if (index) *index = i; });
}
static_assert(sizeof(ab_crop) == sizeof(ab::Engine::CropSpec), "ab_crop mirrors Engine::CropSpec");
ab_status ab_build_crops(ab_net* net, const ab_crop* crops, int n, int dim, double class_weight, double image_weight, const void** imgs_dev,
                         const void** labels_dev)
{
    return guarded([&] {
        const uint8_t* im = nullptr; const void* lb = nullptr;
        E(net).build_crops((const ab::Engine::CropSpec*)crops, n, dim, class_weight, image_weight, &im, &lb);
        if (imgs_dev) *imgs_dev = im;
        if (labels_dev) *labels_dev = lb;
    });
}
ab_status ab_train_step_crops(ab_net* net, const ab_crop* crops, int n, int dim, double class_weight, double image_weight)
{
    return guarded([&] {
        const uint8_t* im = nullptr; const void* lb = nullptr;
        ab::Engine& e = E(net);
        e.build_crops((const ab::Engine::CropSpec*)crops, n, dim, class_weight, image_weight, &im, &lb);
        e.train_step_device(im, lb, n, dim, dim, true);
    });
}

ab_status ab_copy_to_host(ab_net* net, void* dst_host, const void* src_dev, size_t bytes)
{
    return guarded([&] {
        E(net).synchronize();
        if (cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); throw ab::cuda_error("ab_copy_to_host: cudaMemcpy failed"); }
    });
}

ab_status ab_comm_unique_id(void* id128) { return guarded([&] { ab::nccl_unique_id(id128); }); }
ab_status ab_comm_init(ab_net* net, const void* id128, int rank, int world) { return guarded([&] { E(net).comm_init(id128, rank, world); }); }

ab_status ab_set_weights(const uint16_t* labels, int h, int w, double class_weight, double image_weight, float* weights_out)
{
    return guarded([&] {
        if (!labels || !weights_out || h <= 0 || w <= 0) throw std::invalid_argument("bad arguments");
        ab::set_weights(labels, h, w, class_weight, image_weight, weights_out);
    });
}

ab_status ab_profile_enable(int on) { return guarded([&] { ab::k::prof_enable(on != 0); }); }
ab_status ab_profile_read(ab_kernel_stat* out, int cap, int* count)
{
    return guarded([&] {
        ab::k::FamilyStat st[ab::k::F_COUNT];
        ab::k::prof_collect(st);
        if (count) *count = ab::k::F_COUNT;
        for (int i = 0; i < ab::k::F_COUNT && i < cap && out; ++i) {
            std::memset(&out[i], 0, sizeof(out[i]));
            std::strncpy(out[i].name, st[i].name, sizeof(out[i].name) - 1);
            out[i].launches = st[i].launches; out[i].ms = st[i].ms; out[i].flops = st[i].flops; out[i].bytes = st[i].bytes;
        }
    });
}

void* ab_stream(ab_net* net) { return net && net->eng ? (void*)net->eng->stream() : nullptr; }
uint64_t ab_kernel_launches(ab_net*) { return ab::k::g_launches.load(); }
ab_status ab_device_synchronize(ab_net* net) { return guarded([&] { E(net).synchronize(); }); }

}  // extern "C"
