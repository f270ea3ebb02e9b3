// Inference side of the engine: RuntimeNet::Forward / Process (NetPimpl.cpp:287-298) and the tiled
// annonet_infer() (annonet_infer.cpp:32-185) with image, blend canvas and label map resident in HBM.
#include "engine.h"
#include "tiling.h"

#include <cstring>

namespace ab {

void Engine::ensure_infer_ws(int h, int w)
{
    if (iw_.h == h && iw_.w == w) return;
    const LevelDims d = level_dims(h, w, plan_.levels);
    for (int l = 0; l <= plan_.levels; ++l)
        iw_.buf[l].reserve((size_t)plan_.level_channels[l] * d.h[l] * d.w[l] * sizeof(float));
    i_x_.reserve((size_t)3 * h * w * sizeof(float));
    i_logits_.reserve((size_t)plan_.class_count * h * w * sizeof(float));
    iw_.dims = d;
    iw_.n = 1; iw_.h = h; iw_.w = w;
}

// the 49 conv groups, each ONE fused kernel in inference: [max-pool ->] conv -> bias -> mish/affine
void Engine::run_inference_graph(int h, int w)
{
    using namespace k;
    if (model_.is_training) throw state_error("Forward needs a runtime net (anet_type); convert with GetRuntimeNet / as_runtime");
    const LevelDims& d = iw_.dims;
    const float* P = params_.as<float>();
    for (const ConvGroup& g : plan_.groups) {
        if (g.dead) continue;
        ConvArgs a;
        a.N = 1;
        if (g.in_level < 0) { a.in.p = i_x_.as<float>(); a.in.sstride = (long long)3 * h * w; a.in.c0 = 0; a.in.C = 3; }
        else a.in = level_slice(iw_, g.in_level, g.in_c0, g.cin);
        if (g.out_level == -2) { a.out.p = i_logits_.as<float>(); a.out.sstride = (long long)g.cout * h * w; a.out.c0 = 0; a.out.C = g.cout; }
        else a.out = level_slice(iw_, g.out_level, g.out_c0, g.cout);
        const int ol = g.out_level == -2 ? 0 : g.out_level;
        a.H = d.h[ol]; a.W = d.w[ol];
        a.bias = P + g.b_off;
        if (g.dlib_bn >= 0) {
            a.epi = g.bn_first ? EPI_AFFINE_MISH : EPI_MISH_AFFINE;
            a.scale = P + g.bn_off;
            a.shift = P + g.bn_off + g.cout;
        }
        a.wv.w = P + g.w_off;
        if (g.is_cont && g.stride == 2) {
            up_forward(a.in, a.out, 1, d.h[g.in_level], d.w[g.in_level], P + g.w_off, P + g.b_off, stream_);
        } else if (g.ksize == 3) {
            if (!g.is_cont) { a.wv.s_out = g.cin * 9; a.wv.s_in = 9; a.wv.flip = 0; }
            else { a.wv.s_out = 9; a.wv.s_in = g.cout * 9; a.wv.flip = 1; }
            conv3x3(a, stream_);
        } else {
            a.wv.s_out = g.cin; a.wv.s_in = 1;
            a.pool = g.pool_input;
            if (g.pool_input) { a.Hin = d.h[g.in_level]; a.Win = d.w[g.in_level]; }
            conv1x1(a, stream_);
        }
    }
    AB_CUDA(cudaGetLastError());
}

const float* Engine::forward_device(const uint8_t* img_dev, int h, int w)
{
    if (!loaded_) throw state_error("no net loaded");
    AB_CUDA(cudaSetDevice(device_));
    if (precision_ == 1) {  // AB_PRECISION_F16: tcgen05 path
        tc_input(img_dev, h, w, 0, 0, h, w);
        tc_run_inference(h, w);
        return i_logits_.as<float>();
    }
    ensure_infer_ws(h, w);
    k::to_tensor(img_dev, i_x_.as<float>(), 1, h, w, model_.avg[0], model_.avg[1], model_.avg[2], stream_);
    run_inference_graph(h, w);
    return i_logits_.as<float>();
}

const float* Engine::forward_host(const uint8_t* img_host, int h, int w)
{
    if (!loaded_) throw state_error("no net loaded");
    AB_CUDA(cudaSetDevice(device_));
    level_dims(h, w, plan_.levels);  // validate before touching memory
    const size_t bytes = (size_t)h * w * 3;
    void* stage = pinned(bytes);
    std::memcpy(stage, img_host, bytes);
    i_img_.reserve(bytes);
    AB_CUDA(cudaMemcpyAsync(i_img_.p, stage, bytes, cudaMemcpyHostToDevice, stream_));
    return forward_device(i_img_.as<uint8_t>(), h, w);
}

static const double* upload_gains(DevBuf& buf, const double* gains, int ngains, int K, cudaStream_t st)
{
    if (!gains || ngains == 0) return nullptr;
    if (ngains < K) throw std::invalid_argument("fewer gains than classes");
    buf.reserve(sizeof(double) * K);
    AB_CUDA(cudaMemcpyAsync(buf.p, gains, sizeof(double) * K, cudaMemcpyHostToDevice, st));
    return buf.as<double>();
}

void Engine::process_host(const uint8_t* img_host, int h, int w, const double* gains, int ngains, uint16_t* labels_host)
{
    const float* logits = forward_host(img_host, h, w);
    const long long npix = (long long)h * w;
    i_labels_.reserve((size_t)npix * 2);
    const double* gd = upload_gains(gains_dev_, gains, ngains, plan_.class_count, stream_);
    k::argmax_labels(logits, npix, plan_.class_count, npix, gd, i_labels_.as<uint16_t>(), stream_);
    AB_CUDA(cudaMemcpyAsync(labels_host, i_labels_.p, (size_t)npix * 2, cudaMemcpyDeviceToHost, stream_));
    synchronize();
}

void Engine::infer_tiled_device(const uint8_t* img_dev, int h, int w, int max_tile_w, int max_tile_h, int rank, int world)
{
    if (!loaded_) throw state_error("no net loaded");
    AB_CUDA(cudaSetDevice(device_));
    const int K = plan_.class_count;
    const int ov = (1 << (plan_.levels + 1)) - 1;  // TrainingNet::GetRequiredInputDimension, annonet_infer_main.cpp:421-427
    const std::vector<Tile> tiles = get_tiles(w, h, max_tile_w, max_tile_h, ov, ov);
    canvas_.reserve((size_t)K * h * w * sizeof(float));
    AB_CUDA(cudaMemsetAsync(canvas_.p, 0, (size_t)K * h * w * sizeof(float), stream_));
    int row = -1, last_top = INT32_MIN;
    for (const Tile& t : tiles) {
        if (t.full[1] != last_top) { ++row; last_top = t.full[1]; }
        if (world > 1 && row % world != rank) continue;
        const int fw = t.full[2] - t.full[0] + 1, fh = t.full[3] - t.full[1] + 1;
        const int cx = t.full[0] + fw / 2, cy = t.full[1] + fh / 2;
        const int rw = recommended_input_dimension(fw, plan_.levels), rh = recommended_input_dimension(fh, plan_.levels);
        const int left = cx - rw / 2, top = cy - rh / 2;
        if (precision_ == 1) {
            tc_input(img_dev, h, w, left, top, rh, rw);
            tc_run_inference(rh, rw);
        } else {
            ensure_infer_ws(rh, rw);
            k::extract_tile(img_dev, h, w, left, top, i_x_.as<float>(), rh, rw, model_.avg[0], model_.avg[1], model_.avg[2], stream_);
            run_inference_graph(rh, rw);
        }
        k::BlendArgs b;
        b.logits = i_logits_.as<float>(); b.K = K; b.th = rh; b.tw = rw;
        b.canvas = canvas_.as<float>(); b.H = h; b.W = w; b.left = left; b.top = top;
        for (int i = 0; i < 4; ++i) { b.full[i] = t.full[i]; b.unique[i] = t.unique[i]; }
        k::blend_tile(b, stream_);
    }
    AB_CUDA(cudaGetLastError());
}

void Engine::infer_finish_device(int h, int w, const double* gains, int ngains, uint16_t* labels_dev)
{
    const long long npix = (long long)h * w;
    const double* gd = upload_gains(gains_dev_, gains, ngains, plan_.class_count, stream_);
    k::argmax_labels(canvas_.as<float>(), npix, plan_.class_count, npix, gd, labels_dev, stream_);
    AB_CUDA(cudaGetLastError());
}

void Engine::infer_tiled_host(const uint8_t* img_host, int h, int w, int max_tile_w, int max_tile_h, const double* gains, int ngains,
                              uint16_t* labels_host, float* blended_host)
{
    AB_CUDA(cudaSetDevice(device_));
    const size_t bytes = (size_t)h * w * 3;
    i_img_.reserve(bytes);
    // large images: copy straight from the caller's memory (the driver stages pageable memory itself)
    AB_CUDA(cudaMemcpyAsync(i_img_.p, img_host, bytes, cudaMemcpyHostToDevice, stream_));
    infer_tiled_device(i_img_.as<uint8_t>(), h, w, max_tile_w, max_tile_h, 0, 1);
    i_labels_.reserve((size_t)h * w * 2);
    infer_finish_device(h, w, gains, ngains, i_labels_.as<uint16_t>());
    AB_CUDA(cudaMemcpyAsync(labels_host, i_labels_.p, (size_t)h * w * 2, cudaMemcpyDeviceToHost, stream_));
    if (blended_host)
        AB_CUDA(cudaMemcpyAsync(blended_host, canvas_.p, (size_t)plan_.class_count * h * w * sizeof(float), cudaMemcpyDeviceToHost, stream_));
    synchronize();
}

}  // namespace ab
