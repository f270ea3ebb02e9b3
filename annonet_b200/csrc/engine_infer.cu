// Inference side of the engine: RuntimeNet::Forward / Process (NetPimpl.cpp:287-298) and the tiled
// annonet_infer() (annonet_infer.cpp:32-185) with image, blend canvas and label map resident in HBM.
#include "engine.h"
#include "nccl_api.h"
#include "tiling.h"

#include <cmath>
#include <cstring>

namespace ab {

void Engine::ensure_infer_ws(int h, int w)
{
    if (iw_.h == h && iw_.w == w) return;
    const LevelDims d = level_dims(h, w, plan_.levels);
    for (int l = 0; l <= plan_.levels; ++l)
        iw_.buf[l].reserve((size_t)plan_.level_channels[l] * d.h[l] * d.w[l] * sizeof(float));
    size_t up_max = 0;
    for (const ConvGroup& g : plan_.groups)
        if (g.is_cont && g.stride == 2 && d.resized[g.out_level]) up_max = std::max(up_max, (size_t)g.cout * d.up_h[g.out_level] * d.up_w[g.out_level]);
    if (up_max) i_up_.reserve(up_max * sizeof(float));
    iw_.resize.build(d, plan_.levels, stream_);
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
            if (d.resized[ol]) {   // concat_prev_ resizes the up-sampled tensor to the skip's size (layers.h:2438,2494-2521)
                Slice tmp;
                tmp.p = i_up_.as<float>(); tmp.sstride = (long long)g.cout * d.up_h[ol] * d.up_w[ol]; tmp.c0 = 0; tmp.C = g.cout;
                up_forward(a.in, tmp, 1, d.h[g.in_level], d.w[g.in_level], P + g.w_off, P + g.b_off, stream_);
                resize_bilinear_nchw(tmp, d.up_h[ol], d.up_w[ol], a.out, 1, a.H, a.W, iw_.resize.tab[ol], stream_);
            } else
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
        tc_run_inference(1, h, w);
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

// annonet_infer() (annonet_infer.cpp:32-185).  One process per GPU: after ab_comm_init every rank calls this with the same
// image; rank r runs the tiles of band r into a PRIVATE canvas that spans only the rows its tiles touch, the ranks exchange
// the seam strips with their two neighbours (each <= 63 overlap rows x K planes; ncclSend/Recv), every rank runs the argmax of the
// rows it owns and the label strips (2 bytes per pixel) are broadcast so that every rank holds the whole label map.
// No all-reduce, no full-size canvas per rank.  world = 1 is the same code with one band.
void Engine::infer_tiled_device(const uint8_t* img_dev, int h, int w, int max_tile_w, int max_tile_h, const double* gains, int ngains,
                                uint16_t* labels_dev)
{
    if (!loaded_) throw state_error("no net loaded");
    AB_CUDA(cudaSetDevice(device_));
    const int K = plan_.class_count;
    const int ov = (1 << (plan_.levels + 1)) - 1;  // TrainingNet::GetRequiredInputDimension, annonet_infer_main.cpp:421-427
    const std::vector<Tile> tiles = get_tiles(w, h, max_tile_w, max_tile_h, ov, ov);
    const std::vector<Band> bands = plan_bands(tiles, world_, plan_.levels);
    const Band& me = bands[rank_];
    const long long W = w;
    const int rows = me.empty() ? 0 : me.f1 - me.f0 + 1;
    canvas_rows_ = rows; canvas_row0_ = me.empty() ? 0 : me.f0;
    const size_t plane = (size_t)rows * W;
    if (rows) {
        canvas_.reserve((size_t)K * plane * sizeof(float));
        AB_CUDA(cudaMemsetAsync(canvas_.p, 0, (size_t)K * plane * sizeof(float), stream_));
    }
    struct Job { int ti, left, top, rw, rh; };
    std::vector<Job> jobs;
    for (int ti = me.t0; ti < me.t1; ++ti) {
        const Tile& t = tiles[ti];
        const int fw = t.full[2] - t.full[0] + 1, fh = t.full[3] - t.full[1] + 1;
        const int cx = t.full[0] + fw / 2, cy = t.full[1] + fh / 2;
        const int rw = recommended_input_dimension(fw, plan_.levels), rh = recommended_input_dimension(fh, plan_.levels);
        jobs.push_back(Job{ti, cx - rw / 2, cy - rh / 2, rw, rh});
    }
    auto blend = [&](const Job& j, const float* logits) {
        const Tile& t = tiles[j.ti];
        k::BlendArgs b;   // canvas coordinates: image rows shifted by the band's first row
        b.logits = logits; b.K = K; b.th = j.rh; b.tw = j.rw;
        b.canvas = canvas_.as<float>(); b.H = rows; b.W = w; b.left = j.left; b.top = j.top - me.f0;
        for (int i = 0; i < 4; ++i) { b.full[i] = t.full[i]; b.unique[i] = t.unique[i]; }
        b.full[1] -= me.f0; b.full[3] -= me.f0; b.unique[1] -= me.f0; b.unique[3] -= me.f0;
        k::blend_tile(b, stream_);
    };
    if (precision_ == 1) {
        // Tensor-core path: tiles of equal size run in batches -- one tile alone leaves the coarse levels of the net with a
        // handful of CTAs per launch.  Batches are formed in tile order among equal sizes (interior tiles all share one size), so
        // the workspace geometry changes a few times per image; the blend adds commute up to float rounding at 4-tile corners.
        static const int max_batch = getenv("AB_INFER_BATCH") ? std::max(1, atoi(getenv("AB_INFER_BATCH"))) : 8;
        std::vector<char> done(jobs.size(), 0);
        for (size_t i = 0; i < jobs.size(); ++i) {
            if (done[i]) continue;
            std::vector<size_t> same;
            for (size_t j = i; j < jobs.size(); ++j)
                if (!done[j] && jobs[j].rw == jobs[i].rw && jobs[j].rh == jobs[i].rh) same.push_back(j);
            // memory bound of a batch: ~0.5 KB per pixel of workspace
            const int cap = (int)std::max<long long>(1, std::min<long long>(max_batch, (4ll << 30) / (512ll * jobs[i].rw * jobs[i].rh)));
            for (size_t s0 = 0; s0 < same.size(); s0 += cap) {
                const int nb = (int)std::min<size_t>(cap, same.size() - s0);
                for (int q = 0; q < nb; ++q) {
                    const Job& j = jobs[same[s0 + q]];
                    tc_input(img_dev, h, w, j.left, j.top, j.rh, j.rw, q, nb);
                }
                tc_run_inference(nb, jobs[i].rh, jobs[i].rw);
                for (int q = 0; q < nb; ++q) {
                    blend(jobs[same[s0 + q]], i_logits_.as<float>() + (size_t)q * K * jobs[i].rh * jobs[i].rw);
                    done[same[s0 + q]] = 1;
                }
            }
        }
    } else {
        for (const Job& j : jobs) {
            ensure_infer_ws(j.rh, j.rw);
            k::extract_tile(img_dev, h, w, j.left, j.top, i_x_.as<float>(), j.rh, j.rw, model_.avg[0], model_.avg[1], model_.avg[2], stream_);
            run_inference_graph(j.rh, j.rw);
            blend(j, i_logits_.as<float>());
        }
    }
    // ---- seams: what my tiles added to rows a neighbour owns goes there, what the neighbours' tiles added to my rows comes here
    if (world_ > 1) {
        int prev = -1, next = -1;
        for (int r = rank_ - 1; r >= 0 && prev < 0; --r) if (!bands[r].empty()) prev = r;
        for (int r = rank_ + 1; r < world_ && next < 0; ++r) if (!bands[r].empty()) next = r;
        struct Strip { int peer, y0, y1; };   // image rows [y0, y1]
        std::vector<Strip> out, in;
        if (!me.empty()) {
            if (prev >= 0 && me.f0 < me.own0) out.push_back({prev, me.f0, me.own0 - 1});
            if (next >= 0 && me.f1 > me.own1) out.push_back({next, me.own1 + 1, me.f1});
            if (prev >= 0 && bands[prev].f1 >= me.own0) in.push_back({prev, me.own0, bands[prev].f1});
            if (next >= 0 && bands[next].f0 <= me.own1) in.push_back({next, bands[next].f0, me.own1});
        }
        size_t in_floats = 0;
        for (const Strip& s : in) in_floats += (size_t)K * (s.y1 - s.y0 + 1) * W;
        seam_.reserve(std::max<size_t>(in_floats, 1) * sizeof(float));
        AB_CUDA(cudaStreamSynchronize(stream_));   // (first use: NCCL sets up its connections on its own streams)
        nccl_->check(nccl_->GroupStart(), "ncclGroupStart");
        for (const Strip& s : out)
            for (int k = 0; k < K; ++k)
                nccl_->check(nccl_->Send(canvas_.as<float>() + k * plane + (size_t)(s.y0 - me.f0) * W, (size_t)(s.y1 - s.y0 + 1) * W, /*ncclFloat32*/ 7,
                                         s.peer, comm_, stream_), "ncclSend");
        size_t off = 0;
        for (const Strip& s : in) {   // one receive per plane: sends and receives pair up one to one, in order
            const size_t n = (size_t)(s.y1 - s.y0 + 1) * W;
            for (int k = 0; k < K; ++k)
                nccl_->check(nccl_->Recv(seam_.as<float>() + off + k * n, n, 7, s.peer, comm_, stream_), "ncclRecv");
            off += (size_t)K * n;
        }
        nccl_->check(nccl_->GroupEnd(), "ncclGroupEnd");
        off = 0;
        for (const Strip& s : in) {
            const size_t n = (size_t)(s.y1 - s.y0 + 1) * W;
            for (int k = 0; k < K; ++k)
                k::add_inplace(canvas_.as<float>() + k * plane + (size_t)(s.y0 - me.f0) * W, seam_.as<float>() + off + k * n, (long long)n, stream_);
            off += (size_t)K * n;
        }
    }
    // ---- argmax of the rows I own, straight into the caller's label map; then every rank broadcasts its strip
    const double* gd = upload_gains(gains_dev_, gains, ngains, K, stream_);
    // detection levels (annonet_infer.cpp:187-239): seeds are found with the argmax, the blob filter runs on the whole map
    bool detect = false;
    for (double v : detection_levels_) detect = detect || v > 0.0;
    const double* dd = nullptr;
    if (detect) {
        if ((int)detection_levels_.size() < K) throw std::invalid_argument("fewer detection levels than classes");
        dd = upload_gains(det_dev_, detection_levels_.data(), (int)detection_levels_.size(), K, stream_);
        seeds_.reserve((size_t)h * W);
    }
    if (!me.empty()) {
        const long long n = (long long)(me.own1 - me.own0 + 1) * W;
        k::argmax_labels(canvas_.as<float>() + (size_t)(me.own0 - me.f0) * W, (long long)plane, K, n, gd, labels_dev + (size_t)me.own0 * W, stream_,
                         dd, detect ? seeds_.as<uint8_t>() + (size_t)me.own0 * W : nullptr);
    }
    if (world_ > 1) {
        nccl_->check(nccl_->GroupStart(), "ncclGroupStart");
        for (int r = 0; r < world_; ++r) {
            if (bands[r].empty()) continue;
            uint16_t* p = labels_dev + (size_t)bands[r].own0 * W;
            nccl_->check(nccl_->Broadcast(p, p, (size_t)(bands[r].own1 - bands[r].own0 + 1) * W * 2, /*ncclUint8*/ 1, r, comm_, stream_), "ncclBroadcast");
            if (detect) {
                uint8_t* s = seeds_.as<uint8_t>() + (size_t)bands[r].own0 * W;
                nccl_->check(nccl_->Broadcast(s, s, (size_t)(bands[r].own1 - bands[r].own0 + 1) * W, 1, r, comm_, stream_), "ncclBroadcast");
            }
        }
        nccl_->check(nccl_->GroupEnd(), "ncclGroupEnd");
    }
    if (detect) {   // blobs cross band boundaries: every rank filters the whole (2 bytes per pixel) map it now holds
        blob_parent_.reserve((size_t)h * W * sizeof(int));
        blob_keep_.reserve((size_t)h * W);
        // default: the reference's own seed lookup, quirk included (kernels_blobs.cu); AB_DETECTION_SEEDS=intended: same-pixel lookup
        static const bool intended = getenv("AB_DETECTION_SEEDS") && std::string(getenv("AB_DETECTION_SEEDS")) == "intended";
        k::filter_detections(labels_dev, seeds_.as<uint8_t>(), h, w, blob_parent_.as<int>(), blob_keep_.as<uint8_t>(), stream_, intended ? 0 : 1);
    }
    AB_CUDA(cudaGetLastError());
}

// confusion_matrix_per_pixel (annonet_infer_main.cpp:482-489) on the device; update_confusion_matrix_per_region (:202-272) with
// the connected blobs of both label images found on the device (union-find, kernels_blobs.cu) and the per-blob class votes
// counted on the host from the blob ids -- one linear pass, the expensive part of the reference (two flood fills) is gone.
void Engine::evaluate_labels(const uint16_t* truth, const uint16_t* predicted, int h, int w, int K, unsigned long long* per_pixel,
                             unsigned long long* per_region)
{
    AB_CUDA(cudaSetDevice(device_));
    if (K < 1 || K > 64) throw std::invalid_argument("evaluate_labels: 1..64 classes");
    const size_t n = (size_t)h * w;
    e_truth_.reserve(n * 2); e_pred_.reserve(n * 2); e_roots_.reserve(n * 4); blob_keep_.reserve(n); e_counts_.reserve(sizeof(unsigned long long) * K * K);
    AB_CUDA(cudaMemcpyAsync(e_truth_.p, truth, n * 2, cudaMemcpyHostToDevice, stream_));
    AB_CUDA(cudaMemcpyAsync(e_pred_.p, predicted, n * 2, cudaMemcpyHostToDevice, stream_));
    k::confusion_per_pixel(e_truth_.as<uint16_t>(), e_pred_.as<uint16_t>(), (long long)n, K, e_counts_.as<unsigned long long>(), stream_);
    AB_CUDA(cudaMemcpyAsync(per_pixel, e_counts_.p, sizeof(unsigned long long) * K * K, cudaMemcpyDeviceToHost, stream_));
    if (!per_region) { synchronize(); return; }
    std::fill(per_region, per_region + (size_t)K * K, 0ull);
    bool any_labelled = false;
    for (size_t i = 0; i < n && !any_labelled; ++i) any_labelled = truth[i] != 65535;
    if (!any_labelled) { synchronize(); return; }   // labeled_points_by_class.empty() (:210-212)
    std::vector<int> roots(n);
    const uint16_t ignore = 65535;
    for (int pass = 0; pass < 2; ++pass) {   // blobs of the ground truth, then of the result (:269-270)
        k::label_blobs((pass == 0 ? e_truth_ : e_pred_).as<uint16_t>(), h, w, e_roots_.as<int>(), blob_keep_.as<uint8_t>(), stream_);
        AB_CUDA(cudaMemcpyAsync(roots.data(), e_roots_.p, n * 4, cudaMemcpyDeviceToHost, stream_));
        synchronize();
        // blob numbers: background (-1) is blob 0 like dlib's, every other blob gets the next free number
        std::vector<int> number(n, -1);
        int blobs = 1;
        std::vector<unsigned> votes_gt(K), votes_pred(K);   // grown below: [blob][class]
        for (size_t i = 0; i < n; ++i)
            if (roots[i] >= 0 && number[roots[i]] < 0) number[roots[i]] = blobs++;
        votes_gt.assign((size_t)blobs * K, 0); votes_pred.assign((size_t)blobs * K, 0);
        std::vector<unsigned> pred_other((size_t)blobs, 0);   // votes for predictions outside [0, K)
        for (size_t i = 0; i < n; ++i) {
            if (truth[i] == ignore) continue;          // not a labelled point
            const int b = roots[i] < 0 ? 0 : number[roots[i]];
            if (truth[i] < K) ++votes_gt[(size_t)b * K + truth[i]];
            if (predicted[i] < K) ++votes_pred[(size_t)b * K + predicted[i]]; else ++pred_other[b];
        }
        auto winner = [&](const unsigned* v, int from) -> int {   // most votes, lowest class among ties; -1: no votes at all
            int best = -1; unsigned most = 0;
            for (int c = from; c < K; ++c) if (v[c] > most) { most = v[c]; best = c; }
            return best;
        };
        for (int b = 0; b < blobs; ++b) {
            const unsigned* gt = &votes_gt[(size_t)b * K];
            unsigned* pr = &votes_pred[(size_t)b * K];
            const int wgt = winner(gt, 0);
            if (wgt < 0) continue;                                       // winner == label_to_ignore: no labelled point in the blob
            bool only_background = pr[0] > 0 && pred_other[b] == 0;
            for (int c = 1; c < K && only_background; ++c) only_background = pr[c] == 0;
            if (wgt != 0 && !only_background) pr[0] = 0;                 // detections count, however small (:249-256)
            const int wpr = winner(pr, 0);
            if (wpr >= 0) ++per_region[(size_t)wgt * K + wpr];
        }
    }
}

void Engine::render_labels(const uint16_t* labels, int h, int w, int out_h, int out_w, const uint8_t* palette_rgba, int K, uint8_t* rgba_out)
{
    AB_CUDA(cudaSetDevice(device_));
    if (h < 1 || w < 1 || out_h < 1 || out_w < 1 || K < 1) throw std::invalid_argument("render_labels: bad size");
    const size_t n = (size_t)h * w, on = (size_t)out_h * out_w;
    e_pred_.reserve(n * 2); e_rgba_.reserve(on * 4); e_counts_.reserve(std::max<size_t>((size_t)K * 4, 64 * 64 * 8));
    AB_CUDA(cudaMemcpyAsync(e_pred_.p, labels, n * 2, cudaMemcpyHostToDevice, stream_));
    AB_CUDA(cudaMemcpyAsync(e_counts_.p, palette_rgba, (size_t)K * 4, cudaMemcpyHostToDevice, stream_));
    k::render_labels(e_pred_.as<uint16_t>(), h, w, out_h, out_w, e_counts_.as<uint8_t>(), K, e_rgba_.as<uint8_t>(), stream_);
    AB_CUDA(cudaMemcpyAsync(rgba_out, e_rgba_.p, on * 4, cudaMemcpyDeviceToHost, stream_));
    synchronize();
}

void Engine::infer_tiled_host(const uint8_t* img_host, int h, int w, int max_tile_w, int max_tile_h, const double* gains, int ngains,
                              uint16_t* labels_host, float* blended_host)
{
    AB_CUDA(cudaSetDevice(device_));
    const size_t bytes = (size_t)h * w * 3;
    i_img_.reserve(bytes);
    // large images: copy straight from the caller's memory (the driver stages pageable memory itself)
    AB_CUDA(cudaMemcpyAsync(i_img_.p, img_host, bytes, cudaMemcpyHostToDevice, stream_));
    i_labels_.reserve((size_t)h * w * 2);
    infer_tiled_device(i_img_.as<uint8_t>(), h, w, max_tile_w, max_tile_h, gains, ngains, i_labels_.as<uint16_t>());
    AB_CUDA(cudaMemcpyAsync(labels_host, i_labels_.p, (size_t)h * w * 2, cudaMemcpyDeviceToHost, stream_));
    if (blended_host) {   // the blended logits of the rows this rank's canvas holds (one process: the whole image), other rows zero
        const size_t hw = (size_t)h * w;
        std::memset(blended_host, 0, (size_t)plan_.class_count * hw * sizeof(float));
        for (int k = 0; k < plan_.class_count && canvas_rows_; ++k)
            AB_CUDA(cudaMemcpyAsync(blended_host + k * hw + (size_t)canvas_row0_ * w, canvas_.as<float>() + (size_t)k * canvas_rows_ * w,
                                    (size_t)canvas_rows_ * w * sizeof(float), cudaMemcpyDeviceToHost, stream_));
    }
    synchronize();
}

}  // namespace ab
