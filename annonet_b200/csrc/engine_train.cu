// Training side of the engine: one dnn_trainer step on one device (trainer.h:631-671,743-828):
// forward in batch-norm training mode, weighted per-pixel loss, backward, [gradient all-reduce,] update.
#include "engine.h"

#include <cstring>

namespace ab {

using namespace k;

void Engine::ensure_train_ws(int n, int h, int w)
{
    if (tn_ == n && th_ == h && tw_ == w && tw_act_.n == n) return;
    synchronize();
    const LevelDims d = level_dims(h, w, plan_.levels);
    const int K = plan_.class_count;
    for (int l = 0; l <= plan_.levels; ++l) {
        const size_t bytes = (size_t)n * plan_.level_channels[l] * d.h[l] * d.w[l] * sizeof(float);
        tw_act_.buf[l].reserve(bytes);
        tw_grad_.buf[l].reserve(bytes);
    }
    tw_act_.dims = tw_grad_.dims = d;
    size_t up_max = 0;
    for (const ConvGroup& g : plan_.groups)
        if (g.is_cont && g.stride == 2 && d.resized[g.out_level]) up_max = std::max(up_max, (size_t)n * g.cout * d.up_h[g.out_level] * d.up_w[g.out_level]);
    if (up_max) t_up_.reserve(up_max * sizeof(float));
    tw_act_.resize.build(d, plan_.levels, stream_);
    tw_act_.n = tw_grad_.n = n; tw_act_.h = tw_grad_.h = h; tw_act_.w = tw_grad_.w = w;
    t_x_.reserve((size_t)n * 3 * h * w * 4);
    t_logits_.reserve((size_t)n * K * h * w * 4);
    t_glogits_.reserve((size_t)n * K * h * w * 4);
    // per group: pre-activation conv output z (groups with bn), pooled input (transition_down), bn batch statistics
    z_off_.assign(plan_.groups.size(), 0);
    pool_off_.assign(plan_.groups.size(), 0);
    stats_off_.assign(plan_.groups.size(), 0);
    size_t zt = 0, pt = 0, stt = 0, dz_max = 0, dp_max = 0, dead_max = 0;
    int cmax = 1;
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        const int ol = g.out_level == -2 ? 0 : g.out_level;
        const size_t out_elems = (size_t)n * g.cout * d.h[ol] * d.w[ol];
        cmax = std::max(cmax, g.cout);
        if (g.dlib_bn >= 0) {
            stats_off_[gi] = stt; stt += 2 * (size_t)g.cout;
            if (g.dead) dead_max = std::max(dead_max, out_elems);
            else { z_off_[gi] = zt; zt += out_elems; dz_max = std::max(dz_max, out_elems); }
        }
        if (g.pool_input) {
            const size_t pe = (size_t)n * g.cin * d.h[ol] * d.w[ol];
            pool_off_[gi] = pt; pt += pe;
            dp_max = std::max(dp_max, pe);
        }
    }
    t_z_.reserve(zt * 4); t_pool_.reserve(pt * 4); t_dz_.reserve(dz_max * 4); t_dpool_.reserve(dp_max * 4);
    t_dead_.reserve(std::max<size_t>(dead_max, 1) * 4);
    t_stats_.reserve(stt * 4);
    t_sums_.reserve(sizeof(double) * 2 * cmax);
    t_scratch_.reserve(wgrad_scratch_floats(0, 0) * 4);
    t_loss_.reserve(sizeof(double));
    tn_ = n; th_ = h; tw_ = w;
    stats_valid_ = false;
}

static Slice dense_slice(float* p, int C, long long hw)
{
    Slice s; s.p = p; s.sstride = (long long)C * hw; s.c0 = 0; s.C = C; return s;
}

void Engine::forward_train(int n, int h, int w)
{
    const LevelDims& d = tw_act_.dims;
    const float* P = params_.as<float>();
    if (n < 2) throw std::invalid_argument("training needs a mini-batch of at least 2 (bn_ runs in batch mode only for num_samples > 1, layers.h:1392)");
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        if (g.dead && hyper.skip_dead_layers) continue;
        const int ol = g.out_level == -2 ? 0 : g.out_level;
        const int H = d.h[ol], W = d.w[ol];
        const long long hw = (long long)H * W;
        Slice in, out, y;
        if (g.in_level < 0) in = dense_slice(t_x_.as<float>(), 3, (long long)h * w);
        else in = level_slice(tw_act_, g.in_level, g.in_c0, g.cin);
        if (g.out_level == -2) y = dense_slice(t_logits_.as<float>(), g.cout, hw);
        else if (!g.dead) y = level_slice(tw_act_, g.out_level, g.out_c0, g.cout);
        // conv output: z (when a bn follows) or the final destination
        if (g.dlib_bn >= 0) out = dense_slice(g.dead ? t_dead_.as<float>() : t_z_.as<float>() + z_off_[gi], g.cout, hw);
        else out = y;

        if (g.is_cont && g.stride == 2) {
            if (d.resized[ol]) {   // concat_prev_ resizes the up-sampled tensor to the skip's size (layers.h:2438,2494-2521)
                Slice tmp = dense_slice(t_up_.as<float>(), g.cout, (long long)d.up_h[ol] * d.up_w[ol]);
                up_forward(in, tmp, n, d.h[g.in_level], d.w[g.in_level], P + g.w_off, P + g.b_off, stream_);
                resize_bilinear_nchw(tmp, d.up_h[ol], d.up_w[ol], out, n, H, W, tw_act_.resize.tab[ol], stream_);
            } else
                up_forward(in, out, n, d.h[g.in_level], d.w[g.in_level], P + g.w_off, P + g.b_off, stream_);
        } else {
            ConvArgs a;
            a.N = n; a.H = H; a.W = W; a.out = out; a.bias = P + g.b_off; a.wv.w = P + g.w_off;
            if (g.ksize == 3) {
                a.in = in;
                if (!g.is_cont) { a.wv.s_out = g.cin * 9; a.wv.s_in = 9; a.wv.flip = 0; }
                else { a.wv.s_out = 9; a.wv.s_in = g.cout * 9; a.wv.flip = 1; }
                conv3x3(a, stream_);
            } else {
                a.wv.s_out = g.cin; a.wv.s_in = 1;
                if (g.pool_input) {
                    Slice pooled = dense_slice(t_pool_.as<float>() + pool_off_[gi], g.cin, hw);
                    maxpool_forward(in, pooled, n, d.h[g.in_level], d.w[g.in_level], stream_);
                    a.in = pooled;
                } else a.in = in;
                conv1x1(a, stream_);
            }
        }
        if (g.dlib_bn >= 0) {
            const int mish_first = g.bn_first ? 0 : 1;
            float* stats = t_stats_.as<float>() + stats_off_[gi];
            float* rmean = running_.as<float>() + running_off_[gi];
            bn_stats(out, n, H, W, g.has_mish && mish_first, t_sums_.as<double>(), stream_);
            bn_finalize(t_sums_.as<double>(), g.cout, (long long)n * hw, (float)model_.layers[g.dlib_bn].eps, nu_dev_.as<unsigned long long>() + gi, window_[gi], rmean, rmean + g.cout,
                        stats, stream_);
            if (!g.dead)
                bn_apply(out, y, n, H, W, mish_first, g.has_mish, P + g.bn_off, P + g.bn_off + g.cout, stats, stream_);
        }
    }
    stats_valid_ = true;
}

void Engine::backward_train(int n, int h, int w)
{
    const LevelDims& d = tw_act_.dims;
    const float* P = params_.as<float>();
    float* G = grads_.as<float>();
    float* scratch = t_scratch_.as<float>();
    for (int l = 0; l <= plan_.levels; ++l)
        fill_zero(tw_grad_.buf[l].p, (size_t)n * plan_.level_channels[l] * d.h[l] * d.w[l] * sizeof(float), stream_);
    // gradient buckets for the all-reduce: the arena is in forward order, backward completes it from the tail
    const size_t np = plan_.num_params;
    const int nbuckets = 4;
    size_t bucket_hi = np;
    if (world_ > 1) { loss_to_slot(t_loss_.as<double>(), grads_.as<float>() + np, stream_); bucket_hi = np + 1; }

    for (int gi = (int)plan_.groups.size() - 1; gi >= 0; --gi) {
        const ConvGroup& g = plan_.groups[gi];
        if (g.dead || grad_blocked(d, g)) {
            // never consumed -- or reached only through a resized join, which the reference's backward pass does not cross
            // (plan.h): dlib back-propagates an all-zero gradient, every parameter gradient is exactly 0
            fill_zero(G + g.w_off, (g.nweights() + g.cout + (g.dlib_bn >= 0 ? 2 * g.cout : 0)) * sizeof(float), stream_);
        } else {
            const int ol = g.out_level == -2 ? 0 : g.out_level;
            const int H = d.h[ol], W = d.w[ol];
            const long long hw = (long long)H * W;
            Slice in, gin, gy, dz;
            if (g.in_level < 0) in = dense_slice(t_x_.as<float>(), 3, (long long)h * w);
            else { in = level_slice(tw_act_, g.in_level, g.in_c0, g.cin); gin = level_slice(tw_grad_, g.in_level, g.in_c0, g.cin); }
            if (g.out_level == -2) gy = dense_slice(t_glogits_.as<float>(), g.cout, hw);
            else gy = level_slice(tw_grad_, g.out_level, g.out_c0, g.cout);
            if (g.dlib_bn >= 0) {
                const int mish_first = g.bn_first ? 0 : 1;
                const float* stats = t_stats_.as<float>() + stats_off_[gi];
                Slice z = dense_slice(t_z_.as<float>() + z_off_[gi], g.cout, hw);
                bn_backward_reduce(z, gy, n, H, W, mish_first, g.has_mish, P + g.bn_off, P + g.bn_off + g.cout, stats, t_sums_.as<double>(), stream_);
                bn_backward_apply(z, gy, t_dz_.as<float>(), n, H, W, mish_first, g.has_mish, P + g.bn_off, P + g.bn_off + g.cout, stats,
                                  t_sums_.as<double>(), G + g.bn_off, G + g.bn_off + g.cout, stream_);
                dz = dense_slice(t_dz_.as<float>(), g.cout, hw);
            } else dz = gy;

            channel_sum(dz, n, H, W, G + g.b_off, scratch, stream_);  // layers.h:226, cpu_dlib.cpp:470-504

            if (g.is_cont && g.stride == 2) {
                const int hi = d.h[g.in_level], wi = d.w[g.in_level];
                wgrad3x3(in, dz, n, hi, wi, H, W, 2, 0, G + g.w_off, scratch, stream_);
                up_backward_data(dz, gin, n, hi, wi, P + g.w_off, stream_);
            } else if (g.ksize == 3) {
                if (!g.is_cont) wgrad3x3(dz, in, n, H, W, H, W, 1, 1, G + g.w_off, scratch, stream_);
                else wgrad3x3(in, dz, n, H, W, H, W, 1, 1, G + g.w_off, scratch, stream_);
                if (g.in_level >= 0) {
                    ConvArgs a;
                    a.N = n; a.H = H; a.W = W; a.in = dz; a.out = gin; a.accumulate = 1; a.wv.w = P + g.w_off;
                    if (!g.is_cont) { a.wv.s_out = 9; a.wv.s_in = g.cin * 9; a.wv.flip = 1; }
                    else { a.wv.s_out = g.cout * 9; a.wv.s_in = 9; a.wv.flip = 0; }
                    conv3x3(a, stream_);
                }
            } else {
                Slice b = g.pool_input ? dense_slice(t_pool_.as<float>() + pool_off_[gi], g.cin, hw) : in;
                wgrad1x1(dz, b, n, H, W, G + g.w_off, scratch, stream_);
                ConvArgs a;
                a.N = n; a.H = H; a.W = W; a.in = dz; a.wv.w = P + g.w_off; a.wv.s_out = 1; a.wv.s_in = g.cin;
                if (g.pool_input) {
                    Slice dp = dense_slice(t_dpool_.as<float>(), g.cin, hw);
                    a.out = dp;
                    conv1x1(a, stream_);
                    maxpool_backward(in, dp, gin, n, d.h[g.in_level], d.w[g.in_level], stream_);
                } else {
                    a.out = gin; a.accumulate = 1;
                    conv1x1(a, stream_);
                }
            }
        }
        // everything at and above this group's offset is final: hand finished buckets to the all-reduce
        if (world_ > 1 && bucket_hi - g.w_off >= np / nbuckets) {
            allreduce_range(g.w_off, bucket_hi);
            bucket_hi = g.w_off;
        }
    }
    if (world_ > 1) {
        allreduce_range(0, bucket_hi);
        AB_CUDA(cudaEventRecord(ev_comm_done_, comm_stream_));
        AB_CUDA(cudaStreamWaitEvent(stream_, ev_comm_done_, 0));
        slot_to_loss(grads_.as<float>() + np, t_loss_.as<double>(), world_, stream_);   // mean of the ranks' losses
    }
}

// adam's second-moment arena and its device-side step counter; allocated outside any stream capture (tc_train_step calls this
// before it starts capturing)
void Engine::ensure_adam_state()
{
    if (adam_v_.p) return;
    const size_t np = plan_.num_params;
    adam_v_.reserve(np * 4);
    adam_t_dev_.reserve(256);
    AB_CUDA(cudaMemsetAsync(adam_v_.p, 0, np * 4, stream_));
    const float t0 = (float)adam_t_;
    AB_CUDA(cudaMemcpyAsync(adam_t_dev_.p, &t0, 4, cudaMemcpyHostToDevice, stream_));
    AB_CUDA(cudaStreamSynchronize(stream_));
}

void Engine::optimizer_step()
{
    const long long np = (long long)plan_.num_params;
    const float gscale = world_ > 1 ? 1.0f / (float)world_ : 1.0f;
    if (hyper.solver == 0) {
        sgd_step(params_.as<float>(), grads_.as<float>(), mom_.as<float>(), lr_mul_.as<float>(), wd_mul_.as<float>(), np,
                 (float)hyper.learning_rate, (float)hyper.weight_decay, (float)hyper.momentum, gscale, stream_);
    } else {
        ensure_adam_state();
        ++adam_t_;
        adam_step(params_.as<float>(), grads_.as<float>(), mom_.as<float>(), adam_v_.as<float>(), lr_mul_.as<float>(), wd_mul_.as<float>(),
                  np, adam_t_dev_.as<float>(), (float)hyper.learning_rate, (float)hyper.weight_decay, (float)hyper.adam_beta1,
                  (float)hyper.adam_beta2, gscale, stream_);
    }
}

void Engine::train_step_device(const uint8_t* imgs_dev, const void* labels_dev, int n, int h, int w, bool update)
{
    if (!loaded_) throw state_error("no net loaded");
    if (!model_.is_training) throw state_error("StartTraining needs a training net (net_type with bn_ layers)");
    AB_CUDA(cudaSetDevice(device_));
    level_dims(h, w, plan_.levels);   // rejects inputs below the required input dimension
    if (precision_ == 1) { tc_train_step(imgs_dev, labels_dev, n, h, w, update); return; }   // AB_PRECISION_F16: tensor-core path
    ensure_train_ws(n, h, w);
    to_tensor(imgs_dev, t_x_.as<float>(), n, h, w, model_.avg[0], model_.avg[1], model_.avg[2], stream_);
    forward_train(n, h, w);
    loss_forward_backward(t_logits_.as<float>(), labels_dev, t_glogits_.as<float>(), n, plan_.class_count, h, w, t_loss_.as<double>(), stream_);
    backward_train(n, h, w);
    if (update) optimizer_step();
    AB_CUDA(cudaMemcpyAsync(loss_host_, t_loss_.p, sizeof(double), cudaMemcpyDeviceToHost, stream_));
    AB_CUDA(cudaGetLastError());
}

void Engine::train_step_host(const uint8_t* imgs, const void* labels, int n, int h, int w, bool update)
{
    if (!loaded_) throw state_error("no net loaded");
    AB_CUDA(cudaSetDevice(device_));
    level_dims(h, w, plan_.levels);
    const size_t ib = (size_t)n * h * w * 3, lb = (size_t)n * h * w * 8;
    // depth-1 asynchrony (trainer.h:87,1252).  The host->device copy of step i+1 runs on its own stream WHILE step i computes:
    // pinned staging -> device staging (copy stream), then a device-to-device copy into the step's fixed input buffers (the
    // CUDA graph of the step is keyed on those pointers) at the head of the compute stream.
    AB_CUDA(cudaEventSynchronize(ev_stage_));            // the previous step's copy has left the pinned buffer
    char* stage = (char*)pinned(ib + lb);
    std::memcpy(stage, imgs, ib);
    std::memcpy(stage + ib, labels, lb);
    t_imgs_.reserve(ib);
    t_labels_.reserve(lb);
    t_stage_dev_.reserve(ib + lb);
    AB_CUDA(cudaStreamWaitEvent(copy_stream_, ev_unstaged_, 0));   // the previous step has emptied the device staging buffer
    AB_CUDA(cudaMemcpyAsync(t_stage_dev_.p, stage, ib + lb, cudaMemcpyHostToDevice, copy_stream_));
    AB_CUDA(cudaEventRecord(ev_stage_, copy_stream_));
    AB_CUDA(cudaStreamWaitEvent(stream_, ev_stage_, 0));
    AB_CUDA(cudaMemcpyAsync(t_imgs_.p, t_stage_dev_.p, ib, cudaMemcpyDeviceToDevice, stream_));
    AB_CUDA(cudaMemcpyAsync(t_labels_.p, (const char*)t_stage_dev_.p + ib, lb, cudaMemcpyDeviceToDevice, stream_));
    AB_CUDA(cudaEventRecord(ev_unstaged_, stream_));
    train_step_device(t_imgs_.as<uint8_t>(), t_labels_.p, n, h, w, update);
}

void Engine::dataset_clear()
{
    synchronize();
    dataset_.clear();
}

int Engine::dataset_add(const uint8_t* rgb, const uint16_t* labels, int h, int w)
{
    if (!rgb || !labels || h < 1 || w < 1) throw std::invalid_argument("dataset_add: bad image");
    AB_CUDA(cudaSetDevice(device_));
    DatasetImage d;
    d.rgb = std::make_unique<DevBuf>(); d.labels = std::make_unique<DevBuf>();
    d.h = h; d.w = w;
    d.rgb->reserve((size_t)h * w * 3);
    d.labels->reserve((size_t)h * w * 2);
    AB_CUDA(cudaMemcpyAsync(d.rgb->p, rgb, (size_t)h * w * 3, cudaMemcpyHostToDevice, stream_));
    AB_CUDA(cudaMemcpyAsync(d.labels->p, labels, (size_t)h * w * 2, cudaMemcpyHostToDevice, stream_));
    AB_CUDA(cudaStreamSynchronize(stream_));
    dataset_.push_back(std::move(d));
    return (int)dataset_.size() - 1;
}

void Engine::build_crops(const CropSpec* crops, int n, int dim, double class_weight, double image_weight, const uint8_t** imgs_dev, const void** labels_dev)
{
    if (n < 1 || dim < 1 || !crops) throw std::invalid_argument("build_crops: bad arguments");
    AB_CUDA(cudaSetDevice(device_));
    const size_t jb = sizeof(k::CropJob) * (size_t)n;
    if (crop_jobs_host_bytes_ < jb) {
        AB_CUDA(cudaStreamSynchronize(stream_));
        if (crop_jobs_host_) cudaFreeHost(crop_jobs_host_);
        AB_CUDA(cudaMallocHost(&crop_jobs_host_, jb));
        crop_jobs_host_bytes_ = jb;
    }
    AB_CUDA(cudaEventSynchronize(ev_stage_));   // the previous batch's job table has left the pinned buffer
    k::CropJob* J = (k::CropJob*)crop_jobs_host_;
    for (int i = 0; i < n; ++i) {
        const CropSpec& c = crops[i];
        if (c.image < 0 || c.image >= (int)dataset_.size()) throw std::invalid_argument("build_crops: no such dataset image");
        if (c.noise_level < 0 || c.noise_level > 255 || !(c.brightness > 0)) throw std::invalid_argument("build_crops: bad augmentation parameters");
        const DatasetImage& d = dataset_[c.image];
        k::CropJob& j = J[i];
        j.image = d.rgb->as<uint8_t>(); j.labels = d.labels->as<uint16_t>(); j.h = d.h; j.w = d.w;
        j.left = c.left; j.top = c.top; j.flip_lr = c.flip_lr; j.flip_ud = c.flip_ud;
        j.brightness = c.brightness; j.noise_level = c.noise_level; j.seed = c.seed;
        for (int q = 0; q < 3; ++q) j.color_offset[q] = c.color_offset[q];
    }
    crop_jobs_.reserve(jb);
    crop_scratch_.reserve(k::crop_scratch_bytes(n));
    crop_overflow_.reserve(256);
    t_imgs_.reserve((size_t)n * dim * dim * 3);
    t_labels_.reserve((size_t)n * dim * dim * 8);
    AB_CUDA(cudaMemcpyAsync(crop_jobs_.p, J, jb, cudaMemcpyHostToDevice, stream_));
    AB_CUDA(cudaEventRecord(ev_stage_, stream_));
    k::build_crops(crop_jobs_.as<k::CropJob>(), n, dim, class_weight, image_weight, crop_scratch_.p, t_imgs_.as<uint8_t>(), t_labels_.p,
                   crop_overflow_.as<unsigned int>(), stream_);
    AB_CUDA(cudaGetLastError());
    *imgs_dev = t_imgs_.as<uint8_t>();
    *labels_dev = t_labels_.p;
}

double Engine::train_sync()
{
    synchronize();
    return *loss_host_;
}

}  // namespace ab
