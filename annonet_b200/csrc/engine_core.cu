#include "engine.h"
#include "nccl_api.h"
#include "tc_state.h"

#include <cstring>
#include <dlfcn.h>

namespace ab {

void cuda_check(cudaError_t e, const char* what, const char* file, int line)
{
    if (e == cudaSuccess) return;
    std::string msg = std::string("CUDA error: ") + cudaGetErrorString(e) + " (" + cudaGetErrorName(e) + ") at " + file + ":" +
                      std::to_string(line) + " in " + what;
    cudaGetLastError();  // clear the sticky flag of non-fatal errors (e.g. out of memory) so that a caller can retry smaller
    throw cuda_error(msg);
}

void DevBuf::release()
{
    if (p) cudaFree(p);
    p = nullptr;
    bytes = 0;
}

void DevBuf::reserve(size_t n)
{
    if (n <= bytes) return;
    release();
    AB_CUDA(cudaMalloc(&p, n));
    bytes = n;
}

void ResizeTables::build(const LevelDims& d, int levels, cudaStream_t st)
{
    std::vector<int> ints;
    std::vector<float> floats;
    struct Off { size_t i, f; int oh, ow; } off[8] = {};
    for (int l = 0; l < levels; ++l) {
        if (!d.resized[l]) continue;
        const ResizeAxis r = resize_axis(d.up_h[l], d.h[l], false), c = resize_axis(d.up_w[l], d.w[l], true);
        off[l] = Off{ints.size(), floats.size(), d.h[l], d.w[l]};
        ints.insert(ints.end(), r.lo.begin(), r.lo.end()); ints.insert(ints.end(), r.hi.begin(), r.hi.end());
        ints.insert(ints.end(), c.lo.begin(), c.lo.end()); ints.insert(ints.end(), c.hi.begin(), c.hi.end());
        floats.insert(floats.end(), r.frac.begin(), r.frac.end()); floats.insert(floats.end(), r.ifrac.begin(), r.ifrac.end());
        floats.insert(floats.end(), c.frac.begin(), c.frac.end()); floats.insert(floats.end(), c.ifrac.begin(), c.ifrac.end());
    }
    if (ints.empty()) return;
    buf.reserve(ints.size() * 4 + floats.size() * 4);
    AB_CUDA(cudaMemcpyAsync(buf.p, ints.data(), ints.size() * 4, cudaMemcpyHostToDevice, st));
    AB_CUDA(cudaMemcpyAsync((char*)buf.p + ints.size() * 4, floats.data(), floats.size() * 4, cudaMemcpyHostToDevice, st));
    AB_CUDA(cudaStreamSynchronize(st));   // (the host vectors are locals)
    const int* ip = buf.as<int>();
    const float* fp = (const float*)((const char*)buf.p + ints.size() * 4);
    for (int l = 0; l < levels; ++l) {
        if (!d.resized[l]) continue;
        const Off& o = off[l];
        tab[l] = k::ResizeTab{ip + o.i, ip + o.i + o.oh, ip + o.i + 2 * o.oh, ip + o.i + 2 * o.oh + o.ow,
                              fp + o.f, fp + o.f + o.oh, fp + o.f + 2 * o.oh, fp + o.f + 2 * o.oh + o.ow};
    }
}

// ---- NCCL, resolved at run time from the process (torch bundles libnccl.so.2) -----------------

void nccl_unique_id(void* id128)
{
    NcclApi api;
    NcclApi::UniqueId id;
    api.check(api.GetUniqueId(&id), "ncclGetUniqueId");
    std::memcpy(id128, &id, 128);
}

Engine::Engine(int device, int precision) : device_(device), precision_(precision), requested_precision_(precision)
{
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        throw cuda_error("no CUDA device available: this engine has no CPU fallback");
    }
    if (device < 0 || device >= count) throw std::invalid_argument("device index out of range");
    AB_CUDA(cudaSetDevice(device_));
    AB_CUDA(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
    AB_CUDA(cudaStreamCreateWithFlags(&comm_stream_, cudaStreamNonBlocking));
    AB_CUDA(cudaStreamCreateWithFlags(&copy_stream_, cudaStreamNonBlocking));
    AB_CUDA(cudaStreamCreateWithFlags(&side_stream_, cudaStreamNonBlocking));
    for (auto& ev : ev_fork_) AB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    AB_CUDA(cudaEventCreateWithFlags(&ev_join_, cudaEventDisableTiming));
    for (auto& ev : ev_bucket_) AB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    AB_CUDA(cudaEventCreateWithFlags(&ev_comm_done_, cudaEventDisableTiming));
    AB_CUDA(cudaEventCreateWithFlags(&ev_stage_, cudaEventDisableTiming));
    AB_CUDA(cudaEventCreateWithFlags(&ev_unstaged_, cudaEventDisableTiming));
    AB_CUDA(cudaMallocHost((void**)&loss_host_, sizeof(double)));
    *loss_host_ = 0;
}

Engine::~Engine()
{
    cudaSetDevice(device_);
    if (stream_) cudaStreamSynchronize(stream_);
    if (comm_ && nccl_) nccl_->CommDestroy(comm_);
    for (auto& ev : ev_bucket_) if (ev) cudaEventDestroy(ev);
    if (ev_comm_done_) cudaEventDestroy(ev_comm_done_);
    if (ev_stage_) cudaEventDestroy(ev_stage_);
    if (ev_unstaged_) cudaEventDestroy(ev_unstaged_);
    if (pinned_) cudaFreeHost(pinned_);
    if (loss_host_) cudaFreeHost(loss_host_);
    if (crop_jobs_host_) cudaFreeHost(crop_jobs_host_);
    if (comm_stream_) cudaStreamDestroy(comm_stream_);
    if (copy_stream_) cudaStreamDestroy(copy_stream_);
    if (side_stream_) cudaStreamDestroy(side_stream_);
    for (auto& ev : ev_fork_) if (ev) cudaEventDestroy(ev);
    if (ev_join_) cudaEventDestroy(ev_join_);
    if (stream_) cudaStreamDestroy(stream_);
}

void Engine::synchronize()
{
    AB_CUDA(cudaSetDevice(device_));
    AB_CUDA(cudaStreamSynchronize(stream_));
}

void* Engine::pinned(size_t bytes)
{
    if (bytes > pinned_bytes_) {
        if (pinned_) { AB_CUDA(cudaStreamSynchronize(stream_)); cudaFreeHost(pinned_); pinned_ = nullptr; }
        AB_CUDA(cudaMallocHost(&pinned_, bytes));
        pinned_bytes_ = bytes;
    }
    return pinned_;
}

void Engine::load(const NetModel& m)
{
    if (!m.params_allocated()) throw state_error("the net's parameters are not allocated (a stream written before the first forward pass)");
    AB_CUDA(cudaSetDevice(device_));
    synchronize();
    model_ = m;
    plan_ = build_plan(model_);
    tc_.reset();  // filter packs and launch descriptors of the tensor-core path belong to the old parameters
    // The tcgen05 kernels move channels in 16-byte (8 x fp16) vectors and keep a filter-gradient tile of at most 64 output
    // channels; a width-scaled net (annonet_train --net-width-scaler, NetPimpl.cpp:73-130) with 4 / 6 / 10-channel layers or a
    // net with more than 64 classes runs on the fp32 kernels of the same engine instead (same results within the tighter
    // 1e-4 tolerance, slower) -- decided once per net, reported once.
    precision_ = requested_precision_;
    if (precision_ == 1) {
        std::string why;
        for (const ConvGroup& g : plan_.groups) {
            if (g.dead) continue;
            if (g.in_level >= 0 && (g.cin % 8 || g.in_c0 % 8)) why = "an input slice of " + std::to_string(g.cin) + " channels";
            if (g.out_level >= 0 && (g.cout % 8 || g.out_c0 % 8 || g.cout > 64)) why = "a layer with " + std::to_string(g.cout) + " output channels";
        }
        for (int l = 0; l <= plan_.levels; ++l)
            if (plan_.level_channels[l] % 8) why = "a level of " + std::to_string(plan_.level_channels[l]) + " channels";
        if (plan_.class_count > 64) why = std::to_string(plan_.class_count) + " classes";
        if (!why.empty()) {
            precision_ = 0;
            fprintf(stderr, "annonet_b200: this net has %s: running it on the fp32 kernels (the tensor-core path needs channel counts that are "
                            "multiples of 8, at most 64 per layer, and at most 64 classes)\n", why.c_str());
        }
    }
    const size_t np = plan_.num_params;
    std::vector<float> hp(np, 0.f), lrm(np, 1.f), wdm(np, 1.f);
    running_off_.assign(plan_.groups.size(), 0);
    num_updates_.assign(plan_.groups.size(), 0);
    window_.assign(plan_.groups.size(), 100);
    size_t roff = 0;
    std::vector<float> hr;
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        const Layer& lc = model_.layers[g.dlib_conv];
        if (lc.params.size() != g.nweights() + (size_t)g.cout) throw serialization_error("layer<" + std::to_string(g.dlib_conv) + ">: parameter count does not match the graph");
        pack_filters(g, lc.params.v.data(), hp.data() + g.w_off, hp.data() + g.b_off);
        for (size_t i = 0; i < g.nweights(); ++i) { lrm[g.w_off + i] = (float)lc.mult[0]; wdm[g.w_off + i] = (float)lc.mult[1]; }
        for (int i = 0; i < g.cout; ++i) { lrm[g.b_off + i] = (float)(lc.mult[0] * lc.mult[2]); wdm[g.b_off + i] = (float)(lc.mult[1] * lc.mult[3]); }
        if (g.dlib_bn >= 0) {
            const Layer& lb = model_.layers[g.dlib_bn];
            if (lb.params.size() != 2 * (size_t)g.cout) throw serialization_error("layer<" + std::to_string(g.dlib_bn) + ">: parameter count does not match the graph");
            std::memcpy(hp.data() + g.bn_off, lb.params.v.data(), sizeof(float) * 2 * g.cout);
            for (int i = 0; i < g.cout; ++i) {
                lrm[g.bn_off + i] = (float)lb.mult[0]; wdm[g.bn_off + i] = (float)lb.mult[1];
                lrm[g.bn_off + g.cout + i] = (float)(lb.mult[0] * lb.mult[2]); wdm[g.bn_off + g.cout + i] = (float)(lb.mult[1] * lb.mult[3]);
            }
            running_off_[gi] = roff;
            roff += 2 * (size_t)g.cout;
            hr.resize(roff, 0.f);
            if (lb.type == L_BN) {
                if (lb.running_means.size() != (size_t)g.cout || lb.running_variances.size() != (size_t)g.cout) throw serialization_error("bn_: running statistics missing");
                std::memcpy(hr.data() + running_off_[gi], lb.running_means.v.data(), sizeof(float) * g.cout);
                std::memcpy(hr.data() + running_off_[gi] + g.cout, lb.running_variances.v.data(), sizeof(float) * g.cout);
                num_updates_[gi] = lb.num_updates;
                window_[gi] = lb.window;
            }
        }
    }
    // grads_[np]: the step loss of a data-parallel run rides behind the gradients through the all-reduce
    params_.reserve(np * 4); grads_.reserve((np + 1) * 4); mom_.reserve(np * 4); lr_mul_.reserve(np * 4); wd_mul_.reserve(np * 4);
    running_.reserve(std::max<size_t>(roff, 1) * 4);
    AB_CUDA(cudaMemcpy(params_.p, hp.data(), np * 4, cudaMemcpyHostToDevice));
    AB_CUDA(cudaMemcpy(lr_mul_.p, lrm.data(), np * 4, cudaMemcpyHostToDevice));
    AB_CUDA(cudaMemcpy(wd_mul_.p, wdm.data(), np * 4, cudaMemcpyHostToDevice));
    if (roff) AB_CUDA(cudaMemcpy(running_.p, hr.data(), roff * 4, cudaMemcpyHostToDevice));
    nu_dev_.reserve(sizeof(unsigned long long) * num_updates_.size());
    AB_CUDA(cudaMemcpy(nu_dev_.p, num_updates_.data(), sizeof(unsigned long long) * num_updates_.size(), cudaMemcpyHostToDevice));
    AB_CUDA(cudaMemset(grads_.p, 0, np * 4));
    AB_CUDA(cudaMemset(mom_.p, 0, np * 4));
    if (adam_v_.p) AB_CUDA(cudaMemset(adam_v_.p, 0, np * 4));
    adam_t_ = 0;
    stats_valid_ = false;
    loaded_ = true;
    // free the per-layer host parameter copies: the device arena is authoritative from here on
    for (Layer& l : model_.layers) { std::vector<float>().swap(l.params.v); }
}

NetModel Engine::snapshot(bool as_runtime)
{
    if (!loaded_) throw state_error("no net loaded");
    synchronize();
    const size_t np = plan_.num_params;
    std::vector<float> hp(np), hr(running_.bytes / 4);
    AB_CUDA(cudaMemcpy(hp.data(), params_.p, np * 4, cudaMemcpyDeviceToHost));
    if (!hr.empty()) AB_CUDA(cudaMemcpy(hr.data(), running_.p, hr.size() * 4, cudaMemcpyDeviceToHost));
    if (!num_updates_.empty()) AB_CUDA(cudaMemcpy(num_updates_.data(), nu_dev_.p, sizeof(unsigned long long) * num_updates_.size(), cudaMemcpyDeviceToHost));
    // the batch statistics of the last forward pass: bn_ serializes its `means` / `invstds` members (layers.h:1428-1448), and
    // dlib's next training pass asserts that they have the running statistics' dimensions (cpu_dlib.cpp:1087)
    std::vector<float> hs;
    if (stats_valid_ && t_stats_.p && !stats_off_.empty()) {
        size_t n = 0;
        for (size_t gi = 0; gi < plan_.groups.size(); ++gi)
            if (plan_.groups[gi].dlib_bn >= 0) n = std::max(n, stats_off_[gi] + 2 * (size_t)plan_.groups[gi].cout);
        hs.resize(n);
        AB_CUDA(cudaMemcpy(hs.data(), t_stats_.p, n * 4, cudaMemcpyDeviceToHost));
    }
    NetModel out = model_;
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        Layer& lc = out.layers[g.dlib_conv];
        lc.params.v.resize(g.nweights() + g.cout);
        unpack_filters(g, hp.data() + g.w_off, hp.data() + g.b_off, lc.params.v.data());
        if (g.dlib_bn >= 0) {
            Layer& lb = out.layers[g.dlib_bn];
            lb.params.v.assign(hp.begin() + g.bn_off, hp.begin() + g.bn_off + 2 * g.cout);
            if (lb.type == L_BN) {
                lb.running_means.v.assign(hr.begin() + running_off_[gi], hr.begin() + running_off_[gi] + g.cout);
                lb.running_variances.v.assign(hr.begin() + running_off_[gi] + g.cout, hr.begin() + running_off_[gi] + 2 * g.cout);
                lb.num_updates = num_updates_[gi];
                lb.window = window_[gi];
                if (!hs.empty()) {   // [mean(C) | invstd(C)] of the last training forward pass, as [1, C, 1, 1] tensors
                    for (HostTensor* t : {&lb.means, &lb.invstds}) { t->d[0] = 1; t->d[1] = g.cout; t->d[2] = 1; t->d[3] = 1; }
                    lb.means.v.assign(hs.begin() + stats_off_[gi], hs.begin() + stats_off_[gi] + g.cout);
                    lb.invstds.v.assign(hs.begin() + stats_off_[gi] + g.cout, hs.begin() + stats_off_[gi] + 2 * g.cout);
                }
            }
        }
    }
    if (as_runtime && out.is_training) return to_runtime(out);
    return out;
}

void Engine::set_layer_momentum(int dlib_layer, const float* v, size_t n)
{
    if (!loaded_) throw state_error("no net loaded");
    if (dlib_layer < 0 || dlib_layer >= (int)plan_.layer_size.size() || plan_.layer_size[dlib_layer] != n)
        throw std::invalid_argument("layer<" + std::to_string(dlib_layer) + ">: solver state size does not match the layer's parameters");
    synchronize();
    std::vector<float> raw(n);
    for (const ConvGroup& g : plan_.groups) {
        if (g.dlib_conv == dlib_layer) pack_filters(g, v, raw.data(), raw.data() + g.nweights());
        else if (g.dlib_bn == dlib_layer) raw.assign(v, v + n);
        else continue;
        AB_CUDA(cudaMemcpy(mom_.as<float>() + plan_.layer_off[dlib_layer], raw.data(), n * 4, cudaMemcpyHostToDevice));
        return;
    }
}

void Engine::layer_values(int dlib_layer, int which, std::vector<float>& out)
{
    if (!loaded_) throw state_error("no net loaded");
    if (dlib_layer < 0 || dlib_layer >= (int)plan_.layer_size.size() || plan_.layer_size[dlib_layer] == 0)
        throw std::invalid_argument("layer<" + std::to_string(dlib_layer) + "> has no parameters");
    synchronize();
    const size_t off = plan_.layer_off[dlib_layer], n = plan_.layer_size[dlib_layer];
    std::vector<float> raw(n);
    const DevBuf& arena = which == 0 ? params_ : which == 1 ? grads_ : mom_;
    AB_CUDA(cudaMemcpy(raw.data(), arena.as<float>() + off, n * 4, cudaMemcpyDeviceToHost));
    out.resize(n);
    for (const ConvGroup& g : plan_.groups) {
        if (g.dlib_conv == dlib_layer) { unpack_filters(g, raw.data(), raw.data() + g.nweights(), out.data()); return; }
        if (g.dlib_bn == dlib_layer) { out = raw; return; }
    }
}

void Engine::set_bn_window(unsigned long long w)
{
    if (w == 0) throw std::invalid_argument("The batch normalization running stats window size can't be 0.");
    for (auto& x : window_) x = w;
}

void Engine::comm_init(const void* id128, int rank, int world)
{
    if (world < 1 || rank < 0 || rank >= world) throw std::invalid_argument("bad rank/world");
    AB_CUDA(cudaSetDevice(device_));
    rank_ = rank;
    world_ = world;
    if (world == 1) return;
    nccl_ = std::make_unique<NcclApi>();
    NcclApi::UniqueId id;
    std::memcpy(&id, id128, 128);
    nccl_->check(nccl_->CommInitRank(&comm_, world, id, rank), "ncclCommInitRank");
    // a step captured before the communicator existed would replay without its all-reduces and without the 1/world factor
    if (tc_) { tc_->drop_graph(); tc_->seen_key = TcState::GraphKey(); }
    // every replica starts from rank 0's state, as dnn_trainer copies the master net to its devices (trainer.h:1206-1253)
    if (loaded_) {
        synchronize();
        auto bcast = [&](DevBuf& b, size_t bytes) {
            if (b.p && bytes) nccl_->check(nccl_->Broadcast(b.p, b.p, bytes, /*ncclUint8*/ 1, 0, comm_, comm_stream_), "ncclBroadcast");
        };
        const size_t np = plan_.num_params;
        bcast(params_, np * 4);
        bcast(mom_, np * 4);
        bcast(running_, running_.bytes);
        bcast(nu_dev_, sizeof(unsigned long long) * num_updates_.size());
        if (adam_v_.p) { bcast(adam_v_, np * 4); bcast(adam_t_dev_, 4); }
        AB_CUDA(cudaStreamSynchronize(comm_stream_));
        if (tc_) tc_->fwd_packs_valid = false;
    }
}

// sum-all-reduce of grads[lo, hi) on the communication stream, ordered after everything enqueued so far
// on the compute stream (multi_device_tensor_averager, tensor_tools.h:1690-1736; the 1/world factor is
// folded into the optimizer step)
void Engine::allreduce_range(size_t lo, size_t hi)
{
    if (world_ <= 1 || hi <= lo) return;
    cudaEvent_t ev = ev_bucket_[bucket_slot_++ % 8];
    AB_CUDA(cudaEventRecord(ev, stream_));
    AB_CUDA(cudaStreamWaitEvent(comm_stream_, ev, 0));
    float* g = grads_.as<float>() + lo;   // (hi may be num_params + 1: the loss slot rides with the topmost bucket)
    nccl_->check(nccl_->AllReduce(g, g, hi - lo, /*ncclFloat32*/ 7, /*ncclSum*/ 0, comm_, comm_stream_), "ncclAllReduce");
}

k::Slice Engine::level_slice(Levels& L, int level, int c0, int C) const
{
    k::Slice s;
    s.p = L.buf[level].as<float>();
    s.sstride = (long long)plan_.level_channels[level] * L.dims.h[level] * L.dims.w[level];
    s.c0 = c0;
    s.C = C;
    return s;
}

}  // namespace ab
