// Training step of the fp16 tensor-core path (AB_PRECISION_F16): the dnn_trainer step of engine_train.cu
// (trainer.h:631-671,743-828) over padded flat NHWC fp16 tensors (tc.h).
//   forward : conv (tcgen05, kernels_tc.cu; split operands: activations and filters as fp16 hi + lo pairs, three MMAs per
//             K chunk) -> z [fp32] -> batch statistics -> bn/mish -> level slice [hi | lo]
//   loss    : fp32 logits planes (the head conv writes them directly) -> loss + gradient (kernels_misc.cu), times the loss scale
//   backward: bn/mish backward (z, gy -> dz [fp16], + bias gradient) -> filter gradient (kernels_wgrad_tc.cu, hi halves of
//             the activations) -> backward-data as a tcgen05 conv with transposed filters accumulating into the gradient slice
//   update  : fp32 master parameters / fp32 gradient arena (unscaled), the same fused SGD/Adam kernel as the fp32 path;
//             the fp16 filter packs of both directions are rebuilt from the master copy by one batched kernel.
#include "engine.h"
#include "kcommon.cuh"
#include "tc_state.h"

#include <cstdio>
#include <cstring>

namespace ab {

using tc::h16;
using tc::ConvLaunch;
using tc::ConvSetup;
using tc::Geo;
using tc::View;

namespace {

// N of a backward-data GEMM (= cin of the group) padded so that the <= 256 wide N blocks of setup_conv tile it
int tiled_n(int c)
{
    const int npad = tc::padded_n(c);
    const int nblocks = (npad + 255) / 256;
    const int nblk = ((npad + nblocks - 1) / nblocks + 15) / 16 * 16;
    return nblocks * nblk;
}

// taps of the backward-data convolution of a group: gin[pos][ci] += sum_t dz[pos + (dy, dx)_t][co] * w(t)[ci][co]
void backward_taps(const ConvGroup& g, TcGroup& t)
{
    if (g.ksize == 1) {
        t.bwd_ntaps = 1;
        t.bwd_tapmap[0] = 0; t.bwd_tapdy[0] = 0; t.bwd_tapdx[0] = 0;
        return;
    }
    t.bwd_ntaps = 9;
    for (int dx = -1; dx <= 1; ++dx)
        for (int dy = -1; dy <= 1; ++dy) {
            const int i = (dx + 1) * 3 + (dy + 1);   // dx-major: the order the merged-dy-tap mode of the kernel wants
            t.bwd_tapdy[i] = dy; t.bwd_tapdx[i] = dx;
            if (!g.is_cont)               // con_: gin[y][x] += dz[y - ky + 1][x - kx + 1] * w[co][ci][ky][kx]   (layers.h:217)
                t.bwd_tapmap[i] = (1 - dy) * 3 + (1 - dx);
            else                          // cont_ s1: gin[y][x] += dz[y + ky - 1][x + kx - 1] * w[ci][co][ky][kx] (layers.h:542)
                t.bwd_tapmap[i] = (dy + 1) * 3 + (dx + 1);   // cont_ s2 on the fine grid, evaluated at pixel (2i + 1, 2j + 1): the same
        }
}

}  // namespace

void Engine::tc_ensure_train(int n, int h, int w)
{
    if (!tc_) tc_ = std::make_unique<TcState>();
    TcState& T = *tc_;
    // Loss scale: per-pixel gradients are ~1/(n*h*w) times O(1) (loss.h:3158) -- below fp16's normal range.  A power of two near
    // n*h*w/8 puts them around 2^-3 with ~2^18 of headroom (measured spread over the whole net: 2e-4 .. 2e2 x 1/(n*h*w),
    // experiments/bf16_emulation.py); conversions saturate instead of overflowing.  hyper.loss_scale overrides.
    double sc = hyper.loss_scale;
    if (!(sc > 0)) { sc = 1; while (sc * 8 <= (double)n * h * w) sc *= 2; }
    if (T.train_ready && T.tw.n == n && T.tw.h == h && T.tw.w == w && T.loss_scale == (float)sc) return;
    synchronize();
    T.loss_scale = (float)sc;
    const float unscale = 1.f / T.loss_scale;
    T.train_ready = false;
    // group table + forward pack offsets (shared with the inference path)
    T.fwd_packs_valid = false;
    tc_pack_forward_filters();
    const int K = plan_.class_count;
    if (K > 64) throw std::invalid_argument("the tensor-core training path supports at most 64 classes (one 64-channel filter-gradient tile)");
    const int Kp = (K + 7) / 8 * 8;   // the loss gradient as a [positions][Kp] fp16 tensor
    ensure_workspace(T.tw, plan_, n, h, w, true, true, true, stream_);     // activations: hi + lo halves
    ensure_workspace(T.tg, plan_, n, h, w, false, false, false, stream_);  // gradients: single fp16 values, loss-scaled
    T.tw.pool_arg.reserve(std::max<size_t>(T.tw.pool.bytes / 2, 256));
    t_logits_.reserve((size_t)n * K * h * w * 4);
    t_glogits_.reserve((size_t)n * K * h * w * 4);
    t_loss_.reserve(sizeof(double));

    // per group buffers
    size_t z_elems = 0, dpool_elems = 0, upx_elems = 0, bwd_elems = 0, stt = 0, tp_elems = 0;
    int cmax = 8;
    stats_off_.assign(plan_.groups.size(), 0);
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        TcGroup& t = T.groups[gi];
        const int ol = g.out_level == -2 ? 0 : g.out_level;
        cmax = std::max(cmax, g.cout);
        for (int l = 0; l < t.nlaunch; ++l) {   // split forward packs: hi and lo filters side by side along N
            t.train_pack_off[l] = tp_elems;
            tp_elems += (2 * tc::packed_filter_elems(g.in_level < 0 ? 8 : g.cin, t.npad, t.fwd_ntaps[l]) + 511) / 512 * 512;
        }
        if (g.dlib_bn >= 0) {
            stats_off_[gi] = stt; stt += 2 * (size_t)g.cout;
            t.z_off = z_elems;
            z_elems += ((size_t)T.tw.geo[ol].positions() * g.cout + 63) / 64 * 64;
        }
        if (g.dead) continue;
        if (g.pool_input) { t.dpool_off = dpool_elems; dpool_elems += ((size_t)T.tw.geo[ol].positions() * g.cin + 63) / 64 * 64; }
        if (g.is_cont && g.stride == 2) { t.upx_off = upx_elems; upx_elems += ((size_t)T.tw.geo[ol].positions() * g.cin + 63) / 64 * 64; }
        if (g.in_level >= 0) {
            backward_taps(g, t);
            t.bwd_npad = tiled_n(g.cin);
            t.bwd_pack_off = bwd_elems;
            bwd_elems += (tc::packed_filter_elems(std::max(g.cout, 8), t.bwd_npad, t.bwd_ntaps) + 511) / 512 * 512;
        }
    }
    auto zero_reserve = [&](DevBuf& b, size_t bytes) {
        b.reserve(std::max<size_t>(bytes, 256));
        AB_CUDA(cudaMemsetAsync(b.p, 0, std::max<size_t>(bytes, 256), stream_));
    };
    zero_reserve(T.z, z_elems * 4);
    zero_reserve(T.dz, z_elems * 2);
    zero_reserve(T.dpool, dpool_elems * 2);
    zero_reserve(T.upx, upx_elems * 2);
    zero_reserve(T.glogits16, (size_t)T.tw.geo[0].positions() * Kp * 2);
    zero_reserve(T.bwd_packs, bwd_elems * 2);
    zero_reserve(T.train_packs, tp_elems * 2);
    // batch-statistics sums (forward) and backward reduction sums: one slot per layer and pass, all cleared by ONE memset at the
    // start of a step
    T.sums_slot = (size_t)2 * cmax;
    T.sums.reserve(sizeof(double) * T.sums_slot * 2 * plan_.groups.size());
    T.sums_side.reserve(sizeof(double) * 2 * cmax);
    T.bias_tmp.reserve(sizeof(float) * 256);
    T.wscratch.reserve(tc::wgrad_tc_scratch_floats() * 4);
    t_stats_.reserve(std::max<size_t>(stt, 1) * 4);

    // filter-pack jobs: every forward launch and every backward-data conv, rebuilt from the fp32 arena each step
    std::vector<tc::PackJob> jobs;
    T.max_pack_elems = 0;
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        const TcGroup& t = T.groups[gi];
        const int kk = g.ksize * g.ksize;
        for (int l = 0; l < t.nlaunch; ++l) {
            tc::PackJob j{};
            j.w_off = (long long)g.w_off; j.packed_off = (long long)t.train_pack_off[l];
            j.sn = g.is_cont ? kk : (long long)g.cin * kk; j.sk = g.is_cont ? (long long)g.cout * kk : kk;
            j.nreal = g.cout; j.npad = t.npad; j.kreal = g.cin; j.kp = tc::padded_k(g.in_level < 0 ? 8 : g.cin); j.ntaps = t.fwd_ntaps[l];
            j.parts = 2;
            for (int i = 0; i < j.ntaps; ++i) j.tapmap[i] = t.fwd_tapmap[l][i];
            T.max_pack_elems = std::max(T.max_pack_elems, j.ntaps * j.npad * j.kp * 2);
            jobs.push_back(j);
        }
    }
    T.n_fwd_jobs = (int)jobs.size();
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        const TcGroup& t = T.groups[gi];
        const int kk = g.ksize * g.ksize;
        if (g.dead || g.in_level < 0) continue;
        tc::PackJob j{};
        j.w_off = (long long)g.w_off; j.packed_off = (long long)t.bwd_pack_off;
        // n = ci, k = co:  con_ [co][ci][kk], cont_ [ci][co][kk]
        j.sn = g.is_cont ? (long long)g.cout * kk : kk; j.sk = g.is_cont ? kk : (long long)g.cin * kk;
        j.nreal = g.cin; j.npad = t.bwd_npad; j.kreal = g.cout; j.kp = tc::padded_k(std::max(g.cout, 8)); j.ntaps = t.bwd_ntaps;
        j.parts = 1;
        for (int i = 0; i < j.ntaps; ++i) j.tapmap[i] = t.bwd_tapmap[i];
        T.max_pack_elems = std::max(T.max_pack_elems, j.ntaps * j.npad * j.kp);
        jobs.push_back(j);
    }
    T.n_pack_jobs = (int)jobs.size();
    T.pack_jobs.reserve(jobs.size() * sizeof(tc::PackJob));
    AB_CUDA(cudaMemcpyAsync(T.pack_jobs.p, jobs.data(), jobs.size() * sizeof(tc::PackJob), cudaMemcpyHostToDevice, stream_));
    AB_CUDA(cudaStreamSynchronize(stream_));   // `jobs` is a local

    // launches
    const float* P = params_.as<float>();
    float* G = grads_.as<float>();
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        TcGroup& t = T.groups[gi];
        const int ol = g.out_level == -2 ? 0 : g.out_level;
        const bool up = g.is_cont && g.stride == 2;
        float* z = T.z.as<float>() + t.z_off;
        // forward: conv (+ bias) -> z, the logits planes, or (`up`) the destination slice
        const h16* tp = T.train_packs.as<h16>();
        if (g.dlib_bn >= 0)
            setup_forward_launches(plan_, g, t, T.tw, tp, t.train_pack_off, P, t.fwd, z, g.cout, 0, 0, tc::OUT_F32_NHWC, tc::ACT_NONE);
        else if (g.out_level == -2)
            setup_forward_launches(plan_, g, t, T.tw, tp, t.train_pack_off, P, t.fwd, t_logits_.p, 0, 0, 0, tc::OUT_F32_NCHW, tc::ACT_NONE);
        else
            setup_forward_launches(plan_, g, t, T.tw, tp, t.train_pack_off, P, t.fwd, T.tw.level[g.out_level].p,
                                   plan_.level_channels[g.out_level], g.out_c0, T.tw.level_lo[g.out_level], tc::OUT_BF16, tc::ACT_NONE);
        t.dgrad.clear();
        if (g.dead) continue;

        // gradient at the conv output: dz (in z's place), the loss gradient, or the gradient slice itself
        const h16* dz; int dz_ctot, dz_c0, dz_c;
        if (g.dlib_bn >= 0) { dz = T.dz.as<h16>() + t.z_off; dz_ctot = g.cout; dz_c0 = 0; dz_c = g.cout; }
        else if (g.out_level == -2) { dz = T.glogits16.as<h16>(); dz_ctot = Kp; dz_c0 = 0; dz_c = Kp; }
        else { dz = T.tg.level[ol].as<h16>(); dz_ctot = plan_.level_channels[ol]; dz_c0 = g.out_c0; dz_c = g.cout; }

        // filter gradient
        tc::WgradJob& wj = t.wgrad;
        wj = tc::WgradJob();
        wj.dy = dz; wj.dy_ctot = dz_ctot; wj.dy_c0 = dz_c0; wj.cy = dz_c; wj.cy_real = g.cout;
        wj.g = T.tw.geo[ol];
        wj.out = G + g.w_off;
        wj.unscale = unscale;
        const int kk = g.ksize * g.ksize;
        if (g.in_level < 0) { wj.x = T.tw.img.as<h16>(); wj.x_ctot = 8; wj.x_c0 = 0; wj.cx = 8; }
        else if (g.pool_input) { wj.x = T.tw.pool.as<h16>() + T.tw.pool_off[gi]; wj.x_ctot = g.cin; wj.x_c0 = 0; wj.cx = g.cin; }
        else if (up) { wj.x = T.upx.as<h16>() + t.upx_off; wj.x_ctot = g.cin; wj.x_c0 = 0; wj.cx = g.cin; }
        else { wj.x = T.tw.level[g.in_level].as<h16>(); wj.x_ctot = plan_.level_channels[g.in_level]; wj.x_c0 = g.in_c0; wj.cx = g.cin; }
        wj.cx_real = g.cin;
        wj.ntaps = kk;
        if (!g.is_cont) { wj.sy = (long long)g.cin * kk; wj.sx = kk; } else { wj.sy = kk; wj.sx = (long long)g.cout * kk; }
        for (int ky = 0; ky < g.ksize; ++ky)
            for (int kx = 0; kx < g.ksize; ++kx) {
                const int i = ky * g.ksize + kx;
                wj.tapmap[i] = i;
                if (g.ksize == 1) { wj.tap_dy[i] = 0; wj.tap_dx[i] = 0; }
                else if (!g.is_cont) { wj.tap_dy[i] = ky - 1; wj.tap_dx[i] = kx - 1; }   // dW[co][ci][ky][kx] = sum dz[p] x[p + (ky-1, kx-1)]
                else if (!up) { wj.tap_dy[i] = 1 - ky; wj.tap_dx[i] = 1 - kx; }          // cont_ s1: x[p + (1-ky, 1-kx)]
                else { wj.tap_dy[i] = -ky; wj.tap_dx[i] = -kx; }                         // cont_ s2: zero-stuffed x at p - (ky, kx)
            }

        tc::setup_wgrad_tc(t.wgrad_tc, wj, T.wscratch.as<float>());
        if (up) {   // algorithmic work of the stride-2 cont_: its input lives on the COARSE grid (the zero-stuffed copy is an implementation detail)
            const double pxf = (double)wj.g.N * wj.g.H * wj.g.W;
            t.wgrad_tc.flops *= 0.25;
            t.wgrad_tc.bytes = 2.0 * pxf * (g.cin * 0.25 + g.cout);
        }

        // backward-data
        if (g.in_level < 0) continue;
        ConvSetup s;
        s.src = dz; s.src_positions = T.tw.geo[ol].positions(); s.src_ctot = dz_ctot; s.a_c0 = dz_c0; s.cin = dz_c;
        s.packed = T.bwd_packs.as<h16>() + t.bwd_pack_off;
        s.ntaps = t.bwd_ntaps;
        for (int i = 0; i < s.ntaps; ++i) { s.tap_dy[i] = t.bwd_tapdy[i]; s.tap_dx[i] = t.bwd_tapdx[i]; }
        s.npad_total = t.bwd_npad;
        s.grid = T.tw.geo[ol];
        t.dgrad.assign(1, ConvLaunch());
        ConvLaunch& L = t.dgrad[0];
        tc::setup_conv(L, s);
        L.act = tc::ACT_NONE; L.bias = nullptr;
        L.cout = g.cin;
        if (g.pool_input) { L.mode = tc::OUT_BF16; L.out = T.dpool.as<h16>() + t.dpool_off; L.out_ctot = g.cin; L.out_c0 = 0; }
        else {
            L.mode = up ? tc::OUT_BF16_DOWN_ACCUM : tc::OUT_BF16_ACCUM;
            L.out = T.tg.level[g.in_level].p; L.out_ctot = plan_.level_channels[g.in_level]; L.out_c0 = g.in_c0;
            if (up) { L.out_S = T.tw.geo[g.in_level].S(); L.out_P = T.tw.geo[g.in_level].P(); }
        }
        const double px = (double)s.grid.N * s.grid.H * s.grid.W;
        L.flops = 2.0 * px * g.cin * g.cout * s.ntaps / (up ? 4.0 : 1.0);
        L.bytes = 2.0 * px * (g.cin / (up ? 4.0 : 1.0) + g.cout);
    }
    // Gradient slices: the first writer of a channel range in backward order ASSIGNS (no read-modify-write on the epilogue's
    // critical path, no memset of the level), later writers accumulate (dnn/core.h:958-974 zeroes x_grad and lets every
    // consumer add; the static plan knows who comes first).  A level still needs zeroing when a writer would accumulate
    // into channels nobody has assigned (max-pool gradient first, partial overlaps) or when channels are never written.
    {
        std::vector<std::vector<char>> written(plan_.levels + 1);
        for (int l = 0; l <= plan_.levels; ++l) written[l].assign(plan_.level_channels[l], 0);
        T.grad_level_needs_zero.assign(plan_.levels + 1, 0);
        static const bool always_zero = getenv("AB_GRAD_MEMSET") != nullptr;
        for (int gi = (int)plan_.groups.size() - 1; gi >= 0; --gi) {
            const ConvGroup& g = plan_.groups[gi];
            if (g.dead || g.in_level < 0) continue;
            std::vector<char>& w = written[g.in_level];
            int nw = 0;
            for (int c = g.in_c0; c < g.in_c0 + g.cin; ++c) nw += w[c];
            const bool none = nw == 0, all = nw == g.cin;
            if (g.pool_input || always_zero) { if (!all) T.grad_level_needs_zero[g.in_level] = 1; }
            else if (none) {
                ConvLaunch& L = T.groups[gi].dgrad[0];
                L.mode = L.mode == tc::OUT_BF16_DOWN_ACCUM ? tc::OUT_BF16_DOWN : tc::OUT_BF16;
            } else if (!all) T.grad_level_needs_zero[g.in_level] = 1;
            for (int c = g.in_c0; c < g.in_c0 + g.cin; ++c) w[c] = 1;
        }
        for (int l = 0; l <= plan_.levels; ++l)
            for (char c : written[l])
                if (!c) T.grad_level_needs_zero[l] = 1;
        // one zeroing now: assigned slices keep their (never written) pad positions at zero from here on
        for (int l = 0; l <= plan_.levels; ++l)
            k::fill_zero(T.tg.level[l].p, (size_t)T.tg.geo[l].positions() * plan_.level_channels[l] * sizeof(h16), stream_);
        static const bool debug = getenv("AB_TC_DEBUG") != nullptr;
        if (debug)
            for (int l = 0; l <= plan_.levels; ++l) fprintf(stderr, "gradient level %d: memset per step %d\n", l, (int)T.grad_level_needs_zero[l]);
    }
    tn_ = n; th_ = h; tw_ = w;
    stats_valid_ = false;
    T.train_ready = true;
}

void Engine::level_values(int level, bool gradient, std::vector<float>& out, int dims[4])
{
    if (tn_ == 0) throw state_error("no training step has run");
    if (level < 0 || level > plan_.levels) throw std::invalid_argument("no such level");
    synchronize();
    const int C = plan_.level_channels[level];
    if (precision_ == 1) {
        TcState& T = *tc_;
        const Geo& g = T.tw.geo[level];
        dims[0] = g.N; dims[1] = C; dims[2] = g.H; dims[3] = g.W;
        std::vector<uint16_t> raw((size_t)g.positions() * C);
        AB_CUDA(cudaMemcpy(raw.data(), (gradient ? T.tg : T.tw).level[level].p, raw.size() * 2, cudaMemcpyDeviceToHost));
        out.resize((size_t)g.N * C * g.H * g.W);
        for (int n = 0; n < g.N; ++n)
            for (int c = 0; c < C; ++c)
                for (int y = 0; y < g.H; ++y)
                    for (int x = 0; x < g.W; ++x) {
                        __half_raw hr;
                        hr.x = raw[(size_t)g.pos(n, y, x) * C + c];
                        const float f = __half2float(__half(hr));   // (activations: the hi halves; gradients: loss-scaled)
                        out[(((size_t)n * C + c) * g.H + y) * g.W + x] = gradient ? f / T.loss_scale : f;
                    }
    } else {
        Levels& L = gradient ? tw_grad_ : tw_act_;
        dims[0] = L.n; dims[1] = C; dims[2] = L.dims.h[level]; dims[3] = L.dims.w[level];
        out.resize((size_t)dims[0] * dims[1] * dims[2] * dims[3]);
        AB_CUDA(cudaMemcpy(out.data(), L.buf[level].p, out.size() * 4, cudaMemcpyDeviceToHost));
    }
}

// One step = ~390 kernels, many of them a few microseconds long (the coarse levels): launch-bound when issued one by
// one.  The step is static (shapes, pointers, the bn counters live on the device), so it is captured once into a CUDA
// graph and replayed while nothing it bakes in changes (TcState::GraphKey).  Not captured: the first step of a
// configuration (lazy module loading, cudaFuncSetAttribute) and per-kernel profiling (Adam's step counter lives on the
// device, so Adam steps are captured like SGD steps).  Data-parallel runs capture the step as segments split at the gradient-bucket boundaries and issue the NCCL
// all-reduces eagerly between them (TcState::graph_segs).
void Engine::tc_train_step(const uint8_t* imgs_dev, const void* labels_dev, int n, int h, int w, bool update)
{
    if (n < 2) throw std::invalid_argument("training needs a mini-batch of at least 2 (bn_ runs in batch mode only for num_samples > 1, layers.h:1392)");
    tc_ensure_train(n, h, w);
    TcState& T = *tc_;
    static const bool no_graph = getenv("AB_NO_GRAPH") != nullptr;
    if (hyper.solver != 0 && update) ensure_adam_state();   // allocations must happen outside the capture
    if (no_graph || k::prof_enabled()) { tc_train_body(imgs_dev, labels_dev, n, h, w, update, false); return; }
    TcState::GraphKey key;
    key.imgs = imgs_dev; key.labels = labels_dev; key.n = n; key.h = h; key.w = w; key.update = update; key.skip_dead = hyper.skip_dead_layers;
    key.world = world_; key.solver = hyper.solver; key.loss_scale = hyper.loss_scale;
    key.lr = hyper.learning_rate; key.wd = hyper.weight_decay; key.momentum = hyper.momentum; key.window = window_.empty() ? 0 : window_[0];
    if (!(T.graph_segs.size() && T.graph_key == key)) {
        if (!(T.seen_key == key)) {   // first time: run eagerly
            T.seen_key = key;
            tc_train_body(imgs_dev, labels_dev, n, h, w, update, false);
            return;
        }
        T.drop_graph();
        const unsigned long long k0 = k::g_launches.load();
        AB_CUDA(cudaStreamBeginCapture(stream_, cudaStreamCaptureModeThreadLocal));
        try {
            tc_train_body(imgs_dev, labels_dev, n, h, w, update, true);   // closes a segment at every bucket boundary
            tc_close_segment();
        } catch (...) {
            cudaGraph_t graph = nullptr;
            cudaStreamEndCapture(stream_, &graph);
            if (graph) cudaGraphDestroy(graph);
            T.drop_graph();
            throw;
        }
        T.graph_kernels = k::g_launches.load() - k0;
        T.graph_key = key;
    }
    // replay: segment, bucket all-reduce (eager, communication stream), ..., join, last segment (optimizer step)
    for (size_t i = 0; i < T.graph_segs.size(); ++i) {
        AB_CUDA(cudaGraphLaunch(T.graph_segs[i], stream_));
        if (i < T.graph_buckets.size()) {
            allreduce_range(T.graph_buckets[i].first, T.graph_buckets[i].second);
            if (i + 1 == T.graph_buckets.size()) {
                AB_CUDA(cudaEventRecord(ev_comm_done_, comm_stream_));
                AB_CUDA(cudaStreamWaitEvent(stream_, ev_comm_done_, 0));
            }
        }
    }
    k::count_launch((int)T.graph_kernels);
}

// ends the capture running on the compute stream and keeps it as the next segment of the step
void Engine::tc_close_segment()
{
    cudaGraph_t graph = nullptr;
    AB_CUDA(cudaStreamEndCapture(stream_, &graph));
    cudaGraphExec_t exec = nullptr;
    const cudaError_t e = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    AB_CUDA(e);
    tc_->graph_segs.push_back(exec);
}

void Engine::tc_train_body(const uint8_t* imgs_dev, const void* labels_dev, int n, int h, int w, bool update, bool capture)
{
    TcState& T = *tc_;
    const float* P = params_.as<float>();
    float* G = grads_.as<float>();
    const int K = plan_.class_count;

    // h16 filter packs of both directions from the fp32 master parameters
    const tc::PackJob* jobs = T.pack_jobs.as<tc::PackJob>();
    AB_CUDA(cudaMemsetAsync(T.sums.p, 0, sizeof(double) * T.sums_slot * 2 * plan_.groups.size(), stream_));
    tc::pack_filters_batched(P, T.train_packs.as<h16>(), jobs, T.n_fwd_jobs, T.max_pack_elems, stream_);
    tc::pack_filters_batched(P, T.bwd_packs.as<h16>(), jobs + T.n_fwd_jobs, T.n_pack_jobs - T.n_fwd_jobs, T.max_pack_elems, stream_);
    T.fwd_packs_valid = false;   // (the inference packs no longer match the parameters once the optimizer has run)
    const float unscale = 1.f / T.loss_scale;

    tc::image_to_nhwc(imgs_dev, h, w, 0, 0, T.tw.img.as<h16>(), T.tw.geo[0], model_.avg[0], model_.avg[1], model_.avg[2], stream_, T.tw.img_lo);

    // Side stream: work off the critical path of the layer chain -- filter gradients (only the optimizer needs them), the dead
    // layers -- runs next to it; the chain's launches at the coarse levels occupy a few SMs each.  Forks and joins are event
    // dependencies, captured into the step's CUDA graph as parallel branches.
    static const bool no_side = getenv("AB_NO_SIDE") != nullptr;
    bool side_busy = false;
    auto fork_side = [&]() -> cudaStream_t {
        if (no_side || k::prof_enabled()) return stream_;
        cudaEvent_t ev = ev_fork_[fork_slot_++ % 32];
        AB_CUDA(cudaEventRecord(ev, stream_));
        AB_CUDA(cudaStreamWaitEvent(side_stream_, ev, 0));
        side_busy = true;
        return side_stream_;
    };
    auto join_side = [&]() {
        if (!side_busy) return;
        AB_CUDA(cudaEventRecord(ev_join_, side_stream_));
        AB_CUDA(cudaStreamWaitEvent(stream_, ev_join_, 0));
        side_busy = false;
    };

    // ---- forward (batch-norm training mode)
    // The two reduction passes (bn statistics, bn backward sums) walk the positions downwards: their producers walked upwards, so
    // the last ~100 MB they wrote are still in the 126 MB L2, and the apply pass that follows (upwards) starts where the
    // reduction stopped.  Measured: 8.67 -> 8.57 ms per step; alternating the convolutions as well brought nothing (8.59).
    static const int down = getenv("AB_WALK") && atoi(getenv("AB_WALK")) == 0 ? 0 : 1;
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        TcGroup& t = T.groups[gi];
        if (g.dead && hyper.skip_dead_layers) continue;
        const int ol = g.out_level == -2 ? 0 : g.out_level;
        const Geo& geo = T.tw.geo[ol];
        // dead layers (computed, never consumed: NetStructure.h:41-55) only feed their own running statistics: side stream
        cudaStream_t fs = g.dead ? fork_side() : stream_;
        double* sums = T.sums.as<double>() + T.sums_slot * (2 * gi);
        if (g.pool_input)
            tc::maxpool_nhwc(T.tw.level[g.in_level].as<h16>(), plan_.level_channels[g.in_level], g.in_c0, g.cin, T.tw.geo[g.in_level],
                             T.tw.pool.as<h16>() + T.tw.pool_off[gi], geo, fs, T.tw.pool_arg.as<uint8_t>() + T.tw.pool_off[gi],
                             T.tw.level_lo[g.in_level], geo.positions() * g.cin);
        // batch statistics ride in the conv epilogue where one 16-channel block is the whole output (the full-resolution layers):
        // saves the pass that re-reads z
        static const bool no_fused_stats = getenv("AB_NO_FUSED_STATS") != nullptr;
        const bool fused_stats = g.dlib_bn >= 0 && !no_fused_stats && t.fwd.size() == 1 && t.fwd[0].mode == tc::OUT_F32_NHWC &&
                                 t.fwd[0].nblk == 16 * t.fwd[0].nsum && t.fwd[0].npad_total == 16 * t.fwd[0].nsum;
        if (fused_stats) {
            t.fwd[0].stat_sums = sums;
            t.fwd[0].stat_mish = g.has_mish && !g.bn_first;
        }
        for (const ConvLaunch& L : t.fwd) tc::launch_conv(L, fs);
        if (g.out_level >= 0) run_up_resize(plan_, g, T.tw, fs);
        if (fused_stats) t.fwd[0].stat_sums = nullptr;   // (the launch took its copy)
        if (g.dlib_bn >= 0) {
            const int mish_first = g.bn_first ? 0 : 1;
            float* stats = t_stats_.as<float>() + stats_off_[gi];
            float* rmean = running_.as<float>() + running_off_[gi];
            const tc::ViewF z{T.z.as<float>() + t.z_off, g.cout, 0, g.cout};
            if (!fused_stats) tc::bn_stats_nhwc(z, geo, g.has_mish && mish_first, sums, fs, down, 1);
            const double m = (double)((long long)n * geo.H * geo.W);
            const float eps = (float)model_.layers[g.dlib_bn].eps;
            if (g.dead) {
                k::bn_finalize(sums, g.cout, (long long)m, eps, nu_dev_.as<unsigned long long>() + gi, window_[gi], rmean, rmean + g.cout, stats,
                               fs);
            } else {
                // bn_finalize folded into the apply kernel: one launch less per layer
                tc::BnFinalize fin;
                fin.sums = sums; fin.m = m; fin.eps = (double)eps;
                fin.num_updates = nu_dev_.as<unsigned long long>() + gi; fin.window = window_[gi];
                fin.rmean = rmean; fin.rvar = rmean + g.cout; fin.stats_out = stats;
                const View y{T.tw.level[g.out_level].as<h16>(), plan_.level_channels[g.out_level], g.out_c0, g.cout, T.tw.level_lo[g.out_level]};
                tc::bn_apply_nhwc(z, y, geo, mish_first, g.has_mish, P + g.bn_off, P + g.bn_off + g.cout, stats, fs, &fin);
            }
        }
    }
    join_side();
    stats_valid_ = true;

    // ---- loss + gradient at the logits (fp32 planes), then h16 [positions][8] for the tensor-core backward
    k::loss_forward_backward(t_logits_.as<float>(), labels_dev, t_glogits_.as<float>(), n, K, h, w, t_loss_.as<double>(), stream_);
    tc::planes_to_nhwc8(t_glogits_.as<float>(), K, T.glogits16.as<h16>(), T.tw.geo[0], stream_, T.loss_scale);

    // ---- backward
    for (int l = 0; l <= plan_.levels; ++l)
        if (T.grad_level_needs_zero[l])
            k::fill_zero(T.tg.level[l].p, (size_t)T.tg.geo[l].positions() * plan_.level_channels[l] * sizeof(h16), stream_);
    const size_t np = plan_.num_params;
    k::fill_zero(G, np * sizeof(float), stream_);
    static const int nbuckets = getenv("AB_BUCKETS") ? std::max(1, atoi(getenv("AB_BUCKETS"))) : 2;
    // eager: all-reduce the bucket now;  capturing: the bucket ends a graph segment (tc_train_step issues the all-reduce at replay)
    auto bucket_ready = [&](size_t lo, size_t hi) {
        join_side();   // the bucket's filter gradients are complete
        if (!capture) { allreduce_range(lo, hi); return; }
        if (world_ <= 1 || hi <= lo) return;
        tc_close_segment();
        T.graph_buckets.emplace_back(lo, hi);
        AB_CUDA(cudaStreamBeginCapture(stream_, cudaStreamCaptureModeThreadLocal));
    };
    size_t bucket_hi = np;
    if (world_ > 1) { k::loss_to_slot(t_loss_.as<double>(), G + np, stream_); bucket_hi = np + 1; }   // the loss rides with the first bucket
    for (int gi = (int)plan_.groups.size() - 1; gi >= 0; --gi) {
        const ConvGroup& g = plan_.groups[gi];
        TcGroup& t = T.groups[gi];
        // (a group reached only through a resized join gets no gradient in the reference, plan.h grad_blocked: its parameter
        // gradients stay at the zeros the arena was filled with)
        if (!g.dead && !grad_blocked(T.tw.dims, g)) {
            const int ol = g.out_level == -2 ? 0 : g.out_level;
            const Geo& geo = T.tw.geo[ol];
            const bool up = g.is_cont && g.stride == 2;
            if (g.dlib_bn >= 0) {
                const int mish_first = g.bn_first ? 0 : 1;
                const float* stats = t_stats_.as<float>() + stats_off_[gi];
                const tc::ViewF z{T.z.as<float>() + t.z_off, g.cout, 0, g.cout};
                const View dzv{T.dz.as<h16>() + t.z_off, g.cout, 0, g.cout};
                const View gy{T.tg.level[g.out_level].as<h16>(), plan_.level_channels[g.out_level], g.out_c0, g.cout};
                double* red = T.sums.as<double>() + T.sums_slot * (2 * gi + 1);
                tc::bn_backward_reduce_nhwc(z, gy, geo, mish_first, g.has_mish, P + g.bn_off, P + g.bn_off + g.cout, stats, red, stream_, down, 1);
                tc::bn_backward_apply_nhwc(z, gy, dzv, geo, mish_first, g.has_mish, P + g.bn_off, P + g.bn_off + g.cout, stats, red,
                                           G + g.bn_off, G + g.bn_off + g.cout, G + g.b_off, stream_, 0, unscale);
            } else {
                // bias gradient = channel sums of the gradient at the conv output (layers.h:226, cpu_dlib.cpp:470-504)
                const tc::WgradJob& wj = t.wgrad;
                const View dzv{const_cast<h16*>(wj.dy), wj.dy_ctot, wj.dy_c0, wj.cy};
                k::fill_zero(T.bias_tmp.p, sizeof(float) * 64, stream_);
                tc::channel_sum_nhwc(dzv, geo, T.bias_tmp.as<float>(), stream_, unscale);
                AB_CUDA(cudaMemcpyAsync(G + g.b_off, T.bias_tmp.p, sizeof(float) * g.cout, cudaMemcpyDeviceToDevice, stream_));
            }
            {   // the filter gradient is needed by the optimizer only: off the chain, onto the side stream
                cudaStream_t ws = fork_side();
                if (up) {
                    const View xin{T.tw.level[g.in_level].as<h16>(), plan_.level_channels[g.in_level], g.in_c0, g.cin};
                    tc::upsample_zero_nhwc(xin, T.tw.geo[g.in_level], T.upx.as<h16>() + t.upx_off, geo, ws);
                }
                tc::launch_wgrad_tc(t.wgrad_tc, ws);
            }
            for (const ConvLaunch& L : t.dgrad) tc::launch_conv(L, stream_);
            if (g.pool_input) {
                const View gin{T.tg.level[g.in_level].as<h16>(), plan_.level_channels[g.in_level], g.in_c0, g.cin};
                tc::maxpool_backward_nhwc(T.tw.pool_arg.as<uint8_t>() + T.tw.pool_off[gi], g.cin, T.tw.geo[g.in_level], T.dpool.as<h16>() + t.dpool_off,
                                          geo, gin, stream_);
            }
        }
        if (world_ > 1 && bucket_hi - g.w_off >= np / nbuckets) {
            bucket_ready(g.w_off, bucket_hi);
            bucket_hi = g.w_off;
        }
    }
    if (world_ > 1) {
        bucket_ready(0, bucket_hi);
        if (!capture) {
            AB_CUDA(cudaEventRecord(ev_comm_done_, comm_stream_));
            AB_CUDA(cudaStreamWaitEvent(stream_, ev_comm_done_, 0));
        }
        k::slot_to_loss(G + np, t_loss_.as<double>(), world_, stream_);   // mean of the ranks' losses (trainer.h:1206-1253)
    }
    join_side();
    if (update) optimizer_step();
    AB_CUDA(cudaMemcpyAsync(loss_host_, t_loss_.p, sizeof(double), cudaMemcpyDeviceToHost, stream_));
    AB_CUDA(cudaGetLastError());
}

}  // namespace ab
