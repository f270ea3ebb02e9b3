// fp16 tensor-core path of the engine (AB_PRECISION_F16): the same conv groups as the fp32 path
// (plan.h), executed as tcgen05 implicit GEMMs over padded flat NHWC fp16 tensors (tc.h).
#include "engine.h"
#include "tc_state.h"

#include <cstring>

namespace ab {

using tc::h16;
using tc::ConvLaunch;
using tc::ConvSetup;
using tc::Geo;

namespace {

// taps of one launch of a group in the forward direction: (dy, dx) offsets on the M grid and the index of
// the filter tap in the arena ([..][ky*3 + kx])
void forward_taps(const ConvGroup& g, TcGroup& t)
{
    t.nlaunch = 1;
    if (g.ksize == 1) {
        t.fwd_ntaps[0] = 1;
        t.fwd_tapmap[0][0] = 0; t.fwd_tapdy[0][0] = 0; t.fwd_tapdx[0][0] = 0;
    } else if (!g.is_cont) {           // con_ 3x3 s1 p1: out[y][x] += w[ky][kx] * in[y+ky-1][x+kx-1]; taps dx-major (tc.h "kym")
        t.fwd_ntaps[0] = 9;
        for (int dx = -1; dx <= 1; ++dx)
            for (int dy = -1; dy <= 1; ++dy) {
                const int i = (dx + 1) * 3 + (dy + 1);
                t.fwd_tapmap[0][i] = (dy + 1) * 3 + (dx + 1); t.fwd_tapdy[0][i] = dy; t.fwd_tapdx[0][i] = dx;
            }
    } else if (g.stride == 1) {        // cont_ 3x3 s1 p1: out[Y][X] += w[ky][kx] * in[Y-ky+1][X-kx+1]   (layers.h:531)
        t.fwd_ntaps[0] = 9;
        for (int dx = -1; dx <= 1; ++dx)
            for (int dy = -1; dy <= 1; ++dy) {
                const int i = (dx + 1) * 3 + (dy + 1);
                t.fwd_tapmap[0][i] = (1 - dy) * 3 + (1 - dx); t.fwd_tapdy[0][i] = dy; t.fwd_tapdx[0][i] = dx;
            }
    } else {
        // cont_ 3x3 s2 p0 ("up"): out[2r+a][2c+b] += w[ky][kx] * in[r + (a-ky)/2][c + (b-kx)/2] for ky = a (mod 2), kx = b (mod 2).
        // M position (r, c) of the coarse grid sits one row and one column before pixel (r, c): pixel (i, j) is at
        // offset (i - r + 1)*P + (j - c + 1).
        t.nlaunch = 4;
        for (int a = 0; a < 2; ++a)
            for (int b = 0; b < 2; ++b) {
                const int l = a * 2 + b;
                int n = 0;
                for (int ky = a; ky < 3; ky += 2)
                    for (int kx = b; kx < 3; kx += 2) {
                        t.fwd_tapmap[l][n] = ky * 3 + kx;
                        t.fwd_tapdy[l][n] = (a - ky) / 2 + 1;
                        t.fwd_tapdx[l][n] = (b - kx) / 2 + 1;
                        ++n;
                    }
                t.fwd_ntaps[l] = n;
            }
    }
}

}  // namespace

// (re)build the h16 filter packs of the forward direction from the fp32 parameter arena
void Engine::tc_pack_forward_filters()
{
    if (!tc_) tc_ = std::make_unique<TcState>();
    TcState& T = *tc_;
    if (T.groups.size() != plan_.groups.size()) {
        T.groups.assign(plan_.groups.size(), TcGroup());
        size_t off = 0;
        for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
            const ConvGroup& g = plan_.groups[gi];
            TcGroup& t = T.groups[gi];
            forward_taps(g, t);
            t.npad = tc::padded_n(g.cout);
            const int cin = g.in_level < 0 ? 8 : g.cin;
            for (int l = 0; l < t.nlaunch; ++l) {
                t.fwd_pack_off[l] = off;
                off += (tc::packed_filter_elems(cin, t.npad, t.fwd_ntaps[l]) + 511) / 512 * 512;  // keep 1 KB alignment
            }
        }
        T.fwd_packs.reserve(off * sizeof(h16));
        T.fwd_packs_valid = false;
    }
    if (T.fwd_packs_valid) return;
    const float* P = params_.as<float>();
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        const TcGroup& t = T.groups[gi];
        const int kk = g.ksize * g.ksize;
        // arena: con_ [cout][cin][kk], cont_ [cin][cout][kk]
        const long long sn = g.is_cont ? kk : (long long)g.cin * kk, sk = g.is_cont ? (long long)g.cout * kk : kk;
        for (int l = 0; l < t.nlaunch; ++l)
            tc::pack_filters_bf16(P + g.w_off, T.fwd_packs.as<h16>() + t.fwd_pack_off[l], g.cout, t.npad, g.cin, t.fwd_ntaps[l],
                                  t.fwd_tapmap[l], sn, sk, stream_);
    }
    AB_CUDA(cudaGetLastError());
    T.fwd_packs_valid = true;
}

// workspace of padded flat NHWC level buffers for n samples of h x w
void ensure_workspace(TcWorkspace& ws, const Plan& plan, int n, int h, int w, bool with_image, bool with_pool, bool split, cudaStream_t st)
{
    if (ws.n == n && ws.h == h && ws.w == w && ws.split == split) return;
    const LevelDims d = level_dims(h, w, plan.levels);
    const size_t halves = split ? 2 : 1;
    ws.dims = d;
    size_t up_elems = 0;
    for (const ConvGroup& g : plan.groups)
        if (g.is_cont && g.stride == 2 && d.resized[g.out_level]) {
            tc::Geo& ug = ws.up_geo[g.out_level];
            ug.N = n; ug.H = d.up_h[g.out_level]; ug.W = d.up_w[g.out_level];
            up_elems = std::max(up_elems, (size_t)ug.positions() * g.cout);
        }
    if (up_elems) ws.up_tmp.reserve(up_elems * halves * sizeof(h16));
    ws.resize.build(d, plan.levels, st);
    for (int l = 0; l <= plan.levels; ++l) {
        ws.geo[l].N = n; ws.geo[l].H = d.h[l]; ws.geo[l].W = d.w[l];
        const size_t elems = (size_t)ws.geo[l].positions() * plan.level_channels[l];
        ws.level_lo[l] = split ? (long long)elems : 0;
        ws.level[l].reserve(elems * halves * sizeof(h16));
        AB_CUDA(cudaMemsetAsync(ws.level[l].p, 0, elems * halves * sizeof(h16), st));  // pad positions must be zero and stay zero
    }
    if (with_image) {
        const size_t elems = (size_t)ws.geo[0].positions() * 8;
        ws.img_lo = split ? (long long)elems : 0;
        ws.img.reserve(elems * halves * sizeof(h16));
        AB_CUDA(cudaMemsetAsync(ws.img.p, 0, elems * halves * sizeof(h16), st));
    }
    if (with_pool) {
        ws.pool_off.assign(plan.groups.size(), 0);
        size_t off = 0;
        for (size_t gi = 0; gi < plan.groups.size(); ++gi) {
            const ConvGroup& g = plan.groups[gi];
            if (!g.pool_input) continue;
            ws.pool_off[gi] = off;
            // split: the group's lo half follows its hi half directly (positions * cin elements behind)
            off += ((size_t)ws.geo[g.out_level].positions() * g.cin * halves + 63) / 64 * 64;
        }
        ws.pool.reserve(off * sizeof(h16));
        AB_CUDA(cudaMemsetAsync(ws.pool.p, 0, off * sizeof(h16), st));
    }
    ws.n = n; ws.h = h; ws.w = w; ws.split = split;
}

// prepared launch of one conv group in the forward direction.  `dst`/`dst_ctot`/`dst_c0`: where the result goes
// (a level slice, a dense z tensor, or the f32 logits planes), `mode`/`act` as in tc.h.
void setup_forward_launches(const Plan& plan, const ConvGroup& g, const TcGroup& t, const TcWorkspace& ws, const h16* packs,
                            const size_t* pack_off, const float* P, std::vector<ConvLaunch>& out, void* dst, int dst_ctot, int dst_c0,
                            long long dst_lo_off, int mode, int act)
{
    out.assign(t.nlaunch, ConvLaunch());
    const bool up = g.is_cont && g.stride == 2;
    const int ol = g.out_level == -2 ? 0 : g.out_level;
    // the M grid walks the SOURCE tensor's positions: same geometry as the output except for `up`
    const Geo& grid = up ? ws.geo[g.in_level] : ws.geo[ol];
    for (int l = 0; l < t.nlaunch; ++l) {
        ConvSetup s;
        if (g.in_level < 0) { s.src = ws.img.as<h16>(); s.src_ctot = 8; s.a_c0 = 0; s.cin = 8; }
        else if (g.pool_input) { s.src = ws.pool.as<h16>() + ws.pool_off[&g - plan.groups.data()]; s.src_ctot = g.cin; s.a_c0 = 0; s.cin = g.cin; }
        else { s.src = ws.level[g.in_level].as<h16>(); s.src_ctot = plan.level_channels[g.in_level]; s.a_c0 = g.in_c0; s.cin = g.cin; }
        s.src_positions = grid.positions();
        // split workspace: the lo halves sit exactly src_positions rows behind the hi halves (one tensor map covers both)
        s.split = ws.split ? 1 : 0;
        s.packed = packs + pack_off[l];
        s.ntaps = t.fwd_ntaps[l];
        for (int i = 0; i < s.ntaps; ++i) { s.tap_dy[i] = t.fwd_tapdy[l][i]; s.tap_dx[i] = t.fwd_tapdx[l][i]; }
        s.npad_total = t.npad;
        s.grid = grid;
        ConvLaunch& L = out[l];
        tc::setup_conv(L, s);
        L.mode = mode; L.act = act;
        L.cout = g.cout;
        L.out = dst; L.out_ctot = dst_ctot; L.out_c0 = dst_c0; L.out_lo_off = dst_lo_off;
        L.bias = P + g.b_off;
        if (act != tc::ACT_NONE) { L.scale = P + g.bn_off; L.shift = P + g.bn_off + g.cout; }
        if (up) {
            L.mode = (mode == tc::OUT_BF16_ACCUM) ? tc::OUT_BF16_UP_ACCUM : tc::OUT_BF16_UP;
            L.par_a = l / 2; L.par_b = l % 2;
            L.out_S = ws.geo[ol].S(); L.out_P = ws.geo[ol].P();
            if (ws.dims.resized[ol]) {   // into the dense temporary of the up-sampled size; run_up_resize() brings it to the level
                const tc::Geo& ug = ws.up_geo[ol];
                L.out = ws.up_tmp.p; L.out_ctot = g.cout; L.out_c0 = 0;
                L.out_lo_off = ws.split ? (long long)ug.positions() * g.cout : 0;
                L.out_S = ug.S(); L.out_P = ug.P();
            }
        }
        // algorithmic work: the convolution's FLOPs (the three MMAs of a split chunk are an implementation cost, not work) and one
        // read of the input + one write of the output in their stored widths
        const double px = (double)grid.N * grid.H * grid.W;
        L.flops = 2.0 * px * g.cin * g.cout * s.ntaps;
        const double out_b = (mode == tc::OUT_F32_NHWC || mode == tc::OUT_F32_NCHW) ? 4.0 : (dst_lo_off ? 4.0 : 2.0);
        L.bytes = px * ((ws.split ? 4.0 : 2.0) * g.cin + out_b * g.cout * (up ? 1.0 : 1.0));
    }
}

// after the launches of an `up` group whose join needs the bilinear resize: temporary -> the group's level slice
void run_up_resize(const Plan& plan, const ConvGroup& g, const TcWorkspace& ws, cudaStream_t st)
{
    const int ol = g.out_level;
    if (!(g.is_cont && g.stride == 2) || !ws.dims.resized[ol]) return;
    const tc::Geo& ug = ws.up_geo[ol];
    const tc::View out{ws.level[ol].as<h16>(), plan.level_channels[ol], g.out_c0, g.cout, ws.level_lo[ol]};
    const k::ResizeTab& t = ws.resize.tab[ol];
    const tc::ResizeTabDev td{t.top, t.bottom, t.left, t.right, t.tb, t.itb, t.lr, t.ilr};
    tc::resize_bilinear_nhwc(ws.up_tmp.as<h16>(), ws.split ? (long long)ug.positions() * g.cout : 0, g.cout, ug, out, ws.geo[ol], td, st);
}

void Engine::tc_ensure_infer(int n, int h, int w)
{
    tc_pack_forward_filters();
    TcState& T = *tc_;
    if (T.iw.h == h && T.iw.w == w && T.iw.n == n) return;
    ensure_workspace(T.iw, plan_, n, h, w, true, true, false, stream_);
    i_logits_.reserve((size_t)n * plan_.class_count * h * w * sizeof(float));
    const float* P = params_.as<float>();
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        TcGroup& t = T.groups[gi];
        t.infer.clear();
        if (g.dead) continue;
        int act = tc::ACT_NONE;
        if (g.dlib_bn >= 0) act = g.bn_first ? tc::ACT_AFFINE_MISH : tc::ACT_MISH_AFFINE;
        if (g.out_level == -2)
            setup_forward_launches(plan_, g, t, T.iw, T.fwd_packs.as<h16>(), t.fwd_pack_off, P, t.infer, i_logits_.p, 0, 0, 0, tc::OUT_F32_NCHW, act);
        else
            setup_forward_launches(plan_, g, t, T.iw, T.fwd_packs.as<h16>(), t.fwd_pack_off, P, t.infer, T.iw.level[g.out_level].p,
                                   plan_.level_channels[g.out_level], g.out_c0, 0, tc::OUT_BF16, act);
    }
}

void Engine::tc_input(const uint8_t* img_dev, int IH, int IW, int left, int top, int h, int w, int slot, int n)
{
    tc_ensure_infer(n, h, w);
    tc::Geo one = tc_->iw.geo[0];   // sample `slot` of the batch: positions slot*S ... (tc.h)
    one.N = 1;
    tc::image_to_nhwc(img_dev, IH, IW, left, top, tc_->iw.img.as<h16>() + (size_t)slot * one.S() * 8, one, model_.avg[0], model_.avg[1], model_.avg[2], stream_);
}

void Engine::tc_run_inference(int n, int h, int w)
{
    tc_ensure_infer(n, h, w);
    if (model_.is_training) throw state_error("Forward needs a runtime net (anet_type); convert with GetRuntimeNet / as_runtime");
    TcState& T = *tc_;
    for (size_t gi = 0; gi < plan_.groups.size(); ++gi) {
        const ConvGroup& g = plan_.groups[gi];
        if (g.dead) continue;
        if (g.pool_input)
            tc::maxpool_nhwc(T.iw.level[g.in_level].as<h16>(), plan_.level_channels[g.in_level], g.in_c0, g.cin, T.iw.geo[g.in_level],
                             T.iw.pool.as<h16>() + T.iw.pool_off[gi], T.iw.geo[g.out_level], stream_);
        for (const ConvLaunch& L : T.groups[gi].infer) tc::launch_conv(L, stream_);
        run_up_resize(plan_, g, T.iw, stream_);
    }
    AB_CUDA(cudaGetLastError());
}

}  // namespace ab
