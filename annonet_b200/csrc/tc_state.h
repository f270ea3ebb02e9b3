// State of the fp16 tensor-core path of one Engine (AB_PRECISION_F16): padded flat NHWC workspaces,
// fp16 filter packs and the prepared launch descriptors (tensor maps) of every conv group.
#pragma once
#include "engine.h"
#include "tc.h"

#include <vector>

namespace ab {

struct TcGroup {
    // forward filters: one pack per launch (1, or the 4 parity classes of a stride-2 cont_)
    size_t fwd_pack_off[4] = {0, 0, 0, 0};
    int fwd_ntaps[4] = {0, 0, 0, 0};
    int fwd_tapmap[4][9] = {};
    int fwd_tapdy[4][9] = {}, fwd_tapdx[4][9] = {};  // tap offset = dy*P + dx of the M grid
    int nlaunch = 1;
    int npad = 0;
    size_t train_pack_off[4] = {0, 0, 0, 0};   // training forward: split packs (hi and lo filters along N: 2 x the rows), own buffer
    std::vector<tc::ConvLaunch> infer;   // prepared for the current inference geometry
    // training
    std::vector<tc::ConvLaunch> fwd;     // conv -> z (or destination)
    size_t bwd_pack_off = 0;             // backward-data filters [tap][ci][co]
    int bwd_npad = 0, bwd_ntaps = 0;
    int bwd_tapmap[9] = {}, bwd_tapdy[9] = {}, bwd_tapdx[9] = {};
    std::vector<tc::ConvLaunch> dgrad;
    tc::WgradJob wgrad;
    tc::WgradTcLaunch wgrad_tc;          // the same job prepared for the tcgen05 kernel
    size_t z_off = 0, dpool_off = 0, upx_off = 0;   // elements into TcState::z / dpool / upx
};

struct TcWorkspace {
    int n = 0, h = 0, w = 0;
    // split (training activations): every tensor is stored twice, hi halves then lo halves (tc.h); the lo half of a level / the
    // image starts level_lo[l] / img_lo elements behind the hi half
    bool split = false;
    tc::Geo geo[8];
    DevBuf level[8];           // h16 [positions][level_channels]  (x 2 when split)
    DevBuf img;                // h16 [positions][8]
    DevBuf pool;               // pooled inputs of the transition_down groups, dense
    long long level_lo[8] = {0}, img_lo = 0;   // (a pooled tensor's lo half: positions * cin elements behind its hi half)
    DevBuf pool_arg;           // training: u8 index of the winning window element, same element offsets as `pool`
    // non-canonical input sizes: a join whose up-sampled tensor comes back one pixel short (plan.h LevelDims::resized) gets it
    // through this dense temporary [positions of up_geo][cout] (+ lo half when split) and the bilinear resize kernel
    LevelDims dims;
    tc::Geo up_geo[8];
    DevBuf up_tmp;
    ResizeTables resize;
    std::vector<size_t> pool_off;  // per group, in elements
};

struct TcState {
    std::vector<TcGroup> groups;
    DevBuf fwd_packs, bwd_packs, train_packs;
    bool fwd_packs_valid = false, bwd_packs_valid = false;
    float loss_scale = 1.f;    // power of two folded into the loss gradient; parameter gradients are divided by it again
    TcWorkspace iw;            // inference
    // training
    TcWorkspace tw, tg;        // activations / gradients
    DevBuf z;                  // per bn group: conv output z, fp32 [positions][cout]
    DevBuf dz;                 // per bn group: gradient at the conv output, h16 [positions][cout] (same element offsets as z)
    DevBuf dpool;              // per transition_down group: gradient at the pooled tensor              [positions][cin]
    DevBuf upx;                // per `up` group: zero-stuffed input on the fine grid (filter gradient)  [positions][cin]
    DevBuf glogits16;          // loss gradient as h16 [positions][8]
    size_t sums_slot = 0;   // doubles per (layer, pass) slot of `sums`
    DevBuf sums, sums_side, bias_tmp, wscratch, pack_jobs;   // sums_side: batch-statistics scratch of the side stream (dead layers)
    int n_pack_jobs = 0, n_fwd_jobs = 0, max_pack_elems = 0;
    bool train_ready = false;
    std::vector<char> grad_level_needs_zero;   // per level: some writer accumulates into channels nobody assigned before it
    // CUDA graph of one whole training step (static shapes and pointers): replayed while the key stays the same
    struct GraphKey {
        const void* imgs = nullptr; const void* labels = nullptr;
        int n = 0, h = 0, w = 0, update = 0, skip_dead = 0, world = 1, solver = 0;
        double lr = 0, wd = 0, momentum = 0, loss_scale = 0;
        unsigned long long window = 0;
        bool operator==(const GraphKey& o) const
        {
            return imgs == o.imgs && labels == o.labels && n == o.n && h == o.h && w == o.w && update == o.update && skip_dead == o.skip_dead &&
                   world == o.world && solver == o.solver && loss_scale == o.loss_scale && lr == o.lr && wd == o.wd && momentum == o.momentum && window == o.window;
        }
    };
    GraphKey graph_key, seen_key;
    // Data parallel: the step is captured as SEGMENTS split at the gradient-bucket boundaries; the NCCL all-reduce of bucket i
    // is issued eagerly on the communication stream between segment i and i+1 (NCCL inside a capture hung on the 2-GPU box).
    // One process: a single segment.
    std::vector<cudaGraphExec_t> graph_segs;
    std::vector<std::pair<size_t, size_t>> graph_buckets;   // arena range all-reduced after segment i (segs.size() - 1 entries)
    unsigned long long graph_kernels = 0;   // kernels inside the graph (for ab_kernel_launches)
    void drop_graph() { for (cudaGraphExec_t e : graph_segs) cudaGraphExecDestroy(e); graph_segs.clear(); graph_buckets.clear(); }
    ~TcState() { drop_graph(); }
};

// prepared launches of one conv group in the forward direction (engine_bf16.cu)
// packs / pack_off: the group's filter packs (split packs when ws.split); dst_lo_off: element offset of the destination's lo half
void setup_forward_launches(const Plan& plan, const ConvGroup& g, const TcGroup& t, const TcWorkspace& ws, const tc::h16* packs,
                            const size_t* pack_off, const float* P, std::vector<tc::ConvLaunch>& out, void* dst, int dst_ctot, int dst_c0,
                            long long dst_lo_off, int mode, int act);
void run_up_resize(const Plan& plan, const ConvGroup& g, const TcWorkspace& ws, cudaStream_t st);
void ensure_workspace(TcWorkspace& ws, const Plan& plan, int n, int h, int w, bool with_image, bool with_pool, bool split, cudaStream_t st);


}  // namespace ab
