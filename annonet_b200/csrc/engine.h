// The engine behind the C ABI: owns the device-resident parameters of one net (net_type or
// anet_type), the static plan, and the activation workspaces; sequences the kernels of
//   * RuntimeNet::Forward / annonet_infer()            (dnn/core.h:932-948, annonet_infer.cpp:32-185)
//   * dnn_trainer's step: forward, loss, backward, update (trainer.h:631-671,743-828; core.h:998-1023)
// on one CUDA stream (plus a side stream for the NCCL gradient all-reduce).
#pragma once
#include "kernels.h"
#include "net_model.h"
#include "plan.h"
#include "tc.h"

#include <cuda_runtime.h>
#include <memory>
#include <string>
#include <vector>

namespace ab {

struct cuda_error : std::runtime_error { using std::runtime_error::runtime_error; };
struct state_error : std::runtime_error { using std::runtime_error::runtime_error; };

void cuda_check(cudaError_t e, const char* what, const char* file, int line);
#define AB_CUDA(x) ::ab::cuda_check((x), #x, __FILE__, __LINE__)

// cudaMalloc'ed block with RAII
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }
    void release();
    void reserve(size_t n);  // grows (never shrinks); contents are lost on growth
    template <typename T> T* as() const { return (T*)p; }
};

struct Hyper {
    double learning_rate = 1e-2;   // dnn_trainer default (trainer.h:78)
    double weight_decay = 0.001;   // NetPimpl.h:22
    double momentum = 0.9;
    double lr_shrink_factor = 0.1;
    double iters_without_progress = 2000;
    int solver = 0;                // 0 sgd, 1 adam
    double adam_beta1 = 0.9, adam_beta2 = 0.999;
    int skip_dead_layers = 0;
    double loss_scale = 0;         // fp16 path: 0 = automatic (a power of two near n*h*w/4)
};

// device tables of the bilinear resizes a geometry needs (one per resized join; plan.h resize_axis)
struct ResizeTables {
    DevBuf buf;
    k::ResizeTab tab[8];
    void build(const LevelDims& d, int levels, cudaStream_t st);
};

struct NcclApi;
struct TcState;  // h16 tensor-core path: workspaces, packed filters, launch descriptors (engine_bf16.cu)

class Engine {
public:
    Engine(int device, int precision);
    ~Engine();

    void load(const NetModel& m);
    NetModel snapshot(bool as_runtime);
    bool loaded() const { return loaded_; }
    bool is_training() const { return model_.is_training; }
    const Plan& plan() const { return plan_; }
    const NetModel& model() const { return model_; }
    int device() const { return device_; }
    int precision() const { return precision_; }
    cudaStream_t stream() const { return stream_; }
    void synchronize();

    // ---- inference (runtime nets) ----
    // one image already on the device as u8 HWC; returns device logits [K][h][w] (valid until the next call)
    const float* forward_device(const uint8_t* img_dev, int h, int w);
    const float* forward_host(const uint8_t* img_host, int h, int w);
    void process_host(const uint8_t* img_host, int h, int w, const double* gains, int ngains, uint16_t* labels_host);
    // annonet_infer(); image/labels on the device.  After comm_init the tile rows are sharded over the ranks as contiguous bands,
    // seams are exchanged with the neighbours and every rank ends up with the whole label map (engine_infer.cu)
    void infer_tiled_device(const uint8_t* img_dev, int h, int w, int max_tile_w, int max_tile_h, const double* gains, int ngains,
                            uint16_t* labels_dev);
    // annonet_infer_main.cpp's evaluation (confusion matrices :482-491, per region :202-272) and result image (:408-418); host buffers
    void evaluate_labels(const uint16_t* truth, const uint16_t* predicted, int h, int w, int K, unsigned long long* per_pixel,
                         unsigned long long* per_region);
    void render_labels(const uint16_t* labels, int h, int w, int out_h, int out_w, const uint8_t* palette_rgba, int K, uint8_t* rgba_out);
    void set_detection_levels(const double* levels, int n) { detection_levels_.assign(levels, levels + (levels ? n : 0)); }
    void infer_tiled_host(const uint8_t* img_host, int h, int w, int max_tile_w, int max_tile_h, const double* gains, int ngains,
                          uint16_t* labels_host, float* blended_host);

    // ---- training (training nets) ----
    Hyper hyper;
    void set_bn_window(unsigned long long w);
    void train_step_device(const uint8_t* imgs_dev, const void* labels_dev, int n, int h, int w, bool update);
    void train_step_host(const uint8_t* imgs, const void* labels, int n, int h, int w, bool update);
    double train_sync();  // waits; returns the loss of the last step
    const float* train_logits() const { return t_logits_.as<float>(); }
    int last_n() const { return tn_; } int last_h() const { return th_; } int last_w() const { return tw_; }
    float* grads() const { return grads_.as<float>(); }
    float* params() const { return params_.as<float>(); }
    void layer_values(int dlib_layer, int which, std::vector<float>& out);  // dlib layout; which: 0 parameters, 1 gradients, 2 sgd momentum
    void set_layer_momentum(int dlib_layer, const float* v, size_t n);        // the solver state of a dnn_trainer sync file
    // all tensors of one resolution level after the last training step (activations or their gradients), NCHW fp32
    void level_values(int level, bool gradient, std::vector<float>& out, int dims[4]);

    // ---- device-resident dataset + mini-batch assembly (kernels_crop.cu; annonet_train_main.cpp:110-232,520-553)
    struct CropSpec { int image, left, top, flip_lr, flip_ud; double brightness; int noise_level; unsigned seed; int color_offset[3]; };
    void dataset_clear();
    int dataset_add(const uint8_t* rgb, const uint16_t* labels, int h, int w);
    // crops + weighted labels into the step's input buffers (returned), ready for train_step_device
    void build_crops(const CropSpec* crops, int n, int dim, double class_weight, double image_weight, const uint8_t** imgs_dev, const void** labels_dev);

    void comm_init(const void* id128, int rank, int world);
    int world() const { return world_; }

private:
    // workspaces
    struct Levels { DevBuf buf[8]; LevelDims dims; ResizeTables resize; int n = 0, h = 0, w = 0; };
    void ensure_infer_ws(int h, int w);
    void ensure_train_ws(int n, int h, int w);
    void run_inference_graph(int h, int w);
    void forward_train(int n, int h, int w);
    void backward_train(int n, int h, int w);
    void optimizer_step();
    void ensure_adam_state();
    void allreduce_range(size_t lo, size_t hi);
    k::Slice level_slice(Levels& L, int level, int c0, int C) const;
    // h16 tensor-core path (precision_ == AB_PRECISION_F16)
    // n: tiles of equal size run as one batch (the coarse levels of a single tile are a handful of CTAs per launch)
    void tc_ensure_infer(int n, int h, int w);
    void tc_input(const uint8_t* img_dev, int IH, int IW, int left, int top, int h, int w, int slot = 0, int n = 1);
    void tc_run_inference(int n, int h, int w);
    void tc_pack_forward_filters();
    void tc_ensure_train(int n, int h, int w);
    void tc_train_step(const uint8_t* imgs_dev, const void* labels_dev, int n, int h, int w, bool update);
    void tc_train_body(const uint8_t* imgs_dev, const void* labels_dev, int n, int h, int w, bool update, bool capture);
    void tc_close_segment();
    std::unique_ptr<TcState> tc_;

    int device_, precision_, requested_precision_;   // precision_: what the loaded net actually runs on (load() may fall back to fp32)
    cudaStream_t stream_ = nullptr, comm_stream_ = nullptr, copy_stream_ = nullptr;
    // side stream of the training step: work that is off the critical path of the layer chain (filter gradients, the never
    // consumed "dead" layers) runs there, next to the small launches of the coarse levels that leave most SMs idle
    cudaStream_t side_stream_ = nullptr;
    cudaEvent_t ev_fork_[32] = {}, ev_join_ = nullptr;
    int fork_slot_ = 0;
    cudaEvent_t ev_bucket_[8] = {}, ev_comm_done_ = nullptr, ev_stage_ = nullptr, ev_unstaged_ = nullptr;
    bool loaded_ = false;
    NetModel model_;  // host copy of the structure (+ parameters as loaded; device is authoritative after load)
    Plan plan_;

    DevBuf params_, grads_, mom_, adam_v_, lr_mul_, wd_mul_, running_;  // running_: per bn group [mean(C) | var(C)]
    std::vector<size_t> running_off_;   // per group
    std::vector<unsigned long long> num_updates_, window_;  // per group; num_updates_ is a host mirror of nu_dev_ (load/snapshot)
    DevBuf nu_dev_;                     // per group: bn_ num_updates counter, advanced on the device by bn_finalize
    unsigned long long adam_t_ = 0;     // host mirror of the step count (the kernels read the device copy)
    DevBuf adam_t_dev_;
    int bucket_slot_ = 0;

    // inference workspace
    Levels iw_;
    struct DatasetImage { std::unique_ptr<DevBuf> rgb, labels; int h = 0, w = 0; };
    std::vector<DatasetImage> dataset_;
    DevBuf crop_jobs_, crop_scratch_, crop_overflow_;
    void* crop_jobs_host_ = nullptr; size_t crop_jobs_host_bytes_ = 0;   // pinned
    DevBuf e_truth_, e_pred_, e_roots_, e_counts_, e_rgba_;   // evaluation / write-back scratch
    DevBuf i_up_, t_up_;   // the up-sampled tensor of a join that needs the bilinear resize (non-canonical input sizes)
    DevBuf i_x_, i_logits_, i_img_, i_labels_, canvas_, gains_dev_, seam_, det_dev_, seeds_, blob_parent_, blob_keep_;
    std::vector<double> detection_levels_;   // annonet_infer()'s detection_levels; empty or all zero: no blob filter
    int canvas_rows_ = 0, canvas_row0_ = 0;   // the blend canvas holds image rows [canvas_row0_, canvas_row0_ + canvas_rows_)
    void* pinned_ = nullptr; size_t pinned_bytes_ = 0;
    void* pinned(size_t bytes);

    // training workspace
    Levels tw_act_, tw_grad_;
    DevBuf t_x_, t_logits_, t_glogits_, t_z_, t_pool_, t_dz_, t_dpool_, t_dead_, t_stats_, t_sums_, t_scratch_, t_imgs_, t_labels_, t_stage_dev_,
        t_loss_;
    std::vector<size_t> z_off_, pool_off_, stats_off_;
    int tn_ = 0, th_ = 0, tw_ = 0;
    bool stats_valid_ = false;
    double* loss_host_ = nullptr;  // pinned

    // data parallel
    std::unique_ptr<NcclApi> nccl_;
    void* comm_ = nullptr;
    int rank_ = 0, world_ = 1;
};

}  // namespace ab
