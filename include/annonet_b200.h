/* annonet_b200 -- C ABI of the B200-native engine for annonet's segmentation CNN.
 *
 * This is the drop-in boundary: plain C, caller-owned buffers, int status
 * codes.  Each entry point names the reference interface it replaces; paths are relative to
 * the reference checkout (reunanen/annonet).  The reference-side binding (a replacement
 * NetPimpl.cpp that keeps dlib-dnn-pimpl-wrapper/NetPimpl.h unchanged) is shown in
 * INTEGRATION.md and shipped as shim/NetPimpl_b200.cpp.
 *
 * There is no CPU fallback: every compute entry point needs a CUDA device (sm_100a) and
 * returns AB_ERR_CUDA otherwise.
 */
#ifndef ANNONET_B200_H
#define ANNONET_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ab_net ab_net;

typedef enum {
    AB_OK = 0,
    AB_ERR_ARGUMENT = 1,      /* std::invalid_argument / DLIB_CASSERT in the reference            */
    AB_ERR_SERIALIZATION = 2, /* dlib::serialization_error                                        */
    AB_ERR_CUDA = 3,          /* dlib::cuda_error (incl. out of memory -> process exit code != 0) */
    AB_ERR_STATE = 4,         /* call order violated (e.g. training a runtime net)                */
    AB_ERR_TOO_LARGE = 5      /* std::runtime_error of GetRecommendedInputDimension               */
} ab_status;

/* precision of activations/weights inside the convolutions; accumulation is always fp32 */
typedef enum {
    AB_PRECISION_FP32 = 0, /* fp32 FMA path: logits within 1e-4 relative of the dlib CPU build */
    AB_PRECISION_F16 = 1   /* fp16 tcgen05 path (fp32 accumulate; training forward on hi + lo  */
                           /* fp16 pairs): the north star's 2e-2 "bf16 path" tolerance -- bf16 */
                           /* storage cannot meet it on this net (DESIGN.md "Precision")       */
} ab_precision;

/* dlib::loss_multiclass_log_per_pixel_weighted_::weighted_label<uint16_t>
 * (dlib/dlib/dnn/weighted_label.h:8-20): identical memory layout, 8 bytes. */
typedef struct {
    uint16_t label;
    float weight;
} ab_wlabel;

/* inclusive rectangles, as tiling::dlib_tile (tiling/dlib-wrapper.h:11-16) */
typedef struct {
    int32_t full[4];            /* left, top, right, bottom */
    int32_t non_overlapping[4];
    int32_t unique[4];
} ab_tile;

typedef struct {
    int32_t levels;          /* DLIB_DNN_PIMPL_WRAPPER_LEVEL_COUNT (NetStructure.h:103-108); 6 */
    int32_t class_count;     /* num_filters of layer<1> (NetPimpl.cpp:37-41)                   */
    int32_t is_training_net; /* 1: net_type (bn_), 0: anet_type (affine_)                      */
    int32_t precision;
    int64_t num_params;      /* floats in all parameter tensors                                */
    int32_t num_layers;      /* dlib layer count incl. loss and input: 231                     */
    int32_t device;
} ab_net_info_t;

typedef enum {
    AB_HYPER_LEARNING_RATE = 0,          /* TrainingNet::SetLearningRate                  NetPimpl.cpp:43-46   */
    AB_HYPER_WEIGHT_DECAY = 1,           /* dlib::sgd(weight_decay, momentum)             NetPimpl.h:21-22     */
    AB_HYPER_MOMENTUM = 2,
    AB_HYPER_BN_STATS_WINDOW = 3,        /* SetAllBatchNormalizationRunningStatsWindowSizes NetPimpl.cpp:58-61 */
    AB_HYPER_LR_SHRINK_FACTOR = 4,       /* SetLearningRateShrinkFactor                   NetPimpl.cpp:63-66   */
    AB_HYPER_ITERS_WITHOUT_PROGRESS = 5, /* SetIterationsWithoutProgressThreshold         NetPimpl.cpp:48-51   */
    AB_HYPER_SOLVER = 6,                 /* 0 = sgd (compiled-in default), 1 = adam (NetPimpl.h:17-23)        */
    AB_HYPER_ADAM_BETA1 = 7,
    AB_HYPER_ADAM_BETA2 = 8,
    AB_HYPER_SKIP_DEAD_LAYERS = 9,       /* 1: do not run the 7 never-consumed dense layers in training        */
    AB_HYPER_LOSS_SCALE = 10             /* fp16 path: power of two the loss gradient is multiplied by (0: auto) */
} ab_hyper_id;

/* ---- library ------------------------------------------------------------------------------ */
const char* ab_version(void);
/* message of the last failing call on this thread (the what() of the C++ exception the shim rethrows) */
const char* ab_last_error(void);
void ab_free(void* p);

/* ---- net lifetime: NetPimpl::TrainingNet / RuntimeNet constructors, NetPimpl.cpp:18-26,200-224 */
ab_status ab_net_create(int device, int precision, ab_net** out);
void ab_net_destroy(ab_net* net);
/* RuntimeNet copy constructor / operator= (deep copy of the weights), NetPimpl.cpp:210-224 */
ab_status ab_net_clone(const ab_net* net, ab_net** out);
ab_status ab_net_info(const ab_net* net, ab_net_info_t* info);

/* TrainingNet::Initialize + SetNetWidth + SetClassCount (annonet_train_main.cpp:401-405) followed by
 * the parameter allocation dlib does on the first forward pass (layers.h:170-193, Glorot uniform). */
ab_status ab_net_init_random(ab_net* net, int class_count, double width_scaler, int min_filter_count, uint64_t seed);

/* {Training,Runtime}Net::Deserialize (NetPimpl.cpp:170-198,305-320): a dlib stream of net_type
 * (bn_con2 records) or anet_type (affine_ records); which one is detected from the records. */
ab_status ab_net_load_dlib(ab_net* net, const void* bytes, size_t n);
/* Same for a buffer that continues after the net (dlib streams are not length-prefixed; deserialize(net, istream)
 * leaves the stream right behind the net, dnn/core.h:2683-2692): *consumed = bytes of the net. */
ab_status ab_net_load_dlib_prefix(ab_net* net, const void* bytes, size_t n, size_t* consumed);
/* {Training,Runtime}Net::Serialize.  as_runtime=1 on a training net performs
 * RuntimeNet::operator=(TrainingNet) (NetPimpl.cpp:226-232): bn_ -> affine_ with the running
 * statistics folded in (layers.h:2144-2163).  *bytes is released with ab_free. */
ab_status ab_net_save_dlib(ab_net* net, int as_runtime, void** bytes, size_t* n);

/* Host-only codec check (no device needed): deserialize a dlib net stream and serialize it again,
 * optionally converting net_type -> anet_type like RuntimeNet::operator=(TrainingNet).  A clean()ed
 * reference stream transcodes to itself byte for byte. */
ab_status ab_dlib_transcode(const void* bytes, size_t n, int as_runtime, void** out_bytes, size_t* out_n);

/* ---- sizes: NetPimpl.cpp:137-146,246-285 --------------------------------------------------- */
int ab_required_input_dimension(void);
ab_status ab_recommended_input_dimension(int minimum, int* out);

/* ---- inference ---------------------------------------------------------------------------- */
/* RuntimeNet::Forward (NetPimpl.cpp:292-298): u8 HWC image (channels = 3) -> logits, NCHW f32
 * [class_count][h][w] written to logits_host (may be NULL: result stays on the device). */
ab_status ab_forward(ab_net* net, const uint8_t* image_hwc, int h, int w, float* logits_host);
/* RuntimeNet::operator() / Process (NetPimpl.cpp:287-290): Forward + to_label with gains
 * (loss.h:2921-2969). gains may be NULL (ngains = 0). */
ab_status ab_process(ab_net* net, const uint8_t* image_hwc, int h, int w, const double* gains, int ngains,
                     uint16_t* labels_host);

/* tiling::get_tiles (tiling/tiling.cpp:23-87, dlib-wrapper.h:27-39).  Writes at most cap tiles,
 * *count receives the total. */
ab_status ab_get_tiles(int width, int height, int max_tile_w, int max_tile_h, int overlap_x, int overlap_y,
                       ab_tile* out, int cap, int* count);

/* The band plan of a sharded annonet_infer() (host only; what ab_infer_tiled does after ab_comm_init): the tiles of
 * ab_get_tiles(width, height, ...) with overlap 2^(levels+1)-1 are dealt to `world` ranks as contiguous bands of near-equal
 * cost (a band may begin and end inside a tile row where every band still holds a whole row; AB_BAND_ROWS=1: whole rows only).  out: world x 6 ints {first tile, one past the last tile, first/last image row the rank owns (runs argmax on),
 * first/last image row its tiles touch (extent of its private canvas)}; an empty band has first tile == last tile. */
ab_status ab_plan_bands(int width, int height, int max_tile_w, int max_tile_h, int levels, int world, int* out);

/* annonet_infer() (annonet_infer.cpp:32-185; the detection-level blob filter is ab_filter_detections): tiles ->
 * enlarge to 64w+63 -> clamp-to-edge fill -> forward -> blend -> argmax(+gain), all on the device.
 * labels_host: h*w uint16.  blended_host (optional): class_count*h*w f32 (temp.blended_output).
 * Multi-GPU: after ab_comm_init(net, id, rank, world) every rank makes the SAME call; the tile rows are dealt to
 * the ranks as contiguous bands of equal cost, each rank blends its band into a private canvas spanning only the rows its
 * tiles touch, neighbouring ranks exchange the seam strips (ncclSend/Recv), each rank runs the argmax of the rows it owns
 * and the 2-byte label strips are broadcast: every rank returns the whole label map.  (blended_host then holds this
 * rank's rows only, zero elsewhere.) */
ab_status ab_infer_tiled(ab_net* net, const uint8_t* image_hwc, int h, int w, int max_tile_w, int max_tile_h,
                         const double* gains, int ngains, uint16_t* labels_host, float* blended_host);
/* Evaluation of annonet_infer_main.cpp: confusion_matrix_per_pixel (:482-489) and update_confusion_matrix_per_region
 * (:202-272: connected blobs of the ground truth and of the result, one vote per labelled pixel, a non-background region's
 * prediction is background only if nothing else was predicted in it).  truth / predicted: h*w uint16 (host), 65535 = not
 * labelled.  per_pixel / per_region: class_count^2 counters [truth][predicted]; per_region may be NULL. */
ab_status ab_evaluate_labels(ab_net* net, const uint16_t* truth, const uint16_t* predicted, int h, int w, int class_count,
                             uint64_t* per_pixel, uint64_t* per_region);
/* The per-pixel confusion matrix of ab_evaluate_labels again, for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 * The call allocates nothing -- no cudaMalloc / cudaFree, no allocating library routine: the counters it accumulates
 * into live in a static __device__ array of the library (64 x 64 x 8 bytes = 32 KB, held by every GPU that loads the library
 * until the process exits).  truth_dev / predicted_dev: n uint16 each, DEVICE memory of the current device (65535 = not
 * labelled); per_pixel_dev: class_count^2 uint64 [truth][predicted], DEVICE memory; stream: the cudaStream_t to launch on (NULL
 * = the default stream).  Asynchronous: the result is complete when the stream's work is.
 * Limit: class_count <= AB_NOALLOC_MAX_CLASSES (64); anything larger returns AB_ERR_ARGUMENT with the limit in ab_last_error().
 * The array is shared by every call on that GPU and keeps the last call's contents; each call clears what it uses first, on
 * `stream`.  Calls that overlap on the GPU (other streams, unordered host threads) RACE on it: serialise them.  The first
 * call on a device looks the array's address up (cudaGetSymbolAddress); make it outside a stream capture. */
#define AB_NOALLOC_MAX_CLASSES 64
ab_status ab_confusion_per_pixel_noalloc(const uint16_t* truth_dev, const uint16_t* predicted_dev, long long n, int class_count,
                                         uint64_t* per_pixel_dev, void* stream);
/* The result image of annonet_infer_main.cpp:408-418: resize_label_image (nearest neighbour, annonet.cpp:132-141) to
 * out_h x out_w fused with index_label_image_to_rgba_label_image (:74-86); palette_rgba: class_count x 4 bytes,
 * rgba_out: out_h*out_w*4 bytes (host; the caller writes the PNG). */
ab_status ab_render_labels(ab_net* net, const uint16_t* labels, int h, int w, int out_h, int out_w, const uint8_t* palette_rgba,
                           int class_count, uint8_t* rgba_out);
/* annonet_infer()'s detection_levels argument (annonet_infer.cpp:32-41,187-239; the CLI's -d option): with any level > 0
 * the label map of every following ab_infer_tiled* call goes through the blob filter -- connected blobs of equal non-zero
 * labels (dlib::label_connected_blobs, 8-neighbourhood) that contain no pixel with
 * logit[label] - logit[0] > level[label] - level[0] are set to 0.  n = 0 switches it off.
 * The reference looks a seed found at pixel (r, c) up at pixel (row c, col r) (dlib::point(r, c) is x = r, y = c;
 * annonet_infer.cpp:210,221); the engine mirrors that by default so that label maps match the reference's, and does the
 * same-pixel lookup with the environment variable AB_DETECTION_SEEDS=intended. */
ab_status ab_set_detection_levels(ab_net* net, const double* levels, int n);
/* Same with image and labels resident in device memory (image_dev: h*w*3 u8, labels_dev: h*w u16). */
ab_status ab_infer_tiled_device(ab_net* net, const void* image_dev, int h, int w, int max_tile_w, int max_tile_h,
                                const double* gains, int ngains, void* labels_dev);

/* ---- training ----------------------------------------------------------------------------- */
ab_status ab_set_hyper(ab_net* net, int id, double value);
ab_status ab_get_hyper(ab_net* net, int id, double* value);

/* TrainingNet::StartTraining (NetPimpl.cpp:148-151 -> dnn_trainer::train_one_step,
 * trainer.h:197-224): n crops of h x w, images n*h*w*3 u8, labels n*h*w ab_wlabel.  Returns once
 * the inputs are staged and the step is enqueued on the net's stream (depth-1 asynchrony). */
ab_status ab_train_step_async(ab_net* net, const uint8_t* images, const ab_wlabel* labels, int n, int h, int w);
/* Same with inputs already in device memory. */
ab_status ab_train_step_device(ab_net* net, const void* images_dev, const void* labels_dev, int n, int h, int w);
/* Blocks until the enqueued step is finished (dnn_trainer::get_net / wait_for_thread_to_pause,
 * trainer.h:143-151,882-885); *loss = loss of the last step (trainer.h:751-754). */
ab_status ab_train_sync(ab_net* net, double* loss);
/* add_loss_layer::compute_parameter_gradients (core.h:2570-2580): forward + loss + backward, no update. */
ab_status ab_compute_gradients(ab_net* net, const uint8_t* images, const ab_wlabel* labels, int n, int h, int w,
                               double* loss);
/* training-mode logits of the last step, NCHW f32 [n][class_count][h][w] */
ab_status ab_get_logits(ab_net* net, float* out, size_t capacity_floats);
/* every tensor living at one resolution level after the last training step -- the outputs (gradient = 0) or the
 * gradient inputs (gradient = 1) dlib exposes as layer<i>(net).get_output() / get_gradient_input() (core.h:1021-1033) --
 * as channel slices of one NCHW f32 block [n][level channels][h][w]; dims receives {n, c, h, w}; out may be NULL. */
ab_status ab_get_level_tensor(ab_net* net, int level, int gradient, float* out, size_t capacity_floats, int32_t dims[4]);
/* parameter gradients of one dlib layer<idx> in dlib's layout (visit_layer_parameter_gradients, core.h:3806-3816) */
ab_status ab_get_layer_gradient(ab_net* net, int dlib_layer_index, float* out, size_t capacity_floats, size_t* count);
ab_status ab_get_layer_params(ab_net* net, int dlib_layer_index, float* out, size_t capacity_floats, size_t* count);
/* The solver state of one layer: dlib::sgd's momentum tensor v (solvers.h:121-127, "sgd2") in dlib's layout -- what a
 * dnn_trainer synchronization file stores per computational layer (trainer.h:931-964: serialize(devices[0]->solvers)). */
ab_status ab_get_layer_momentum(ab_net* net, int dlib_layer_index, float* out, size_t capacity_floats, size_t* count);
ab_status ab_set_layer_momentum(ab_net* net, int dlib_layer_index, const float* v, size_t count);
/* kind of every dlib layer<i>, i = 0 (loss) .. num_layers-1 (input): the static layer list of NetStructure.h:172-299 as
 * ab_layer_kind values; out may be NULL */
typedef enum ab_layer_kind {
    AB_LAYER_LOSS = 0, AB_LAYER_CON, AB_LAYER_CONT, AB_LAYER_BN, AB_LAYER_AFFINE, AB_LAYER_MISH, AB_LAYER_MAX_POOL,
    AB_LAYER_CONCAT_PREV, AB_LAYER_CONCAT, AB_LAYER_TAG, AB_LAYER_INPUT
} ab_layer_kind;
ab_status ab_get_layer_types(ab_net* net, int32_t* kinds, int capacity, int* count);

/* ---- device-resident dataset and mini-batch assembly ---------------------------------------------
 * The reference's loader threads cut every crop on the CPU (annonet_train_main.cpp:110-232,520-553: extract_image_chip,
 * outpaint, set_weights, flips, brightness, noise, colour offset) and the trainer copies it to the GPU.  Here the full images
 * and label maps are uploaded once; per step the host sends only the random decisions and the crops and their weighted labels
 * are written by kernels straight into the step's input buffers.  Crop windows may leave the image: pixels outside repeat
 * the nearest image pixel (outpaint, annonet.h:74-120), labels outside are ignored (65535, weight 0).  Weights are
 * set_weights(class_weight, image_weight) of each crop (annonet_train.h:20-83; class ids < 256 on this path).  Order of the
 * augmentations as in the reference: crop -> weights -> flips -> brightness -> noise -> colour offset.  Noise is a
 * reproducible function of (seed, pixel, channel), not dlib::rand's stream. */
typedef struct ab_crop {
    int32_t image;            /* index returned by ab_dataset_add */
    int32_t left, top;        /* crop window [left, left + dim) x [top, top + dim) in image coordinates */
    int32_t flip_lr, flip_ud;
    double brightness;        /* multiplicative brightness change, 1.0 = none (annonet_train_main.cpp:198-221) */
    int32_t noise_level;      /* uniform integer noise in [-level, level] per channel, 0 = none (:73-101) */
    uint32_t seed;
    int32_t color_offset[3];  /* per-channel offsets of dlib::apply_random_color_offset, 0 = none */
} ab_crop;
ab_status ab_dataset_clear(ab_net* net);
ab_status ab_dataset_add(ab_net* net, const uint8_t* rgb_hwc, const uint16_t* labels, int h, int w, int* index);
/* device memory handed out by the net (ab_build_crops) -> host, after the net's stream has drained */
ab_status ab_copy_to_host(ab_net* net, void* dst_host, const void* src_dev, size_t bytes);
/* crops -> u8 [n][dim][dim][3] and ab_wlabel [n][dim][dim] in device memory owned by the net (valid until the next call) */
ab_status ab_build_crops(ab_net* net, const ab_crop* crops, int n, int dim, double class_weight, double image_weight,
                         const void** imgs_dev, const void** labels_dev);
/* ab_build_crops + ab_train_step_device: one asynchronous training step on a mini-batch cut on the device */
ab_status ab_train_step_crops(ab_net* net, const ab_crop* crops, int n, int dim, double class_weight, double image_weight);

/* data-parallel training: flat gradient arena for the caller's/engine's all-reduce
 * (multi_device_tensor_averager, tensor_tools.h:1609-1736 -> NCCL) */
ab_status ab_gradient_arena(ab_net* net, void** grads_dev, size_t* count_floats);
/* NCCL communicator owned by the engine: id = 128-byte ncclUniqueId produced on rank 0 */
ab_status ab_comm_unique_id(void* id128);
ab_status ab_comm_init(ab_net* net, const void* id128, int rank, int world);

/* set_weights (annonet_train.h:20-83): per-pixel loss weights from a u16 label crop (host function) */
ab_status ab_set_weights(const uint16_t* labels, int h, int w, double class_weight, double image_weight, float* weights_out);

/* ---- instrumentation ---------------------------------------------------------------------- */
/* Device time per kernel family, measured with CUDA event pairs on the engine's stream around every
 * launch while enabled (bench.py's roofline figures).  flops/bytes are the ALGORITHMIC work of the
 * bracketed launches (DESIGN.md "Algorithmic work").  Process-wide; ab_profile_enable(1) restarts. */
typedef struct {
    char name[24];
    uint64_t launches;
    double ms, flops, bytes;
} ab_kernel_stat;
ab_status ab_profile_enable(int on);
ab_status ab_profile_read(ab_kernel_stat* out, int cap, int* count);
void* ab_stream(ab_net* net);              /* cudaStream_t the engine launches on */
uint64_t ab_kernel_launches(ab_net* net);   /* kernels of this library launched so far on this net */
ab_status ab_device_synchronize(ab_net* net);

#ifdef __cplusplus
}
#endif
#endif /* ANNONET_B200_H */
