"""annonet_b200: ctypes binding of libannonet_b200.so (the C ABI in include/annonet_b200.h).

The product is the shared library (C++ host code + CUDA sm_100a kernels); this module only loads
it for the tests, bench.py and __graft_entry__.py.  The classes mirror the reference's pimpl
surface (dlib-dnn-pimpl-wrapper/NetPimpl.h:29-102) method for method so that the parity tests read
like code written against the reference.  There is no CPU fallback: if the library is missing the
import fails, and every compute call fails with AB_ERR_CUDA when no device is present.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libannonet_b200.so")

AB_PRECISION_FP32, AB_PRECISION_F16 = 0, 1
(HYPER_LEARNING_RATE, HYPER_WEIGHT_DECAY, HYPER_MOMENTUM, HYPER_BN_STATS_WINDOW, HYPER_LR_SHRINK_FACTOR,
 HYPER_ITERS_WITHOUT_PROGRESS, HYPER_SOLVER, HYPER_ADAM_BETA1, HYPER_ADAM_BETA2, HYPER_SKIP_DEAD_LAYERS, HYPER_LOSS_SCALE) = range(11)

WLABEL = np.dtype([("label", "<u2"), ("weight", "<f4")], align=True)  # == ab_wlabel, 8 bytes
assert WLABEL.itemsize == 8


class AbError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"[ab_status {status}] {message}")
        self.status = status


class SerializationError(AbError):
    pass


class CudaError(AbError):
    pass


class _NetInfo(C.Structure):
    _fields_ = [("levels", C.c_int32), ("class_count", C.c_int32), ("is_training_net", C.c_int32), ("precision", C.c_int32),
                ("num_params", C.c_int64), ("num_layers", C.c_int32), ("device", C.c_int32)]


class _KernelStat(C.Structure):
    _fields_ = [("name", C.c_char * 24), ("launches", C.c_uint64), ("ms", C.c_double), ("flops", C.c_double), ("bytes", C.c_double)]


class Crop(C.Structure):
    """ab_crop (include/annonet_b200.h): one crop of a device-resident dataset image and its augmentation"""
    _fields_ = [("image", C.c_int32), ("left", C.c_int32), ("top", C.c_int32), ("flip_lr", C.c_int32), ("flip_ud", C.c_int32),
                ("brightness", C.c_double), ("noise_level", C.c_int32), ("seed", C.c_uint32), ("color_offset", C.c_int32 * 3)]


class _Tile(C.Structure):
    _fields_ = [("full", C.c_int32 * 4), ("non_overlapping", C.c_int32 * 4), ("unique", C.c_int32 * 4)]


# every symbol include/annonet_b200.h declares (checked by tests/test_abi.py)
EXPORTS = [
    "ab_version", "ab_last_error", "ab_free", "ab_net_create", "ab_net_destroy", "ab_net_clone", "ab_net_info",
    "ab_net_init_random", "ab_net_load_dlib", "ab_net_load_dlib_prefix", "ab_net_save_dlib", "ab_dlib_transcode", "ab_required_input_dimension",
    "ab_recommended_input_dimension", "ab_forward", "ab_process", "ab_get_tiles", "ab_plan_bands", "ab_infer_tiled", "ab_infer_tiled_device",
    "ab_dataset_clear", "ab_dataset_add", "ab_build_crops", "ab_copy_to_host", "ab_train_step_crops", "ab_set_detection_levels", "ab_evaluate_labels", "ab_confusion_per_pixel_noalloc", "ab_render_labels", "ab_get_layer_momentum", "ab_set_layer_momentum", "ab_get_layer_types", "ab_set_hyper", "ab_get_hyper", "ab_train_step_async", "ab_train_step_device",
    "ab_train_sync", "ab_compute_gradients", "ab_get_logits", "ab_get_level_tensor", "ab_get_layer_gradient", "ab_get_layer_params",
    "ab_gradient_arena", "ab_comm_unique_id", "ab_comm_init", "ab_set_weights", "ab_stream", "ab_kernel_launches",
    "ab_device_synchronize", "ab_profile_enable", "ab_profile_read",
]


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C annonet_b200/csrc` (or __graft_entry__.build()); "
                          "there is no fallback path")
    lib = C.CDLL(LIB_PATH)
    lib.ab_version.restype = C.c_char_p
    lib.ab_last_error.restype = C.c_char_p
    lib.ab_stream.restype = C.c_void_p
    lib.ab_stream.argtypes = [C.c_void_p]
    lib.ab_kernel_launches.restype = C.c_uint64
    lib.ab_kernel_launches.argtypes = [C.c_void_p]
    lib.ab_free.argtypes = [C.c_void_p]
    lib.ab_net_destroy.argtypes = [C.c_void_p]
    return lib


lib = _load()


def _check(status: int) -> None:
    if status == 0:
        return
    msg = lib.ab_last_error().decode("utf-8", "replace")
    if status == 2:
        raise SerializationError(status, msg)
    if status == 3:
        raise CudaError(status, msg)
    if status == 1:
        raise ValueError(msg)
    raise AbError(status, msg)


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _gains(gains: Optional[Sequence[float]]):
    if gains is None or len(gains) == 0:
        return None, 0
    g = np.ascontiguousarray(gains, dtype=np.float64)
    return g, len(g)


def required_input_dimension() -> int:
    """TrainingNet::GetRequiredInputDimension (NetPimpl.cpp:137-146)."""
    return int(lib.ab_required_input_dimension())


def recommended_input_dimension(minimum: int) -> int:
    """RuntimeNet::GetRecommendedInputDimension (NetPimpl.cpp:271-285)."""
    out = C.c_int(0)
    status = lib.ab_recommended_input_dimension(C.c_int(minimum), C.byref(out))
    if status == 5:
        raise RuntimeError(lib.ab_last_error().decode())
    _check(status)
    return out.value


def get_tiles(width: int, height: int, max_tile_w: int, max_tile_h: int, overlap_x: int, overlap_y: int):
    """tiling::get_tiles (tiling/tiling.cpp:23-87) -> list of dicts of inclusive (l, t, r, b)."""
    n = C.c_int(0)
    _check(lib.ab_get_tiles(width, height, max_tile_w, max_tile_h, overlap_x, overlap_y, None, 0, C.byref(n)))
    arr = (_Tile * max(n.value, 1))()
    _check(lib.ab_get_tiles(width, height, max_tile_w, max_tile_h, overlap_x, overlap_y, arr, n.value, C.byref(n)))
    return [dict(full=tuple(t.full), non_overlapping=tuple(t.non_overlapping), unique=tuple(t.unique)) for t in arr[:n.value]]


def plan_bands(width: int, height: int, max_tile_w: int, max_tile_h: int, world: int, levels: int = 6):
    """The band plan of a sharded annonet_infer() (engine_infer.cu / tiling.cpp plan_bands): per rank a dict with the tile
    range [t0, t1), the image rows it owns (own0..own1, argmax) and the rows its tiles touch (f0..f1, private canvas)."""
    out = (C.c_int * (6 * world))()
    _check(lib.ab_plan_bands(width, height, max_tile_w, max_tile_h, levels, world, out))
    keys = ("t0", "t1", "own0", "own1", "f0", "f1")
    return [dict(zip(keys, out[6 * r:6 * r + 6])) for r in range(world)]


NOALLOC_MAX_CLASSES = 64


def confusion_per_pixel_noalloc(truth_dev: int, predicted_dev: int, n: int, class_count: int, per_pixel_dev: int, stream: int = 0) -> None:
    """ab_confusion_per_pixel_noalloc: the per-pixel confusion matrix of evaluate_labels() on DEVICE pointers (plain ints
    here), accumulating in the library's static device scratch -- the call allocates nothing.  No counterpart in the
    reference.  Asynchronous on `stream` (a cudaStream_t as int)."""
    _check(lib.ab_confusion_per_pixel_noalloc(C.c_void_p(truth_dev), C.c_void_p(predicted_dev), C.c_longlong(n), C.c_int(class_count),
                                              C.c_void_p(per_pixel_dev), C.c_void_p(stream)))


def set_weights(labels: np.ndarray, class_weight: float, image_weight: float) -> np.ndarray:
    """set_weights (annonet_train.h:20-83): u16 [H,W] -> f32 [H,W]."""
    labels = np.ascontiguousarray(labels, dtype=np.uint16)
    out = np.empty(labels.shape, dtype=np.float32)
    _check(lib.ab_set_weights(_ptr(labels), labels.shape[0], labels.shape[1], C.c_double(class_weight), C.c_double(image_weight), _ptr(out)))
    return out


def transcode(data: bytes, as_runtime: bool = False) -> bytes:
    """Host-only: parse a dlib net stream with the C++ codec and write it back (optionally bn_ -> affine_)."""
    p, n = C.c_void_p(), C.c_size_t(0)
    _check(lib.ab_dlib_transcode(data, C.c_size_t(len(data)), int(as_runtime), C.byref(p), C.byref(n)))
    try:
        return C.string_at(p, n.value)
    finally:
        lib.ab_free(p)


def comm_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    _check(lib.ab_comm_unique_id(buf))
    return buf.raw


def pack_labels(labels: np.ndarray, weights: np.ndarray) -> np.ndarray:
    """u16 labels + f32 weights -> array of weighted_label<uint16_t> (dnn/weighted_label.h:8-20)."""
    out = np.zeros(labels.shape, dtype=WLABEL)
    out["label"] = labels
    out["weight"] = weights
    return out


class _Net:
    def __init__(self, device: int = 0, precision: int = AB_PRECISION_FP32):
        self._h = C.c_void_p()
        _check(lib.ab_net_create(device, precision, C.byref(self._h)))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and lib is not None:  # (module globals are already gone at interpreter shutdown)
            lib.ab_net_destroy(h)

    @property
    def handle(self):
        return self._h

    def info(self) -> dict:
        i = _NetInfo()
        _check(lib.ab_net_info(self._h, C.byref(i)))
        return {k: getattr(i, k) for k, _ in _NetInfo._fields_}

    def _serialize(self, as_runtime: bool) -> bytes:
        p, n = C.c_void_p(), C.c_size_t(0)
        _check(lib.ab_net_save_dlib(self._h, int(as_runtime), C.byref(p), C.byref(n)))
        try:
            return C.string_at(p, n.value)
        finally:
            lib.ab_free(p)

    def Deserialize(self, data) -> None:
        if isinstance(data, str):
            with open(data, "rb") as f:
                data = f.read()
        _check(lib.ab_net_load_dlib(self._h, data, len(data)))

    def stream(self) -> int:
        return int(lib.ab_stream(self._h) or 0)

    def kernel_launches(self) -> int:
        return int(lib.ab_kernel_launches(self._h))

    def synchronize(self) -> None:
        _check(lib.ab_device_synchronize(self._h))

    def layer_params(self, idx: int) -> np.ndarray:
        return self._layer(lib.ab_get_layer_params, idx)

    def _layer(self, fn, idx: int) -> np.ndarray:
        n = C.c_size_t(0)
        _check(fn(self._h, idx, None, C.c_size_t(0), C.byref(n)))
        out = np.empty(n.value, dtype=np.float32)
        _check(fn(self._h, idx, _ptr(out), C.c_size_t(out.size), C.byref(n)))
        return out


class RuntimeNet(_Net):
    """NetPimpl::RuntimeNet (NetPimpl.h:73-102)."""

    def Serialize(self) -> bytes:
        return self._serialize(True)

    def Forward(self, image: np.ndarray) -> np.ndarray:
        """u8 [H,W,3] -> logits f32 [K,H,W] (the host-readable tensor of NetPimpl.cpp:292-298)."""
        image = np.ascontiguousarray(image, dtype=np.uint8)
        h, w, _ = image.shape
        out = np.empty((self.info()["class_count"], h, w), dtype=np.float32)
        _check(lib.ab_forward(self._h, _ptr(image), h, w, _ptr(out)))
        return out

    def Process(self, image: np.ndarray, gains: Optional[Sequence[float]] = None) -> np.ndarray:
        image = np.ascontiguousarray(image, dtype=np.uint8)
        h, w, _ = image.shape
        out = np.empty((h, w), dtype=np.uint16)
        g, ng = _gains(gains)
        _check(lib.ab_process(self._h, _ptr(image), h, w, _ptr(g) if g is not None else None, ng, _ptr(out)))
        return out

    __call__ = Process

    @staticmethod
    def GetRecommendedInputDimension(minimum: int) -> int:
        return recommended_input_dimension(minimum)

    def infer(self, image: np.ndarray, max_tile_w: int, max_tile_h: int, gains: Optional[Sequence[float]] = None,
              return_blended: bool = False, labels_out: Optional[np.ndarray] = None):
        """annonet_infer() (annonet_infer.cpp:32-185) without detection levels.  labels_out: an optional contiguous uint16 [h, w]
        array to receive the label map (e.g. a view of pinned memory)."""
        image = np.ascontiguousarray(image, dtype=np.uint8)
        h, w, _ = image.shape
        if labels_out is not None:
            assert labels_out.dtype == np.uint16 and labels_out.shape == (h, w) and labels_out.flags["C_CONTIGUOUS"]
        labels = labels_out if labels_out is not None else np.empty((h, w), dtype=np.uint16)
        blended = np.empty((self.info()["class_count"], h, w), dtype=np.float32) if return_blended else None
        g, ng = _gains(gains)
        _check(lib.ab_infer_tiled(self._h, _ptr(image), h, w, max_tile_w, max_tile_h, _ptr(g) if g is not None else None, ng,
                                  _ptr(labels), _ptr(blended) if blended is not None else None))
        return (labels, blended) if return_blended else labels

    def evaluate_labels(self, truth: np.ndarray, predicted: np.ndarray, class_count: int):
        """confusion matrices of annonet_infer_main.cpp (per pixel :482-489, per region :202-272) -> two [K, K] uint64 arrays"""
        truth = np.ascontiguousarray(truth, dtype=np.uint16)
        predicted = np.ascontiguousarray(predicted, dtype=np.uint16)
        h, w = truth.shape
        assert predicted.shape == (h, w)
        pp = np.zeros((class_count, class_count), np.uint64)
        pr = np.zeros((class_count, class_count), np.uint64)
        _check(lib.ab_evaluate_labels(self._h, _ptr(truth), _ptr(predicted), h, w, class_count, _ptr(pp), _ptr(pr)))
        return pp, pr

    def render_labels(self, labels: np.ndarray, out_h: int, out_w: int, palette_rgba) -> np.ndarray:
        """resize_label_image + index_label_image_to_rgba_label_image (annonet_infer_main.cpp:408-418) -> rgba8 [out_h, out_w, 4]"""
        labels = np.ascontiguousarray(labels, dtype=np.uint16)
        pal = np.ascontiguousarray(palette_rgba, dtype=np.uint8).reshape(-1, 4)
        out = np.empty((out_h, out_w, 4), np.uint8)
        _check(lib.ab_render_labels(self._h, _ptr(labels), labels.shape[0], labels.shape[1], out_h, out_w, _ptr(pal), pal.shape[0], _ptr(out)))
        return out

    def set_detection_levels(self, levels: Optional[Sequence[float]]) -> None:
        """annonet_infer()'s detection_levels (annonet_infer.cpp:187-239); None / empty switches the blob filter off"""
        a = np.ascontiguousarray(levels if levels is not None else [], dtype=np.float64)
        _check(lib.ab_set_detection_levels(self._h, _ptr(a) if a.size else None, int(a.size)))

    def infer_device(self, image_ptr: int, h: int, w: int, max_tile_w: int, max_tile_h: int, labels_ptr: int) -> None:
        _check(lib.ab_infer_tiled_device(self._h, C.c_void_p(image_ptr), h, w, max_tile_w, max_tile_h, None, 0, C.c_void_p(labels_ptr)))


    def comm_init(self, unique_id: bytes, rank: int, world: int) -> None:
        """join an NCCL communicator: infer / infer_device then shard the tile rows over the ranks"""
        _check(lib.ab_comm_init(self._h, unique_id, rank, world))


class TrainingNet(_Net):
    """NetPimpl::TrainingNet (NetPimpl.h:29-71)."""

    def Initialize(self, class_count: int = 2, width_scaler: float = 1.0, min_filter_count: int = 1, seed: int = 0) -> None:
        """Initialize + SetNetWidth + SetClassCount in the CLI's order (annonet_train_main.cpp:401-405)."""
        _check(lib.ab_net_init_random(self._h, class_count, C.c_double(width_scaler), min_filter_count, C.c_uint64(seed)))

    def _set(self, hid: int, v: float) -> None:
        _check(lib.ab_set_hyper(self._h, hid, C.c_double(v)))

    set_hyper = _set

    def SetLearningRate(self, lr: float) -> None:
        self._set(HYPER_LEARNING_RATE, lr)

    def GetLearningRate(self) -> float:
        self.synchronize()
        v = C.c_double(0)
        _check(lib.ab_get_hyper(self._h, HYPER_LEARNING_RATE, C.byref(v)))
        return v.value

    def SetAllBatchNormalizationRunningStatsWindowSizes(self, window: int) -> None:
        self._set(HYPER_BN_STATS_WINDOW, window)

    def SetLearningRateShrinkFactor(self, f: float) -> None:
        self._set(HYPER_LR_SHRINK_FACTOR, f)

    def SetIterationsWithoutProgressThreshold(self, n: int) -> None:
        self._set(HYPER_ITERS_WITHOUT_PROGRESS, n)

    def SetSolver(self, name: str, weight_decay: float, momentum: float = 0.9, beta2: float = 0.999) -> None:
        self._set(HYPER_SOLVER, {"sgd": 0, "adam": 1}[name])
        self._set(HYPER_WEIGHT_DECAY, weight_decay)
        if name == "sgd":
            self._set(HYPER_MOMENTUM, momentum)
        else:
            self._set(HYPER_ADAM_BETA1, momentum)
            self._set(HYPER_ADAM_BETA2, beta2)

    @staticmethod
    def GetRequiredInputDimension() -> int:
        return required_input_dimension()

    def StartTraining(self, images: np.ndarray, labels: np.ndarray) -> None:
        """images u8 [N,H,W,3]; labels WLABEL [N,H,W].  Asynchronous (depth 1) like dnn_trainer::train_one_step."""
        images = np.ascontiguousarray(images, dtype=np.uint8)
        labels = np.ascontiguousarray(labels, dtype=WLABEL)
        n, h, w, _ = images.shape
        assert labels.shape == (n, h, w)
        _check(lib.ab_train_step_async(self._h, _ptr(images), _ptr(labels), n, h, w))

    def train_step_device(self, images_ptr: int, labels_ptr: int, n: int, h: int, w: int) -> None:
        _check(lib.ab_train_step_device(self._h, C.c_void_p(images_ptr), C.c_void_p(labels_ptr), n, h, w))

    # ---- device-resident dataset + mini-batch assembly (annonet_train_main.cpp:110-232 on the GPU)
    def dataset_clear(self) -> None:
        _check(lib.ab_dataset_clear(self._h))

    def dataset_add(self, rgb: np.ndarray, labels: np.ndarray) -> int:
        rgb = np.ascontiguousarray(rgb, dtype=np.uint8)
        labels = np.ascontiguousarray(labels, dtype=np.uint16)
        h, w, _ = rgb.shape
        assert labels.shape == (h, w)
        idx = C.c_int(0)
        _check(lib.ab_dataset_add(self._h, _ptr(rgb), _ptr(labels), h, w, C.byref(idx)))
        return idx.value

    @staticmethod
    def _crop_array(crops):
        arr = (Crop * len(crops))()
        for i, c in enumerate(crops):
            arr[i].image, arr[i].left, arr[i].top = c["image"], c["left"], c["top"]
            arr[i].flip_lr, arr[i].flip_ud = int(c.get("flip_lr", 0)), int(c.get("flip_ud", 0))
            arr[i].brightness, arr[i].noise_level, arr[i].seed = float(c.get("brightness", 1.0)), int(c.get("noise_level", 0)), int(c.get("seed", 0))
            for q, v in enumerate(c.get("color_offset", (0, 0, 0))):
                arr[i].color_offset[q] = int(v)
        return arr

    def build_crops(self, crops, dim: int, class_weight: float, image_weight: float):
        """the mini-batch as the device wrote it: (u8 [n, dim, dim, 3], WLABEL [n, dim, dim]) copied back for inspection"""
        arr = self._crop_array(crops)
        pi, pl = C.c_void_p(), C.c_void_p()
        _check(lib.ab_build_crops(self._h, arr, len(crops), dim, C.c_double(class_weight), C.c_double(image_weight), C.byref(pi), C.byref(pl)))
        self.synchronize()
        n = len(crops)
        imgs = np.empty((n, dim, dim, 3), np.uint8)
        labs = np.empty((n, dim, dim), WLABEL)
        _check(lib.ab_copy_to_host(self._h, _ptr(imgs), pi, C.c_size_t(imgs.nbytes)))
        _check(lib.ab_copy_to_host(self._h, _ptr(labs), pl, C.c_size_t(labs.nbytes)))
        return imgs, labs

    def train_step_crops(self, crops, dim: int, class_weight: float, image_weight: float) -> None:
        arr = self._crop_array(crops)
        _check(lib.ab_train_step_crops(self._h, arr, len(crops), dim, C.c_double(class_weight), C.c_double(image_weight)))

    def sync(self) -> float:
        loss = C.c_double(0)
        _check(lib.ab_train_sync(self._h, C.byref(loss)))
        return loss.value

    def compute_gradients(self, images: np.ndarray, labels: np.ndarray) -> float:
        images = np.ascontiguousarray(images, dtype=np.uint8)
        labels = np.ascontiguousarray(labels, dtype=WLABEL)
        n, h, w, _ = images.shape
        loss = C.c_double(0)
        _check(lib.ab_compute_gradients(self._h, _ptr(images), _ptr(labels), n, h, w, C.byref(loss)))
        self._shape = (n, h, w)
        return loss.value

    def logits(self) -> np.ndarray:
        n, h, w = self._shape
        out = np.empty((n, self.info()["class_count"], h, w), dtype=np.float32)
        _check(lib.ab_get_logits(self._h, _ptr(out), C.c_size_t(out.size)))
        return out

    def level_tensor(self, level: int, gradient: bool = False) -> np.ndarray:
        dims = (C.c_int32 * 4)()
        _check(lib.ab_get_level_tensor(self._h, level, int(gradient), None, C.c_size_t(0), dims))
        out = np.empty(tuple(dims), dtype=np.float32)
        _check(lib.ab_get_level_tensor(self._h, level, int(gradient), _ptr(out), C.c_size_t(out.size), dims))
        return out

    def layer_gradient(self, idx: int) -> np.ndarray:
        return self._layer(lib.ab_get_layer_gradient, idx)

    def gradient_arena(self):
        p, n = C.c_void_p(), C.c_size_t(0)
        _check(lib.ab_gradient_arena(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def comm_init(self, unique_id: bytes, rank: int, world: int) -> None:
        _check(lib.ab_comm_init(self._h, unique_id, rank, world))

    def Serialize(self) -> bytes:
        return self._serialize(False)

    def GetRuntimeNet(self, device: Optional[int] = None, precision: Optional[int] = None) -> RuntimeNet:
        """RuntimeNet::operator=(TrainingNet) (NetPimpl.cpp:226-232): waits for the step, folds bn_ into affine_."""
        info = self.info()
        rt = RuntimeNet(info["device"] if device is None else device, info["precision"] if precision is None else precision)
        rt.Deserialize(self._serialize(True))
        return rt
