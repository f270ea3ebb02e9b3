"""TEST INFRASTRUCTURE ONLY (oracle).  Never imported by the product path.

numpy restatement of the reference's CPU algorithm for the hot path:
forward / backward / loss / argmax / SGD of the segmentation net plus tiling, blend and
set_weights.  It interprets the reference's OWN layer list (oracle/dlib_format.net_layers,
i.e. NetStructure.h) with dlib's executor semantics (tags, concat_prev, concat), so it is
independent of the engine's concat-free channel plan.  Pinned against the compiled reference
(oracle/_ref/annonet_ref) by tests/test_oracle_vs_ref.py and the fixtures in tests/golden/.

Each function cites the reference code it restates (paths under /root/reference).
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from dlib_format import Net

F32 = np.float32
LABEL_TO_IGNORE = 65535

# ----------------------------------------------------------------------------------------------
# elementwise / layer ops


def to_tensor(images: np.ndarray, avg: Sequence[float]) -> np.ndarray:
    """input_rgb_image::to_tensor, dlib/dlib/dnn/input.h:68-117: (p - avg)/256, NCHW f32."""
    assert images.dtype == np.uint8 and images.ndim == 4 and images.shape[3] == 3
    x = images.astype(F32) - np.asarray(avg, dtype=F32).reshape(1, 1, 1, 3)
    x = (x.astype(np.float64) / 256.0).astype(F32)
    return np.ascontiguousarray(x.transpose(0, 3, 1, 2))


def mish(x: np.ndarray) -> np.ndarray:
    """cpu::mish, dlib/dlib/cuda/cpu_dlib.cpp:1510-1523 (fp32 exp, overflow -> y = x)."""
    with np.errstate(over="ignore", invalid="ignore"):
        e = np.exp(x.astype(F32))
        delta = F32(2) * e + e * e + F32(2)
        return (x - F32(2) * x / delta).astype(F32)


def mish_grad(x: np.ndarray) -> np.ndarray:
    """cpu::mish_gradient, cpu_dlib.cpp:1525-1558."""
    with np.errstate(over="ignore", invalid="ignore"):
        xc = np.clip(x, -8, 8).astype(F32)
        e = np.exp(xc)
        delta = F32(2) * e + e * e + F32(2)
        omega = F32(4) * (xc + F32(1)) + F32(4) * e * e + e * e * e + e * (F32(4) * xc + F32(6))
        g = e * omega / (delta * delta)
    g = np.where(x >= 8, F32(1), g)
    g = np.where(x <= -8, F32(0), g)
    return g.astype(F32)


def _im2col(x: np.ndarray, k: int, pad: int) -> np.ndarray:
    """cpu::img2col column order (cin, ky, kx), cpu_dlib.cpp:2057-2202; stride 1.
    Returns [N, Cin*k*k, Hout*Wout]."""
    n, c, h, w = x.shape
    xp = np.pad(x, ((0, 0), (0, 0), (pad, pad), (pad, pad))) if pad else x
    win = np.lib.stride_tricks.sliding_window_view(xp, (k, k), axis=(2, 3))  # N,C,Ho,Wo,k,k
    ho, wo = win.shape[2], win.shape[3]
    return np.ascontiguousarray(win.transpose(0, 1, 4, 5, 2, 3)).reshape(n, c * k * k, ho * wo), ho, wo


def conv2d(x: np.ndarray, filt: np.ndarray, pad: int) -> np.ndarray:
    """tensor_conv::operator(), cpu_dlib.cpp:2250-2296: cross-correlation, stride 1, zero pad."""
    cout, cin, k, _ = filt.shape
    cols, ho, wo = _im2col(x, k, pad)
    out = np.matmul(filt.reshape(cout, -1), cols)
    return out.reshape(x.shape[0], cout, ho, wo).astype(F32)


def conv2d_strided(x: np.ndarray, filt: np.ndarray, stride: int, pad: int) -> np.ndarray:
    """Same with a stride: only used as the backward of the stride-2 cont_ (layers.h:542)."""
    full = conv2d(x, filt, pad)
    return np.ascontiguousarray(full[:, :, ::stride, ::stride])


def conv2d_bwd_data(gy: np.ndarray, filt: np.ndarray, stride: int, pad: int, out_hw: Tuple[int, int]) -> np.ndarray:
    """tensor_conv::get_gradient_for_data, cpu_dlib.cpp:2298-2325 (col2img scatter):
    dx[ci, y*s+ky-p, x*s+kx-p] += gy[co, y, x] * filt[co, ci, ky, kx]."""
    n, cout, hy, wy = gy.shape
    _, cin, k, _ = filt.shape
    h, w = out_hw
    dx = np.zeros((n, cin, h + 2 * pad, w + 2 * pad), dtype=F32)
    g2 = gy.reshape(n, cout, hy * wy)
    for ky in range(k):
        for kx in range(k):
            contrib = np.matmul(filt[:, :, ky, kx].T, g2).reshape(n, cin, hy, wy)
            dx[:, :, ky:ky + stride * (hy - 1) + 1:stride, kx:kx + stride * (wy - 1) + 1:stride] += contrib
    if pad:
        dx = dx[:, :, pad:pad + h, pad:pad + w]
    return np.ascontiguousarray(dx).astype(F32)


def conv2d_bwd_filter(gy: np.ndarray, x: np.ndarray, k: int, stride: int, pad: int) -> np.ndarray:
    """tensor_conv::get_gradient_for_filters, cpu_dlib.cpp:2327-2355:
    dW[co, ci, ky, kx] = sum_{n,y,x} gy[n,co,y,x] * xpad[n,ci,y*s+ky,x*s+kx]."""
    n, cout, hy, wy = gy.shape
    cin = x.shape[1]
    xp = np.pad(x, ((0, 0), (0, 0), (pad, pad), (pad, pad))) if pad else x
    dw = np.zeros((cout, cin, k, k), dtype=np.float64)
    g2 = gy.reshape(n, cout, hy * wy)
    for ky in range(k):
        for kx in range(k):
            xs = xp[:, :, ky:ky + stride * (hy - 1) + 1:stride, kx:kx + stride * (wy - 1) + 1:stride].reshape(n, cin, hy * wy)
            dw[:, :, ky, kx] = np.einsum("nop,ncp->oc", g2, xs, optimize=True)
    return dw.astype(F32)


def max_pool_3x3s2(x: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """pooling::operator() for max_pool<3,3,2,2> pad 0, cpu_dlib.cpp:1888-1943.
    Returns (out, argmax index 0..8 of the FIRST maximum in row-major window order, :2011)."""
    n, c, h, w = x.shape
    ho, wo = 1 + (h - 3) // 2, 1 + (w - 3) // 2
    win = np.lib.stride_tricks.sliding_window_view(x, (3, 3), axis=(2, 3))[:, :, ::2, ::2][:, :, :ho, :wo]
    flat = win.reshape(n, c, ho, wo, 9)
    return flat.max(axis=-1).astype(F32), flat.argmax(axis=-1)


def max_pool_bwd(gy: np.ndarray, arg: np.ndarray, in_hw: Tuple[int, int]) -> np.ndarray:
    """pooling::get_gradient (max), cpu_dlib.cpp:1971-2017."""
    n, c, ho, wo = gy.shape
    h, w = in_hw
    dx = np.zeros((n, c, h, w), dtype=F32)
    ys = (2 * np.arange(ho)).reshape(1, 1, ho, 1) + arg // 3
    xs = (2 * np.arange(wo)).reshape(1, 1, 1, wo) + arg % 3
    ni = np.arange(n).reshape(n, 1, 1, 1)
    ci = np.arange(c).reshape(1, c, 1, 1)
    np.add.at(dx, (np.broadcast_to(ni, arg.shape), np.broadcast_to(ci, arg.shape), ys, xs), gy)
    return dx


def _resize_axis_tables(n_in: int, n_out: int, simd: bool):
    """source index pairs and the fraction of dlib's bilinear resize_image along one axis (image_transforms/interpolation.h:
    722-806).  Rows: y accumulates in double (y += y_scale), the fraction is later rounded to float.  Columns: groups of four
    run in SSE floats -- x_j starts at float(-4*xs + j*xs) and is advanced by float(4*xs) per group, left = trunc(x) -- until a
    group's right-most `right` index would leave the row; the remaining columns accumulate in double again."""
    scale = (n_in - 1) / float(max(n_out - 1, 1))
    lo, hi, frac = np.zeros(n_out, np.int64), np.zeros(n_out, np.int64), np.zeros(n_out, np.float64)
    c = 0
    if simd:
        x0 = -4 * scale
        xv = np.array([x0, x0 + scale, x0 + 2 * scale, x0 + 3 * scale], dtype=F32)
        step = F32(4 * scale)
        while True:
            xv = (xv + step).astype(F32)
            left = xv.astype(np.int32)            # simd4i(_x): cvttps -> truncation
            right = left + 1
            if right[3] >= n_in or c + 3 >= n_out + 3:   # (the reference's loop has no c bound of its own: `right` leaves the row first)
                break
            if c + 4 > n_out:
                break
            lr = (xv - left.astype(F32)).astype(F32)
            lo[c:c + 4], hi[c:c + 4], frac[c:c + 4] = left, right, lr
            c += 4
    x = -scale + c * scale
    for cc in range(c, n_out):
        x += scale
        left = int(math.floor(x))
        lo[cc], hi[cc] = left, min(left + 1, n_in - 1)
        frac[cc] = float(F32(x - left)) if simd else x - left
    return lo, hi, frac


def resize_bilinear(x: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """tt::resize_bilinear -> cpu::resize_bilinear -> dlib::resize_image(interpolate_bilinear) per channel plane,
    cpu_dlib.cpp:1735-1762 and image_transforms/interpolation.h:700-806 (float pixels)."""
    n, c, h, w = x.shape
    top, bottom, tb = _resize_axis_tables(h, out_h, simd=False)
    left, right, lr = _resize_axis_tables(w, out_w, simd=True)
    tbf, lrf = tb.astype(F32), lr.astype(F32)
    itb, ilr = (F32(1) - tbf).astype(F32), (F32(1) - lrf).astype(F32)
    tl, tr = x[:, :, top][:, :, :, left], x[:, :, top][:, :, :, right]
    bl, br = x[:, :, bottom][:, :, :, left], x[:, :, bottom][:, :, :, right]
    wtl = (itb[:, None] * ilr[None, :]).astype(F32)
    wtr = (itb[:, None] * lrf[None, :]).astype(F32)
    wbl = (tbf[:, None] * ilr[None, :]).astype(F32)
    wbr = (tbf[:, None] * lrf[None, :]).astype(F32)
    return (((wtl * tl + wtr * tr).astype(F32) + wbl * bl).astype(F32) + wbr * br).astype(F32)


def bn_inference(x, gamma, beta, rm, rv, eps):
    """batch_normalize_conv_inference, cpu_dlib.cpp:1008-1071."""
    invstd = (F32(1) / np.sqrt(rv + F32(eps))).astype(F32)
    sh = (1, -1, 1, 1)
    return (gamma.reshape(sh) * (x - rm.reshape(sh)) * invstd.reshape(sh) + beta.reshape(sh)).astype(F32)


# How the per-channel batch-norm reductions are accumulated:
#   "dlib_fp32": sequentially in fp32 in dlib's loop order (sample, channel, pixel) -- the faithful
#                restatement of cpu_dlib.cpp:1127-1140,1228-1270.  This is what pins the oracle to the
#                compiled reference (gradients agree to ~4e-5; tests/test_oracle_golden.py).
#   "fp64":      exact sums.  dlib's fp32 running sums carry rounding noise that the ~40 stacked
#                bn_ layers amplify to ~1e-2 on the deepest gradients of the narrow golden net; the CUDA
#                engine accumulates in fp64, so it is compared with this mode at 1e-4 and with the
#                reference's own vectors at the reference's noise floor (DESIGN.md "Parity").
SUM_MODE = "dlib_fp32"


def _channel_sum(a: np.ndarray) -> np.ndarray:
    """sum over (n, y, x) per channel of an NCHW array, accumulated according to SUM_MODE."""
    if SUM_MODE == "fp64":
        return a.sum(axis=(0, 2, 3), dtype=np.float64)
    c = a.shape[1]
    flat = np.ascontiguousarray(a.transpose(1, 0, 2, 3)).reshape(c, -1).astype(F32)
    return np.cumsum(flat, axis=1, dtype=F32)[:, -1]


def bn_train(x, gamma, beta, eps):
    """batch_normalize_conv, cpu_dlib.cpp:1073-1182 (biased variance for normalisation)."""
    m = x.shape[0] * x.shape[2] * x.shape[3]
    if SUM_MODE == "fp64":
        mean = (_channel_sum(x) / m).astype(F32)
        sq = (_channel_sum(np.square(x, dtype=np.float64)) / m).astype(F32)
    else:
        mean = (_channel_sum(x) / F32(m)).astype(F32)
        sq = (_channel_sum(x * x) / F32(m)).astype(F32)
    var = (sq - mean * mean).astype(F32)
    invstd = (1.0 / np.sqrt(var.astype(np.float64) + eps)).astype(F32)
    sh = (1, -1, 1, 1)
    y = ((x - mean.reshape(sh)) * invstd.reshape(sh)) * gamma.reshape(sh) + beta.reshape(sh)
    return y.astype(F32), mean, invstd, var


def _bn_backward_fp32(gy, x, gamma, mean, invstd):
    """cpu_dlib.cpp:1228-1288 statement by statement, in fp32 with sequential sums."""
    sh = (1, -1, 1, 1)
    m = x.shape[0] * x.shape[2] * x.shape[3]
    xc = (x - mean.reshape(sh)).astype(F32)
    xhat = (xc * invstd.reshape(sh)).astype(F32)
    dbeta = _channel_sum(gy)
    dgamma = _channel_sum(gy * xhat)
    dx = (gy * gamma.reshape(sh)).astype(F32)
    invstd_pow = (F32(-0.5) * invstd ** 3).astype(F32)
    dvars = _channel_sum(dx * xc * invstd_pow.reshape(sh))
    invnum = F32(1.0 / m)
    dmeans = _channel_sum((-dx * invstd.reshape(sh) + dvars.reshape(sh) * F32(-2) * xc * invnum).astype(F32))
    out = dx * invstd.reshape(sh) + dvars.reshape(sh) * 2 * xc * invnum + dmeans.reshape(sh) * invnum
    return out.astype(F32), dgamma.astype(F32), dbeta.astype(F32)


def bn_backward(gy, x, gamma, mean, invstd):
    """batch_normalize_conv_gradient, cpu_dlib.cpp:1184-1289; returns (dx, dgamma, dbeta)."""
    if SUM_MODE != "fp64":
        return _bn_backward_fp32(gy, x, gamma, mean, invstd)
    sh = (1, -1, 1, 1)
    m = x.shape[0] * x.shape[2] * x.shape[3]
    gy64 = gy.astype(np.float64)
    xc = x.astype(np.float64) - mean.reshape(sh).astype(np.float64)
    istd = invstd.astype(np.float64)
    xhat = xc * istd.reshape(sh)
    dbeta = gy64.sum(axis=(0, 2, 3))
    dgamma = (gy64 * xhat).sum(axis=(0, 2, 3))
    dxh = gy64 * gamma.reshape(sh).astype(np.float64)
    dvar = (dxh * xc).sum(axis=(0, 2, 3)) * (-0.5 * istd ** 3)
    dmean = (-dxh * istd.reshape(sh)).sum(axis=(0, 2, 3)) + dvar * (-2.0 * xc.sum(axis=(0, 2, 3)) / m)
    dx = dxh * istd.reshape(sh) + dvar.reshape(sh) * 2.0 * xc / m + dmean.reshape(sh) / m
    return dx.astype(F32), dgamma.astype(F32), dbeta.astype(F32)


def affine(x, gamma, beta):
    """affine_transform_conv, cpu_dlib.cpp:629-660."""
    sh = (1, -1, 1, 1)
    return (x * gamma.reshape(sh) + beta.reshape(sh)).astype(F32)


# ----------------------------------------------------------------------------------------------
# the executor (dlib/dlib/dnn/core.h:932-948 forward, :998-1009 backward)


def _split_conv_params(layer) -> Tuple[np.ndarray, np.ndarray]:
    p = layer["params"].reshape(-1)
    fa = layer["filters_alias"]
    nf = fa[0] * fa[1] * fa[2] * fa[3]
    return p[:nf].reshape(fa), p[nf:]


def _tagged(layers, i, tag_id):
    for j in range(i + 1, len(layers)):
        if layers[j]["type"] == "tag" and layers[j]["id"] == tag_id:
            return j
    raise ValueError(f"tag {tag_id} not found below layer {i}")


def _res(layers, i):
    """tags forward their subnetwork's tensors (core.h:1555-1731)."""
    while layers[i]["type"] == "tag":
        i += 1
    return i


class Context:
    def __init__(self, n):
        self.out: List[Optional[np.ndarray]] = [None] * n
        self.aux: Dict[int, object] = {}
        self.training = False


def forward(net: Net, x: np.ndarray, training: bool, update_running_stats: bool = True, skip_dead: bool = False) -> Context:
    """x = NCHW f32 tensor (to_tensor output).  training=True runs bn_ in batch mode
    (layers.h:1387-1408, only legal for batch > 1) and moves the running statistics."""
    L = net.layers
    n = len(L)
    ctx = Context(n)
    ctx.training = training
    ctx.out[n - 1] = x
    for i in range(n - 2, 0, -1):
        l = L[i]
        t = l["type"]
        if t == "tag":
            continue
        s = _res(L, i + 1)
        xin = ctx.out[s]
        if skip_dead and l.get("dead") and t == "con":
            ctx.out[i] = None
            continue
        if xin is None and t != "concat":  # downstream of a skipped dead conv (concat_ reads tags only)
            ctx.out[i] = None
            continue
        if t == "con":
            w, b = _split_conv_params(l)
            ctx.out[i] = conv2d(xin, w, l["pad"]) + b.reshape(1, -1, 1, 1)
        elif t == "cont":
            w, b = _split_conv_params(l)  # [Cin, Cout, k, k]
            st, pd = l["stride"], l["pad"]
            ho = st * (xin.shape[2] - 1) + l["nr"] - 2 * pd
            wo = st * (xin.shape[3] - 1) + l["nc"] - 2 * pd
            ctx.out[i] = conv2d_bwd_data(xin, w, st, pd, (ho, wo)) + b.reshape(1, -1, 1, 1)
        elif t == "mish":
            ctx.out[i] = mish(xin)
        elif t == "bn":
            c = l["gamma_alias"][1]
            p = l["params"].reshape(-1)
            gamma, beta = p[:c], p[c:]
            if training:
                assert xin.shape[0] > 1, "bn_ runs in batch mode only for batch > 1 (layers.h:1392)"
                y, mean, invstd, var = bn_train(xin, gamma, beta, l["eps"])
                ctx.aux[i] = (mean, invstd)
                if update_running_stats:
                    m = xin.shape[0] * xin.shape[2] * xin.shape[3]
                    nu = l["num_updates"]
                    decay = 1.0 - nu / (nu + 1.0)
                    l["num_updates"] = min(nu + 1, l["window"])
                    scale = m / (m - 1.0)
                    rv = l["running_variances"].reshape(-1)
                    rm = l["running_means"].reshape(-1)
                    if decay == 1:
                        rv_new = (scale * var.astype(np.float64)).astype(F32)
                        rm_new = mean.copy()
                    else:
                        rv_new = ((1 - decay) * rv.astype(np.float64) + scale * decay * var.astype(np.float64)).astype(F32)
                        rm_new = ((1 - decay) * rm.astype(np.float64) + decay * mean.astype(np.float64)).astype(F32)
                    l["running_variances"] = rv_new.reshape(1, c, 1, 1)
                    l["running_means"] = rm_new.reshape(1, c, 1, 1)
                    l["means"] = mean.reshape(1, c, 1, 1).copy()
                    l["invstds"] = invstd.reshape(1, c, 1, 1).copy()
                ctx.out[i] = y
            else:
                ctx.out[i] = bn_inference(xin, gamma, beta, l["running_means"].reshape(-1), l["running_variances"].reshape(-1), l["eps"])
        elif t == "affine":
            c = l["gamma_alias"][1]
            p = l["params"].reshape(-1)
            ctx.out[i] = affine(xin, p[:c], p[c:])
        elif t == "max_pool":
            y, arg = max_pool_3x3s2(xin)
            ctx.out[i] = y
            ctx.aux[i] = arg
        elif t == "concat_prev":
            t2 = ctx.out[_res(L, _tagged(L, i, l["tag"]))]
            if xin.shape[2:] != t2.shape[2:]:
                # concat_prev_::resize_if_needed (layers.h:2438,2494-2521): the sub-network output is bilinearly resized to the
                # tagged tensor's size (they differ by at most one pixel: an even-sized level pools to (s-1)/2 and comes back as s-1)
                assert abs(xin.shape[2] - t2.shape[2]) <= 1 and abs(xin.shape[3] - t2.shape[3]) <= 1
                xin = resize_bilinear(xin, t2.shape[2], t2.shape[3])
                ctx.aux[("resized", i)] = True
            ctx.out[i] = np.concatenate([xin, t2], axis=1)
        elif t == "concat":
            ctx.out[i] = np.concatenate([ctx.out[_res(L, _tagged(L, i, tid))] for tid in l["tags"]], axis=1)
        else:
            raise ValueError(t)
    return ctx


def logits_of(ctx: Context) -> np.ndarray:
    return ctx.out[1]


def loss_and_grad(logits: np.ndarray, labels: np.ndarray, weights: np.ndarray) -> Tuple[float, np.ndarray]:
    """loss_multiclass_log_per_pixel_weighted_::compute_loss_value_and_gradient, loss.h:3121-3192
    with cpu::softmax (cpu_dlib.cpp:1326-1383).  labels u16 [N,H,W], weights f32 [N,H,W]."""
    n, k, h, w = logits.shape
    mx = logits.max(axis=1, keepdims=True)
    e = np.exp((logits - mx).astype(F32))
    p = (e / e.sum(axis=1, keepdims=True, dtype=F32)).astype(F32)
    scale = 1.0 / (n * h * w)
    valid = weights != 0
    if np.any((labels >= k) & valid):
        raise ValueError("label out of range with non-zero weight (loss.h:3173)")
    y = np.where(labels < k, labels, 0).astype(np.int64)
    py = np.take_along_axis(p, y[:, None], axis=1)[:, 0]
    wl = weights.astype(np.float64) * scale
    # NB the reference multiplies float(weight) by double(scale), then by float g: (weight*scale)*g in double
    loss = float(np.sum(np.where(labels < k, wl * -np.log(np.maximum(py.astype(F32), F32(1e-10))).astype(np.float64), 0.0)))
    onehot = np.zeros_like(p)
    np.put_along_axis(onehot, y[:, None], 1.0, axis=1)
    onehot *= (labels < k)[:, None]
    grad = (wl[:, None] * (p.astype(np.float64) - onehot)).astype(F32)
    return loss, grad


def backward(net: Net, ctx: Context, grad_logits: np.ndarray) -> Dict[int, np.ndarray]:
    """back_propagate_error, core.h:998-1009 + each layer's backward (layers.h).  Returns the
    parameter gradient of every parameter-bearing layer keyed by dlib layer index, in dlib's
    flat layout (filters then biases; gamma then beta)."""
    L = net.layers
    n = len(L)
    g: List[Optional[np.ndarray]] = [None] * n
    g[1] = grad_logits
    pg: Dict[int, np.ndarray] = {}

    def add(j, val):
        j = _res(L, j)
        if j == n - 1:
            return
        g[j] = val.copy() if g[j] is None else g[j] + val

    for i in range(1, n - 1):
        l = L[i]
        t = l["type"]
        if t == "tag":
            continue
        s = _res(L, i + 1)
        xin = ctx.out[s]
        gin = g[i]
        if gin is None:
            # never consumed (the "dead" dense layers): dlib back-propagates an all-zero
            # gradient, so every parameter gradient below is exactly zero
            if t in ("con", "cont", "bn"):
                pg[i] = np.zeros(l["params"].size, dtype=F32)
            if xin is not None and ctx.out[i] is not None:
                gin = np.zeros_like(ctx.out[i])
            else:
                continue
        if t == "con":
            w, _ = _split_conv_params(l)
            add(i + 1, conv2d_bwd_data(gin, w, 1, l["pad"], xin.shape[2:]))
            dw = conv2d_bwd_filter(gin, xin, l["nr"], 1, l["pad"])
            db = gin.sum(axis=(0, 2, 3), dtype=np.float64).astype(F32)
            pg[i] = np.concatenate([dw.reshape(-1), db])
        elif t == "cont":
            w, _ = _split_conv_params(l)  # [Cin, Cout, k, k] == a conv filter bank mapping Cout -> Cin
            add(i + 1, conv2d_strided(gin, w, l["stride"], l["pad"]))
            dw = conv2d_bwd_filter(xin, gin, l["nr"], l["stride"], l["pad"])  # roles swapped, layers.h:547
            db = gin.sum(axis=(0, 2, 3), dtype=np.float64).astype(F32)
            pg[i] = np.concatenate([dw.reshape(-1), db])
        elif t == "mish":
            add(i + 1, gin * mish_grad(xin))
        elif t == "bn":
            c = l["gamma_alias"][1]
            gamma = l["params"].reshape(-1)[:c]
            mean, invstd = ctx.aux[i]
            dx, dgamma, dbeta = bn_backward(gin, xin, gamma, mean, invstd)
            add(i + 1, dx)
            pg[i] = np.concatenate([dgamma, dbeta])
        elif t == "max_pool":
            add(i + 1, max_pool_bwd(gin, ctx.aux[i], xin.shape[2:]))
        elif t == "concat_prev":
            k1 = xin.shape[1]
            # Where the forward pass resized, the reference's backward pass adds this slice into a TEMPORARY
            # (resize_if_needed returns resize_temp, layers.h:2452,2494-2518) that nobody reads: the sub-network's gradient input
            # stays zero.  Restated as is -- the compiled reference's gradients confirm it (tests/test_gpu_resize.py).
            if ctx.aux.get(("resized", i)):
                add(i + 1, np.zeros_like(xin))
            else:
                add(i + 1, gin[:, :k1])
            add(_tagged(L, i, l["tag"]), gin[:, k1:])
        elif t == "concat":
            off = 0
            for tid in l["tags"]:
                j = _tagged(L, i, tid)
                kk = ctx.out[_res(L, j)].shape[1]
                add(j, gin[:, off:off + kk])
                off += kk
        elif t == "affine":
            raise ValueError("anet_type is not trainable")
    return pg


def sgd_step(net: Net, grads: Dict[int, np.ndarray], momentum_state: Dict[int, np.ndarray], lr: float,
             weight_decay: float = 0.001, momentum: float = 0.9) -> None:
    """sgd::operator() + update_considering_bias, dlib/dlib/dnn/solvers.h:37-119,155-185, and
    add_layer::update_parameters, core.h:1011-1023.  In place on net and momentum_state."""
    for i, l in net.param_layers():
        if l["type"] == "affine":
            continue
        p = l["params"].reshape(-1)
        gr = grads[i].reshape(-1)
        v = momentum_state.setdefault(i, np.zeros_like(p))
        lrm, wdm, blrm, bwdm = l["multipliers"]
        if l["type"] in ("con", "cont"):
            bias_off = p.size - l["num_filters"]
        else:
            bias_off = p.size // 2
        lr_w = float(F32(lr)) * lrm
        wd_w = float(F32(weight_decay)) * wdm
        for lo, hi, lr_, wd_ in ((0, bias_off, lr_w, wd_w), (bias_off, p.size, lr_w * blrm, wd_w * bwdm)):
            a, b, c = F32(float(F32(momentum))), F32(-wd_ * lr_), F32(-lr_)
            v[lo:hi] = a * v[lo:hi] + b * p[lo:hi] + c * gr[lo:hi]
        p += v
        l["params"] = p.reshape(l["params"].shape)


def adam_step(net: Net, grads: Dict[int, np.ndarray], state: Dict[int, Dict[str, np.ndarray]], lr: float,
              weight_decay: float = 0.001, momentum1: float = 0.9, momentum2: float = 0.999) -> None:
    """adam::operator() + update_considering_bias, dlib/dlib/dnn/solvers.h:225-372, with cpu::compute_adam_update,
    dlib/dlib/cuda/cpu_dlib.cpp:704-743 (fp32 arithmetic, eps = 1e-8, bias-corrected step size
    alpha = lr * sqrt(1 - b2^t) / (1 - b1^t)) and add_layer::update_parameters, core.h:1011-1023.  In place on net and state."""
    for i, l in net.param_layers():
        if l["type"] == "affine":
            continue
        p = l["params"].reshape(-1)
        gr = grads[i].reshape(-1).astype(F32)
        st = state.setdefault(i, {"m": np.zeros_like(p), "v": np.zeros_like(p), "t": 0})
        st["t"] += 1
        t = F32(st["t"])
        lrm, wdm, blrm, bwdm = l["multipliers"]
        bias_off = p.size - l["num_filters"] if l["type"] in ("con", "cont") else p.size // 2
        lr_w = F32(float(F32(lr)) * lrm)
        wd_w = F32(float(F32(weight_decay)) * wdm)
        b1, b2 = F32(momentum1), F32(momentum2)
        s = np.empty_like(p)
        for lo, hi, lr_, wd_ in ((0, bias_off, lr_w, wd_w), (bias_off, p.size, F32(lr_w * F32(blrm)), F32(wd_w * F32(bwdm)))):
            alpha = F32(lr_ * F32(np.sqrt(F32(1) - F32(np.power(b2, t)))) / (F32(1) - F32(np.power(b1, t))))
            g = (wd_ * p[lo:hi] + gr[lo:hi]).astype(F32)
            st["m"][lo:hi] = b1 * st["m"][lo:hi] + (F32(1) - b1) * g
            st["v"][lo:hi] = b2 * st["v"][lo:hi] + (F32(1) - b2) * g * g
            s[lo:hi] = -alpha * st["m"][lo:hi] / (np.sqrt(st["v"][lo:hi]) + F32(1e-8))
        p += s
        l["params"] = p.reshape(l["params"].shape)


def to_label(logits_khw: np.ndarray, gains: Sequence[float] = ()) -> np.ndarray:
    """loss_multiclass_log_per_pixel_::to_label, loss.h:2921-2969 / annonet_infer.cpp:170-185:
    argmax_k float(logit_k + gain_k), strict '>' from -inf: lowest index wins ties, all-NaN -> 65535."""
    k = logits_khw.shape[0]
    v = logits_khw.astype(np.float64)
    if len(gains):
        v = v + np.asarray(gains, dtype=np.float64).reshape(k, 1, 1)
    v = v.astype(F32)
    best = np.full(v.shape[1:], -np.inf, dtype=F32)
    label = np.full(v.shape[1:], LABEL_TO_IGNORE, dtype=np.uint16)
    for c in range(k):
        better = v[c] > best
        label[better] = c
        best = np.where(better, v[c], best)
    return label


# ----------------------------------------------------------------------------------------------
# tiling + tiled inference


def recommended_input_dimension(minimum: int, levels: int = 6) -> int:
    """RuntimeNet::GetRecommendedInputDimension, NetPimpl.cpp:271-285: smallest 2^L*(w+1)-1 >= minimum, w in 1..499."""
    for w in range(1, 500):
        d = (w + 1) * (1 << levels) - 1
        if d >= minimum:
            return d
    raise ValueError(f"Requested minimum input dimension {minimum} is too large")


def find_starting_center(full: int, tile: int, overlap: int) -> int:
    """tiling/tiling.cpp:7-21."""
    stride = tile - overlap
    c = full // 2
    while c - stride >= 0:
        c -= stride
    if c > tile // 2:
        c -= stride // 2
    return c


def get_tiles(width: int, height: int, max_w: int, max_h: int, ov_x: int, ov_y: int):
    """tiling::get_tiles, tiling/tiling.cpp:23-87 + to_dlib_rectangle (dlib-wrapper.h:18-25).
    Returns a list of dicts with inclusive (l, t, r, b) rects: full, non_overlapping, unique."""
    sx = find_starting_center(width, max_w, ov_x)
    sy = find_starting_center(height, max_h, ov_y)
    stx, sty = max_w - ov_x, max_h - ov_y
    tiles = []
    cy = sy
    while cy < height:
        top_row, bottom_row = cy == sy, cy + sty >= height
        cx = sx
        while cx < width:
            left_col, right_col = cx == sx, cx + stx >= width
            dl, dt = cx - max_w // 2, cy - max_h // 2
            left, top = max(0, dl), max(0, dt)
            right, bottom = min(dl + max_w, width), min(dt + max_h, height)
            tw, th = right - left, bottom - top
            no_l = left if left_col else left + (ov_x + 1) // 2
            no_t = top if top_row else top + (ov_y + 1) // 2
            un_l = left if left_col else left + ov_x
            un_t = top if top_row else top + ov_y
            fr, fb = left + tw, top + th
            no_r = fr if right_col else fr - ov_x // 2
            no_b = fb if bottom_row else fb - ov_y // 2
            un_r = fr if right_col else fr - ov_x
            un_b = fb if bottom_row else fb - ov_y
            tiles.append(dict(full=(left, top, fr - 1, fb - 1),
                              non_overlapping=(no_l, no_t, no_r - 1, no_b - 1),
                              unique=(un_l, un_t, un_r - 1, un_b - 1)))
            cx += stx
        cy += sty
    return tiles


def _get_t(c, first_possible, first_in, last_in, last_possible):
    """annonet_infer.cpp:102-114."""
    if c < first_in:
        return (c - first_possible) / float(first_in - first_possible)
    if c > last_in:
        return (last_possible - c) / float(last_possible - last_in)
    return 1.0


def annonet_infer(anet: Net, image: np.ndarray, max_tile_w: int, max_tile_h: int, gains: Sequence[float] = (),
                  overlap: Optional[int] = None, return_blended: bool = False, forward_fn=None, tile_range: Optional[Tuple[int, int]] = None):
    """annonet_infer(), annonet_infer.cpp:32-185 (without the detection-level blob filter):
    tiles -> enlarge to the recommended size around the centre -> clamp-to-edge fill (outpaint,
    annonet.h:74-120) -> Forward -> blend (double ramps, float accumulate) -> argmax."""
    h, w, _ = image.shape
    levels = anet.levels
    ov = overlap if overlap is not None else (1 << (levels + 1)) - 1
    tiles = get_tiles(w, h, max_tile_w, max_tile_h, ov, ov)
    blended = None
    for ti, tile in enumerate(tiles):
        fl, ft, fr, fb = tile["full"]
        # multi-GPU sharding of the engine: a rank runs the tiles [t0, t1) of its band into a private canvas
        if tile_range is not None and not (tile_range[0] <= ti < tile_range[1]):
            continue
        ul, ut, ur, ub = tile["unique"]
        fw, fh = fr - fl + 1, fb - ft + 1
        cx, cy = fl + fw // 2, ft + fh // 2
        rw, rh = recommended_input_dimension(fw, levels), recommended_input_dimension(fh, levels)
        left, top = cx - rw // 2, cy - rh // 2
        ys = np.clip(np.arange(top, top + rh), 0, h - 1)
        xs = np.clip(np.arange(left, left + rw), 0, w - 1)
        chip = image[ys][:, xs]
        if forward_fn is not None:   # tests: check the scheduler / blend / argmax at sizes the numpy net cannot reach
            out = forward_fn(np.ascontiguousarray(chip))
        else:
            x = to_tensor(chip[None], anet.layers[-1]["avg"])
            out = logits_of(forward(anet, x, training=False, skip_dead=True))[0]
        k = out.shape[0]
        if blended is None:
            blended = np.zeros((k, h, w), dtype=F32)
        y0, y1 = max(ft, 0, top), min(fb, h - 1, top + rh - 1)
        x0, x1 = max(fl, 0, left), min(fr, w - 1, left + rw - 1)
        yy = np.arange(y0, y1 + 1)
        xx = np.arange(x0, x1 + 1)
        tv = np.array([_get_t(y, ft, ut, ub, fb) for y in yy], dtype=np.float64)
        th = np.array([_get_t(x_, fl, ul, ur, fr) for x_ in xx], dtype=np.float64)
        inside_v = (yy >= ut) & (yy <= ub)
        inside_h = (xx >= ul) & (xx <= ur)
        unique = inside_v[:, None] & inside_h[None, :]
        t = th[None, :] * tv[:, None]
        src = out[:, y0 - top:y1 - top + 1, x0 - left:x1 - left + 1]
        dst = blended[:, y0:y1 + 1, x0:x1 + 1]
        acc = (dst.astype(np.float64) + t[None] * src.astype(np.float64)).astype(F32)  # `out += t * in` with float out
        dst[...] = np.where(unique[None], src, acc)
    if blended is None:
        return (None, None) if return_blended else None
    labels = to_label(blended, gains)
    return (labels, blended) if return_blended else labels


def filter_detections(labels: np.ndarray, blended: np.ndarray, detection_levels: Sequence[float], transposed_seeds: bool = True) -> np.ndarray:
    """the detection-level blob filter of annonet_infer(), annonet_infer.cpp:187-239: a pixel is a SEED when its label is
    non-zero and blended[label] - blended[0] (float) > detection_levels[label] - detection_levels[0] (double); connected blobs
    of equal non-zero labels (dlib::label_connected_blobs with zero_pixels_are_background, neighbors_8, connected_if_equal)
    that hold no seed are set to 0.  Connected components by scipy.ndimage.label with the full 3x3 structure, per label.

    Reference quirk, restated as is: the seeds are stored as dlib::point(r, c) -- i.e. x = r, y = c (annonet_infer.cpp:210) -- and
    looked up as connected_blobs(point.y(), point.x()) (:221), so the seed found at pixel (r, c) marks the blob at pixel
    (row = c, col = r).  Out-of-range lookups (c >= rows or r >= cols, possible on non-square images) read past the matrix in
    the reference; here they mark nothing.  transposed_seeds=False gives the evidently intended behaviour."""
    from scipy import ndimage
    det = np.asarray(detection_levels, dtype=np.float64)
    if not np.any(det > 0):
        return labels.copy()
    k = blended.shape[0]
    lab = labels.astype(np.int64)
    mine = np.take_along_axis(blended, np.minimum(lab, k - 1)[None], axis=0)[0]
    diff = (mine - blended[0]).astype(F32).astype(np.float64)
    seeds = (labels > 0) & (labels < k) & (diff > det[np.minimum(lab, k - 1)] - det[0])
    if transposed_seeds:
        h, w = labels.shape
        m = min(h, w)
        marked = np.zeros_like(seeds)
        marked[:min(w, h), :min(h, w)] = False
        t = seeds.T            # t[c, r] = seeds[r, c]
        marked[:min(h, t.shape[0]), :min(w, t.shape[1])] = t[:h, :w]
        seeds = marked
    out = labels.copy()
    for value in np.unique(labels):
        if value == 0:
            continue
        comp, n = ndimage.label(labels == value, structure=np.ones((3, 3), dtype=bool))
        if n == 0:
            continue
        has_seed = np.zeros(n + 1, dtype=bool)
        has_seed[np.unique(comp[seeds & (comp > 0)])] = True
        out[(comp > 0) & ~has_seed[comp]] = 0
    return out


def crop_and_augment(image: np.ndarray, labels: np.ndarray, left: int, top: int, dim: int, class_weight: float, image_weight: float,
                     flip_lr: bool = False, flip_ud: bool = False, brightness: float = 1.0, color_offset: Sequence[int] = (0, 0, 0)):
    """randomly_crop_image, annonet_train_main.cpp:110-232, for given random decisions and further_downscaling_factor 1
    (extract_image_chip with an unscaled dim x dim rectangle copies pixels; outside the image the chip is black and outpaint,
    annonet.h:74-120, then repeats the nearest valid pixel; labels outside become label_to_ignore): crop -> set_weights ->
    fliplr / flipud -> multiplicative brightness (tuc::round of the clamped product) -> colour offset
    (random_color_transform.h:98-150: put_in_range(0, 255, v + offset)).  Noise (:73-101) is left out: it is a random stream.
    Returns (u8 [dim, dim, 3], u16 labels [dim, dim], f32 weights [dim, dim])."""
    h, w = labels.shape
    ys, xs = top + np.arange(dim), left + np.arange(dim)
    inside = ((ys >= 0) & (ys < h))[:, None] & ((xs >= 0) & (xs < w))[None, :]
    cy, cx = np.clip(ys, 0, h - 1), np.clip(xs, 0, w - 1)
    img = image[cy][:, cx].copy()
    lab = np.where(inside, labels[cy][:, cx], LABEL_TO_IGNORE).astype(np.uint16)
    wts = set_weights(lab, class_weight, image_weight)
    if flip_lr:
        img, lab, wts = img[:, ::-1], lab[:, ::-1], wts[:, ::-1]
    if flip_ud:
        img, lab, wts = img[::-1], lab[::-1], wts[::-1]
    v = img.astype(np.float64)
    if brightness != 1.0:
        v = np.floor(np.clip(v * brightness, 0.0, 255.0) + 0.5)
    v = np.clip(v + np.asarray(color_offset, dtype=np.float64).reshape(1, 1, 3), 0, 255)
    return v.astype(np.uint8), np.ascontiguousarray(lab), np.ascontiguousarray(wts)


def confusion_per_pixel(truth: np.ndarray, predicted: np.ndarray, k: int) -> np.ndarray:
    """confusion_matrix_per_pixel, annonet_infer_main.cpp:482-489: [truth][predicted] over the labelled pixels (truth != 65535)."""
    m = (truth < k) & (predicted < k)
    out = np.zeros((k, k), dtype=np.uint64)
    np.add.at(out, (truth[m].astype(np.int64), predicted[m].astype(np.int64)), 1)
    return out


def confusion_per_region(truth: np.ndarray, predicted: np.ndarray, k: int) -> np.ndarray:
    """update_confusion_matrix_per_region, annonet_infer_main.cpp:202-272: for the connected blobs (equal non-zero labels,
    8-neighbourhood; all zero pixels together are blob 0) of the ground truth and then of the result, every labelled pixel
    votes for its truth and its prediction; in a blob whose ground truth is predominantly non-background the background
    prediction is dropped unless it is the only one; the blob adds one count at [truth winner][prediction winner].
    (The reference picks the winner with max_element over an unordered_map: ties are iteration-order dependent there, the
    lowest class here.)"""
    from scipy import ndimage
    out = np.zeros((k, k), dtype=np.uint64)
    labelled = truth != LABEL_TO_IGNORE
    if not labelled.any():
        return out
    for img in (truth, predicted):
        comp = np.zeros(img.shape, dtype=np.int64)
        nxt = 1
        for value in np.unique(img):
            if value == 0:
                continue
            c, n = ndimage.label(img == value, structure=np.ones((3, 3), dtype=bool))
            comp[c > 0] = c[c > 0] + (nxt - 1)
            nxt += n
        vg = np.zeros((nxt, k), dtype=np.int64)
        vp = np.zeros((nxt, k), dtype=np.int64)
        b = comp[labelled]
        np.add.at(vg, (b, truth[labelled].astype(np.int64)), 1)
        np.add.at(vp, (b, predicted[labelled].astype(np.int64)), 1)
        for blob in range(nxt):
            if vg[blob].sum() == 0:
                continue
            wgt = int(np.argmax(vg[blob]))
            only_background = vp[blob, 0] > 0 and vp[blob, 1:].sum() == 0
            if wgt != 0 and not only_background:
                vp[blob, 0] = 0
            out[wgt, int(np.argmax(vp[blob]))] += 1
    return out


def render_labels(labels: np.ndarray, out_h: int, out_w: int, palette_rgba: np.ndarray) -> np.ndarray:
    """resize_label_image (annonet.cpp:132-141: dlib::resize_image + interpolate_nearest_neighbor, i.e. source pixel
    floor(c * (iw-1)/max(ow-1,1) + 0.5), image_transforms/interpolation.h:559-593 and dlib's vector<double> -> point rounding)
    followed by index_label_image_to_rgba_label_image (annonet_infer_main.cpp:74-86)."""
    ih, iw = labels.shape
    ys = (ih - 1) / float(max(out_h - 1, 1))
    xs = (iw - 1) / float(max(out_w - 1, 1))
    sy = np.floor(np.arange(out_h) * ys + 0.5).astype(np.int64)
    sx = np.floor(np.arange(out_w) * xs + 0.5).astype(np.int64)
    idx = labels[sy][:, sx].astype(np.int64)
    pal = np.asarray(palette_rgba, dtype=np.uint8).reshape(-1, 4)
    out = np.zeros((out_h, out_w, 4), dtype=np.uint8)
    ok = idx < pal.shape[0]
    out[ok] = pal[idx[ok]]
    return out


def set_weights(labels: np.ndarray, class_weight: float, image_weight: float) -> np.ndarray:
    """set_weights, annonet_train.h:20-83.  labels u16 [H,W] -> f32 weights [H,W]."""
    nr, nc = labels.shape
    valid = labels != LABEL_TO_IGNORE
    if not valid.any():
        return np.zeros((nr, nc), dtype=F32)
    counts = np.bincount(labels[valid].astype(np.int64))
    # the reference's histogram vector grows as index*2+16 (make_sure_vector_is_large_enough,
    # annonet_train.h:30-35) and average_count divides by that *allocated* size, zeros
    # included; growth only depends on the order in which labels are first seen
    flat = labels.reshape(-1)
    flat = flat[flat != LABEL_TO_IGNORE]
    _, first = np.unique(flat, return_index=True)
    size = 0
    for lab in flat[np.sort(first)].tolist():
        if lab >= size:
            size = lab * 2 + 16
    total = int(counts.sum())
    average = total / float(size)
    lw = np.zeros(max(size, counts.size), dtype=np.float64)
    total_unnorm = 0.0
    for lab, cnt in enumerate(counts.tolist()):
        if cnt > 0:
            u = math.pow(average / cnt, class_weight)
            lw[lab] = u
            total_unnorm += cnt * u
    target = total * math.pow(nr * nc / float(total), image_weight)
    lw *= target / total_unnorm
    out = np.where(valid, lw[np.where(valid, labels, 0).astype(np.int64)], 0.0)
    return out.astype(F32)
