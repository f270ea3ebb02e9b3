"""CPU-side attribution of the bf16 path's gradient error: runs the numpy oracle (oracle/annonet_numpy.py) with bf16
rounding injected where the engine's tensor-core path stores bf16 (activations in the level buffers, filters, gradients),
selectable per resolution, and prints each layer's parameter-gradient error against the unrounded run.

  python experiments/bf16_emulation.py N H W  [act=MINH] [wt=MINH] [grad=MINH] [pooltie=exact|bf16]

MINH: round only tensors whose height is >= MINH (0 = everywhere, 99999 = nowhere).  Test infrastructure only."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import annonet_numpy as on  # noqa: E402
import dlib_format as df  # noqa: E402
from conftest import synth_image, synth_labels  # noqa: E402

F32 = np.float32


def rnd(a, bits=7):
    """round-to-nearest-even to `bits` explicit mantissa bits (7 = bfloat16, 10 = fp16 / tf32 without the range limits)"""
    a = np.ascontiguousarray(a, dtype=F32)
    u = a.view(np.uint32)
    sh = np.uint32(23 - bits)
    r = ((u >> sh) & np.uint32(1)) + np.uint32((1 << (23 - bits - 1)) - 1)
    return ((u + r) & np.uint32((0xFFFFFFFF << (23 - bits)) & 0xFFFFFFFF)).view(F32)


class Cfg:
    act = 99999      # level activations stored as bf16 where H >= act
    wt = 99999       # conv filters rounded where the conv's OUTPUT height >= wt
    grad = 99999     # gradients (dz, level gradient buffers) rounded where H >= grad
    pooltie = "bf16"  # "exact": break max-pool ties on the unrounded values
    abits = 7        # mantissa bits of the stored activations / filters / gradients
    wbits = 7
    gbits = 7
    w2 = 0           # filters as a hi + lo pair (two MMAs): residual rounded to wbits as well
    wbbits = 0       # > 0: the backward-data convolutions use the filters rounded to this many bits instead of the forward ones
    xgbits = 0       # > 0: the filter-gradient GEMMs read the activations rounded to this many bits (the hi half only)
    poolarg = 0      # 1: route the max-pool gradient with the argmax of the unrounded baseline run


def run(net, x, labels, weights, cfg, base_aux=None):
    """forward (training) + loss + backward with rounding hooks; returns (loss, logits, param grads)"""
    L = net.layers
    n = len(L)
    out = [None] * n
    hi = [None] * n   # unrounded values of the same tensors (for pool tie-breaking)
    aux = {}
    ra = lambda t: rnd(t, cfg.abits) if t.shape[2] >= cfg.act else t
    out[n - 1] = ra(x)
    hi[n - 1] = x
    # which layer ends a conv group (its output is what the engine stores in a level buffer)
    def is_store(i):
        t = L[i]["type"]
        if t == "cont":
            return L[i - 1]["type"] != "bn"          # `up`: straight into the level
        if t == "bn":
            return L[i + 1]["type"] == "mish"        # con -> mish -> bn
        if t == "mish":
            return L[i + 1]["type"] == "bn"          # con -> bn -> mish
        return False
    def rw(w):
        hi_ = rnd(w, cfg.wbits)
        return hi_ + rnd(w - hi_, cfg.wbits) if cfg.w2 else hi_
    for i in range(n - 2, 0, -1):
        l = L[i]
        t = l["type"]
        if t == "tag":
            continue
        s = on._res(L, i + 1)
        xin = out[s]
        if l.get("dead") and t == "con":
            continue
        if xin is None and t != "concat":
            continue
        if t in ("con", "cont"):
            w, b = on._split_conv_params(l)
            if t == "con":
                wr = rw(w) if xin.shape[2] >= cfg.wt else w
                y = on.conv2d(xin, wr, l["pad"]) + b.reshape(1, -1, 1, 1)
            else:
                st, pd = l["stride"], l["pad"]
                ho = st * (xin.shape[2] - 1) + l["nr"] - 2 * pd
                wo = st * (xin.shape[3] - 1) + l["nc"] - 2 * pd
                wr = rw(w) if ho >= cfg.wt else w
                y = on.conv2d_bwd_data(xin, wr, st, pd, (ho, wo)) + b.reshape(1, -1, 1, 1)
            aux[("w", i)] = rnd(w, cfg.wbbits) if cfg.wbbits else wr
        elif t == "mish":
            y = on.mish(xin)
        elif t == "bn":
            c = l["gamma_alias"][1]
            p = l["params"].reshape(-1)
            y, mean, invstd, var = on.bn_train(xin, p[:c], p[c:], l["eps"])
            aux[i] = (mean, invstd)
        elif t == "max_pool":
            src = hi[s] if cfg.pooltie == "exact" else xin
            _, arg = on.max_pool_3x3s2(src)
            nn, cc, ho, wo = arg.shape
            win = np.lib.stride_tricks.sliding_window_view(xin, (3, 3), axis=(2, 3))[:, :, ::2, ::2][:, :, :ho, :wo].reshape(nn, cc, ho, wo, 9)
            y = np.take_along_axis(win, arg[..., None], axis=-1)[..., 0]
            if cfg.pooltie != "exact":
                y = win.max(axis=-1)
            aux[i] = base_aux[i] if (cfg.poolarg and base_aux is not None) else arg
            hi[i] = y
        elif t == "concat_prev":
            j = on._res(L, on._tagged(L, i, l["tag"]))
            y = np.concatenate([xin, out[j]], axis=1)
            hi[i] = np.concatenate([hi[s], hi[j]], axis=1)
        elif t == "concat":
            js = [on._res(L, on._tagged(L, i, tid)) for tid in l["tags"]]
            y = np.concatenate([out[j] for j in js], axis=1)
            hi[i] = np.concatenate([hi[j] for j in js], axis=1)
        else:
            raise ValueError(t)
        if is_store(i):
            hi[i] = y
            y = ra(y)
        elif hi[i] is None:
            hi[i] = y
        out[i] = y
    logits = out[1]
    loss, gl = on.loss_and_grad(logits, labels, weights)

    rg = lambda t: rnd(t, cfg.gbits) if t.shape[2] >= cfg.grad else t
    g = [None] * n
    g[1] = rg(gl)
    pg = {}

    def add(j, val):
        j = on._res(L, j)
        if j == n - 1:
            return
        g[j] = rg(val.copy() if g[j] is None else g[j] + val)

    for i in range(1, n - 1):
        l = L[i]
        t = l["type"]
        if t == "tag":
            continue
        s = on._res(L, i + 1)
        xin = out[s]
        gin = g[i]
        if gin is None:
            continue
        if t == "con":
            add(i + 1, on.conv2d_bwd_data(gin, aux[("w", i)], 1, l["pad"], xin.shape[2:]))
            dw = on.conv2d_bwd_filter(gin, rnd(xin, cfg.xgbits) if cfg.xgbits else xin, l["nr"], 1, l["pad"])
            pg[i] = np.concatenate([dw.reshape(-1), gin.sum(axis=(0, 2, 3), dtype=np.float64).astype(F32)])
        elif t == "cont":
            add(i + 1, on.conv2d_strided(gin, aux[("w", i)], l["stride"], l["pad"]))
            dw = on.conv2d_bwd_filter(rnd(xin, cfg.xgbits) if cfg.xgbits else xin, gin, l["nr"], l["stride"], l["pad"])
            pg[i] = np.concatenate([dw.reshape(-1), gin.sum(axis=(0, 2, 3), dtype=np.float64).astype(F32)])
        elif t == "mish":
            add(i + 1, gin * on.mish_grad(xin))
        elif t == "bn":
            c = l["gamma_alias"][1]
            gamma = l["params"].reshape(-1)[:c]
            mean, invstd = aux[i]
            dx, dgamma, dbeta = on.bn_backward(gin, xin, gamma, mean, invstd)
            add(i + 1, dx)
            pg[i] = np.concatenate([dgamma, dbeta])
        elif t == "max_pool":
            add(i + 1, on.max_pool_bwd(gin, aux[i], xin.shape[2:]))
        elif t == "concat_prev":
            k1 = xin.shape[1]
            add(i + 1, gin[:, :k1])
            add(on._tagged(L, i, l["tag"]), gin[:, k1:])
        elif t == "concat":
            off = 0
            for tid in l["tags"]:
                j = on._tagged(L, i, tid)
                kk = out[on._res(L, j)].shape[1]
                add(j, gin[:, off:off + kk])
                off += kk
    return loss, logits, pg, aux


def compare(net, base, test, label):
    l0, lg0, g0, _ = base
    l1, lg1, g1, _ = test
    print(f"== {label}: loss {l1:.6f} vs {l0:.6f}; logits max-rel {np.abs(lg1 - lg0).max() / np.abs(lg0).max():.3e}")
    worst_l2 = worst_mx = 0
    rows = []
    for i, layer in net.param_layers():
        if i not in g0:
            continue
        a, b = g1[i], g0[i]
        if np.abs(b).max() == 0:
            continue
        nb = layer["num_filters"] if layer["type"] in ("con", "cont") else b.size // 2
        L = net.layers
        feeds_bn = layer["type"] in ("con", "cont") and (L[i - 1]["type"] == "bn" or (L[i - 1]["type"] == "mish" and L[i - 2]["type"] == "bn"))
        for part, sl in (("w", slice(0, b.size - nb)), ("b", slice(b.size - nb, None))):
            if part == "b" and feeds_bn and L[i - 1]["type"] == "bn":
                continue    # bias in front of a batch norm: its exact gradient is zero, the reference value is rounding noise
            l2 = float(np.linalg.norm(a[sl] - b[sl]) / np.linalg.norm(b[sl]))
            mx = float(np.abs(a[sl] - b[sl]).max() / np.abs(b[sl]).max())
            rows.append((i, layer["type"], part, l2, mx))
            worst_l2, worst_mx = max(worst_l2, l2), max(worst_mx, mx)
    for r in rows:
        if r[2] == "w":
            print("   layer<%d> %-5s %s relL2 %.3e maxrel %.3e" % r)
    print(f"   WORST relL2 {worst_l2:.3e}  maxrel {worst_mx:.3e}   (# > 2e-2 relL2: {sum(r[3] > 2e-2 for r in rows)} of {len(rows)})")


def varied_image(rng, h, w, labels, k):
    """a synthetic 'scene': per-image brightness / contrast / colour cast, per-class colour and texture inside the label blocks,
    sinusoidal shading with a per-image frequency band, noise with a per-image sigma"""
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    base = rng.uniform(40, 215, size=3)
    contrast = rng.uniform(0.15, 1.6)
    img = np.zeros((h, w, 3))
    cls_col = rng.uniform(-70, 70, size=(k + 1, 3))
    cls_freq = rng.uniform(0.05, 1.5, size=(k + 1, 2))
    lab = np.where(labels < k, labels, k).astype(np.int64)
    for c in range(3):
        img[:, :, c] = cls_col[lab, c] + 25 * np.sin(cls_freq[lab, 0] * xx + cls_freq[lab, 1] * yy + rng.uniform(0, 6.28))
        fmax = rng.choice([0.02, 0.08, 0.3])
        for _ in range(3):
            fx, fy, ph = rng.uniform(0.2 * fmax, fmax), rng.uniform(0.2 * fmax, fmax), rng.uniform(0, 6.28)
            img[:, :, c] += rng.uniform(10, 50) * np.sin(fx * xx + fy * yy + ph)
    img = base + contrast * img + rng.normal(0, rng.uniform(1, 25), size=img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def main():
    nb, h, w = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    on.SUM_MODE = "fp64"
    rng = np.random.default_rng(19)
    net = df.random_net(class_count=3, seed=21)
    labels = np.stack([synth_labels(rng, h, w, 3) for _ in range(nb)])
    if os.environ.get("DATA") == "varied":
        imgs = np.stack([varied_image(rng, h, w, labels[i], 3) for i in range(nb)])
    elif os.environ.get("DATA") == "uniform":
        imgs = rng.integers(0, 256, size=(nb, h, w, 3), dtype=np.uint8)
    else:
        imgs = np.stack([synth_image(rng, h, w) for _ in range(nb)])
    weights = np.stack([on.set_weights(l, 0.5, 0.5) for l in labels])
    x = on.to_tensor(imgs, net.layers[-1]["avg"])
    t0 = time.time()
    base = run(net, x, labels, weights, Cfg())
    print(f"baseline run: {time.time() - t0:.1f} s")
    for spec in sys.argv[4:]:
        cfg = Cfg()
        for kv in spec.split(","):
            k, v = kv.split("=")
            setattr(cfg, k, v if k == "pooltie" else int(v))
        compare(net, base, run(net, x, labels, weights, cfg, base[3]), spec)


if __name__ == "__main__":
    main()
