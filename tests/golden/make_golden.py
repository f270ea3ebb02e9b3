"""Generates the committed golden fixtures in tests/golden/ by running the UNMODIFIED reference
(oracle/_ref/annonet_ref, built from /root/reference by oracle/Makefile) on seeded synthetic inputs.

    make -C oracle -j8 && python tests/golden/make_golden.py

Only runnable where /root/reference exists (the build container).  Everything written here is an
output of the reference's own code; nothing is computed by the engine or by the numpy oracle.
The nets are narrow (SetNetWidth 0.25, min 4 filters) so that the fixtures stay small; full-width
parity is checked on the GPU box against the same binary (tests/test_gpu_*.py).
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.path.join(ROOT, "oracle", "_ref", "annonet_ref")


def run(*args):
    r = subprocess.run([REF, *map(str, args)], check=True, capture_output=True, text=True)
    return r.stdout


def synth_image(rng, h, w):
    """smooth low-frequency colour field + noise, so that bn statistics are not degenerate"""
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    img = np.zeros((h, w, 3))
    for c in range(3):
        for _ in range(4):
            fx, fy, ph = rng.uniform(0.01, 0.08), rng.uniform(0.01, 0.08), rng.uniform(0, 6.28)
            img[:, :, c] += rng.uniform(20, 50) * np.sin(fx * xx + fy * yy + ph)
    img += 128 + rng.normal(0, 12, size=img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def synth_labels(rng, h, w, k):
    yy, xx = np.mgrid[0:h, 0:w]
    lab = (((yy // 16) + (xx // 16)) % k).astype(np.uint16)
    lab[rng.random((h, w)) < 0.05] = 65535
    return lab


def main():
    if not os.path.exists(REF):
        sys.exit("build the reference first: make -C oracle -j8")
    rng = np.random.default_rng(20261104)
    K = 3
    with tempfile.TemporaryDirectory() as tmp:
        p = lambda name: os.path.join(tmp, name)
        g = lambda name: os.path.join(HERE, name)

        # ---- nets: configure (SetNetWidth, SetClassCount), materialise with srand(1), fold to a runtime net
        run("config", K, 0.25, 4, p("cfg.net"))
        run("materialize", p("cfg.net"), g("net_small.dnn"), 1)

        # ---- training fixtures: 2 steps x 3 crops of 191x191 (deepest level 2x2: with 127x127 crops the
        # deepest bn_ layers would normalise over 2 values per channel, which amplifies fp32 noise to percents)
        n, h, w, steps = 3, 191, 191, 2
        imgs = np.stack([synth_image(rng, h, w) for _ in range(n * steps)])
        labs = np.stack([synth_labels(rng, h, w, K) for _ in range(n * steps)])
        wts = np.zeros(labs.shape, np.float32)
        for i in range(n * steps):
            labs[i].tofile(p("l1.u16"))
            run("set_weights", p("l1.u16"), h, w, 0.5, 0.5, p("w1.f32"))
            wts[i] = np.fromfile(p("w1.f32"), np.float32).reshape(h, w)
        imgs.tofile(p("imgs.u8")); labs.tofile(p("labs.u16")); wts.tofile(p("wts.f32"))
        imgs[:n].tofile(p("imgs0.u8")); labs[:n].tofile(p("labs0.u16")); wts[:n].tofile(p("wts0.f32"))
        run("grads", g("net_small.dnn"), p("imgs0.u8"), p("labs0.u16"), p("wts0.f32"), n, h, w, p("g"))
        out = run("train", g("net_small.dnn"), p("imgs.u8"), p("labs.u16"), p("wts.f32"), n, h, w, 0.1, steps,
                  g("net_small_trained.dnn"), 50, g("anet_small_trained.dnn"))
        np.savez_compressed(g("train_inputs.npz"), images=imgs, labels=labs, weights=wts)
        loss = float(np.fromfile(p("g.loss"), np.float64)[0])
        # logits record = 36-byte header (int32 idx, 4 x int64 dims) + f32 data; keep every 2nd pixel to stay small
        logits = np.fromfile(p("g.logits"), np.float32, offset=36).reshape(n, K, h, w)[:, :, ::2, ::2].copy()
        grads = np.fromfile(p("g.grads"), np.uint8)
        np.savez_compressed(g("train_grads.npz"), loss=loss, logits_sub2=logits, grads=grads)

        # ---- inference fixtures (on the trained runtime net: non-trivial affine parameters)
        ih, iw = 127, 191
        img = synth_image(rng, ih, iw)
        img.tofile(p("img.u8"))
        run("forward", g("anet_small_trained.dnn"), p("img.u8"), ih, iw, p("fwd.bin"))
        fwd = np.fromfile(p("fwd.bin"), np.uint8)
        run("process", g("anet_small_trained.dnn"), p("img.u8"), ih, iw, p("proc.u16"), "0.0,0.05,-0.05")
        proc = np.fromfile(p("proc.u16"), np.uint16).reshape(ih, iw)
        np.savez_compressed(g("forward.npz"), image=img, forward=fwd, process_labels=proc, gains=np.array([0.0, 0.05, -0.05]))

        # ---- tiled inference: 260x230 image, 192x192 tiles -> 2x2 tiles with seams
        th, tw, tile = 230, 260, 192
        big = synth_image(rng, th, tw)
        big.tofile(p("big.u8"))
        run("infer", g("anet_small_trained.dnn"), p("big.u8"), th, tw, tile, tile, p("big.u16"), "-", "-", p("big.blend"))
        np.savez_compressed(g("infer_tiled.npz"), image=big, tile=tile,
                            labels=np.fromfile(p("big.u16"), np.uint16).reshape(th, tw),
                            blended=np.fromfile(p("big.blend"), np.float32).reshape(K, th, tw))

        # ---- set_weights on random label crops + tiling + recommended sizes
        sw = {}
        for i, (cw, iw_) in enumerate([(0.0, 0.0), (1.0, 0.0), (0.5, 0.5), (0.25, 1.0)]):
            lab = synth_labels(rng, 40, 56, 5)
            if i == 3:
                lab[:, :20] = 9  # sparse label ids: exercises the index*2+16 histogram growth
            lab.tofile(p("sw.u16"))
            run("set_weights", p("sw.u16"), 40, 56, cw, iw_, p("sw.f32"))
            sw[f"labels{i}"] = lab
            sw[f"weights{i}"] = np.fromfile(p("sw.f32"), np.float32).reshape(40, 56)
            sw[f"params{i}"] = np.array([cw, iw_])
        np.savez_compressed(g("set_weights.npz"), **sw)

        tiles = {}
        for (W, H, mw, mh, ox, oy) in [(1024, 1024, 1024, 1024, 127, 127), (3000, 3000, 2048, 2048, 227, 227),
                                       (5000, 5000, 2048, 2048, 227, 227), (5000, 5000, 2048, 2048, 226, 226),
                                       (16384, 16384, 1024, 1024, 127, 127), (16384, 16384, 4096, 4096, 127, 127),
                                       (260, 230, 192, 192, 127, 127), (1000, 700, 512, 384, 127, 127), (100, 90, 1024, 1024, 127, 127)]:
            rows = [list(map(int, line.split())) for line in run("tiles", W, H, mw, mh, ox, oy).strip().splitlines()]
            tiles[f"{W}x{H}_{mw}x{mh}_{ox}_{oy}"] = rows
        rec = {str(m): int(run("recommended", m).split()[1]) for m in (1, 127, 128, 191, 192, 256, 1024, 1087, 1088, 4096, 31999)}
        rec["required"] = int(run("recommended", 1).split()[0])
        with open(g("tiles.json"), "w") as f:
            json.dump(dict(tiles=tiles, recommended=rec), f)
        print("loss", loss, out.strip())

        # ---- dlib's own Adam arithmetic (tt::compute_adam_update) on a seeded vector: pins oracle adam_step
        arng = np.random.default_rng(5)
        n, steps = 1000, 4
        params = arng.normal(0, 0.3, n).astype(np.float32)
        grads = arng.normal(0, 1e-3, (steps, n)).astype(np.float32)
        grads[:, :50] = 0   # exact zeros: m / (sqrt(v) + eps) with v = 0
        lr, wd, b1, b2 = 0.01, 0.0005, 0.9, 0.999
        params.tofile(p("adam_p.f32")); grads.tofile(p("adam_g.f32"))
        run("adam", n, steps, repr(lr), repr(wd), repr(b1), repr(b2), p("adam_p.f32"), p("adam_g.f32"), p("adam_o.f32"))
        np.savez(g("adam.npz"), params=params, grads=grads, out=np.fromfile(p("adam_o.f32"), np.float32), lr=lr, wd=wd, b1=b1, b2=b2)


if __name__ == "__main__":
    main()
