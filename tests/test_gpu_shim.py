"""The drop-in boundary exercised the way a user of the reference would: the reference's OWN, unmodified front ends
(annonet_train_main.cpp; annonet_infer.cpp behind a headless driver) compiled against shim/NetPimpl_b200.cpp and linked to
libannonet_b200.so (shim/Makefile), run on the GPU, and cross-checked with the reference build (oracle/_ref/annonet_ref):
  * annonet_train on a synthetic anno dataset (PNG images + *_mask.png) -> annonet.dnn + dnn_trainer state file
  * the reference's annonet_infer() over RuntimeNet::Forward of the engine == the reference's own result on the same net
  * the dnn_trainer synchronization file (trainer.h:931-964) is interchangeable in both directions, solver state included"""
import os
import subprocess
import tempfile

import numpy as np
import pytest

import dlib_format as df
from conftest import REF_BIN, ROOT, golden_bytes, golden_npz, have_ref_binary, rel_err, run_ref, synth_image
from png_util import write_anno_dataset

pytestmark = pytest.mark.gpu

TRAIN_BIN = os.path.join(ROOT, "shim", "_build", "annonet_train")
CLI_BIN = os.path.join(ROOT, "shim", "_build", "annonet_cli")
needs_bins = pytest.mark.skipif(not (os.access(TRAIN_BIN, os.X_OK) and os.access(CLI_BIN, os.X_OK) and have_ref_binary()),
                                reason="shim/_build/* or oracle/_ref/annonet_ref not built (make -C shim; make -C oracle)")


def _run(cmd, cwd=None, env=None, timeout=900):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([str(c) for c in cmd], cwd=cwd, env=e, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, f"{cmd[0]} failed ({r.returncode}):\n{r.stdout[-2000:]}\n{r.stderr[-2000:]}"
    return r.stdout


@needs_bins
def test_reference_train_cli_over_the_engine_then_reference_infer_over_the_engine():
    rng = np.random.default_rng(11)
    with tempfile.TemporaryDirectory() as tmp:
        data = os.path.join(tmp, "data")
        os.makedirs(data)
        write_anno_dataset(data, rng, 4, 300, 320, synth_image)
        out = _run([TRAIN_BIN, data, "--minibatch-size", 4, "--max-total-steps", 8, "--input-dimension-multiplier", 1.0,
                    "--initial-learning-rate", 0.05, "--save-interval", 4, "--data-loader-thread-count", 2, "--cached-image-count", 4],
                   cwd=tmp, env={"AB_PRECISION": "1"})
        assert "Now training" in out
        assert os.path.exists(os.path.join(tmp, "annonet.dnn")) and os.path.exists(os.path.join(tmp, "annonet_trainer_state_file.dat"))
        # annonet.dnn = serialize(json) << factor << serialize(anet bytes)   (annonet_train_main.cpp:557-565)
        r = df.Reader(open(os.path.join(tmp, "annonet.dnn"), "rb").read())
        json_classes, factor, anet_len = r.string(), r.float(), r.int()
        anet = bytes(r.d[r.pos:r.pos + anet_len])
        assert factor == 1.0 and len(anet) == anet_len and len(anet) > 9_000_000
        p = lambda n: os.path.join(tmp, n)
        open(p("anet.dnn"), "wb").write(anet)
        img = synth_image(rng, 330, 290)
        img.tofile(p("img.u8"))
        # the reference's annonet_infer() (unmodified annonet_infer.cpp) tile by tile over the ENGINE's RuntimeNet::Forward ...
        _run([CLI_BIN, "infer", p("anet.dnn"), p("img.u8"), 330, 290, 192, 192, p("mine.u16"), "-", "-", p("mine.f32")], env={"AB_PRECISION": "1"})
        # ... and over the reference's own dlib CPU net
        run_ref("infer", p("anet.dnn"), p("img.u8"), 330, 290, 192, 192, p("ref.u16"), "-", "-", p("ref.f32"))
        mine, ref = np.fromfile(p("mine.u16"), np.uint16), np.fromfile(p("ref.u16"), np.uint16)
        bm, br = np.fromfile(p("mine.f32"), np.float32).reshape(3, -1), np.fromfile(p("ref.f32"), np.float32).reshape(3, -1)
        err, agree = rel_err(bm, br), float(np.mean(mine == ref))
        print(f"annonet_train over the engine: 8 steps; annonet_infer() over the engine vs the reference: blended logits max-rel {err:.3e}, labels agree {agree:.5f}")
        assert err <= 2e-2
        top2 = np.sort(br, axis=0)[-2:]
        assert np.all((top2[1] - top2[0])[mine != ref] < 4e-2 * np.abs(br).max())


def _nets_equal(a, b, tol=0.0):
    na, nb = df.parse_net(open(a, "rb").read()), df.parse_net(open(b, "rb").read())
    worst = 0.0
    for (i, la), (_, lb) in zip(na.param_layers(), nb.param_layers()):
        worst = max(worst, float(np.abs(la["params"] - lb["params"]).max()))
        if la["type"] == "bn":
            worst = max(worst, float(np.abs(la["running_means"] - lb["running_means"]).max()))
            assert la["num_updates"] == lb["num_updates"]
    assert worst <= tol, worst
    return na, nb


@needs_bins
def test_trainer_state_file_is_interchangeable_with_the_reference():
    """dnn_trainer version-13 synchronization file: written by the reference and resumed by the engine, and the other way
    round.  Resuming + 0 steps must give back the stored net exactly; resuming + 1 step must move the parameters like the
    other implementation's resumed step does (which needs the stored sgd momentum)."""
    t = golden_npz("train_inputs.npz")
    with tempfile.TemporaryDirectory() as tmp:
        p = lambda n: os.path.join(tmp, n)
        open(p("net.dnn"), "wb").write(golden_bytes("net_small.dnn"))
        n, hw = 2, 191
        t["images"][:2 * n].tofile(p("i.u8")); t["labels"][:2 * n].tofile(p("l.u16")); t["weights"][:2 * n].astype(np.float32).tofile(p("w.f32"))
        t["images"][2 * n:3 * n].tofile(p("i3.u8")); t["labels"][2 * n:3 * n].tofile(p("l3.u16")); t["weights"][2 * n:3 * n].astype(np.float32).tofile(p("w3.f32"))
        for f in ("i0.u8", "l0.u16", "w0.f32"):
            open(p(f), "wb").close()   # resume + 0 steps: no samples
        eng = lambda *a: _run([CLI_BIN, *a], env={"AB_PRECISION": "0"})
        ref = lambda *a: _run([REF_BIN, *a])
        for writer, reader, tag in ((ref, eng, "reference -> engine"), (eng, ref, "engine -> reference")):
            sync = p("state_" + tag[:3])
            # two steps, state file written by the forced flush of GetRuntimeNet
            writer("train", p("net.dnn"), p("i.u8"), p("l.u16"), p("w.f32"), n, hw, hw, 0.1, 2, p("w2.dnn"), 50, p("w2a.dnn"), sync)
            assert os.path.exists(sync)
            for who, name in ((writer, "ww"), (reader, "rr")):
                # resume + 0 steps: the stored net comes back
                who("train", p("net.dnn"), p("i0.u8"), p("l0.u16"), p("w0.f32"), n, hw, hw, 0.1, 0, p(name + "0.dnn"), "-", p(name + "0a.dnn"), sync)
            _nets_equal(p("ww0.dnn"), p("w2.dnn"))
            _nets_equal(p("rr0.dnn"), p("w2.dnn"))
            # resume + 1 step on both sides from COPIES of the state (the step rewrites the file)
            for who, name in ((writer, "ww"), (reader, "rr")):
                s2 = sync + "_" + name
                open(s2, "wb").write(open(sync, "rb").read())
                who("train", p("net.dnn"), p("i3.u8"), p("l3.u16"), p("w3.f32"), n, hw, hw, 0.1, 1, p(name + "1.dnn"), "-", p(name + "1a.dnn"), s2)
            base = df.parse_net(open(p("w2.dnn"), "rb").read())
            a, b = df.parse_net(open(p("ww1.dnn"), "rb").read()), df.parse_net(open(p("rr1.dnn"), "rb").read())
            worst = 0.0
            for (i, l0), (_, la), (_, lb) in zip(base.param_layers(), a.param_layers(), b.param_layers()):
                da, db = la["params"] - l0["params"], lb["params"] - l0["params"]
                if np.abs(da).max() > 0:
                    worst = max(worst, rel_err(db, da))
            print(f"state file {tag}: resumed nets identical; one resumed step moves the parameters alike (max-rel {worst:.2e})")
            assert worst < 2e-2


def _label_maps(rng, h, w, k):
    yy, xx = np.mgrid[0:h, 0:w]
    truth = ((((yy // 23) + (xx // 31)) % (k + 1))).astype(np.uint16)
    truth[truth == k] = 0                                          # plenty of background
    truth[rng.random((h, w)) < 0.03] = 65535                       # unlabelled pixels
    pred = truth.copy()
    pred[pred == 65535] = 0
    flip = rng.random((h, w)) < 0.08
    pred[flip] = rng.integers(0, k, size=int(flip.sum()), dtype=np.uint16)
    pred[40:60, 10:90] = 0                                         # a missed region
    return truth, pred


def test_confusion_matrices_and_result_image_match_oracle(ab):
    """annonet_infer_main.cpp's evaluation: per-pixel and per-region confusion matrices (:202-272,482-491) and the result image
    (:408-418: nearest-neighbour label upscale + class colours) on the device against the oracle's restatement"""
    import annonet_numpy as on
    rng = np.random.default_rng(31)
    rt = ab.RuntimeNet(0, ab.AB_PRECISION_FP32)
    for (h, w, k) in ((200, 260, 4), (97, 131, 3), (64, 64, 8)):
        truth, pred = _label_maps(rng, h, w, k)
        pp, pr = rt.evaluate_labels(truth, pred, k)
        assert np.array_equal(pp, on.confusion_per_pixel(truth, pred, k))
        ref_pr = on.confusion_per_region(truth, pred, k)
        assert np.array_equal(pr, ref_pr), (pr, ref_pr)
        assert pr.sum() > 4
        palette = rng.integers(0, 256, size=(k, 4), dtype=np.uint8)
        for (oh, ow) in ((h, w), (2 * h + 3, 3 * w - 1), (h // 2 + 1, w // 3 + 2), (1, 1)):
            assert np.array_equal(rt.render_labels(pred, oh, ow, palette), on.render_labels(pred, oh, ow, palette))
    unlabelled = np.full((50, 60), 65535, np.uint16)
    pp, pr = rt.evaluate_labels(unlabelled, np.zeros((50, 60), np.uint16), 3)
    assert pp.sum() == 0 and pr.sum() == 0


INFER_BIN = os.path.join(ROOT, "shim", "_build", "annonet_infer")
INFER_REF_BIN = os.path.join(ROOT, "shim", "_build", "annonet_infer_reference")


@pytest.mark.skipif(not (os.access(INFER_BIN, os.X_OK) and os.access(INFER_REF_BIN, os.X_OK)), reason="shim/_build/annonet_infer* not built")
def test_reference_infer_cli_over_the_engine_matches_the_reference_cli():
    """the UNMODIFIED annonet_infer_main.cpp over the engine (shim) and over the reference's own NetPimpl.cpp: same annonet.dnn,
    same PNG dataset with ground-truth masks, gains and detection levels on the command line -- the result PNGs and the two
    printed confusion matrices must agree (a handful of near-tie pixels aside)."""
    import annonet_numpy as on
    from conftest import ref_random_net
    from png_util import read_png
    rng = np.random.default_rng(12)
    outs = {}
    with tempfile.TemporaryDirectory() as tmp:
        w = df.Writer()
        anet = ref_random_net(3, 5, training=False)
        w.string(""); w.float(1.0, 53); w.int(len(anet)); w.parts.append(anet)     # annonet_train_main.cpp:557-565
        for who, exe in (("engine", INFER_BIN), ("reference", INFER_REF_BIN)):
            d = os.path.join(tmp, who)
            os.makedirs(os.path.join(d, "data"))
            open(os.path.join(d, "annonet.dnn"), "wb").write(w.bytes())
            write_anno_dataset(os.path.join(d, "data"), np.random.default_rng(5), 2, 200, 330, synth_image)
            out = _run([exe, os.path.join(d, "data"), "-w", 256, "-h", 256, "-g", "1:0.05", "-g", "2:-0.02",
                        "--full-image-reader-thread-count", 1, "--result-image-writer-thread-count", 1], cwd=d, env={"AB_PRECISION": "1"}, timeout=1500)
            assert "Confusion matrix per pixel" in out
            mats = out[out.index("Confusion matrix per pixel"):]
            pngs = {f: read_png(os.path.join(d, "data", f)) for f in sorted(os.listdir(os.path.join(d, "data"))) if f.endswith("_result.png")}
            assert len(pngs) == 2
            outs[who] = (mats, pngs)
    (m_e, p_e), (m_r, p_r) = outs["engine"], outs["reference"]
    agree = min(float(np.mean(np.all(p_e[f] == p_r[f], axis=-1))) for f in p_r)
    print(f"annonet_infer CLI over the engine vs the reference: result PNGs agree on {agree:.5f} of the pixels")
    print(m_e)
    assert agree >= 0.999
    nums = lambda s: [float(t) for t in s.replace("%", " ").split() if t.replace(".", "", 1).isdigit()]
    ne, nr = nums(m_e), nums(m_r)
    assert len(ne) == len(nr) and len(ne) > 20
    assert sum(abs(a - b) for a, b in zip(ne, nr)) <= 0.002 * sum(nr)
