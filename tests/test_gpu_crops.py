"""Mini-batch assembly on the device: randomly_crop_image + set_weights of annonet_train_main.cpp:110-232 /
annonet_train.h:20-83 as kernels over a device-resident dataset, against the oracle's restatement; then training from it."""
import numpy as np
import pytest

import annonet_numpy as on
import dlib_format as df
from conftest import golden_bytes, synth_image, synth_labels

pytestmark = pytest.mark.gpu


def _dataset(rng, count, h, w, k):
    imgs = [synth_image(rng, h + 13 * i, w - 7 * i) for i in range(count)]
    labs = [synth_labels(rng, h + 13 * i, w - 7 * i, k) for i in range(count)]
    return imgs, labs


def test_crops_match_oracle(ab):
    rng = np.random.default_rng(71)
    imgs, labs = _dataset(rng, 3, 300, 340, 5)
    labs[2][:, :50] = 200          # a sparse class id: the growing label_counts vector of set_weights (index*2+16)
    net = ab.TrainingNet(0, ab.AB_PRECISION_FP32)
    for im, lb in zip(imgs, labs):
        net.dataset_add(im, lb)
    dim = 127
    crops = [dict(image=0, left=20, top=30), dict(image=1, left=-40, top=-25, flip_lr=1),                # leaves the image top-left
             dict(image=1, left=260, top=250, flip_ud=1, brightness=1.37),                               # leaves it bottom-right
             dict(image=2, left=10, top=100, flip_lr=1, flip_ud=1, brightness=0.61, color_offset=(9, -14, 3)),
             dict(image=2, left=-100, top=5, color_offset=(-250, 250, 0)), dict(image=0, left=173, top=173, brightness=2.5)]
    for cw, iw in ((0.5, 0.5), (0.0, 0.0), (1.0, 0.25)):
        got_img, got_lab = net.build_crops(crops, dim, cw, iw)
        for i, c in enumerate(crops):
            img, lab, wts = on.crop_and_augment(imgs[c["image"]], labs[c["image"]], c["left"], c["top"], dim, cw, iw, bool(c.get("flip_lr")),
                                                bool(c.get("flip_ud")), c.get("brightness", 1.0), c.get("color_offset", (0, 0, 0)))
            assert np.array_equal(got_img[i], img), f"crop {i}: pixels"
            assert np.array_equal(got_lab[i]["label"], lab), f"crop {i}: labels"
            np.testing.assert_allclose(got_lab[i]["weight"], wts, rtol=2e-6, atol=0, err_msg=f"crop {i}: weights (cw {cw}, iw {iw})")
    # noise: reproducible, bounded by the level, roughly uniform, different per seed
    base, _ = net.build_crops([dict(image=0, left=20, top=30)], dim, 0.5, 0.5)
    a, _ = net.build_crops([dict(image=0, left=20, top=30, noise_level=7, seed=5)], dim, 0.5, 0.5)
    b, _ = net.build_crops([dict(image=0, left=20, top=30, noise_level=7, seed=5)], dim, 0.5, 0.5)
    c2, _ = net.build_crops([dict(image=0, left=20, top=30, noise_level=7, seed=6)], dim, 0.5, 0.5)
    d = a.astype(np.int32) - base.astype(np.int32)
    mid = (base > 7) & (base < 248)
    assert np.array_equal(a, b) and not np.array_equal(a, c2)
    assert np.abs(d).max() <= 7 and set(np.unique(d[mid])) == set(range(-7, 8)) and abs(float(d[mid].mean())) < 0.1


def test_training_from_device_crops_equals_training_from_host_crops(ab):
    """ab_train_step_crops == building the same crops on the host and calling StartTraining"""
    rng = np.random.default_rng(72)
    imgs, labs = _dataset(rng, 2, 260, 280, 3)
    data = golden_bytes("net_small.dnn")
    crops = [dict(image=i % 2, left=int(rng.integers(-30, 150)), top=int(rng.integers(-30, 130)), flip_lr=int(i % 3 == 0), brightness=1.0 + 0.1 * i)
             for i in range(4)]
    dim = 127
    nets = []
    for mode in ("device", "host"):
        net = ab.TrainingNet(0, ab.AB_PRECISION_FP32)
        net.Deserialize(data)
        net.SetLearningRate(0.05)
        if mode == "device":
            for im, lb in zip(imgs, labs):
                net.dataset_add(im, lb)
            for _ in range(3):
                net.train_step_crops(crops, dim, 0.5, 0.5)
        else:
            parts = [on.crop_and_augment(imgs[c["image"]], labs[c["image"]], c["left"], c["top"], dim, 0.5, 0.5, bool(c["flip_lr"]), False, c["brightness"])
                     for c in crops]
            bi = np.stack([p[0] for p in parts])
            bl = ab.pack_labels(np.stack([p[1] for p in parts]), np.stack([p[2] for p in parts]))
            for _ in range(3):
                net.StartTraining(bi, bl)
        nets.append((net.sync(), df.parse_net(net.Serialize())))
    (l0, n0), (l1, n1) = nets
    assert abs(l0 - l1) <= 1e-5 * abs(l1)
    for (i, a), (_, b) in zip(n0.param_layers(), n1.param_layers()):
        assert np.abs(a["params"] - b["params"]).max() <= 1e-5 * max(1.0, float(np.abs(b["params"]).max())), f"layer<{i}>"
