"""Pre-computes the compiled reference's outputs for the BASELINE-sized cases of tests/test_gpu_f16.py into
oracle/_ref/cache/ (see tests/conftest.py: the cache travels to the GPU box with the reference binary, so the GPU tests
spend their box time on the engine, not on minutes of single-socket dlib CPU convolutions).  Run after `make -C oracle`.
Every entry is keyed by a digest of its exact inputs; the tests recompute an entry that is missing or stale."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(HERE), os.path.join(os.path.dirname(os.path.dirname(HERE)), "oracle")]
import conftest as c  # noqa: E402

if __name__ == "__main__":
    net3 = c.ref_random_net(3, 5)
    for n, h, w, seed, name in ((6, 191, 255, 102, "train_6x191x255_k3"), (32, 255, 255, 101, "train_32x255_k3")):
        imgs, labs, wts = c.train_case(n, h, w, 3, seed=seed)
        out = c.ref_train_outputs(net3, imgs, labs, wts, name)
        print(name, "reference loss", float(out["loss"][0]), flush=True)
        out = c.oracle_train_outputs(net3, imgs, labs, wts, name + "_exact")
        print(name, "exact-sum oracle loss", float(out["loss"][0]), flush=True)
    anet3 = c.ref_random_net(3, 5, training=False)
    img = c.synth_image(np.random.default_rng(103), 1024, 1024)
    print("infer_1024_k3", c.ref_infer_outputs(anet3, img, 1024, "infer_1024_k3")["labels"].shape, flush=True)
    anet8 = c.ref_random_net(8, 6, training=False)
    img = c.synth_image(np.random.default_rng(104), 700, 900)
    gains = [0.0, 0.05, -0.05, 0.0, 0.1, 0.0, -0.1, 0.02]
    print("infer_700x900_k8_t512", c.ref_infer_outputs(anet8, img, 512, "infer_700x900_k8_t512", gains=gains)["labels"].shape, flush=True)
    det = [0.0, 0.30, 0.25, 0.35, 0.40, 0.20, 0.30, 0.25]
    img = c.synth_image(np.random.default_rng(105), 640, 640)
    c.ref_infer_outputs(anet8, img, 512, "infer_640_k8_t512", gains=gains)
    out = c.ref_infer_outputs(anet8, img, 512, "infer_640_k8_t512_det", gains=gains, detection_levels=det)
    print("infer_640_k8_t512_det: non-zero labels", int((out["labels"] > 0).sum()), flush=True)
    # non-canonical sizes (tests/test_gpu_resize.py)
    small = c.golden_bytes("net_small.dnn")
    for net, n, h, w, name in ((small, 2, 130, 150, "train_2x130x150_small"), (net3, 2, 256, 256, "train_2x256_k3")):
        imgs, labs, wts = c.train_case(n, h, w, 3, seed=h + w)
        c.ref_train_outputs(net, imgs, labs, wts, name)
        c.oracle_train_outputs(net, imgs, labs, wts, name + "_exact")
        print(name, flush=True)
