"""Non-canonical input sizes: concat_prev_'s +-1 bilinear resize of the up-sampled tensor at the U-Net joins
(layers.h:2438,2494-2521; cpu_dlib.cpp:1735-1762).  BASELINE.json names literal 256 x 256 crops: an even-sized level pools
to (s-1)/2 and comes back as s-1, one pixel short of the skip it is concatenated with.  Forward: bilinear resize with dlib's
own index arithmetic.  Backward: the reference does NOT pass the gradient through a resized join (resize_if_needed hands the
backward pass a temporary) -- every layer behind it has an exactly zero gradient in the compiled reference, and here."""
import numpy as np
import pytest

import annonet_numpy as on
import dlib_format as df
from conftest import (golden_bytes, oracle_train_outputs, ref_grads_by_layer, ref_random_net, ref_train_outputs, rel_err, synth_image, train_case)

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("size", [(130, 150), (256, 256), (200, 127)])
def test_forward_with_resized_joins_matches_oracle(ab, size):
    h, w = size
    anet_bytes = golden_bytes("anet_small_trained.dnn")
    anet = df.parse_net(anet_bytes)
    img = synth_image(np.random.default_rng(h * 1000 + w), h, w)
    ref = on.logits_of(on.forward(anet, on.to_tensor(img[None], anet.layers[-1]["avg"]), training=False))[0]
    rt = ab.RuntimeNet(0, ab.AB_PRECISION_FP32)
    rt.Deserialize(anet_bytes)
    err = rel_err(rt.Forward(img), ref)
    print(f"{h}x{w}: fp32 forward with resized joins vs oracle {err:.2e}")
    assert err < 1e-4


def test_f16_forward_256_matches_oracle(ab):
    full = df.to_runtime(df.random_net(class_count=3, seed=21))
    img = synth_image(np.random.default_rng(256), 256, 256)
    ref = on.logits_of(on.forward(full, on.to_tensor(img[None], full.layers[-1]["avg"]), training=False, skip_dead=True))[0]
    rt = ab.RuntimeNet(0, ab.AB_PRECISION_F16)
    rt.Deserialize(df.serialize_net(full))
    err = rel_err(rt.Forward(img), ref)
    print(f"256x256: fp16 forward with a resized top join vs oracle {err:.2e}")
    assert err < 2e-2


def _check(ab, precision, net_bytes, n, h, w, name, tol):
    imgs, labs, wts = train_case(n, h, w, 3, seed=h + w)
    ref = ref_train_outputs(net_bytes, imgs, labs, wts, name)
    exact = oracle_train_outputs(net_bytes, imgs, labs, wts, name + "_exact")
    onet = df.parse_net(net_bytes)
    net = ab.TrainingNet(0, precision)
    net.Deserialize(net_bytes)
    loss = net.compute_gradients(imgs, ab.pack_labels(labs, wts))
    assert abs(loss - float(ref["loss"][0])) <= tol * abs(loss)
    assert rel_err(net.logits()[:, :, ::2, ::2], ref["logits_sub2"]) <= tol
    blocked = live = 0
    worst = 0.0
    for i, r in sorted(ref_grads_by_layer(onet, bytes(ref["grads"])).items()):
        mine = net.layer_gradient(i)
        if np.abs(r).max() == 0:
            assert np.abs(mine).max() == 0, f"layer<{i}>: zero in the reference (dead, or behind a resized join), non-zero here"
            blocked += 1
            continue
        assert np.abs(mine).max() > 0
        live += 1
        e = exact[f"g{i}"]
        worst = max(worst, float(np.linalg.norm(mine - e) / np.linalg.norm(e)))
    print(f"{name}: {live} layers with a gradient (worst rel-L2 vs exact-sum oracle {worst:.2e}), {blocked} exactly zero as in the reference")
    assert worst <= tol and blocked > 60 and live >= 8


def test_training_gradients_stop_at_a_resized_join_like_the_reference_fp32(ab):
    _check(ab, ab.AB_PRECISION_FP32, golden_bytes("net_small.dnn"), 2, 130, 150, "train_2x130x150_small", 1e-3)


def test_training_gradients_256_crops_f16(ab):
    """BASELINE's literal 256 x 256 crops on the tensor-core path (full-width net)"""
    _check(ab, ab.AB_PRECISION_F16, ref_random_net(3, 5), 2, 256, 256, "train_2x256_k3", 2e-2)
