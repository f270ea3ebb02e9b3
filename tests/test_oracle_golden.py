"""Pins the numpy oracle (oracle/annonet_numpy.py, oracle/dlib_format.py) against the golden vectors
produced by the compiled reference (tests/golden/make_golden.py), and against the known-answer tests
the reference itself holds for this path (test/annonet_test.cpp:54-120, tiling/test/test.cpp:35-163)."""
import math

import numpy as np
import pytest

import annonet_numpy as on
import dlib_format as df
from conftest import golden_bytes, golden_json, golden_npz, parse_ref_tensors, ref_grads_by_layer, rel_err


def test_codec_roundtrip_is_byte_exact():
    for name in ("net_small.dnn", "net_small_trained.dnn", "anet_small_trained.dnn"):
        data = golden_bytes(name)
        assert df.serialize_net(df.parse_net(data)) == data, name


def test_to_runtime_matches_reference_conversion():
    net = df.parse_net(golden_bytes("net_small_trained.dnn"))
    ref = df.parse_net(golden_bytes("anet_small_trained.dnn"))
    mine = df.to_runtime(net)
    for (i, a), (_, b) in zip(mine.param_layers(), ref.param_layers()):
        np.testing.assert_allclose(a["params"], b["params"], rtol=2e-6, atol=1e-7, err_msg=f"layer<{i}>")


def test_forward_and_to_label_match_reference():
    g = golden_npz("forward.npz")
    anet = df.parse_net(golden_bytes("anet_small_trained.dnn"))
    ref = parse_ref_tensors(g["forward"])[1][0]
    x = on.to_tensor(g["image"][None], anet.layers[-1]["avg"])
    logits = on.logits_of(on.forward(anet, x, training=False))[0]
    assert rel_err(logits, ref) < 1e-5
    labels = on.to_label(ref, list(g["gains"]))
    assert np.array_equal(labels, g["process_labels"])


def test_loss_and_gradients_match_reference():
    t = golden_npz("train_inputs.npz")
    gr = golden_npz("train_grads.npz")
    net = df.parse_net(golden_bytes("net_small.dnn"))
    n = 3
    x = on.to_tensor(t["images"][:n], net.layers[-1]["avg"])
    ctx = on.forward(net, x, training=True, update_running_stats=False)
    assert rel_err(on.logits_of(ctx)[:, :, ::2, ::2], gr["logits_sub2"]) < 1e-4  # dlib sums the batch statistics in fp32
    loss, grad = on.loss_and_grad(on.logits_of(ctx), t["labels"][:n], t["weights"][:n])
    assert abs(loss - float(gr["loss"])) < 1e-6 * abs(float(gr["loss"]))
    pg = on.backward(net, ctx, grad)
    ref_grads = ref_grads_by_layer(net, gr["grads"])
    assert set(pg) == set(ref_grads)
    for i, ref in ref_grads.items():
        assert rel_err(pg[i], ref.reshape(-1)) < 2e-4, f"layer<{i}>"


def test_two_sgd_steps_match_reference():
    t = golden_npz("train_inputs.npz")
    net = df.parse_net(golden_bytes("net_small.dnn"))
    ref = df.parse_net(golden_bytes("net_small_trained.dnn"))
    for l in net.layers:
        if l["type"] == "bn":
            l["window"] = 50
    mom = {}
    n = 3
    for s in range(2):
        x = on.to_tensor(t["images"][s * n:(s + 1) * n], net.layers[-1]["avg"])
        ctx = on.forward(net, x, training=True)
        _, grad = on.loss_and_grad(on.logits_of(ctx), t["labels"][s * n:(s + 1) * n], t["weights"][s * n:(s + 1) * n])
        on.sgd_step(net, on.backward(net, ctx, grad), mom, lr=0.1)
    for (i, a), (_, b) in zip(net.param_layers(), ref.param_layers()):
        assert rel_err(a["params"], b["params"]) < 2e-4, f"layer<{i}> params"
        if a["type"] == "bn":
            # dlib accumulates the batch sums in fp32 (cpu_dlib.cpp:1127-1140); the oracle in fp64
            assert rel_err(a["running_means"], b["running_means"]) < 1e-3 or np.max(np.abs(b["running_means"])) < 1e-6, f"layer<{i}>"
            assert rel_err(a["running_variances"], b["running_variances"]) < 1e-3, f"layer<{i}>"
            assert a["num_updates"] == b["num_updates"]


def test_tiled_inference_matches_reference():
    g = golden_npz("infer_tiled.npz")
    anet = df.parse_net(golden_bytes("anet_small_trained.dnn"))
    tile = int(g["tile"])
    labels, blended = on.annonet_infer(anet, g["image"], tile, tile, return_blended=True)
    assert rel_err(blended, g["blended"]) < 1e-5
    assert np.mean(labels == g["labels"]) >= 0.999


def test_set_weights_known_answers():
    """golden values of the reference's own unit test, test/annonet_test.cpp:54-120"""
    lab = np.array([[0, 65535, 1, 0, 0]], dtype=np.uint16)
    w = on.set_weights(lab, 0.0, 0.0)
    assert list(w[0]) == [1.0, 0.0, 1.0, 1.0, 1.0] and w.sum() == 4.0
    w = on.set_weights(lab, 1.0, 0.0)
    assert abs(w[0, 0] - 0.666667) < 1e-6 and w[0, 1] == 0 and w[0, 2] == 2.0 and abs(w.sum() - 4.0) < 1e-6
    w = on.set_weights(lab, 0.5, 0.0)
    assert abs(w[0, 0] - 0.845299) < 1e-6 and abs(w[0, 2] - 0.845299 * math.sqrt(3)) < 1e-6 and abs(w.sum() - 4.0) < 1e-6
    w = on.set_weights(lab, 0.0, 1.0)
    assert list(w[0]) == [1.25, 0.0, 1.25, 1.25, 1.25] and w.sum() == 5.0


def test_set_weights_matches_reference_outputs():
    g = golden_npz("set_weights.npz")
    for i in range(4):
        cw, iw = g[f"params{i}"]
        np.testing.assert_allclose(on.set_weights(g[f"labels{i}"], cw, iw), g[f"weights{i}"], rtol=1e-6)


def test_tiles_match_reference_and_its_identities():
    t = golden_json("tiles.json")
    for key, rows in t["tiles"].items():
        dims, maxs, ox, oy = key.split("_")
        W, H = map(int, dims.split("x"))
        mw, mh = map(int, maxs.split("x"))
        tiles = on.get_tiles(W, H, mw, mh, int(ox), int(oy))
        got = [list(tl["full"]) + list(tl["non_overlapping"]) + list(tl["unique"]) for tl in tiles]
        assert got == rows, key
    # tiling/test/test.cpp: a single tile when the image fits; 3000^2 -> 2x2, 5000^2 -> 3x3 at 2048/227
    assert len(on.get_tiles(2048, 2048, 2048, 2048, 227, 227)) == 1
    assert len(on.get_tiles(3000, 3000, 2048, 2048, 227, 227)) == 4
    tiles = on.get_tiles(5000, 5000, 2048, 2048, 227, 227)
    assert len(tiles) == 9 and tiles[0]["full"][:2] == (0, 0) and tiles[-1]["full"][2:] == (4999, 4999)
    for key in ("required",):
        assert t["recommended"][key] == 127
    for m, d in t["recommended"].items():
        if m != "required":
            assert on.recommended_input_dimension(int(m)) == d
    with pytest.raises(ValueError):
        on.recommended_input_dimension(32000)


def test_adam_restatement_matches_dlib():
    """oracle adam_step against dlib's own compute_adam_update (annonet_ref adam, tests/golden/adam.npz): 4 steps on a
    parameter vector with a bias segment -- bit-exact up to the last fp32 bit of pow/sqrt"""
    g = golden_npz("adam.npz")
    n = g["params"].size
    layer = {"type": "bn", "params": g["params"].copy().reshape(1, n, 1, 1), "multipliers": [1.0, 1.0, 1.0, 1.0]}

    class FakeNet:
        def param_layers(self):
            return [(7, layer)]

    state = {}
    for s in range(g["grads"].shape[0]):
        on.adam_step(FakeNet(), {7: g["grads"][s]}, state, float(g["lr"]), float(g["wd"]), float(g["b1"]), float(g["b2"]))
    np.testing.assert_allclose(layer["params"].reshape(-1), g["out"], rtol=2e-6, atol=1e-9)
