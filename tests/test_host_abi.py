"""CPU-side checks of the product library: it loads, exports every symbol include/annonet_b200.h
declares, its host logic (C++ dlib-stream codec, tiling, set_weights, input-size table) reproduces the
reference's golden vectors, and compute entry points fail loudly (no fallback) without a device."""
import ctypes
import math
import os
import re

import numpy as np
import pytest

import dlib_format as df
from conftest import ROOT, golden_bytes, golden_json, golden_npz


def test_library_exports_every_declared_symbol(ab):
    header = open(os.path.join(ROOT, "include", "annonet_b200.h")).read()
    declared = set(re.findall(r"\b(ab_[a-z0-9_]+)\s*\(", header))
    assert declared == set(ab.EXPORTS), declared ^ set(ab.EXPORTS)
    for name in declared:
        assert hasattr(ab.lib, name), name
    assert b"sm_100a" in ab.lib.ab_version()


def test_cpp_codec_is_byte_exact_on_reference_streams(ab):
    for name in ("net_small.dnn", "net_small_trained.dnn", "anet_small_trained.dnn"):
        data = golden_bytes(name)
        assert ab.transcode(data) == data, name


def test_cpp_runtime_conversion_matches_reference(ab):
    mine = df.parse_net(ab.transcode(golden_bytes("net_small_trained.dnn"), as_runtime=True))
    ref = df.parse_net(golden_bytes("anet_small_trained.dnn"))
    assert not mine.is_training
    for (i, a), (_, b) in zip(mine.param_layers(), ref.param_layers()):
        np.testing.assert_allclose(a["params"], b["params"], rtol=2e-6, atol=1e-7, err_msg=f"layer<{i}>")


def test_cpp_codec_rejects_garbage(ab):
    data = golden_bytes("anet_small_trained.dnn")
    with pytest.raises(ab.SerializationError):
        ab.transcode(data[:1000])
    with pytest.raises(ab.SerializationError):
        ab.transcode(data + b"\x00")
    with pytest.raises(ab.SerializationError):
        ab.transcode(b"\x01\x01\x01\x05hello")


def test_cpp_codec_rejects_hostile_tensor_headers(ab, monkeypatch):
    """a tensor header whose dimensions are negative in pairs, or whose product wraps around 64 bits, must end in a
    serialization error -- not in a huge allocation or a read past the buffer"""
    import dlib_format as df
    net = df.parse_net(golden_bytes("anet_small_trained.dnn"))
    orig = df.Writer.tensor
    for dims in ((-1, -1, 4, 4), (1 << 31, 1 << 31, 4, 1), (1 << 40, 1 << 20, 1, 1), (1 << 62, 1, 1, 1), (3, -2, 1, 1)):
        state = {"done": False}

        def hostile(self, a, dims=dims, state=state):
            if a is not None and a.size and not state["done"]:
                state["done"] = True
                self.int(2)
                for d in dims:
                    self.int(d)
                return None
            return orig(self, a)
        monkeypatch.setattr(df.Writer, "tensor", hostile)
        data = df.serialize_net(net)
        monkeypatch.setattr(df.Writer, "tensor", orig)
        assert state["done"]
        with pytest.raises(ab.SerializationError):
            ab.transcode(data)


def test_tiles_match_reference(ab):
    t = golden_json("tiles.json")
    for key, rows in t["tiles"].items():
        dims, maxs, ox, oy = key.split("_")
        W, H = map(int, dims.split("x"))
        mw, mh = map(int, maxs.split("x"))
        got = [list(tl["full"]) + list(tl["non_overlapping"]) + list(tl["unique"]) for tl in ab.get_tiles(W, H, mw, mh, int(ox), int(oy))]
        assert got == rows, key


def test_tiling_identities(ab):
    """tiling/test/test.cpp:35-163"""
    assert len(ab.get_tiles(2048, 1000, 2048, 2048, 227, 227)) == 1
    for ov in (227, 226):
        for size, n in ((3000, 2), (5000, 3)):
            tiles = ab.get_tiles(size, size, 2048, 2048, ov, ov)
            assert len(tiles) == n * n
            assert tiles[0]["full"][:2] == (0, 0) and tiles[-1]["full"][2:] == (size - 1, size - 1)
            # unique rects are disjoint from every other tile's full rect; non-overlapping rects partition the image
            cover = np.zeros((size, size), np.uint8)
            for tl in tiles:
                l, t_, r, b = tl["non_overlapping"]
                cover[t_:b + 1, l:r + 1] += 1
            assert cover.min() == 1 and cover.max() == 1
            for i, a in enumerate(tiles):
                ul, ut, ur, ub = a["unique"]
                for j, o in enumerate(tiles):
                    if i != j:
                        fl, ft, fr, fb = o["full"]
                        assert ur < fl or fr < ul or ub < ft or fb < ut
    with pytest.raises(ValueError):
        ab.get_tiles(100, 100, 100, 100, 127, 127)


def test_recommended_input_dimension(ab):
    t = golden_json("tiles.json")["recommended"]
    assert ab.required_input_dimension() == t.pop("required") == 127
    for m, d in t.items():
        assert ab.recommended_input_dimension(int(m)) == d
    with pytest.raises(RuntimeError):
        ab.recommended_input_dimension(32000)  # NetPimpl.cpp:282-284


def test_set_weights(ab):
    lab = np.array([[0, 65535, 1, 0, 0]], dtype=np.uint16)  # test/annonet_test.cpp:6-39
    assert list(ab.set_weights(lab, 0.0, 0.0)[0]) == [1.0, 0.0, 1.0, 1.0, 1.0]
    w = ab.set_weights(lab, 1.0, 0.0)
    assert abs(w[0, 0] - 0.666667) < 1e-6 and w[0, 2] == 2.0 and abs(w.sum() - 4) < 1e-6
    w = ab.set_weights(lab, 0.5, 0.0)
    assert abs(w[0, 0] - 0.845299) < 1e-6 and abs(w[0, 2] - 0.845299 * math.sqrt(3)) < 1e-6
    assert list(ab.set_weights(lab, 0.0, 1.0)[0]) == [1.25, 0.0, 1.25, 1.25, 1.25]
    assert ab.set_weights(np.full((3, 3), 65535, np.uint16), 0.5, 0.5).sum() == 0
    g = golden_npz("set_weights.npz")
    for i in range(4):
        cw, iw = g[f"params{i}"]
        np.testing.assert_allclose(ab.set_weights(g[f"labels{i}"], cw, iw), g[f"weights{i}"], rtol=1e-6)


def test_no_cpu_fallback(ab):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    with pytest.raises(ab.CudaError):
        ab.RuntimeNet()
