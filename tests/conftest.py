import json
import os
import struct
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "annonet_ref")
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden_path(name):
    return os.path.join(GOLDEN, name)


def golden_bytes(name):
    with open(golden_path(name), "rb") as f:
        return f.read()


def golden_npz(name):
    return np.load(golden_path(name))


def golden_json(name):
    with open(golden_path(name)) as f:
        return json.load(f)


def parse_ref_tensors(buf):
    """records written by oracle/ref_io.h: int32 layer index, int64 n,k,nr,nc, then f32 data."""
    buf = bytes(buf)
    out, pos = {}, 0
    while pos < len(buf):
        idx, n, k, nr, nc = struct.unpack_from("<iqqqq", buf, pos)
        pos += 36
        cnt = n * k * nr * nc
        out[idx] = np.frombuffer(buf, dtype="<f4", count=cnt, offset=pos).reshape(n, k, nr, nc).copy()
        pos += 4 * cnt
    return out


def ref_grads_by_layer(net, buf):
    """visit_layer_parameter_gradients (dlib/dlib/dnn/core.h:3806-3816) numbers the computational layers
    (add_layer instances: everything but loss, tags and input) from the output side; map that index to
    dlib's layer<i> index and keep the parameter-bearing ones."""
    comp = [i for i, l in enumerate(net.layers) if l["type"] not in ("loss", "tag", "input")]
    out = {}
    for j, t in parse_ref_tensors(buf).items():
        if t.size:
            out[comp[j]] = t
    return out


def rel_err(a, b):
    """max |a-b| relative to the largest reference magnitude (the north-star's "relative" tolerance)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    denom = max(float(np.max(np.abs(b))), 1e-30)
    return float(np.max(np.abs(a - b))) / denom


def have_ref_binary():
    return os.path.exists(REF_BIN) and os.access(REF_BIN, os.X_OK)


def run_ref(*args, timeout=1800):
    r = subprocess.run([REF_BIN, *map(str, args)], check=True, capture_output=True, text=True, timeout=timeout)
    return r.stdout


def synth_image(rng, h, w):
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float64)
    img = np.zeros((h, w, 3))
    for c in range(3):
        for _ in range(4):
            fx, fy, ph = rng.uniform(0.01, 0.08), rng.uniform(0.01, 0.08), rng.uniform(0, 6.28)
            img[:, :, c] += rng.uniform(20, 50) * np.sin(fx * xx + fy * yy + ph)
    img += 128 + rng.normal(0, 12, size=img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def synth_labels(rng, h, w, k, ignore=0.05):
    yy, xx = np.mgrid[0:h, 0:w]
    lab = (((yy // 16) + (xx // 16)) % k).astype(np.uint16)
    lab[rng.random((h, w)) < ignore] = 65535
    return lab


@pytest.fixture(scope="session")
def ab():
    import annonet_b200
    return annonet_b200
