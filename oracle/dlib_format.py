"""TEST INFRASTRUCTURE ONLY (oracle).  Never imported by the product path.

Python restatement of the reference's serialized-net format, used by the tests and the
numpy oracle to read the nets written by the reference (`oracle/_ref/annonet_ref`) and to
hand identical parameters to the reference and to the CUDA engine.  The product's own
reader/writer is the C++ one in annonet_b200/csrc/dlib_stream.cpp; the two are
cross-checked byte-for-byte in tests/.

Follows (all under /root/reference/dlib/dlib):
  integers        serialize.h:297-345   control byte = #bytes | 0x80 if negative, LE magnitude
  bool            serialize.h:130-131   ASCII '1' / '0'
  float/double    float_details.h:86-135, serialize.h:598-632   int64 mantissa, int16 exponent
  std::string     serialize.h:1209-1221  int length + bytes
  tensor          cuda/tensor.h:444-466  int 2, 4 x long long, raw LE f32
  alias_tensor    cuda/tensor.h:657-665  int 1, 4 x long long
  add_loss_layer  dnn/core.h:2675-2681   add_layer core.h:1056-1068 (v2) / :1425-1438 (v3, input side)
  add_tag_layer   dnn/core.h:1662-1667
  con_ layers.h:234-252, cont_ :559-577, max_pool_ :1017-1026, bn_ :1428-1448, affine_ :2231-2239,
  mish_ :3344-3347, concat_prev_ :2463-2466, concat_ :3675-3680, input_rgb_image input.h:119-125
and the layer list of dlib-dnn-pimpl-wrapper/NetStructure.h:25-299 (net_type / anet_type).
"""
from __future__ import annotations

import math
import struct
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

import numpy as np

LOSS_NAME = "loss_multiclass_log_per_pixel_weighted_"

# ----------------------------------------------------------------------------------------------
# primitives


class Reader:
    def __init__(self, data: bytes, pos: int = 0):
        self.d = memoryview(data)
        self.pos = pos

    def int(self) -> int:
        c = self.d[self.pos]
        n = c & 0x0F
        neg = c & 0x80
        if n == 0 or n > 8:
            raise ValueError(f"bad integer control byte {c:#x} at {self.pos}")
        v = int.from_bytes(self.d[self.pos + 1:self.pos + 1 + n], "little")
        self.pos += 1 + n
        return -v if neg else v

    def bool(self) -> bool:
        c = self.d[self.pos]
        self.pos += 1
        if c == ord("1"):
            return True
        if c == ord("0"):
            return False
        raise ValueError(f"bad bool byte {c:#x} at {self.pos - 1}")

    def float(self) -> float:
        m = self.int()
        e = self.int()
        if e == 32000:
            return math.inf
        if e == 32001:
            return -math.inf
        if e == 32002:
            return math.nan
        return math.ldexp(m, e)

    def string(self) -> str:
        n = self.int()
        s = bytes(self.d[self.pos:self.pos + n]).decode("latin-1")
        self.pos += n
        return s

    def tensor(self) -> np.ndarray:
        v = self.int()
        if v != 2:
            raise ValueError(f"tensor version {v} at {self.pos}")
        dims = [self.int() for _ in range(4)]
        n = dims[0] * dims[1] * dims[2] * dims[3]
        a = np.frombuffer(self.d, dtype="<f4", count=n, offset=self.pos).reshape(dims).copy()
        self.pos += 4 * n
        return a

    def alias(self) -> List[int]:
        v = self.int()
        if v != 1:
            raise ValueError(f"alias_tensor version {v} at {self.pos}")
        return [self.int() for _ in range(4)]


class Writer:
    def __init__(self):
        self.parts: List[bytes] = []

    def int(self, v: int) -> None:
        neg = 0x80 if v < 0 else 0
        v = abs(int(v))
        n = max(1, (v.bit_length() + 7) // 8)
        self.parts.append(bytes([n | neg]) + v.to_bytes(n, "little"))

    def bool(self, v: bool) -> None:
        self.parts.append(b"1" if v else b"0")

    def float(self, v: float, digits: int) -> None:
        """digits = 24 for a C++ float, 53 for a double (float_details.h:110-127)."""
        if v == math.inf:
            m, e = 0, 32000
        elif v == -math.inf:
            m, e = 0, 32001
        elif v != v:
            m, e = 0, 32002
        else:
            fr, ex = math.frexp(v)
            m = int(fr * (1 << digits))
            e = ex - digits
            for _ in range(8):
                if m & 0xFF:
                    break
                m >>= 8
                e += 8
        self.int(m)
        self.int(e)

    def string(self, s: str) -> None:
        b = s.encode("latin-1")
        self.int(len(b))
        self.parts.append(b)

    def tensor(self, a: Optional[np.ndarray]) -> None:
        self.int(2)
        if a is None or a.size == 0:
            dims = (0, 0, 0, 0) if a is None else tuple(a.shape)
            for d in dims:
                self.int(d)
            return
        assert a.ndim == 4
        for d in a.shape:
            self.int(d)
        self.parts.append(np.ascontiguousarray(a, dtype="<f4").tobytes())

    def alias(self, dims) -> None:
        self.int(1)
        for d in dims:
            self.int(d)

    def bytes(self) -> bytes:
        return b"".join(self.parts)


# ----------------------------------------------------------------------------------------------
# the static layer list (NetStructure.h), index 0 = loss ... last = input


FEATURES = {0: 16, 1: 16, 2: 24, 3: 32, 4: 40, 5: 48, 6: 56}
DEEPEST = {0: 16, 1: 24, 2: 32, 3: 40, 4: 48, 5: 56, 6: 64}
DENSE_KIND = {1: 1, 2: 2, 3: 2, 4: 3, 5: 3, 6: 4}  # level -> dense_blockN (NetStructure.h:142-147)

# tag ids (NetStructure.h:27-32, 87-92)
DINTAG, DOTAG0, DCTAG0, DOTAG1, DCTAG1, DOTAG2 = 2000, 2001, 2002, 2003, 2004, 2005


def UTAG(n):
    return 2100 + n


def _con(n, k, dead=False):
    return dict(type="con", num_filters=n, nr=k, nc=k, stride=1, pad=k // 2, dead=dead)


def _cont(n, stride):
    return dict(type="cont", num_filters=n, nr=3, nc=3, stride=stride, pad=0 if stride != 1 else 1)


def _bn():
    return dict(type="bn")


def _mish():
    return dict(type="mish")


def _tag(i):
    return dict(type="tag", id=i)


def _cprev(i):
    return dict(type="concat_prev", tag=i)


def _concat(*ids):
    return dict(type="concat", tags=list(ids))


def _dense_layer(n, dead=False):
    return [_bn(), _mish(), _con(n, 3, dead)]


def _dense(kind, n):
    """dense_block{1..4}<N,BN,SUBNET>, outermost first (NetStructure.h:40-66)."""
    if kind == 1:
        return [_cprev(DINTAG)] + _dense_layer(n) + [_tag(DINTAG)]
    if kind == 2:
        return ([_cprev(DOTAG0)] + _dense_layer(n) + [_cprev(DINTAG), _tag(DOTAG0)] + _dense_layer(n) + [_tag(DINTAG)])
    if kind == 3:
        return ([_concat(DOTAG0, DOTAG1)] + _dense_layer(n, dead=True)
                + [_cprev(DCTAG0), _tag(DOTAG1)] + _dense_layer(n)
                + [_tag(DCTAG0), _cprev(DINTAG), _tag(DOTAG0)] + _dense_layer(n) + [_tag(DINTAG)])
    if kind == 4:
        return ([_concat(DOTAG0, DOTAG1, DOTAG2)] + _dense_layer(n, dead=True)
                + [_cprev(DCTAG1), _tag(DOTAG2)] + _dense_layer(n)
                + [_tag(DCTAG1), _cprev(DCTAG0), _tag(DOTAG1)] + _dense_layer(n)
                + [_tag(DCTAG0), _cprev(DINTAG), _tag(DOTAG0)] + _dense_layer(n) + [_tag(DINTAG)])
    raise ValueError(kind)


def net_layers(levels: int = 6, class_count: int = 2) -> List[Dict[str, Any]]:
    """Layer list of net_type/anet_type (NetStructure.h:172-299), loss first, input last."""
    L: List[Dict[str, Any]] = [dict(type="loss")]
    L.append(_con(class_count, 1))
    L += [_mish(), _bn(), _cont(FEATURES[0], 1)]                       # level0t
    for n in range(1, levels + 1):                                      # level1t ... level6t
        L += _dense(DENSE_KIND[n], FEATURES[n]) + [_cprev(UTAG(n)), _cont(FEATURES[n], 2)]
    L += _dense(4, DEEPEST[levels])                                     # deepest
    for n in range(levels, 0, -1):                                      # level6 ... level1
        L += [_bn(), _mish(), _con(FEATURES[n], 1), dict(type="max_pool")]   # down
        L += [_tag(UTAG(n))] + _dense(DENSE_KIND[n], FEATURES[n])
    L += [_mish(), _bn(), _con(FEATURES[0], 3)]                         # level0
    L.append(dict(type="input"))
    return [dict(l) for l in L]


# ----------------------------------------------------------------------------------------------
# whole-net records


@dataclass
class Net:
    """A parsed net_type (training, bn_) or anet_type (runtime, affine_) stream.

    layers[i] mirrors dlib's layer<i>(net): 0 = loss, last = input.  Parameter-bearing layers
    carry their tensors as numpy arrays in dlib's own layout.
    """
    layers: List[Dict[str, Any]] = field(default_factory=list)
    levels: int = 6

    @property
    def is_training(self) -> bool:
        return any(l["type"] == "bn" for l in self.layers)

    @property
    def class_count(self) -> int:
        return self.layers[1]["num_filters"]

    def param_layers(self):
        """(index, layer) of parameter-bearing layers in visit order (output side first)."""
        return [(i, l) for i, l in enumerate(self.layers) if l["type"] in ("con", "cont", "bn", "affine")]

    def num_params(self) -> int:
        return sum(int(l["params"].size) for _, l in self.param_layers())


def _read_trailer(r: Reader, layer: Dict[str, Any], input_side: bool) -> None:
    layer["setup_called"] = r.bool()
    layer["grad_stale"] = r.bool()
    layer["getters_disabled"] = r.bool()
    layer["x_grad"] = r.tensor()
    layer["cached_output"] = r.tensor()
    if input_side:
        layer["grad_final"] = r.tensor()
        layer["sample_expansion_factor"] = r.int()
    else:
        layer["params_grad"] = r.tensor()


def _write_trailer(w: Writer, layer: Dict[str, Any], input_side: bool) -> None:
    w.bool(layer.get("setup_called", True))
    w.bool(layer.get("grad_stale", True))
    w.bool(layer.get("getters_disabled", False))
    w.tensor(layer.get("x_grad"))
    w.tensor(layer.get("cached_output"))
    if input_side:
        w.tensor(layer.get("grad_final"))
        w.int(layer.get("sample_expansion_factor", 1))
    else:
        w.tensor(layer.get("params_grad"))


def _read_details(r: Reader, layer: Dict[str, Any]) -> None:
    t = layer["type"]
    name = r.string()
    if t in ("con", "cont"):
        want = "con_5" if t == "con" else "cont_2"
        if name != want:
            raise ValueError(f"expected {want}, found {name!r} at {r.pos}")
        layer["params"] = r.tensor()
        layer["num_filters"] = r.int()
        nr, nc, sy, sx, py, px = (r.int() for _ in range(6))
        if (nr, nc, sy, sx, py, px) != (layer["nr"], layer["nc"], layer["stride"], layer["stride"], layer["pad"], layer["pad"]):
            raise ValueError(f"{name}: geometry mismatch {(nr, nc, sy, sx, py, px)} vs static {layer}")
        layer["filters_alias"] = r.alias()
        layer["biases_alias"] = r.alias()
        layer["multipliers"] = [r.float() for _ in range(4)]  # lr, wd, bias lr, bias wd
        layer["use_bias"] = r.bool()
    elif t in ("bn", "affine"):
        if name == "bn_con2":
            layer["type"] = "bn"
            layer["params"] = r.tensor()
            layer["gamma_alias"] = r.alias()
            layer["beta_alias"] = r.alias()
            layer["means"] = r.tensor()
            layer["invstds"] = r.tensor()
            layer["running_means"] = r.tensor()
            layer["running_variances"] = r.tensor()
            layer["num_updates"] = r.int()
            layer["window"] = r.int()
            layer["multipliers"] = [r.float() for _ in range(4)]
            layer["eps"] = r.float()
        elif name == "affine_":
            layer["type"] = "affine"
            layer["params"] = r.tensor()
            layer["gamma_alias"] = r.alias()
            layer["beta_alias"] = r.alias()
            layer["mode"] = r.int()
        else:
            raise ValueError(f"expected bn_con2/affine_, found {name!r} at {r.pos}")
    elif t == "mish":
        if name != "mish_":
            raise ValueError(f"expected mish_, found {name!r}")
    elif t == "max_pool":
        if name != "max_pool_2":
            raise ValueError(f"expected max_pool_2, found {name!r}")
        layer["geometry"] = [r.int() for _ in range(6)]
    elif t == "concat_prev":
        if name != "concat_prev_":
            raise ValueError(f"expected concat_prev_, found {name!r}")
    elif t == "concat":
        if name != "concat_":
            raise ValueError(f"expected concat_, found {name!r}")
        n = r.int()
        if n != len(layer["tags"]):
            raise ValueError("concat_ tag count mismatch")
    else:
        raise ValueError(t)


def _write_details(w: Writer, layer: Dict[str, Any]) -> None:
    t = layer["type"]
    if t in ("con", "cont"):
        w.string("con_5" if t == "con" else "cont_2")
        w.tensor(layer.get("params"))
        w.int(layer["num_filters"])
        for v in (layer["nr"], layer["nc"], layer["stride"], layer["stride"], layer["pad"], layer["pad"]):
            w.int(v)
        w.alias(layer.get("filters_alias", [0, 0, 0, 0]))
        w.alias(layer.get("biases_alias", [0, 0, 0, 0]))
        for v in layer.get("multipliers", [1.0, 1.0, 1.0, 0.0]):
            w.float(v, 53)
        w.bool(layer.get("use_bias", True))
    elif t == "bn":
        w.string("bn_con2")
        w.tensor(layer.get("params"))
        w.alias(layer.get("gamma_alias", [0, 0, 0, 0]))
        w.alias(layer.get("beta_alias", [0, 0, 0, 0]))
        w.tensor(layer.get("means"))
        w.tensor(layer.get("invstds"))
        w.tensor(layer.get("running_means"))
        w.tensor(layer.get("running_variances"))
        w.int(layer.get("num_updates", 0))
        w.int(layer.get("window", 100))
        for v in layer.get("multipliers", [1.0, 0.0, 1.0, 1.0]):
            w.float(v, 53)
        w.float(layer.get("eps", 1e-4), 53)
    elif t == "affine":
        w.string("affine_")
        w.tensor(layer.get("params"))
        w.alias(layer.get("gamma_alias", [0, 0, 0, 0]))
        w.alias(layer.get("beta_alias", [0, 0, 0, 0]))
        w.int(layer.get("mode", 0))
    elif t == "mish":
        w.string("mish_")
    elif t == "max_pool":
        w.string("max_pool_2")
        for v in layer.get("geometry", [3, 3, 2, 2, 0, 0]):
            w.int(v)
    elif t == "concat_prev":
        w.string("concat_prev_")
    elif t == "concat":
        w.string("concat_")
        w.int(len(layer["tags"]))
    else:
        raise ValueError(t)


def parse_net(data: bytes, levels: int = 6) -> Net:
    """Recursive-descent parse: versions on the way down, payloads on the way back up."""
    layers = net_layers(levels)
    r = Reader(data)
    n = len(layers)
    # way down
    if r.int() != 1:
        raise ValueError("add_loss_layer version")
    if r.string() != LOSS_NAME:
        raise ValueError("unexpected loss layer name")
    for i in range(1, n - 1):
        v = r.int()
        t = layers[i]["type"]
        want = 1 if t == "tag" else (3 if i == n - 2 else 2)
        if v != want:
            raise ValueError(f"layer<{i}> ({t}): version {v}, expected {want} at byte {r.pos}")
    # input layer
    name = r.string()
    if name != "input_rgb_image":
        raise ValueError(f"unexpected input layer {name!r}")
    layers[-1]["avg"] = [r.float() for _ in range(3)]
    # way back up
    for i in range(n - 2, 0, -1):
        if layers[i]["type"] == "tag":
            continue
        _read_details(r, layers[i])
        _read_trailer(r, layers[i], input_side=(i == n - 2))
    if r.pos != len(data):
        raise ValueError(f"{len(data) - r.pos} trailing bytes after the net")
    return Net(layers=layers, levels=levels)


def serialize_net(net: Net) -> bytes:
    w = Writer()
    L = net.layers
    n = len(L)
    w.int(1)
    w.string(LOSS_NAME)
    for i in range(1, n - 1):
        w.int(1 if L[i]["type"] == "tag" else (3 if i == n - 2 else 2))
    w.string("input_rgb_image")
    for v in L[-1].get("avg", [122.782, 117.001, 104.298]):
        w.float(float(np.float32(v)), 24)
    for i in range(n - 2, 0, -1):
        if L[i]["type"] == "tag":
            continue
        _write_details(w, L[i])
        _write_trailer(w, L[i], input_side=(i == n - 2))
    return w.bytes()


# ----------------------------------------------------------------------------------------------
# construction helpers (used to create full-width nets without the reference binary)


def _sub_channels(layers: List[Dict[str, Any]]) -> List[int]:
    """Output channel count of every layer (index-aligned), following dlib's executor."""
    n = len(layers)
    k = [0] * n
    k[n - 1] = 3

    def tagged(i, tag_id):
        for j in range(i + 1, n):
            if layers[j]["type"] == "tag" and layers[j]["id"] == tag_id:
                return j
        raise ValueError(f"tag {tag_id} not found below layer {i}")

    for i in range(n - 2, 0, -1):
        t = layers[i]["type"]
        if t in ("con", "cont"):
            k[i] = layers[i]["num_filters"]
        elif t == "concat_prev":
            k[i] = k[i + 1] + k[tagged(i, layers[i]["tag"])]
        elif t == "concat":
            k[i] = sum(k[tagged(i, tid)] for tid in layers[i]["tags"])
        else:
            k[i] = k[i + 1]
    return k


def random_net(class_count: int, seed: int, levels: int = 6, scaler: float = 1.0, min_filters: int = 1,
               randomize_bn: bool = True) -> Net:
    """A training net (bn_) with Glorot-uniform filters like `con_::setup` (layers.h:170-193,
    dnn/utilities.h:33-47) but drawn from numpy's RNG; optionally non-trivial bn gamma/beta/
    running stats so that every term of the arithmetic is exercised."""
    rng = np.random.default_rng(seed)
    layers = net_layers(levels, class_count)
    n = len(layers)
    for i, l in enumerate(layers):
        if l["type"] in ("con", "cont") and i != 1:
            l["num_filters"] = max(min_filters, int(round(scaler * l["num_filters"])))
    k = _sub_channels(layers)
    for i in range(n - 2, 0, -1):
        l = layers[i]
        t = l["type"]
        if t in ("con", "cont"):
            cin, cout, ks = k[i + 1], l["num_filters"], l["nr"]
            num_inputs = ks * ks * cin
            lim = math.sqrt(6.0 / (num_inputs + cout))
            p = rng.uniform(-lim, lim, size=num_inputs * cout + cout).astype(np.float32)
            p[num_inputs * cout:] = rng.uniform(-0.1, 0.1, size=cout).astype(np.float32) if randomize_bn else 0
            l["params"] = p.reshape(-1, 1, 1, 1)
            l["filters_alias"] = [cout, cin, ks, ks] if t == "con" else [cin, cout, ks, ks]
            l["biases_alias"] = [1, cout, 1, 1]
            l["multipliers"] = [1.0, 1.0, 1.0, 0.0]
            l["use_bias"] = True
        elif t == "bn":
            c = k[i + 1]
            if randomize_bn:
                gamma = rng.uniform(0.5, 1.5, size=c)
                beta = rng.uniform(-0.2, 0.2, size=c)
                rm = rng.uniform(-0.2, 0.2, size=c)
                rv = rng.uniform(0.5, 1.5, size=c)
            else:
                gamma, beta, rm, rv = np.ones(c), np.zeros(c), np.zeros(c), np.ones(c)
            l["params"] = np.concatenate([gamma, beta]).astype(np.float32).reshape(-1, 1, 1, 1)
            l["gamma_alias"] = [1, c, 1, 1]
            l["beta_alias"] = [1, c, 1, 1]
            l["means"] = np.zeros((0, 0, 0, 0), np.float32)
            l["invstds"] = np.zeros((0, 0, 0, 0), np.float32)
            l["running_means"] = rm.astype(np.float32).reshape(1, c, 1, 1)
            l["running_variances"] = rv.astype(np.float32).reshape(1, c, 1, 1)
            l["num_updates"] = 0
            l["window"] = 100
            l["multipliers"] = [1.0, 0.0, 1.0, 1.0]
            l["eps"] = 1e-4
        elif t == "max_pool":
            l["geometry"] = [3, 3, 2, 2, 0, 0]
        if t != "tag":
            l["setup_called"] = True
            l["grad_stale"] = True
            l["getters_disabled"] = False
            empty = np.zeros((0, 0, 0, 0), np.float32)
            l["x_grad"] = empty
            l["cached_output"] = empty
            if i == n - 2:
                l["grad_final"] = empty
                l["sample_expansion_factor"] = 1
            else:
                l["params_grad"] = empty
    layers[-1]["avg"] = [float(np.float32(122.782)), float(np.float32(117.001)), float(np.float32(104.298))]
    return Net(layers=layers, levels=levels)


def to_runtime(net: Net) -> Net:
    """anet_type from net_type: bn_ -> affine_ with the running statistics folded in
    (layers.h:2144-2163: gamma' = gamma/sqrt(running_var+eps), beta' = beta - gamma'*running_mean)."""
    out = Net(layers=[dict(l) for l in net.layers], levels=net.levels)
    for l in out.layers:
        if l["type"] != "bn":
            continue
        c = l["gamma_alias"][1]
        p = l["params"].reshape(-1)
        gamma, beta = p[:c], p[c:]
        rm = l["running_means"].reshape(-1)
        rv = l["running_variances"].reshape(-1)
        g = (gamma / np.sqrt(rv + np.float32(l["eps"]))).astype(np.float32)
        b = (beta - g * rm).astype(np.float32)
        for key in ("means", "invstds", "running_means", "running_variances", "num_updates", "window", "multipliers", "eps"):
            l.pop(key, None)
        l["type"] = "affine"
        l["params"] = np.concatenate([g, b]).astype(np.float32).reshape(-1, 1, 1, 1)
        l["mode"] = 0  # CONV_MODE
    return out
