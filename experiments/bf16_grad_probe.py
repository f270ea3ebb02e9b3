"""bf16 vs fp32 training path: per-layer gradient error and batch-statistics error (forward activation error per depth)."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import annonet_b200 as ab
import dlib_format as df
from conftest import synth_image, synth_labels

n, h, w = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
rng = np.random.default_rng(19)
net = df.random_net(class_count=3, seed=21)
data = df.serialize_net(net)
imgs = np.stack([synth_image(rng, h, w) for _ in range(n)])
labels = np.stack([synth_labels(rng, h, w, 3) for _ in range(n)])
weights = np.stack([ab.set_weights(l, 0.5, 0.5) for l in labels])
wl = ab.pack_labels(labels, weights)
nets = []
for prec in (ab.AB_PRECISION_FP32, ab.AB_PRECISION_BF16):
    t = ab.TrainingNet(0, prec)
    t.Deserialize(data)
    t.compute_gradients(imgs, wl)
    nets.append(t)
n32, n16 = nets
g = {}
for i, layer in net.param_layers():
    a, b = n16.layer_gradient(i), n32.layer_gradient(i)
    if np.abs(b).max() == 0:
        continue
    nb = layer["num_filters"] if layer["type"] in ("con", "cont") else b.size // 2
    g[i] = float(np.linalg.norm(a[:-nb] - b[:-nb]) / np.linalg.norm(b[:-nb]))
s32, s16 = df.parse_net(n32.Serialize()), df.parse_net(n16.Serialize())
print(f"n {n} h {h} w {w}: layer, type, grad relL2 (weights part), running-mean err / rms(mean), running-var rel err")
for (i, a), (_, b) in zip(s16.param_layers(), s32.param_layers()):
    line = f"layer<{i}> {a['type']:5s} grad {g.get(i, float('nan')):.3e}"
    if a["type"] == "bn":
        rm_a, rm_b = np.asarray(a["running_means"]).ravel(), np.asarray(b["running_means"]).ravel()
        rv_a, rv_b = np.asarray(a["running_variances"]).ravel(), np.asarray(b["running_variances"]).ravel()
        line += f"  mean err/std {np.max(np.abs(rm_a - rm_b) / np.sqrt(rv_b)):.3e}  var rel {np.max(np.abs(rv_a - rv_b) / rv_b):.3e}"
    print(line)

print("per level / channel: relative L2 error of activations | gradients (bf16 vs fp32 path)")
for l in range(7):
    a16, a32 = n16.level_tensor(l), n32.level_tensor(l)
    g16, g32 = n16.level_tensor(l, True), n32.level_tensor(l, True)
    ea = [float(np.linalg.norm(a16[:, c] - a32[:, c]) / max(np.linalg.norm(a32[:, c]), 1e-30)) for c in range(a32.shape[1])]
    eg = [float(np.linalg.norm(g16[:, c] - g32[:, c]) / max(np.linalg.norm(g32[:, c]), 1e-30)) for c in range(g32.shape[1])]
    print(f"level {l} shape {a32.shape}")
    print("  act :", " ".join("%.3f" % v for v in ea))
    print("  grad:", " ".join("%.3f" % v for v in eg))
    # border vs interior for the gradient
    if g32.shape[2] > 8:
        inner = (slice(None), slice(None), slice(3, -3), slice(3, -3))
        d = g16 - g32
        tot = np.linalg.norm(d) / np.linalg.norm(g32)
        inn = np.linalg.norm(d[inner]) / np.linalg.norm(g32[inner])
        print(f"  grad whole {tot:.4f} interior {inn:.4f}")
