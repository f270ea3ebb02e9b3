"""fp16 vs fp32 training path: per level / channel error of the forward activations and of the gradients (debug aid)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import annonet_b200 as ab
import dlib_format as df
from conftest import train_case

n, h, w = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
net = df.random_net(class_count=3, seed=21)
data = df.serialize_net(net)
imgs, labs, wts = train_case(n, h, w, 3, seed=19)
wl = ab.pack_labels(labs, wts)
nets = []
for prec in (ab.AB_PRECISION_FP32, ab.AB_PRECISION_F16):
    t = ab.TrainingNet(0, prec)
    t.Deserialize(data)
    print("loss", t.compute_gradients(imgs, wl))
    nets.append(t)
n32, n16 = nets
print("logits max-rel", np.abs(n16.logits() - n32.logits()).max() / np.abs(n32.logits()).max())
for l in range(7):
    a16, a32 = n16.level_tensor(l), n32.level_tensor(l)
    g16, g32 = n16.level_tensor(l, True), n32.level_tensor(l, True)
    ea = [float(np.linalg.norm(a16[:, c] - a32[:, c]) / max(np.linalg.norm(a32[:, c]), 1e-30)) for c in range(a32.shape[1])]
    eg = [float(np.linalg.norm(g16[:, c] - g32[:, c]) / max(np.linalg.norm(g32[:, c]), 1e-30)) for c in range(g32.shape[1])]
    print(f"level {l} shape {a32.shape}")
    print("  act :", " ".join("%.4f" % v for v in ea))
    print("  grad:", " ".join("%.4f" % v for v in eg))
