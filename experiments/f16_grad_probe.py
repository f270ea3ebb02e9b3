"""Per-layer parameter-gradient error against the exact-sum oracle (cached by tests/conftest.py) of: the fp32 engine, the fp16
tensor-core engine with the automatic loss scale, and with 64 x that scale.  usage: f16_grad_probe.py N H W seed cache_name"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import annonet_b200 as ab
import dlib_format as df
import conftest as c

n, h, w, seed, name = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
net_bytes = c.ref_random_net(3, 5)
onet = df.parse_net(net_bytes)
imgs, labs, wts = c.train_case(n, h, w, 3, seed=seed)
exact = c.oracle_train_outputs(net_bytes, imgs, labs, wts, name + "_exact")
ref = c.ref_grads_by_layer(onet, bytes(c.ref_train_outputs(net_bytes, imgs, labs, wts, name)["grads"]))
wl = ab.pack_labels(labs, wts)
cols = {}
auto = 2.0 ** int(np.floor(np.log2(n * h * w / 8)))
for label, prec, scale in (("fp32", ab.AB_PRECISION_FP32, 0), ("f16", ab.AB_PRECISION_F16, 0), ("f16x64", ab.AB_PRECISION_F16, 64 * auto), ("f16/64", ab.AB_PRECISION_F16, auto / 64)):
    t = ab.TrainingNet(0, prec)
    t.Deserialize(net_bytes)
    if scale:
        t.set_hyper(ab.HYPER_LOSS_SCALE, scale)
    loss = t.compute_gradients(imgs, wl)
    print(label, "loss", loss, "logits max-rel vs exact", np.abs(t.logits()[:, :, ::2, ::2] - exact["logits_sub2"]).max() / np.abs(exact["logits_sub2"]).max())
    cols[label] = {i: t.layer_gradient(i) for i, _ in onet.param_layers()}
rl2 = lambda a, b: float(np.linalg.norm(a.astype(np.float64) - b) / max(np.linalg.norm(b), 1e-300))
mxr = lambda a, b: float(np.abs(a.astype(np.float64) - b).max() / max(np.abs(b).max(), 1e-300))
print("layer type | rel-L2 vs exact: " + " ".join(cols) + " reference | max-rel vs exact: same")
for i, layer in onet.param_layers():
    e = exact[f"g{i}"].astype(np.float64)
    if np.abs(e).max() == 0:
        continue
    r = ref[i].reshape(-1)
    print("layer<%d> %-5s | " % (i, layer["type"]) + " ".join("%.2e" % rl2(cols[k][i], e) for k in cols) + " %.2e | " % rl2(r, e) +
          " ".join("%.2e" % mxr(cols[k][i], e) for k in cols) + " %.2e" % mxr(r, e))
