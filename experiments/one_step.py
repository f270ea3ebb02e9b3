"""two training steps of BASELINE configs[1] (32 x 255 x 255, full-width net) for ncu captures; AB_NO_GRAPH=1 makes every
kernel an individual launch.  usage: one_step.py [n h w]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
os.environ.setdefault("AB_NO_GRAPH", "1")
import annonet_b200 as ab
from conftest import train_case

n, h, w = (int(x) for x in sys.argv[1:4]) if len(sys.argv) > 3 else (32, 255, 255)
imgs, labs, wts = train_case(n, h, w, 3, seed=1)
net = ab.TrainingNet(0, ab.AB_PRECISION_F16)
net.Initialize(3, 1.0, 1, seed=1)
net.SetLearningRate(1e-3)
wl = ab.pack_labels(labs, wts)
for _ in range(2):
    net.StartTraining(imgs, wl)
print("loss", net.sync())
