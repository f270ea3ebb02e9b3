import sys, os
sys.path[:0] = [os.getcwd(), "oracle"]
import numpy as np, torch
import annonet_b200 as ab
t = ab.TrainingNet(0); t.Initialize(3, 1.0, 1, seed=1)
img = np.random.default_rng(0).integers(0, 256, (1024, 1024, 3), dtype=np.uint8)
rt = t.GetRuntimeNet(0, ab.AB_PRECISION_BF16)
d_img = torch.from_numpy(img).cuda(); d_lab = torch.empty((1024, 1024), dtype=torch.int16, device="cuda")
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    rt.infer_device(d_img.data_ptr(), 1024, 1024, 1024, 1024, d_lab.data_ptr())
rt.synchronize()
print("ok")
