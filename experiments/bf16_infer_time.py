import sys, time, os
sys.path[:0] = [os.getcwd(), "oracle"]
import numpy as np, torch
import annonet_b200 as ab
t = ab.TrainingNet(0); t.Initialize(3, 1.0, 1, seed=1)
img = np.random.default_rng(0).integers(0, 256, (1024, 1024, 3), dtype=np.uint8)
for prec in (ab.AB_PRECISION_FP32, ab.AB_PRECISION_BF16):
    rt = t.GetRuntimeNet(0, prec)
    d_img = torch.from_numpy(img).cuda(); d_lab = torch.empty((1024, 1024), dtype=torch.int16, device="cuda")
    for _ in range(3): rt.infer_device(d_img.data_ptr(), 1024, 1024, 1024, 1024, d_lab.data_ptr())
    rt.synchronize(); t0 = time.perf_counter()
    for _ in range(10): rt.infer_device(d_img.data_ptr(), 1024, 1024, 1024, 1024, d_lab.data_ptr())
    rt.synchronize(); dt = (time.perf_counter() - t0) / 10
    print("precision", prec, "ms", dt * 1e3, "Mpix/s", 1.048576 / dt, "TFLOP/s", 170281 * 1087 * 1087 / dt / 1e12)
    if prec == ab.AB_PRECISION_BF16:
        import ctypes as C
        ab._check(ab.lib.ab_profile_enable(1))
        for _ in range(3): rt.infer_device(d_img.data_ptr(), 1024, 1024, 1024, 1024, d_lab.data_ptr())
        stats = (ab._KernelStat * 32)(); cnt = C.c_int(0)
        ab._check(ab.lib.ab_profile_read(stats, 32, C.byref(cnt)))
        for s in stats[:cnt.value]:
            if s.launches: print(s.name.decode(), s.launches, "ms/iter", s.ms / 3, "TF", s.flops / s.ms / 1e9 if s.flops else 0, "GB/s", s.bytes / s.ms / 1e6)
