#!/usr/bin/env python
"""bench.py -- the hot path of annonet on B200, measured.

Headline (BASELINE.json configs[1]): one annonet_train step -- 32 crops of 255 x 255 (the 64w+63 size the
CLI feeds for "256"), 3 classes, weighted per-pixel loss, SGD -- in crops/s.  A "step" is one
forward + loss + backward + update over one synthetic minibatch.

  python bench.py --gpus N --steps K --warmup W            our engine (libannonet_b200.so through the C ABI)
  python bench.py --impl reference ...                      the reference's own dlib CPU build (oracle/_ref/annonet_ref)

`value`  : device-resident inputs, CUDA events on the engine's stream, max over ranks.
`e2e`    : the same steps through ab_train_step_async with HOST buffers (H2D of crops + labels every step,
           8-byte loss read back every step), timed on the host between device synchronizes.
`roofline`: the family of kernels with the largest share of the step, timed live with CUDA event pairs
           (ab_profile_*) during the timed region; algorithmic flops/bytes as DESIGN.md defines them.
Also reported (not the headline): tiled inference of one 1024 x 1024 image (configs[0]) in Mpix/s.
"""
import argparse
import atexit
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CROP, BATCH, CLASSES = 255, 32, 3          # BASELINE configs[1]; --workload configs3 switches to 511 / 64 per GPU
WORKLOAD_TAG = "BASELINE configs[1]"
INFER_SIZE, INFER_TILE = 1024, 1024
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "annonet_ref")
FWD_FLOP_PER_PX = 85140.0  # live layers: SURVEY App. B sums to 42,570 MAC per input pixel (the survey's 170,281 double-counted)


def peaks():
    p = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            m = json.load(f)
        p.update({k: m[k] for k in ("hbm_gbs", "bf16_tflops", "bf16_tflops_sustained") if k in m})
        p["source"] = "measured"
    except Exception:
        pass
    return p


class ClockSampler:
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md's clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines, self.t0, self.t1 = index, None, [], None, None

    def start(self):
        """Started well before the timed region: nvidia-smi needs > 1 s to come up on an 8-GPU box."""
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
            atexit.register(self._kill)   # an early exit must not leave the sampler looping
        except Exception:
            self.proc = None
        return self

    def _kill(self):
        if self.proc and self.proc.poll() is None:
            self.proc.kill()

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def __enter__(self):   # the timed region
        if self.proc is None:
            self.start()
        self.t0 = time.perf_counter()
        return self

    def __exit__(self, *a):
        self.t1 = time.perf_counter()
        if self.proc:
            time.sleep(0.12)
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
            self.thread.join(timeout=2)

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # samples that arrived during the timed region (a sample is up to 20 ms old); if the region was shorter than the
        # sampling could resolve, the samples taken under the warm-up load just before it
        inside = [ln for t, ln in self.lines if self.t0 is not None and self.t0 <= t <= self.t1 + 0.03]
        window = "timed region"
        if not inside:
            inside, window = [ln for t, ln in self.lines if self.t0 is None or t >= self.t0 - 0.5], "warm-up + timed region"
        for ln in inside:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm), "window": window}


def synthetic_batch(n, h, w, k, seed):
    """seeded u8 crops (smooth sinusoids + noise so the batch-norm statistics are not degenerate), blocky
    labels with 5 % ignored pixels, weights from set_weights(0.5, 0.5) like annonet_train_main.cpp:214-232."""
    import numpy as np
    import annonet_b200 as ab
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    imgs = np.empty((n, h, w, 3), np.uint8)
    labels = np.empty((n, h, w), np.uint16)
    weights = np.empty((n, h, w), np.float32)
    for i in range(n):
        img = np.zeros((h, w, 3), np.float32)
        for c in range(3):
            for _ in range(3):
                fx, fy, ph = rng.uniform(0.01, 0.08), rng.uniform(0.01, 0.08), rng.uniform(0, 6.28)
                img[:, :, c] += rng.uniform(20, 50) * np.sin(fx * xx + fy * yy + ph)
        img += 128 + rng.normal(0, 12, size=img.shape)
        imgs[i] = np.clip(img, 0, 255).astype(np.uint8)
        lab = ((((yy.astype(np.int64) + 7 * i) // 16) + (xx.astype(np.int64) // 16)) % k).astype(np.uint16)
        lab[rng.random((h, w)) < 0.05] = 65535
        labels[i] = lab
        weights[i] = ab.set_weights(lab, 0.5, 0.5)
    return imgs, ab.pack_labels(labels, weights)


# ---- the reference's own CPU implementation (oracle/_ref/annonet_ref: unmodified dlib CPU build) --------------
def _ref(*args, timeout=3000):
    env = dict(os.environ)
    # all the host threads it can use: torchrun exports OMP_NUM_THREADS=1 for its workers, which would pin the reference's OpenMP
    # loops to one core (AB_REF_THREADS overrides)
    try:
        avail = len(os.sched_getaffinity(0))
    except AttributeError:
        avail = os.cpu_count() or 1
    env["OMP_NUM_THREADS"] = os.environ.get("AB_REF_THREADS", str(avail))
    r = subprocess.run([REF_BIN, *map(str, args)], check=True, capture_output=True, text=True, timeout=timeout, env=env)
    return r.stdout


def reference_infer(size, tile, reps=1):
    """annonet_infer() of the reference (unmodified annonet_infer.cpp + NetPimpl.cpp, dlib CPU build) on a size x size image"""
    with tempfile.TemporaryDirectory() as tmp:
        cfg, net, anet = (os.path.join(tmp, n) for n in ("cfg.net", "net.dnn", "anet.dnn"))
        _ref("config", CLASSES, 1.0, 1, cfg)
        _ref("materialize", cfg, net, 5)
        _ref("runtime", net, anet)
        return json.loads(_ref("bench_infer", anet, size, size, tile, reps).strip().splitlines()[-1])


def reference_train(sample_crops, steps):
    """StartTraining x steps + sync of the reference (NetPimpl::TrainingNet, dlib CPU backend) on
    `sample_crops` crops of CROP x CROP: a bounded sample of the 32-crop minibatch (per-crop cost is
    independent of the minibatch size on the CPU path: per-sample im2col + sgemm, cpu_dlib.cpp:2250-2355)."""
    with tempfile.TemporaryDirectory() as tmp:
        cfg, net = os.path.join(tmp, "cfg.net"), os.path.join(tmp, "net.dnn")
        _ref("config", CLASSES, 1.0, 1, cfg)
        _ref("materialize", cfg, net, 5)
        out = json.loads(_ref("bench_train", net, sample_crops, CROP, CROP, steps).strip().splitlines()[-1])
    return out


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    if not os.access(REF_BIN, os.X_OK):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/annonet_ref was not built (needs /root/reference at build time)"}))
        return 0
    sample, steps = 2, max(1, min(args.steps, 3))
    out = reference_train(sample, steps)
    v = float(out["crops_per_s"])
    line = {
        "impl": "reference", "metric": "train_crops_per_s", "value": v, "unit": "crops/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": 0, "ms_per_step": 1e3 * float(out["seconds"]) / steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"annonet_train step: {BATCH} crops {CROP}x{CROP} per GPU, {CLASSES} classes, weighted per-pixel loss, SGD ({WORKLOAD_TAG})",
                   "net": "net_type LEVEL_COUNT=6 full width, 2,279,571 params, random init"},
        "cpu_baseline": {"value": v, "unit": "crops/s", "cores": int(out.get("threads", 1)), "kind": "reference",
                         "sample": f"{steps} step(s) of {sample} crops {CROP}x{CROP} (of the {BATCH}-crop minibatch); dlib CPU build, no BLAS, OpenMP im2col"},
        "e2e": {"value": v, "unit": "crops/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ---- our engine -----------------------------------------------------------------------------------------------
def own_arm(args):
    import ctypes as C
    import numpy as np
    import torch
    import torch.distributed as dist
    import annonet_b200 as ab

    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the engine has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    clocks = ClockSampler(local).start()   # long before the timed region (nvidia-smi start-up)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    precision = ab.AB_PRECISION_F16 if args.precision == "f16" else ab.AB_PRECISION_FP32

    net = ab.TrainingNet(local, precision)
    net.Initialize(CLASSES, 1.0, 1, seed=1)         # identical weights on every rank (same seed)
    net.SetLearningRate(1e-3)
    if world > 1:                                    # engine-owned NCCL communicator; id travels through torch.distributed
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(ab.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        net.comm_init(bytes(uid.cpu().numpy().tobytes()), rank, world)

    imgs, wl = synthetic_batch(BATCH, CROP, CROP, CLASSES, seed=1234 + rank)
    h_imgs = torch.from_numpy(imgs).pin_memory()
    h_wl = torch.from_numpy(wl.view(np.uint8).reshape(BATCH, CROP, CROP, 8)).pin_memory()
    d_imgs, d_wl = h_imgs.cuda(), h_wl.cuda()
    stream = torch.cuda.ExternalStream(net.stream(), device=local)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        net.train_step_device(d_imgs.data_ptr(), d_wl.data_ptr(), BATCH, CROP, CROP)

    def step_host():
        ab._check(ab.lib.ab_train_step_async(net.handle, C.c_void_p(h_imgs.data_ptr()), C.c_void_p(h_wl.data_ptr()), BATCH, CROP, CROP))

    for _ in range(max(args.warmup, 3)):
        step_device()
    loss0 = net.sync()

    # ---- value: K steps, inputs resident in HBM; working set per step (~2.3 GB of activations) >> 126 MB L2
    launches0 = net.kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with clocks:
        e0.record(stream)
        for _ in range(args.steps):
            step_device()
        e1.record(stream)
        loss = net.sync()
        barrier()
    ms = e0.elapsed_time(e1)
    launches = net.kernel_launches() - launches0

    # ---- roofline: the same K steps again with a CUDA event pair around every launch (ab_profile_*); kept out of the
    # region above because ~800 event records per step are not free and the profiled step is issued kernel by kernel
    ab._check(ab.lib.ab_profile_enable(1))
    barrier()
    for _ in range(args.steps):
        step_device()
    net.sync()
    stats = (ab._KernelStat * 32)()
    cnt = C.c_int(0)
    ab._check(ab.lib.ab_profile_read(stats, 32, C.byref(cnt)))
    ab._check(ab.lib.ab_profile_enable(0))
    fam = [{"name": s.name.decode(), "launches": int(s.launches), "ms": s.ms, "flops": s.flops, "bytes": s.bytes}
           for s in stats[:cnt.value] if s.launches]

    # ---- e2e: host buffers through the C ABI
    for _ in range(2):
        step_host()
    net.sync()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    net.sync()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0

    t = torch.tensor([ms, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0]), float(t[1])

    # ---- secondary: tiled inference of a 1024 x 1024 image (configs[0]); rank 0 only
    infer = None
    if rank == 0 and not args.no_infer:
        infer = bench_infer(ab, net, local, precision, torch)

    # ---- secondary: tiled inference of a 16384 x 16384 image, 8 classes, tile rows sharded over the ranks (configs[2])
    big = None
    if not args.no_infer16k:
        try:
            big = bench_infer_sharded(ab, local, precision, torch, dist, rank, world)
        except Exception as e:  # secondary figure: never lose the headline line over it
            big = {"error": repr(e)[:300]}

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    pk = peaks()
    total_ms = sum(f["ms"] for f in fam) or 1.0
    top = max(fam, key=lambda f: f["ms"])
    tensor_bound = top["name"] in ("conv3x3", "conv1x1", "up_cont", "wgrad", "conv_tc", "wgrad_tc")
    avg_ms = top["ms"] / top["launches"]
    if tensor_bound:
        achieved = top["flops"] / (top["ms"] * 1e-3) / 1e12
        roof = {"bound": "tensor", "achieved": achieved, "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s",
                "frac": achieved / pk["bf16_tflops_sustained"]}
    else:
        achieved = top["bytes"] / (top["ms"] * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": achieved / pk["hbm_gbs"]}
    # DRAM traffic per launch of the dominant family: ncu dram__bytes_read.sum + dram__bytes_write.sum over every launch of one step,
    # averaged per launch like `achieved` (profiles/r01_traffic.json, written from the committed ncu csv by profiles/traffic_from_ncu.py)
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r02_traffic.json")) as f:
            tj = json.load(f)
        ent = tj.get("f16" if precision != ab.AB_PRECISION_FP32 else "fp32", {}).get(top["name"])
        if ent:
            traffic, traffic_src = ent["dram_bytes_per_launch"], tj.get("source")
    except (OSError, ValueError, KeyError):
        pass
    if tensor_bound and top["bytes"]:
        # the full-resolution layers of this net (8..64 channels) sit below the bf16 ridge (~215 FLOP/B): the same launches against HBM
        gbs = top["bytes"] / (top["ms"] * 1e-3) / 1e9
        roof["hbm_view"] = {"achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"],
                            "note": "algorithmic bytes (operands read once, result written once) / launch time"}
    roof.update({"traffic": traffic, "traffic_source": traffic_src, "kernel": top["name"], "avg_launch_ms": avg_ms, "share_of_step": top["ms"] / total_ms,
                 "peak_source": pk["source"] + (" (sustained bf16 cuBLAS)" if tensor_bound else " (copy)")})
    families = {f["name"]: {"launches_per_step": f["launches"] / args.steps, "ms_per_step": f["ms"] / args.steps,
                            "tflops": f["flops"] / (f["ms"] * 1e-3) / 1e12 if f["flops"] else None,
                            "gbs": f["bytes"] / (f["ms"] * 1e-3) / 1e9} for f in sorted(fam, key=lambda f: -f["ms"])}

    # the constant and the per-launch accounting must describe the same work: forward + backward-data + filter gradients of the
    # live layers = 3 x FWD_FLOP_PER_PX per pixel, minus the first layer's backward-data (the image needs no gradient) and the
    # border effects of the strided levels (< 1 %)
    if precision != ab.AB_PRECISION_FP32 and args.workload != "crops256":
        fam_flops = sum(f["flops"] for f in fam if f["name"] in ("conv_tc", "wgrad_tc")) / args.steps
        want = 3 * FWD_FLOP_PER_PX * CROP * CROP * BATCH
        assert 0.93 * want <= fam_flops <= 1.04 * want, f"kernel families account for {fam_flops:.3e} FLOP per step, the constant says {want:.3e}"

    cpu = None
    if infer is not None and world == 1 and not args.no_cpu and os.access(REF_BIN, os.X_OK):
        # inference cpu_baseline: the reference's annonet_infer() on a 512 x 512 sample (one tile, run at 575 x 575) -- the per-pixel
        # cost of the CPU path does not depend on the image size (per-tile im2col + GEMM), the 1024 x 1024 image would take minutes
        out = reference_infer(512, 1024, 1)
        infer["cpu_baseline"] = {"value": float(out["mpix_per_s"]), "unit": "Mpix/s", "cores": int(out.get("threads", 1)), "kind": "reference",
                                 "sample": "annonet_infer() of one 512x512 image (1 tile at 575x575) by oracle/_ref/annonet_ref: dlib CPU build, no BLAS"}
    if world == 1 and not args.no_cpu and os.access(REF_BIN, os.X_OK):
        out = reference_train(2, 1)
        cpu = {"value": float(out["crops_per_s"]), "unit": "crops/s", "cores": int(out.get("threads", 1)), "kind": "reference",
               "sample": f"1 step of 2 crops {CROP}x{CROP} (of the {BATCH}-crop minibatch) by oracle/_ref/annonet_ref: dlib CPU build, no BLAS, OpenMP im2col"}

    crops = BATCH * world * args.steps
    line = {
        "metric": "train_crops_per_s", "value": crops / (ms * 1e-3), "unit": "crops/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if precision == ab.AB_PRECISION_FP32 else "f16", "data": "synthetic",
        "config": {"workload": f"annonet_train step: {BATCH} crops {CROP}x{CROP} per GPU, {CLASSES} classes, weighted per-pixel loss, SGD ({WORKLOAD_TAG})",
                   "net": "net_type LEVEL_COUNT=6 full width, 2,279,571 params, random init", "parallelism": f"dp{world}",
                   "l2": "no flush: each step streams GBs of activations, far beyond the 126 MB L2",
                   "roofline_timing": "second pass of the same K steps with a CUDA event pair around every launch",
                   "train_tflops_algorithmic": 3 * FWD_FLOP_PER_PX * CROP * CROP * crops / (ms * 1e-3) / 1e12},
        "clocks": clocks.summary(),
        "e2e": {"value": crops / (e2e_ms * 1e-3), "unit": "crops/s", "h2d_bytes_per_step": int(h_imgs.numel() + h_wl.numel()),
                "d2h_bytes_per_step": 8, "timer": "host clock between device synchronizes, max over ranks"},
        "gpu_launches": int(launches),
        "roofline": roof, "kernel_families": families, "cpu_baseline": cpu,
        "loss": {"after_warmup": loss0, "last": loss}, "inference": infer, "inference_16k": big,
    }
    print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def bench_infer(ab, train_net, device, precision, torch, reps=3):
    """annonet_infer() equivalent on one 1024 x 1024 RGB image, 3 classes (BASELINE configs[0]): tiles ->
    forward -> blend -> argmax.  device-resident (image + label map in HBM) and e2e (host image in, host labels out)."""
    import numpy as np
    rt = train_net.GetRuntimeNet(device, precision)
    rng = np.random.default_rng(99)
    img = rng.integers(0, 256, size=(INFER_SIZE, INFER_SIZE, 3), dtype=np.uint8)
    d_img = torch.from_numpy(img).cuda()
    d_lab = torch.empty((INFER_SIZE, INFER_SIZE), dtype=torch.int16, device="cuda")
    stream = torch.cuda.ExternalStream(rt.stream(), device=device)
    for _ in range(2):
        rt.infer_device(d_img.data_ptr(), INFER_SIZE, INFER_SIZE, INFER_TILE, INFER_TILE, d_lab.data_ptr())
    rt.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(reps):
        rt.infer_device(d_img.data_ptr(), INFER_SIZE, INFER_SIZE, INFER_TILE, INFER_TILE, d_lab.data_ptr())
    e1.record(stream)
    rt.synchronize()
    ms = e0.elapsed_time(e1) / reps
    # end to end through the host API: image from pinned host memory, label map back into pinned host memory; one untimed
    # call first (the engine allocates its image / label staging buffers on the first host-side call)
    p_img = torch.from_numpy(img).pin_memory()
    p_lab = torch.empty((INFER_SIZE, INFER_SIZE), dtype=torch.int16).pin_memory()
    h_img, h_lab = p_img.numpy(), p_lab.numpy().view(np.uint16)
    rt.infer(h_img, INFER_TILE, INFER_TILE, labels_out=h_lab)
    e2e_reps = 4 * reps
    t0 = time.perf_counter()
    for _ in range(e2e_reps):
        rt.infer(h_img, INFER_TILE, INFER_TILE, labels_out=h_lab)
    e2e = (time.perf_counter() - t0) / e2e_reps
    assert np.array_equal(h_lab.view(np.int16), d_lab.cpu().numpy()), "host and device-resident inference disagree"
    npix = INFER_SIZE * INFER_SIZE
    tile = ab.recommended_input_dimension(INFER_SIZE)
    return {"workload": f"annonet_infer of one {INFER_SIZE}x{INFER_SIZE} RGB image, {CLASSES} classes, tile {INFER_TILE} (1 tile run at {tile}x{tile})",
            "mpix_per_s": npix / (ms * 1e-3) / 1e6, "ms": ms, "e2e_mpix_per_s": npix / e2e / 1e6,
            "tflops_algorithmic": FWD_FLOP_PER_PX * tile * tile / (ms * 1e-3) / 1e12}


def bench_infer_sharded(ab, device, precision, torch, dist, rank, world, size=16384, tile=1024, classes=8, reps=2):
    """BASELINE configs[2]: annonet_infer() of one size x size image, 8 classes.  Every rank makes the same ab_infer_tiled_device
    call; the engine deals the tile rows to the ranks as contiguous bands of equal cost, exchanges the seam strips between
    neighbours (ncclSend/Recv) and broadcasts the 2-byte label strips.  Device-resident image and label map;
    CUDA events on the engine's stream, max over ranks."""
    net = ab.TrainingNet(device, precision)
    net.Initialize(classes, 1.0, 1, seed=2)
    rt = net.GetRuntimeNet(device, precision)
    del net
    if world > 1:
        uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(ab.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        rt.comm_init(bytes(uid.cpu().numpy().tobytes()), rank, world)
    g = torch.Generator(device="cuda").manual_seed(7)
    d_img = torch.randint(0, 256, (size, size, 3), dtype=torch.uint8, device="cuda", generator=g)
    d_lab = torch.empty((size, size), dtype=torch.int16, device="cuda")
    stream = torch.cuda.ExternalStream(rt.stream(), device=device)

    def once():
        rt.infer_device(d_img.data_ptr(), size, size, tile, tile, d_lab.data_ptr())

    once()
    rt.synchronize()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    e0.record(stream)
    for _ in range(reps):
        once()
    e1.record(stream)
    rt.synchronize()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    return {"workload": f"annonet_infer of one {size}x{size} RGB image, {classes} classes, tile {tile} (19 x 19 tiles), tile rows sharded over {world} GPU(s) "
                        "as contiguous bands, seam strips exchanged between neighbours, label strips broadcast",
            "mpix_per_s": size * size / (ms * 1e-3) / 1e6, "ms": ms, "n_gpus": world}


# ---- BASELINE configs[4]: max-minibatch sweep at 1023 x 1023 crops against the GPU's HBM ----------------------------------
SWEEP_CROP = 1023           # the 64w + 63 size the CLI feeds for "1024" (NetPimpl.cpp:246-285)


def sweep_trial(args):
    """one trial of the sweep, in its own process like the reference's script runs the CLI once per trial
    (find_max_mini-batch_size.cmd:47-50: `%* -b N --max-total-steps=3`): 3 training steps of N crops; exit code 2 when the
    engine reports CUDA out of memory (annonet_train_main.cpp:616-620 exits 2 on any exception), 3 on any other engine limit"""
    import numpy as np
    import torch
    import annonet_b200 as ab
    n, crop = args.sweep_trial, SWEEP_CROP
    torch.cuda.set_device(0)
    free0, total = torch.cuda.mem_get_info()
    precision = ab.AB_PRECISION_F16 if args.precision == "f16" else ab.AB_PRECISION_FP32
    net = ab.TrainingNet(0, precision)
    net.Initialize(CLASSES, 1.0, 1, seed=1)
    net.SetLearningRate(1e-3)
    g = torch.Generator(device="cuda").manual_seed(5)
    try:
        d_imgs = torch.randint(0, 256, (n, crop, crop, 3), dtype=torch.uint8, device="cuda", generator=g)
        wl = np.zeros((crop, crop), dtype=ab.WLABEL)
        yy, xx = np.mgrid[0:crop, 0:crop]
        wl["label"] = ((yy // 16) + (xx // 16)) % CLASSES
        wl["weight"] = 1.0
        d_wl = torch.from_numpy(wl.view(np.uint8).reshape(crop, crop, 8)).cuda().unsqueeze(0).repeat(n, 1, 1, 1).contiguous()
        stream = torch.cuda.ExternalStream(net.stream(), device=0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        net.train_step_device(d_imgs.data_ptr(), d_wl.data_ptr(), n, crop, crop)
        loss = net.sync()
        e0.record(stream)
        for _ in range(2):
            net.train_step_device(d_imgs.data_ptr(), d_wl.data_ptr(), n, crop, crop)
        e1.record(stream)
        loss = net.sync()
        ms = e0.elapsed_time(e1) / 2
    except (ab.CudaError, torch.OutOfMemoryError) as e:
        oom = "out of memory" in str(e).lower()
        print(json.dumps({"n": n, "ok": False, "why": "cuda out of memory" if oom else str(e)[:200]}))
        return 2 if oom else 3
    except ValueError as e:    # an engine limit other than memory (AB_ERR_ARGUMENT)
        print(json.dumps({"n": n, "ok": False, "why": str(e)[:200]}))
        return 3
    free1, _ = torch.cuda.mem_get_info()
    inputs = d_imgs.numel() + d_wl.numel()
    print(json.dumps({"n": n, "ok": True, "ms_per_step": ms, "crops_per_s": n / (ms * 1e-3), "loss": loss, "hbm_used_bytes": int(free0 - free1),
                      "engine_bytes_per_sample": (free0 - free1 - inputs) / n, "hbm_total_bytes": int(total)}))
    return 0


def sweep(args):
    """find_max_mini-batch_size.cmd:1-69 restated: start at 100, double while the trial succeeds, then bisect between the
    largest success and the smallest failure; every trial is its own process"""
    trials = []

    def trial(n):
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--sweep-trial", str(n), "--precision", args.precision],
                           capture_output=True, text=True, timeout=1200)
        line = {"n": n, "ok": False, "why": f"exit {r.returncode}: {r.stderr[-200:]}"}
        for ln in r.stdout.strip().splitlines()[::-1]:
            if ln.startswith("{"):
                line = json.loads(ln)
                break
        line["exit_code"] = r.returncode
        trials.append(line)
        print(f"# trial N = {n}: {'ok' if r.returncode == 0 else 'FAILED (' + str(line.get('why')) + ')'}", file=sys.stderr, flush=True)
        return r.returncode == 0

    lo, hi, cur = 2, 0, 100
    while True:
        if trial(cur):
            lo = cur
            if hi == 0:
                cur *= 2
            elif lo + 0 >= hi:
                break
            else:
                cur = (cur + hi + 1) // 2
        else:
            hi = cur - 1
            cur = (lo + cur) // 2
            if cur == lo or hi < lo:
                break
    best = max((t for t in trials if t.get("ok")), key=lambda t: t["n"], default=None)
    first_fail = min((t for t in trials if not t.get("ok")), key=lambda t: t["n"], default=None)
    ref_bytes_per_sample = 2 * 619.3 * 4 * SWEEP_CROP * SWEEP_CROP   # dlib keeps ~619 floats/px of activations + as many gradients
    total = best["hbm_total_bytes"] if best else 0
    out = {"metric": "max_minibatch_1023x1023", "value": best["n"] if best else 0, "unit": "crops", "n_gpus": 1, "higher_is_better": True,
           "dtype": "f16" if args.precision == "f16" else "f32", "data": "synthetic",
           "config": {"workload": f"annonet_train, largest mini-batch of {SWEEP_CROP}x{SWEEP_CROP} crops that trains (3 steps) on one B200, {CLASSES} classes "
                                  "(BASELINE configs[4]; bisection of find_max_mini-batch_size.cmd, one process per trial)"},
           "best": best, "first_failure": first_fail, "limit": (first_fail or {}).get("why"),
           "reference_equivalent": {"bytes_per_sample": ref_bytes_per_sample, "max_minibatch_at_this_hbm": int(total // ref_bytes_per_sample) if total else None,
                                    "note": "dlib's executor keeps every layer output and gradient in fp32 (619 floats/px each, SURVEY §7): ~5.2 GB per 1023^2 sample"},
           "trials": trials}
    print(json.dumps(out))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default=os.environ.get("AB_PRECISION", "f16"), choices=["fp32", "f16"],
                    help="f16: tcgen05 tensor-core path (fp16 operands, split hi + lo training forward, fp32 accumulate, fp32 master weights); "
                         "fp32: FMA parity path")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-infer", action="store_true", help="skip the secondary inference measurement")
    ap.add_argument("--no-infer16k", action="store_true", help="skip the sharded 16384^2 tiled inference (BASELINE configs[2])")
    ap.add_argument("--sweep", action="store_true", help="BASELINE configs[4]: bisect the largest 1023x1023 mini-batch that fits the GPU")
    ap.add_argument("--sweep-trial", type=int, default=0, help=argparse.SUPPRESS)
    ap.add_argument("--workload", default="configs1", choices=["configs1", "configs3", "crops256"],
                    help="configs1: 32 crops of 255^2 per GPU (headline); configs3: 64 crops of 511^2 per GPU (BASELINE configs[3]); "
                         "crops256: 32 literal 256^2 crops (resized top join; gradients stop there as in the reference)")
    args = ap.parse_args()
    global CROP, BATCH, WORKLOAD_TAG
    if args.workload == "configs3":
        CROP, BATCH, WORKLOAD_TAG = 511, 64, "BASELINE configs[3]"
    elif args.workload == "crops256":
        CROP, BATCH, WORKLOAD_TAG = 256, 32, "BASELINE configs[1] with literal 256x256 crops: bilinear resize at the top join"
    if args.sweep_trial:
        return sweep_trial(args)
    if args.sweep:
        return sweep(args)
    if args.impl == "reference":
        return reference_arm(args)
    return own_arm(args)


if __name__ == "__main__":
    # The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner to stdout on the first
    # communicator): descriptor 1 goes to stderr for the whole run, and only our own print()s reach the real stdout.
    sys.stdout.flush()
    _real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = _real
    try:
        _rc = main()
    finally:
        _real.flush()
    sys.exit(_rc)
