"""One rank of the multi-GPU checks (launched by tests/test_gpu_multi.py under torchrun, one process per GPU).
Every assertion failure exits non-zero; rank 0 prints one OK line per case."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import annonet_b200 as ab  # noqa: E402
import dlib_format as df  # noqa: E402
from conftest import synth_image, train_case  # noqa: E402


def unique_id(rank):
    uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        uid.copy_(torch.frombuffer(bytearray(ab.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(uid, 0)
    return bytes(uid.cpu().numpy().tobytes())


def gather(a):
    """all ranks' copies of a numpy array"""
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [o.cpu().numpy() for o in out]


def case_dp(rank, world, local, precision):
    """data-parallel training (trainer.h:1206-1253, tensor_tools.h:1690-1736): the all-reduced gradient is the SUM of the ranks'
    own gradients (the 1/world factor is folded into the solver), the reported loss is the mean of the ranks' losses, and after
    three steps -- first one eager, then captured graph segments -- every rank holds bit-identical parameters."""
    onet = df.random_net(class_count=3, seed=51)
    data = df.serialize_net(onet)
    imgs, labs, wts = train_case(4, 127, 127, 3, seed=300 + rank)      # a different mini-batch per rank
    wl = ab.pack_labels(labs, wts)
    layers = [i for i, _ in onet.param_layers()]

    solo = ab.TrainingNet(local, precision)
    solo.Deserialize(data)
    loss_solo = solo.compute_gradients(imgs, wl)
    g_solo = np.concatenate([solo.layer_gradient(i) for i in layers])
    g_all = gather(g_solo)
    losses = gather(np.array([loss_solo]))

    net = ab.TrainingNet(local, precision)
    net.Deserialize(data if rank == 0 else df.serialize_net(df.random_net(class_count=3, seed=52)))   # rank 0's parameters must win
    net.comm_init(unique_id(rank), rank, world)
    loss_dp = net.compute_gradients(imgs, wl)
    g_dp = np.concatenate([net.layer_gradient(i) for i in layers])
    want = np.sum(np.stack(g_all).astype(np.float64), axis=0)
    err = np.abs(g_dp - want).max() / np.abs(want).max()
    assert err < 1e-5, f"rank {rank}: all-reduced gradient differs from the sum of the ranks' gradients: {err:.3e}"
    assert abs(loss_dp - float(np.mean(losses))) < 1e-6 * abs(loss_dp), (loss_dp, losses)

    net.SetLearningRate(0.05)
    before = np.concatenate([net.layer_params(i) for i in layers])
    for _ in range(4):
        net.StartTraining(imgs, wl)
    net.sync()
    after = np.concatenate([net.layer_params(i) for i in layers])
    assert np.abs(after - before).max() > 0
    copies = gather(after)
    for r, c in enumerate(copies):
        assert np.array_equal(c, copies[0]), f"parameters of rank {r} differ from rank 0 after 4 data-parallel steps"
    # one SGD step from identical parameters with the mean gradient (solvers.h:155-185): v = -lr*wd*w - lr*g, w += v
    if rank == 0:
        print(f"OK dp precision {precision}: gradient sum err {err:.2e}, loss {loss_dp:.6f} = mean{[float(x) for x in losses]}, "
              f"parameters identical on {world} ranks after 4 steps")


def case_infer(rank, world, local, precision):
    """sharded annonet_infer(): band per rank + seam exchange + label broadcast equals the one-GPU label map"""
    rng = np.random.default_rng(77)
    anet = df.to_runtime(df.random_net(class_count=8, seed=61))
    data = df.serialize_net(anet)
    img = synth_image(rng, 1400, 1100)
    gains = [0.0, 0.05, -0.05, 0.0, 0.1, 0.0, -0.1, 0.02]
    one = ab.RuntimeNet(local, precision)
    one.Deserialize(data)
    rt = ab.RuntimeNet(local, precision)
    rt.Deserialize(data)
    rt.comm_init(unique_id(rank), rank, world)
    # Tiles run alone (AB_INFER_BATCH=1, and always on the fp32 path): a tile's logits do not depend on which rank runs it, only
    # the seams are summed in another order (float addition is not associative), so at most exact ties flip.  With the default
    # batching of equal-sized tiles (tensor-core path) a band forms other batches than the whole image does, the convolution
    # tiling follows the batch size, the fp32 accumulation order follows the tiling and fp16-rounded activations can move by an
    # ulp: labels may then differ where the two best classes are nearly tied, and nowhere else.
    batched = precision == ab.AB_PRECISION_F16 and os.environ.get("AB_INFER_BATCH", "8") != "1"
    for tile in (512, 300):
        ref_labels, ref_blend = one.infer(img, tile, tile, gains=gains, return_blended=True)
        labels = rt.infer(img, tile, tile, gains=gains)
        agree = float(np.mean(labels == ref_labels))
        copies = gather(labels.astype(np.int32))
        for c in copies:
            assert np.array_equal(c, copies[0]), "ranks returned different label maps"
        if not batched:
            assert agree >= 0.99999, f"tile {tile}: sharded labels agree on {agree:.6f} of the pixels"
        else:
            scores = ref_blend + np.asarray(gains, np.float32)[:, None, None]
            top2 = np.sort(scores, axis=0)[-2:]
            margin = (top2[1] - top2[0])[labels != ref_labels]
            bound = 5e-3 * float(np.abs(ref_blend).max())
            assert agree >= 0.999, f"tile {tile}: sharded labels agree on {agree:.6f} of the pixels"
            assert margin.size == 0 or float(margin.max()) < bound, \
                f"tile {tile}: a label differs where the one-GPU margin is {float(margin.max()):.3e} (near-tie bound {bound:.3e})"
        if rank == 0:
            print(f"OK infer precision {precision} tile {tile}: {world} ranks, labels agree with one GPU on {agree:.6f} of {labels.size} pixels")


def main():
    case = sys.argv[1]
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    precision = ab.AB_PRECISION_F16 if (len(sys.argv) < 3 or sys.argv[2] == "f16") else ab.AB_PRECISION_FP32
    {"dp": case_dp, "infer": case_infer}[case](rank, world, local, precision)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
