"""world_size-2 checks of the multi-GPU host logic on CPU (gloo): the tile-row sharding of tiled inference
(rank r blends the tile rows r, r + world, ... into a private canvas, canvases are summed by one
all-reduce, every rank runs the argmax) must reproduce the single-process annonet_infer() result, and the data-parallel
gradient exchange (sum all-reduce of the flat arena, 1/world folded into the update) must equal the reference's
per-device mean (trainer.h:797-801)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import annonet_numpy as on
import dlib_format as df

H, W, TILE, K = 700, 900, 256, 3


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_forward(chip):
    """a cheap deterministic stand-in for the net: K planes that depend on the pixel values and on the position inside the tile"""
    c = chip.astype(np.float32)
    yy, xx = np.mgrid[0:c.shape[0], 0:c.shape[1]].astype(np.float32)
    return np.stack([c[..., 0] * 0.01 + 0.001 * yy, c[..., 1] * 0.01 - 0.002 * xx, c[..., 2] * 0.01 + 0.0005 * (xx - yy)]).astype(np.float32)


def _image():
    return np.random.default_rng(5).integers(0, 256, size=(H, W, 3), dtype=np.uint8)


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        anet = df.Net(layers=[{"type": "loss"}, {"type": "con", "num_filters": K}], levels=6)
        _, canvas = on.annonet_infer(anet, _image(), TILE, TILE, return_blended=True, forward_fn=_fake_forward, rank=rank, world=world)
        t = torch.from_numpy(canvas.copy())
        dist.all_reduce(t, op=dist.ReduceOp.SUM)          # the one exchange step of sharded inference
        labels = on.to_label(t.numpy(), ())
        # data-parallel gradient exchange on a fake arena: sum all-reduce, then 1/world inside the update
        g = torch.from_numpy(np.random.default_rng(100 + rank).normal(size=1000).astype(np.float32))
        gs = g.clone()
        dist.all_reduce(gs, op=dist.ReduceOp.SUM)
        gathered = [torch.empty_like(g) for _ in range(world)]
        dist.all_gather(gathered, g)
        if rank == 0:
            np.savez(out, canvas=t.numpy(), labels=labels, gsum=gs.numpy(), gmean=torch.stack(gathered).mean(0).numpy())
    finally:
        dist.destroy_process_group()


def test_sharded_inference_and_gradient_exchange_world2(tmp_path):
    out = str(tmp_path / "rank0.npz")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = np.load(out)
    anet = df.Net(layers=[{"type": "loss"}, {"type": "con", "num_filters": K}], levels=6)
    labels, canvas = on.annonet_infer(anet, _image(), TILE, TILE, return_blended=True, forward_fn=_fake_forward)
    # seams are sums of up to four tile contributions: the all-reduce may add them in another order (fp32 rounding only)
    assert np.max(np.abs(got["canvas"] - canvas)) <= 1e-5 * np.abs(canvas).max()
    assert np.mean(got["labels"] != labels) < 1e-4
    assert np.allclose(got["gsum"] / 2, got["gmean"], rtol=0, atol=1e-6)


def test_tile_rows_partition_the_tiles():
    tiles = on.get_tiles(W, H, TILE, TILE, 127, 127)
    tops = sorted({t["full"][1] for t in tiles})
    for world in (2, 4, 8):
        owner = [tops.index(t["full"][1]) % world for t in tiles]
        assert sorted(set(owner)) == list(range(min(world, len(tops))))
        assert sum(owner.count(r) for r in range(world)) == len(tiles)
