"""world_size-2 checks of the multi-GPU host logic on CPU (gloo), following the engine's own protocol (DESIGN.md §6,
engine_infer.cu): the band plan comes from the library (ab_plan_bands, host only), every rank blends the tiles of its band
into a private canvas, the seam strips go to the rank that owns the rows (send/recv, added there), every rank runs the argmax
of its rows and the label strips are gathered -- the result must be the single-process annonet_infer() label map.  And the
data-parallel gradient exchange (sum all-reduce of the flat arena, 1/world folded into the update) must equal the reference's
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


def _strips(bands, rank):
    """(outgoing, incoming) seam strips of a rank as (peer, y0, y1), image rows inclusive -- engine_infer.cu's Strip lists"""
    me = bands[rank]
    live = [r for r, b in enumerate(bands) if b["t1"] > b["t0"]]
    prev = max([r for r in live if r < rank], default=-1)
    nxt = min([r for r in live if r > rank], default=-1)
    out, inc = [], []
    if me["t1"] > me["t0"]:
        if prev >= 0 and me["f0"] < me["own0"]: out.append((prev, me["f0"], me["own0"] - 1))
        if nxt >= 0 and me["f1"] > me["own1"]: out.append((nxt, me["own1"] + 1, me["f1"]))
        if prev >= 0 and bands[prev]["f1"] >= me["own0"]: inc.append((prev, me["own0"], bands[prev]["f1"]))
        if nxt >= 0 and bands[nxt]["f0"] <= me["own1"]: inc.append((nxt, bands[nxt]["f0"], me["own1"]))
    return out, inc


def _worker(rank, world, port, out):
    import annonet_b200 as ab
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        anet = df.Net(layers=[{"type": "loss"}, {"type": "con", "num_filters": K}], levels=6)
        bands = ab.plan_bands(W, H, TILE, TILE, world)
        me = bands[rank]
        _, canvas = on.annonet_infer(anet, _image(), TILE, TILE, return_blended=True, forward_fn=_fake_forward, tile_range=(me["t0"], me["t1"]))
        assert not canvas[:, :me["f0"]].any() and not canvas[:, me["f1"] + 1:].any()      # a band canvas spans f0..f1 only
        sends, recvs = _strips(bands, rank)
        reqs = [dist.isend(torch.from_numpy(np.ascontiguousarray(canvas[:, y0:y1 + 1])), peer) for peer, y0, y1 in sends]
        for peer, y0, y1 in recvs:
            t = torch.empty((K, y1 - y0 + 1, W), dtype=torch.float32)
            dist.recv(t, peer)
            canvas[:, y0:y1 + 1] += t.numpy()
        for r in reqs:
            r.wait()
        labels = np.zeros((H, W), dtype=np.uint16)
        labels[me["own0"]:me["own1"] + 1] = on.to_label(canvas[:, me["own0"]:me["own1"] + 1], ())
        for r in range(world):     # the label strips are broadcast by their owners
            t = torch.from_numpy(labels[bands[r]["own0"]:bands[r]["own1"] + 1].astype(np.int32))
            dist.broadcast(t, r)
            labels[bands[r]["own0"]:bands[r]["own1"] + 1] = t.numpy().astype(np.uint16)
        own = np.zeros_like(canvas)
        own[:, me["own0"]:me["own1"] + 1] = canvas[:, me["own0"]:me["own1"] + 1]
        t = torch.from_numpy(own)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)          # (test only: assemble the owned rows to compare the canvas too)
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


def _tile_cost(t):
    fw, fh = t["full"][2] - t["full"][0] + 1, t["full"][3] - t["full"][1] + 1
    return on.recommended_input_dimension(fw, 6) * on.recommended_input_dimension(fh, 6)


def test_bands_partition_the_tiles_and_the_rows():
    import annonet_b200 as ab
    for (w, h, tile) in ((W, H, TILE), (16384, 16384, 1024), (300, 5000, 256), (200, 200, 512), (2200, 2200, 256)):
        tiles = on.get_tiles(w, h, tile, tile, 127, 127)
        assert tiles == ab.get_tiles(w, h, tile, tile, 127, 127)
        tops = sorted({t["full"][1] for t in tiles})
        cost = [_tile_cost(t) for t in tiles]
        row_cost = {}
        for t, c in zip(tiles, cost):
            row_cost[t["full"][1]] = row_cost.get(t["full"][1], 0) + c
        for world in (1, 2, 3, 4, 8):
            bands = ab.plan_bands(w, h, tile, tile, world)
            live = [b for b in bands if b["t1"] > b["t0"]]
            assert len(live) == min(world, len(tops)) or len(live) == world      # nobody idles while there is work to deal
            assert live[0]["t0"] == 0 and live[-1]["t1"] == len(tiles)
            assert all(a["t1"] == b["t0"] for a, b in zip(live, live[1:]))  # contiguous, in order
            assert live[0]["own0"] == 0 and live[-1]["own1"] == h - 1
            assert all(a["own1"] + 1 == b["own0"] for a, b in zip(live, live[1:]))   # owned rows tile the image
            by_row = all(b["t0"] == 0 or tiles[b["t0"]]["full"][1] != tiles[b["t0"] - 1]["full"][1] for b in live)
            for b in live:
                assert b["f0"] <= b["own0"] <= b["own1"] <= b["f1"]
                assert b["f0"] == tiles[b["t0"]]["full"][1] and b["f1"] == tiles[b["t1"] - 1]["full"][3]
                # a split row belongs to the rank that holds its first tile; every band holds a whole row it owns
                first_own = next(t for t in tiles[b["t0"]:b["t1"]] if t["non_overlapping"][1] == b["own0"])
                assert first_own["full"][0] == tiles[0]["full"][0]                     # ... which starts at the left edge
                assert tiles[b["t1"] - 1]["non_overlapping"][3] == b["own1"]
                # strips only ever go to the two neighbours: nothing of this band lies above the previous band's rows or below the next's
                i = live.index(b)
                if i > 0: assert b["f0"] >= live[i - 1]["own0"]
                if i + 1 < len(live): assert b["f1"] <= live[i + 1]["own1"]
            # balance: no band costs more than the ideal share plus one unit of the plan's granularity (a tile, or a tile row)
            per = [sum(cost[b["t0"]:b["t1"]]) for b in live]
            unit = max(row_cost.values()) if by_row else max(cost)
            assert max(per) <= sum(cost) / len(live) + unit + 1e-9
        if (w, h, tile) in ((16384, 16384, 1024), (2200, 2200, 256)):     # enough rows per rank: the tile-granular plan applies
            bands = ab.plan_bands(w, h, tile, tile, 8)
            per = [sum(cost[b["t0"]:b["t1"]]) for b in bands]
            assert max(per) <= 1.03 * sum(cost) / 8, "8 ranks: busiest band within 3 % of the mean"


@pytest.mark.parametrize("world", [2, 3, 5, 8])
def test_strip_protocol_reproduces_the_one_process_label_map(world):
    """the engine's protocol for `world` ranks, run in ONE process: private band canvases, strips added at their owners, argmax of
    the owned rows -- on an image with enough tile rows that bands begin and end in the middle of a row"""
    import annonet_b200 as ab
    h = w = 2200
    img = np.random.default_rng(11).integers(0, 256, size=(h, w, 3), dtype=np.uint8)
    anet = df.Net(layers=[{"type": "loss"}, {"type": "con", "num_filters": K}], levels=6)
    ref_labels, ref_canvas = on.annonet_infer(anet, img, TILE, TILE, return_blended=True, forward_fn=_fake_forward)
    bands = ab.plan_bands(w, h, TILE, TILE, world)
    tiles = on.get_tiles(w, h, TILE, TILE, 127, 127)
    if world == 8:
        assert any(tiles[b["t0"]]["full"][0] != tiles[0]["full"][0] for b in bands[1:]), "expected bands that start inside a tile row"
    private = []
    for b in bands:
        _, c = on.annonet_infer(anet, img, TILE, TILE, return_blended=True, forward_fn=_fake_forward, tile_range=(b["t0"], b["t1"]))
        assert not c[:, :b["f0"]].any() and not c[:, b["f1"] + 1:].any()
        private.append(c)
    final = [c.copy() for c in private]
    for r in range(world):
        sends, recvs = _strips(bands, r)
        for peer, y0, y1 in sends:
            assert (r, y0, y1) in _strips(bands, peer)[1], "every send has its matching receive"
            final[peer][:, y0:y1 + 1] += private[r][:, y0:y1 + 1]
        for peer, y0, y1 in recvs:
            assert (r, y0, y1) in _strips(bands, peer)[0], "every receive has its matching send"
    labels = np.zeros((h, w), np.uint16)
    for r, b in enumerate(bands):
        rows = slice(b["own0"], b["own1"] + 1)
        assert np.max(np.abs(final[r][:, rows] - ref_canvas[:, rows])) <= 1e-5 * np.abs(ref_canvas).max()
        labels[rows] = on.to_label(final[r][:, rows], ())
    assert np.mean(labels != ref_labels) < 1e-4
