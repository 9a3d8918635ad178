"""Host-side multi-GPU logic: band balancing and the band exchange, the latter with world_size 2 over gloo on CPU."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_balanced_row_bands_cover_and_balance():
    from volumerenderer_b200.tiles import balanced_row_bands, band_tile
    rng = np.random.default_rng(0)
    cost = np.concatenate([np.zeros(100), rng.integers(1000, 90000, 800), np.zeros(180)]).astype(np.float64)
    for n in (1, 2, 3, 4, 8):
        bands = balanced_row_bands(cost, n)
        assert len(bands) == n and bands[0][0] == 0 and sum(h for _, h in bands) == len(cost)
        for (y0, h), (y1, _) in zip(bands, bands[1:]):
            assert y0 + h == y1 and h >= 1
        loads = [cost[y:y + h].sum() for y, h in bands]
        assert max(loads) <= cost.sum() / n + cost.max() + 1e-9
    assert band_tile((10, 5), 1920) == (0, 10, 1920, 5)
    assert balanced_row_bands(np.zeros(16), 4) == [(0, 4), (4, 4), (8, 4), (12, 4)]
    with pytest.raises(ValueError):
        balanced_row_bands(np.ones(3), 4)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, H, W):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    from volumerenderer_b200.tiles import balanced_row_bands, exchange_bands, exchange_rows_p2p, reduce_progressive_sums
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cost = np.arange(H, dtype=np.float64) + 1.0
    bands = balanced_row_bands(cost, world)
    full = torch.arange(H * W * 4, dtype=torch.float32).reshape(H, W, 4)
    img = torch.full((H, W, 4), -1.0)
    y0, h = bands[rank]
    img[y0:y0 + h] = full[y0:y0 + h]          # this rank "rendered" only its band
    exchange_bands(img, bands)
    assert torch.equal(img, full), f"rank {rank}: band exchange did not reassemble the frame"
    # the grouped send/recv exchange of the 1-channel frame colour gives the same reassembly
    fc_full = torch.arange(H * W, dtype=torch.float32).reshape(H, W)
    fc = torch.full((H, W), -1.0)
    fc[y0:y0 + h] = fc_full[y0:y0 + h]
    exchange_rows_p2p(fc, bands)
    assert torch.equal(fc, fc_full), f"rank {rank}: p2p row exchange did not reassemble the frame"
    s = torch.full((H, W, 4), float(rank + 1))
    reduce_progressive_sums(s, dst=0)
    if rank == 0:
        assert torch.equal(s, torch.full((H, W, 4), float(sum(range(1, world + 1)))))
    dist.barrier()
    dist.destroy_process_group()


def test_band_exchange_gloo_world2():
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), 37, 16), nprocs=2, join=True)


def test_band_exchange_gloo_world3():
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(3, _free_port(), 29, 8), nprocs=3, join=True)


def test_complement_tiles_and_row_costs():
    from volumerenderer_b200.tiles import balanced_row_bands, complement_tiles, row_costs
    assert complement_tiles((0, 10), 64, 40) == [(0, 10, 64, 30)]
    assert complement_tiles((10, 5), 64, 40) == [(0, 0, 64, 10), (0, 15, 64, 25)]
    assert complement_tiles((30, 10), 64, 40) == [(0, 0, 64, 30)]
    assert complement_tiles((0, 40), 64, 40) == []
    # rows without samples still cost march time: the split moves towards equal row counts as the march share grows
    samples = np.zeros(100)
    samples[40:60] = 1000.0
    b_eval = balanced_row_bands(row_costs(samples, march_ms=0.0, eval_ms=10.0) + 1e-9, 2)
    b_mix = balanced_row_bands(row_costs(samples, march_ms=10.0, eval_ms=10.0), 2)
    assert b_eval[0][1] == 50 and b_mix[0][1] == 50  # symmetric case stays in the middle
    samples[:] = 0.0
    samples[80:] = 1000.0
    heavy_only = balanced_row_bands(row_costs(samples, 0.0, 10.0) + 1e-9, 2)
    mixed = balanced_row_bands(row_costs(samples, 10.0, 10.0), 2)
    assert heavy_only[0][1] == 90 and 70 <= mixed[0][1] < 90
    c = row_costs(samples, 5.0, 10.0)
    assert abs(c.sum() - samples.sum() * 1.5) < 1e-6 * c.sum()


def test_refine_row_costs_converges():
    """feedback balancing against a hidden true cost the model does not know"""
    from volumerenderer_b200.tiles import balanced_row_bands, refine_row_costs
    rng = np.random.default_rng(3)
    H, world = 270, 8
    true = 1.0 + 40.0 * np.exp(-((np.arange(H) - 150.0) / 40.0) ** 2) + rng.random(H)
    model = np.ones(H)                      # a poor first guess: all rows equal
    bands = balanced_row_bands(model, world)
    for _ in range(4):
        rank_ms = [true[y0:y0 + h].sum() for y0, h in bands]
        model = refine_row_costs(model, bands, rank_ms)
        bands = balanced_row_bands(model, world)
    rank_ms = np.array([true[y0:y0 + h].sum() for y0, h in bands])
    assert rank_ms.max() / rank_ms.mean() < 1.08
    assert sum(h for _, h in bands) == H


def test_polish_move_takes_one_row_from_the_slowest_band():
    from volumerenderer_b200 import tiles
    bands = [(0, 10), (10, 5), (15, 5), (20, 12)]
    # slowest in the middle: the row goes to the faster neighbour, on the right side of the boundary
    assert tiles.polish_move(bands, [1.0, 1.1, 1.3, 1.2]) == [(0, 10), (10, 6), (16, 4), (20, 12)]
    assert tiles.polish_move(bands, [1.0, 1.2, 1.3, 1.1]) == [(0, 10), (10, 5), (15, 4), (19, 13)]
    # slowest at an end: only one neighbour
    assert tiles.polish_move(bands, [1.5, 1.1, 1.3, 1.2]) == [(0, 9), (9, 6), (15, 5), (20, 12)]
    assert tiles.polish_move(bands, [1.0, 1.1, 1.3, 1.6]) == [(0, 10), (10, 5), (15, 6), (21, 11)]
    for out in (tiles.polish_move(bands, [1.0, 1.1, 1.3, 1.6]), tiles.polish_move(bands, [1.5, 1.1, 1.3, 1.2])):
        assert sum(h for _, h in out) == 32 and all(out[i][0] + out[i][1] == out[i + 1][0] for i in range(3))
    assert tiles.polish_move([(0, 1), (1, 9)], [2.0, 1.0]) is None       # a one-row band cannot give a row away
    assert tiles.polish_move([(0, 10)], [1.0]) is None
