"""Offline statistic for DESIGN.md section 10 (CPU only, oracle lookups): which share of a frame's camera samples lies in aligned chunks of 32 / 16
consecutive samples that are also consecutive march steps (no zero-density gap inside)? C2 geometry, every 16th x 9th pixel."""
import sys, math
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
import numpy as np
from oracle import pyoracle
import volumerenderer_b200 as vr
o = pyoracle.Oracle()
g = vr.build_cloud_grid(256)
info = o.grid_info(g)
bmin, bmax = info['bbox_min'].astype(np.float64), info['bbox_max'].astype(np.float64) + 1.0
W, H = 1920, 1080
sx, sy = 16, 9     # every 16th / 9th pixel -> 120 x 120 rays
tan = math.tan(45.0 * 0.5 * math.pi / 180.0)
cam = np.array([0.0, 20.0, -75.0])
o_idx = o.world_to_index(g, cam.astype(np.float32)).astype(np.float64)
# linear part of world->index from two probes
e = np.eye(3, dtype=np.float32)
Minv = np.stack([o.world_to_index(g, (cam + e[i]).astype(np.float32)).astype(np.float64) - o_idx for i in range(3)], axis=1)
tot = 0; hist = np.zeros(5, np.int64)  # samples in: leftovers(<8), 8-runs only, 16-chunks contiguous, 32-chunks contiguous
n_samples = 0; in32 = 0; in16 = 0; in32_any = 0
for py in range(sy // 2, H, sy):
    for px in range(sx // 2, W, sx):
        x = ((2.0 * (px + 0.5)) / W - 1.0) * (W / H) * tan
        y = (1.0 - (2.0 * (py + 0.5)) / H) * tan
        d = np.array([x, y, 1.0]); d /= np.linalg.norm(d)
        di = Minv @ d; di /= np.linalg.norm(di)
        with np.errstate(divide='ignore', invalid='ignore'):
            t0 = (bmin - o_idx) / di; t1 = (bmax - o_idx) / di
        tn = np.nanmax(np.minimum(t0, t1)); tf = np.nanmin(np.maximum(t0, t1))
        tn = max(tn, 0.0); tf = min(tf, 2500.0)
        if not (tf > tn): continue
        ts = tn + np.arange(int((tf - tn)) + 1, dtype=np.float64)
        ts = ts[ts < tf]
        p = o_idx[None, :] + ts[:, None] * di[None, :]
        ijk = np.floor(p).astype(np.int32)
        val = o.lookup(g, ijk)
        # transmittance cut-off
        T = np.exp(-np.cumsum(np.concatenate([[0.0], val[:-1]])) * 0.05 * 1.0)
        live = T > 0.001
        idx = np.nonzero((val != 0) & live)[0]
        c = len(idx); n_samples += c
        # entries: consecutive samples in chunks of 32 / 16 from the start of the pixel's list
        for k in range(0, c - 31, 32):
            if idx[k + 31] - idx[k] == 31: in32 += 32
        for k in range(0, c - 15, 16):
            if idx[k + 15] - idx[k] == 15: in16 += 16
print('samples', n_samples, 'in contiguous aligned 32-chunks %.3f' % (in32 / n_samples), 'in contiguous aligned 16-chunks %.3f' % (in16 / n_samples))
