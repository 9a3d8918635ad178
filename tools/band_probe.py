"""GPU experiment (one GPU): what each of N GPUs would spend on its row band of a workload's frame.

    python tools/band_probe.py [workload=C2] [N=8] [rounds=6] [side=1] [check=0]

The frame is split into N row bands exactly like bench.py does at N > 1 (tiles.row_costs from one full frame, then measured
feedback), but every band is rendered on THIS GPU, one after the other, and timed with the per-pass CUDA events. The
slowest band plus the exchange is what an N-GPU step costs; the spread between bands is the imbalance. Prints one JSON
line per feedback round: per-band march/eval/total ms, their max and mean, and the full frame's time for the ratio.
check=1: afterwards the N bands of one frame (same light list) are compared with the whole-frame render bit for bit.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import volumerenderer_b200 as vr
from volumerenderer_b200 import tiles

WORKLOADS = {"C2": dict(n=256, algo="point", ml=1000, gx=32), "C3": dict(n=256, algo="ray", ml=4096, gx=128),
             # BASELINE.json's C5: 4K frame through the 1024^3 cloud, "image-tiled across 8 B200"
             "C5S": dict(n=1024, algo="sphere", ml=1000, gx=32, W=3840, H=2160),
             "C5P": dict(n=1024, algo="path", ml=1000, gx=32, W=3840, H=2160)}


def band_ms(r, algo, tile, frames=3):
    r.reset_accumulation()
    for _ in range(2):
        r.draw_frame(algo, tile)
    r.sync()
    r.timings_sum(reset=True)
    r.timer_start()
    for _ in range(frames):
        r.draw_frame(algo, tile)
    dev = r.timer_stop() / frames
    ts, nf = r.timings_sum(reset=True)
    return dict(march=ts["march_ms"] / nf, eval=ts["eval_ms"] / nf, resolve=ts["resolve_ms"] / nf, total=ts["total_ms"] / nf, step=dev)


def main():
    wl = WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "C2"]
    N = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    rounds = int(sys.argv[3]) if len(sys.argv) > 3 else 6
    side = int(sys.argv[4]) if len(sys.argv) > 4 else 1
    check = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    W, H, algo = wl.get("W", 1920), wl.get("H", 1080), wl["algo"]
    r = vr.VolumeRenderer(0)
    r.set_eval_side_stream(side)
    r.set_grid(vr.build_cloud_grid(wl["n"]))
    r.resize(W, H)
    r.set_max_lights_capacity(max(87500, wl["ml"]))
    r.ubo = r.default_ubo()
    r.update_ubo(maxLights=wl["ml"])
    r.set_light_gen_groups(wl["gx"], 1, 1)
    full = band_ms(r, algo, None)
    print(json.dumps({"full_frame": {k: round(v, 4) for k, v in full.items()}, "ideal_per_band": round(full["step"] / N, 4)}), flush=True)
    cost = tiles.row_costs(r.row_sample_counts(), full["march"], full["eval"])
    bands = tiles.balanced_row_bands(cost, N)
    for rnd in range(rounds):
        res = [band_ms(r, algo, tiles.band_tile(b, W)) for b in bands]
        steps = [x["step"] for x in res]
        print(json.dumps({"round": rnd, "bands": bands, "step_ms": [round(s, 4) for s in steps],
                          "eval_ms": [round(x["eval"], 4) for x in res], "march_ms": [round(x["march"], 4) for x in res],
                          "max": round(max(steps), 4), "mean": round(float(np.mean(steps)), 4),
                          "speedup_at_max": round(full["step"] / max(steps), 3)}), flush=True)
        best = (max(steps), bands) if rnd == 0 or max(steps) < best[0] else best
        cost = tiles.refine_row_costs(cost, bands, steps)
        bands = tiles.balanced_row_bands(cost, N)
    if check:
        bands = best[1]
        r.update_ubo(frameCount=1)
        if algo != "path":
            r.reset_lights()
            r.dispatch_light_gen(algo, (wl["gx"], 1, 1))
        r.clear_image()
        r.dispatch_compute_color(algo)
        whole = (r.read_frame_color(), r.read_accum(), r.read_image_rgba8())
        info = r.march_info()
        r.clear_image()
        for b in bands:
            r.dispatch_compute_color(algo, tiles.band_tile(b, W))
        tiled = (r.read_frame_color(), r.read_accum(), r.read_image_rgba8())
        print(json.dumps({"check": "bands of one frame vs the whole-frame render", "bands": bands,
                          "whole_frame_march": {k: (int(v) if not isinstance(v, bool) else v) for k, v in info.items()},
                          "frame_color_equal": bool(np.array_equal(whole[0], tiled[0])), "accum_equal": bool(np.array_equal(whole[1], tiled[1])),
                          "rgba8_equal": bool(np.array_equal(whole[2], tiled[2])), "mean": float(whole[0].mean())}), flush=True)
    r.close()


if __name__ == "__main__":
    main()
