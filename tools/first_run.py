"""Exploratory GPU run: microbenchmarks + one C2-shaped frame with per-pass timings (not the bench)."""
import json
import sys
import time
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import volumerenderer_b200 as vr

r = vr.VolumeRenderer(0)
print("ubench fma TFLOP/s", r.ubench(0), "rcp T/s", r.ubench(1), "point Teval/s", r.ubench(2), "ffma2 TFLOP/s", r.ubench(3),
      "packed point Teval/s", r.ubench(4), flush=True)
r.set_eval_variant(int(os.environ.get("VR_EVAL_VARIANT", "1")))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
algo = sys.argv[2] if len(sys.argv) > 2 else "point"
ml = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
gx = int(sys.argv[4]) if len(sys.argv) > 4 else 32
t = time.time()
grid = vr.build_cloud_grid(n)
print("grid", n, grid.nbytes, "built in", round(time.time() - t, 2), "s", flush=True)
r.set_grid(grid)
r.resize(1920, 1080)
r.set_max_lights_capacity(max(87500, ml))
r.ubo = r.default_ubo()
r.update_ubo(maxLights=ml)
r.set_light_gen_groups(gx, 1, 1)
for f in range(4):
    t = time.time()
    r.draw_frame(algo)
    r.sync()
    wall = time.time() - t
    tm = r.timings()
    c = r.counters()
    print(json.dumps(dict(frame=f + 1, wall_ms=round(wall * 1e3, 3), **{k: round(v, 3) for k, v in tm.items()}, **c,
                          teval_per_s=round(c["E"] / (tm["eval_ms"] * 1e-3) / 1e12, 3))), flush=True)
fc = r.read_frame_color()
print("frame colour mean", float(fc.mean()), "max", float(fc.max()))
