"""GPU experiment: time every evaluation-kernel variant on the C2 frame and compare the frame colour with variant 1."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import volumerenderer_b200 as vr

algo = sys.argv[1] if len(sys.argv) > 1 else "point"
ml = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
gx = int(sys.argv[3]) if len(sys.argv) > 3 else 32
variants = [int(v) for v in sys.argv[4].split(",")] if len(sys.argv) > 4 else [2, 8, 1, 3, 0]
near = float(sys.argv[5]) if len(sys.argv) > 5 else None
r = vr.VolumeRenderer(0)
grid = vr.build_cloud_grid(256)
r.set_grid(grid)
r.resize(1920, 1080)
r.set_max_lights_capacity(max(87500, ml))
r.ubo = r.default_ubo()
r.update_ubo(maxLights=ml)
r.set_light_gen_groups(gx, 1, 1)
ref = None
for v in variants:
    r.set_eval_variant(v)
    if near is not None:
        r.set_eval_near_distance(near)
    r.reset_accumulation()
    ms = []
    for f in range(5):
        r.reset_accumulation()          # same frame (frameCount 1) every time
        r.draw_frame(algo)
        r.sync()
        ms.append(r.timings()["eval_ms"])
    fc = r.read_frame_color()
    c = r.counters()
    if ref is None:
        ref = fc
    denom = np.maximum(np.abs(ref), 1e-6)
    rel = float(np.max(np.abs(fc - ref) / denom))
    extra = r.eval_segment_counts() if v == 8 else {}
    rmse = float(np.sqrt(np.mean((fc.astype(np.float64) - ref) ** 2)))
    print(json.dumps(dict(variant=v, rmse_vs_first=rmse, **extra, eval_ms=[round(x, 3) for x in ms], best=round(min(ms), 3),
                          teval_per_s=round(c["E"] / (min(ms) * 1e-3) / 1e12, 3), E=c["E"], bitwise_equal=bool(np.array_equal(fc, ref)),
                          max_rel_vs_first=rel, redo_ctas=r.eval_redo_count())), flush=True)
