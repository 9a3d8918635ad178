"""GPU experiment: one row band of the C2 frame (what one of N GPUs renders) on a single GPU: per-pass times."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import volumerenderer_b200 as vr

y0 = int(sys.argv[1]) if len(sys.argv) > 1 else 465
h = int(sys.argv[2]) if len(sys.argv) > 2 else 69
variant = int(sys.argv[3]) if len(sys.argv) > 3 else 8
r = vr.VolumeRenderer(0)
r.set_grid(vr.build_cloud_grid(256))
r.resize(1920, 1080)
r.ubo = r.default_ubo()
r.update_ubo(maxLights=1000)
r.set_light_gen_groups(32, 1, 1)
r.set_eval_variant(variant)
for f in range(6):
    r.draw_frame("point", (0, y0, 1920, h))
    r.sync()
    t = r.timings()
    print(json.dumps({k: round(v, 4) for k, v in t.items()}), r.counters()["S_eff"], flush=True)
