"""GPU experiment: path mode (k_path) frame times, whole frame and one row band of eight.

    python tools/path_probe.py [grid N=256] [W=1920] [H=1080] [frames=4]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import volumerenderer_b200 as vr

n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
W = int(sys.argv[2]) if len(sys.argv) > 2 else 1920
H = int(sys.argv[3]) if len(sys.argv) > 3 else 1080
frames = int(sys.argv[4]) if len(sys.argv) > 4 else 4
r = vr.VolumeRenderer(0)
r.set_grid(vr.build_cloud_grid(n))
r.resize(W, H)
r.ubo = r.default_ubo()
r.update_ubo(maxLights=1000)
out = {"version": vr.load().vr_version().decode(), "grid": n, "size": [W, H]}
for name, tile in (("whole", None), ("band_3_of_8", (0, 3 * H // 8, W, H // 8)), ("band_0_of_8", (0, 0, W, H // 8))):
    r.reset_accumulation()
    r.draw_frame("path", tile)
    r.sync()
    r.timer_start()
    for _ in range(frames):
        r.draw_frame("path", tile)
    out[name + "_ms"] = round(r.timer_stop() / frames, 4)
    out[name + "_S_eff"] = r.counters()["S_eff"]
print(json.dumps(out), flush=True)
r.close()
