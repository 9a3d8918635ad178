"""Workload written for compute-sanitizer, run on the VR_CHECKED build by tools/gpu_checked.sh: __graft_entry__.smoke(), then one 160x96 frame sequence per
algorithm on the 64^3 cloud (both streams of the evaluation pass, the light-generation prefetch, tiles), the exact
evaluation variants, and a 3-band vr_group run with the peer-store exchange on one device. Results are checked against
each other (tiled == whole), so a sanitizer-clean run is also a correct one."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import __graft_entry__ as entry
import volumerenderer_b200 as vr

entry.smoke()
grid = vr.build_cloud_grid(64)
W, H = 160, 96
for algo in ("point", "sphere", "ray", "beam", "path"):
    with vr.VolumeRenderer(0) as r:
        r.set_grid(grid)
        r.resize(W, H)
        r.ubo = r.default_ubo()
        r.update_ubo(maxLights=90)
        r.set_light_gen_groups(6, 1, 1)
        for _ in range(4):  # frames 3 and 4 run with the prefetched light list
            r.draw_frame(algo)
        whole = r.read_accum()
        # one more frame as three row bands (same lights, tile arguments): tiled == whole, bit for bit
        r.update_ubo(frameCount=1)
        r.reset_lights()
        r.dispatch_light_gen(algo, (6, 1, 1))
        r.clear_image()
        r.dispatch_compute_color(algo)
        one = r.read_accum()
        r.clear_image()
        for y0, h in ((0, 40), (40, 32), (72, 24)):
            r.dispatch_compute_color(algo, (0, y0, W, h))
        assert np.array_equal(r.read_accum(), one), algo
        assert np.isfinite(whole).all()
    print("sanitize:", algo, "ok", flush=True)
for variant in (0, 1, 2, 3):
    with vr.VolumeRenderer(0) as r:
        r.set_grid(grid)
        r.resize(W, H)
        r.ubo = r.default_ubo()
        r.update_ubo(maxLights=90)
        r.set_light_gen_groups(6, 1, 1)
        r.set_eval_variant(variant)
        r.draw_frame("point")
        r.draw_frame("sphere")
        r.sync()
    print("sanitize: variant", variant, "ok", flush=True)
for exchange in ("none", "allgather"):
    with vr.RendererGroup([0, 0, 0]) as g:
        g.set_grid(grid)
        g.resize(W, H)
        g.ubo = g.default_ubo()
        g.update_ubo(maxLights=90)
        g.set_light_gen_groups(6, 1, 1)
        g.set_exchange(exchange)
        g.balance("point", rounds=1)
        for _ in range(3):
            g.draw_frame("point")
        g.sync()
        got = g.read_accum()
    with vr.VolumeRenderer(0) as r:
        r.set_grid(grid)
        r.resize(W, H)
        r.ubo = r.default_ubo()
        r.update_ubo(maxLights=90)
        r.set_light_gen_groups(6, 1, 1)
        for _ in range(3):
            r.draw_frame("point")
        assert np.array_equal(r.read_accum(), got), exchange
    print("sanitize: group", exchange, "ok", flush=True)
print("sanitize: all ok")
