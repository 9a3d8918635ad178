#!/usr/bin/env python3
"""Every BASELINE.json configuration on one GPU, next to the CPU paths, with parity on a pixel stride.

    python tools/run_configs.py [C1 C2 C3 C4A C4B C5S C5P ...] > gpurun_out/configs.jsonl

One JSON line per configuration: ms/frame (CUDA events over the timed frames), light-pixel contributions/s, point or
sphere evaluations/s and the FP32 fraction (SURVEY 8d flop figures over the FFMA peak measured in the same run),
per-pass times, the oracle port on all host threads ("fast": exact zero-density skip) and the reference shader source
("faithful": every in-bbox sample) on a pixel stride, and max-relative-error / RMSE of the GPU frame against the
oracle on that stride. Not the bench (bench.py is); this fills BASELINE.md section 4.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import volumerenderer_b200 as vr
from oracle import pyoracle

CONFIGS = {
    # name: grid N, W, H, algo, maxLights, light-gen groups, frames timed, oracle stride, faithful stride
    "C2": dict(n=256, W=1920, H=1080, algo="point", ml=1000, gx=32, frames=10, stride=8, fstride=16),
    "C3": dict(n=256, W=1920, H=1080, algo="ray", ml=4096, gx=128, frames=3, stride=40, fstride=80),
    "C4A": dict(n=512, W=1920, H=1080, algo="beam", ml=256, gx=8, frames=64, stride=24, fstride=60),
    "C4B": dict(n=512, W=1920, H=1080, algo="beam", ml=16384, gx=512, frames=1, stride=120, fstride=240),
    "C5S": dict(n=1024, W=3840, H=2160, algo="sphere", ml=1000, gx=32, frames=3, stride=48, fstride=96),
    "C5P": dict(n=1024, W=3840, H=2160, algo="path", ml=1000, gx=32, frames=3, stride=16, fstride=32),
}
FLOP = {"point": 11, "sphere": 29, "ray": 19, "beam": 37}


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def rel_err(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    denom = np.where(np.abs(b) < 1e-6, 1.0, np.abs(b))
    return float(np.max(np.abs(a - b) / denom)), float(np.sqrt(np.mean((a - b) ** 2)))


def run_c1(r, orc):
    grid = vr.build_fog_sphere_grid(32.0, center=(0, 250, -200))
    r.set_grid(grid)
    for _ in range(3):
        d = r.cpu_test_render(256, 256)
    ms = []
    for _ in range(10):
        d = r.cpu_test_render(256, 256)
        ms.append(r.timings()["eval_ms"])
    t0 = time.perf_counter()
    od, _ = orc.cpu_test(grid, 256, 256, nthreads=1)
    cpu_ms = (time.perf_counter() - t0) * 1e3
    lookups = 256 * 256 * 1200
    return dict(config="C1", desc="CPU_test density ray-march, 64^3-extent fog sphere (r=32) at 256x256, 1200 steps/pixel",
                ms_per_frame=float(np.median(ms)), lookups_per_s=lookups / (np.median(ms) * 1e-3), cpu_ms_1thread=cpu_ms,
                cpu_lookups_per_s=lookups / (cpu_ms * 1e-3), speedup=cpu_ms / float(np.median(ms)),
                bit_exact=bool(np.array_equal(d, od)), algorithmic_bytes=lookups * 4,
                hbm_gbs_algorithmic=lookups * 4 / (np.median(ms) * 1e-3) / 1e9)


def run_cfg(name, r, orc, ref, peak_fma, grids):
    c = CONFIGS[name]
    n, W, H, algo = c["n"], c["W"], c["H"], c["algo"]
    if n not in grids:
        t0 = time.perf_counter()
        grids.clear()
        grids[n] = vr.build_cloud_grid(n)
        log(f"[{name}] cloud {n}^3: {grids[n].nbytes / 1e6:.1f} MB built in {time.perf_counter() - t0:.1f} s")
    grid = grids[n]
    assert grid.nbytes < (1 << 32), "beyond the reference's 32-bit GLSL addressing"
    r.set_grid(grid)
    r.resize(W, H)
    r.set_max_lights_capacity(max(87500, c["ml"]))
    r.ubo = r.default_ubo()
    r.update_ubo(maxLights=c["ml"])
    r.set_light_gen_groups(c["gx"], 1, 1)
    # frame 1: parity against the oracle on a pixel stride
    r.reset_accumulation()
    r.draw_frame(algo)
    fc = r.read_frame_color()
    lights, count, emitted = r.read_lights()
    cnt1 = r.counters()
    ou = pyoracle.default_ubo(W, H, frameCount=1, maxLights=c["ml"])
    ol, oc, oem, _ = orc.light_gen(grid, ou, (c["gx"], 1, 1), 0) if algo != "path" else (lights, 0, 0, 0)
    lights_exact = bool(algo == "path" or ((count, emitted) == (oc, oem) and lights.tobytes() == ol.tobytes()))
    s = c["stride"]
    t0 = time.perf_counter()
    ofc, ocnt = orc.gather(grid, ou, algo, ol, stride=s, skip_zero=True)
    fast_s = time.perf_counter() - t0
    rel, rmse = rel_err(fc[::s, ::s], ofc[::s, ::s])
    nl = max(1, len(ol)) if algo != "path" else 1
    fast_units = ocnt["S_eff"] * nl            # contributions in the strided CPU sample
    fs = c["fstride"]
    faithful = None
    if ref is not None:
        _, fcnt = orc.gather(grid, ou, algo, np.zeros(0, pyoracle.LIGHT_DTYPE), stride=fs, skip_zero=True)
        t0 = time.perf_counter()
        ref.gather(grid, ou, algo, ol, stride=fs, nthreads=1)
        faithful = dict(seconds=time.perf_counter() - t0, stride=fs, threads=1, contributions=fcnt["S_eff"] * nl,
                        full_frame_ms_equiv=(time.perf_counter() - t0) * 1e3 * cnt1["S_eff"] / max(1, fcnt["S_eff"]))
    # timed frames (fresh accumulation, frames 1..K as the reference would run them)
    for _ in range(2):
        r.draw_frame(algo)
    r.sync()
    r.reset_accumulation()
    r.timings_sum(reset=True)
    ceff = evals = 0
    r.timer_start()
    for _ in range(c["frames"]):
        r.draw_frame(algo)
        k = r.counters()
        ceff += k["S_eff"] * (k["light_count"] if algo != "path" else 1)
        evals += k["E"]
    dev_ms = r.timer_stop()
    tsum, nfr = r.timings_sum(reset=True)
    ms = dev_ms / c["frames"]
    out = dict(config=name, algo=algo, grid=f"{n}^3 cloud, {grid.nbytes / 1e6:.1f} MB", image=f"{W}x{H}", max_lights=c["ml"],
               light_gen_groups=c["gx"], frames=c["frames"], ms_per_frame=ms, contributions_per_s=ceff / (dev_ms * 1e-3),
               evals_per_s=evals / (dev_ms * 1e-3), S_ref=cnt1["S_ref"], S_eff=cnt1["S_eff"], light_count=cnt1["light_count"],
               eval_lights=cnt1["eval_lights"], passes_ms={k2: v / nfr for k2, v in tsum.items()},
               lights_bit_exact=lights_exact, parity=dict(stride=s, max_rel=rel, rmse=rmse, mean=float(ofc[::s, ::s].mean())),
               cpu_fast=dict(seconds=fast_s, stride=s, threads=os.cpu_count(), contributions=fast_units,
                             contributions_per_s=fast_units / fast_s,
                             full_frame_ms_equiv=fast_s * 1e3 * cnt1["S_eff"] / max(1, ocnt["S_eff"])),
               cpu_faithful=faithful)
    if algo in FLOP:
        eval_s = tsum["eval_ms"] * 1e-3
        out["fp32_tflops"] = evals * FLOP[algo] / eval_s / 1e12
        out["fp32_frac_of_measured_ffma_peak"] = out["fp32_tflops"] / peak_fma
        out["eval_only_evals_per_s"] = evals / eval_s
    out["speedup_vs_cpu_fast"] = out["cpu_fast"]["full_frame_ms_equiv"] / ms
    return out


def main():
    names = sys.argv[1:] or ["C1", "C2", "C3", "C4A", "C4B", "C5S", "C5P"]
    r = vr.VolumeRenderer(0)
    orc = pyoracle.Oracle()
    ref = pyoracle.Ref() if pyoracle.have_ref() else None
    peak_fma = r.ubench("fma")
    log("FFMA peak", peak_fma, "TFLOP/s; host threads", os.cpu_count())
    grids = {}
    for name in names:
        t0 = time.perf_counter()
        res = run_c1(r, orc) if name == "C1" else run_cfg(name, r, orc, ref, peak_fma, grids)
        res["ffma_peak_tflops"] = peak_fma
        res["host_threads"] = os.cpu_count()
        print(json.dumps(res), flush=True)
        log(f"[{name}] done in {time.perf_counter() - t0:.1f} s")
    r.close()


if __name__ == "__main__":
    main()
