"""The BASELINE.json configurations other than the headline one (C2), measured for bench.py's `configs` array.

bench.py times C2 (Point VPLs, 1080p) as the headline line and then calls run_all() here, so that the driver's record also
carries, per configuration: device time per step, the configuration's throughput, a roofline entry for its dominant
kernel, the oracle port timed on the host cores on a pixel stride of the same frame (cpu_baseline), an end-to-end figure
through the C ABI with host buffers (volume upload and float image read-back inside the timed region, every step), and a
parity object computed in this same process against the oracle (oracle/pyoracle.py: the checker, never the thing timed).

  C1   CPU_test/main.cpp:25-43, 89-126          density ray-march 256x256, 1200 steps per pixel, 1 CPU thread beside it
  C3   shaders/ray_compute_color.comp:9-26      Ray lights, maxLights 4096, 256^3 cloud, 1920x1080
  C4A  shaders/beam_compute_color.comp:9-26     Beam lights, maxLights 256 per frame, 512^3 cloud, 1920x1080, ALL 64 frames of
       + point_compute_color.comp:97-105        the progressive average: float accumulator and re-quantised rgba8 image
  C5S  shaders/sphere_compute_color.comp:9-15   Sphere lights, maxLights 1000, 1024^3-extent cloud, 3840x2160
  C5P  shaders/path_compute_color.comp:33-128   Path mode, same volume and image

tools/run_configs.py is the command-line front end (one JSON line per configuration).
"""
import os
import time

import numpy as np

CONFIGS = {
    # n: cloud extent; ml: maxLights; gx: light-gen work-groups; steps: timed frames; stride: pixel stride of the CPU sample
    "C3": dict(n=256, W=1920, H=1080, algo="ray", ml=4096, gx=128, steps=3, stride=20, e2e_steps=3,
               desc="C3: Ray lights, maxLights 4096, 256^3 procedural cloud, 1920x1080"),
    "C4A": dict(n=512, W=1920, H=1080, algo="beam", ml=256, gx=8, steps=64, stride=20, e2e_steps=8, accumulate=True,
                desc="C4-A: Beam lights, maxLights 256 per frame, 512^3 procedural cloud, 1920x1080, 64-frame progressive accumulation"),
    "C5S": dict(n=1024, W=3840, H=2160, algo="sphere", ml=1000, gx=32, steps=5, stride=40, e2e_steps=3,
                desc="C5 sphere: Sphere lights, maxLights 1000, 1024^3-extent procedural cloud, 3840x2160"),
    "C5P": dict(n=1024, W=3840, H=2160, algo="path", ml=1000, gx=32, steps=5, stride=32, e2e_steps=3,
                desc="C5 path: Path mode, 1024^3-extent procedural cloud, 3840x2160"),
}
ORDER = ["C1", "C3", "C4A", "C5S", "C5P"]
FLOP_PER_EVAL = {"point": 11, "sphere": 29, "ray": 19, "beam": 37}  # flops per evaluation as the shaders write them
# Flops an evaluation NEEDS once the sub-point of a ray/beam light (8 flops, the same for every camera sample) is computed
# once per light, as the expanded list does: the roofline fraction uses these, so it cannot exceed 1
FLOP_PERFORMED = {"point": 11, "sphere": 29, "ray": 11, "beam": 29}
REL_TOL, RMSE_TOL = 1e-3, 1e-4  # BASELINE.json: max relative error / RMSE on linear float radiance
KERNELS = {
    "ray": "k_eval_seg<256,2,point,16,full> over the expanded sub-point list (+ slice units, k_eval_point3<8,3,2,list> on leftovers and near-light segments)",
    "beam": "k_eval_seg<256,2,sphere,8,full> over the expanded sub-point list (+ slice units, k_eval_sphere2<8,list>)",
    "sphere": "k_eval_seg<256,2,sphere,8,full> (+ slice units, k_eval_sphere2<8,list>)",
}


def rel_err(a, b):
    """max relative error with the |ref| < 1e-6 -> absolute rule of BASELINE.json, and RMSE"""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    denom = np.where(np.abs(b) < 1e-6, 1.0, np.abs(b))
    return float(np.max(np.abs(a - b) / denom)), float(np.sqrt(np.mean((a - b) ** 2)))


def _pinned(nbytes_or_shape, dtype):
    import torch
    t = torch.empty(nbytes_or_shape, dtype=dtype, pin_memory=True)
    return t, t.numpy()


def run_c1(vr, r, orc, peaks, log):
    """CPU_test's loop as a kernel (k_cpu_test) against the oracle's restatement on one host thread, bit for bit."""
    import torch
    grid = vr.build_fog_sphere_grid(32.0, center=(0, 250, -200))
    W = H = 256
    r.set_grid(grid)
    for _ in range(3):
        d = r.cpu_test_render(W, H)
    ms = []
    for _ in range(20):
        d = r.cpu_test_render(W, H)
        ms.append(r.timings()["eval_ms"])
    ms_dev = float(np.median(ms))
    t0 = time.perf_counter()
    od, _ = orc.cpu_test(grid, W, H, nthreads=1)
    cpu_s = time.perf_counter() - t0
    lookups = W * H * 1200
    # end to end: volume upload + render + density image back, per step (vr_cpu_test_render returns host data itself)
    k = 10
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(k):
        r.set_grid(grid)
        r.cpu_test_render(W, H)
    e2e_s = (time.perf_counter() - t0) / k
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    ach = lookups * 4 / (ms_dev * 1e-3) / 1e9
    return {
        "name": "C1", "workload": "C1: CPU_test density ray-march (CPU_test/main.cpp:25-43, 89-126), fog sphere r=32 in front of the fixed camera, 256x256, 1200 unit steps per pixel",
        "metric": "grid_lookups_per_s", "unit": "lookups/s", "value": lookups / (ms_dev * 1e-3), "ms_per_step": ms_dev, "steps": 20,
        "roofline": {"bound": "hbm", "kernel": "k_cpu_test", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                     "note": "algorithmic bytes = 4 B per lookup; the 0.8 MB grid stays in L1/L2, so this is a cache-bandwidth figure "
                             "set against the HBM peak, and the launch (65 536 threads) is too small to fill 148 SMs"},
        "cpu_baseline": {"value": lookups / cpu_s, "unit": "lookups/s", "cores": 1, "kind": "port",
                         "sample": f"oracle restatement of CPU_test's loop, whole 256x256 image, 1 thread, {cpu_s:.2f} s"},
        "e2e": {"value": lookups / e2e_s, "unit": "lookups/s", "ms_per_step": e2e_s * 1e3, "steps": k,
                "h2d_bytes_per_step": int(grid.nbytes), "d2h_bytes_per_step": W * H * 4},
        "parity": {"bit_exact": bool(np.array_equal(d, od)), "pixels": W * H, "against": "oracle cpu_test (pinned to oracle/_ref's tree().getValue golden)"},
        "speedup_vs_cpu_baseline": cpu_s * 1e3 / ms_dev,
    }


def run_gather(name, vr, r, orc, peaks, grids, log):
    from oracle import pyoracle
    import torch
    c = CONFIGS[name]
    n, W, H, algo, K, s = c["n"], c["W"], c["H"], c["algo"], c["steps"], c["stride"]
    threads = os.cpu_count() or 1
    if n not in grids:
        t0 = time.perf_counter()
        grids.clear()
        grids[n] = vr.build_cloud_grid(n)
        log(f"[{name}] cloud {n}^3: {grids[n].nbytes / 1e6:.1f} MB built in {time.perf_counter() - t0:.1f} s")
    grid = grids[n]
    r.set_grid(grid)
    r.resize(W, H)
    r.set_max_lights_capacity(max(87500, c["ml"]))
    r.ubo = r.default_ubo()
    r.update_ubo(maxLights=c["ml"])
    r.set_light_gen_groups(c["gx"], 1, 1)
    sub = (slice(None, None, s), slice(None, None, s))
    npx = len(range(0, H, s)) * len(range(0, W, s))

    # ---- warm-up (buffers sized, light prefetch settled), then the timed frames 1..K of a fresh accumulation
    for _ in range(2):
        r.draw_frame(algo)
    r.sync()
    r.reset_accumulation()
    r.timings_sum(reset=True)
    ceff = evals = 0
    r.timer_start()
    for _ in range(K):
        r.draw_frame(algo)
        k = r.counters()
        ceff += k["S_eff"] * (k["light_count"] if algo != "path" else 1)
        evals += k["E"]
    dev_ms = r.timer_stop()
    tsum, nfr = r.timings_sum(reset=True)
    ms = dev_ms / K
    got_acc = r.read_accum()          # after K frames
    got_img = r.read_image_rgba8()

    # ---- parity in this process: frames re-rendered one by one, each against the oracle on the pixel stride
    # (accumulate: all K frames with the running average and the rgba8 re-quantisation; otherwise frame 1 only)
    check_frames = K if c.get("accumulate") else 1
    r.reset_accumulation()
    acc = np.zeros((H, W, 4), np.float32)
    acc[..., 3] = 1.0
    img8 = np.zeros((H, W, 4), np.uint8)
    img8[..., 3] = 255
    fc_o = np.zeros((H, W), np.float32)
    lights_exact, worst_rel, worst_rmse, cpu_s, cpu_units, S_eff_sample = True, 0.0, 0.0, 0.0, 0, 0
    for frame in range(1, check_frames + 1):
        r.draw_frame(algo)
        ou = pyoracle.default_ubo(W, H, frameCount=frame, maxLights=c["ml"])
        if algo != "path":
            ol, oc, oem, _ = orc.light_gen(grid, ou, (c["gx"], 1, 1), 0)
            gl, gc, gem = r.read_lights()
            lights_exact = lights_exact and (gc, gem) == (oc, oem) and gl.tobytes() == ol.tobytes()
        else:
            ol = np.zeros(0, pyoracle.LIGHT_DTYPE)
        t0 = time.perf_counter()
        _, ocnt = orc.gather(grid, ou, algo, ol, stride=s, skip_zero=True, nthreads=threads, out=fc_o)
        cpu_s += time.perf_counter() - t0
        S_eff_sample = ocnt["S_eff"]
        cpu_units += ocnt["S_eff"] * (max(1, len(ol)) if algo != "path" else 1)
        orc.accumulate(fc_o, frame, acc, img8, path_mode=(algo == "path"))
        if frame == 1:
            fc = r.read_frame_color()
            worst_rel, worst_rmse = rel_err(fc[sub], fc_o[sub])
            frame1_equal = bool(np.array_equal(fc[sub], fc_o[sub]))
    parity = {"lights_bit_exact": bool(lights_exact) if algo != "path" else None, "stride": s, "pixels": npx, "frames_checked": check_frames,
              "max_rel": worst_rel, "rmse": worst_rmse, "tolerance": {"max_rel": REL_TOL, "rmse": RMSE_TOL}, "against": "oracle port, frame 1 colour"}
    if algo == "path":
        parity["bit_exact"] = frame1_equal
    if c.get("accumulate"):
        a_rel, a_rmse = rel_err(got_acc[sub][..., 0], acc[sub][..., 0])
        again = r.read_accum()
        parity.update({"accum_f32_max_rel": a_rel, "accum_f32_rmse": a_rmse,
                       "rgba8_max_abs_lsb": int(np.max(np.abs(got_img[sub].astype(np.int32) - img8[sub].astype(np.int32)))),
                       "timed_run_equals_checked_run": bool(np.array_equal(again, got_acc)),
                       "against": f"oracle port: light generation + gather + running average + rgba8 re-quantisation over all {K} frames"})
    parity["ok"] = bool((lights_exact or algo == "path") and worst_rel <= REL_TOL and worst_rmse <= RMSE_TOL
                        and (algo != "path" or frame1_equal)
                        and (not c.get("accumulate") or (parity["accum_f32_max_rel"] <= REL_TOL and parity["accum_f32_rmse"] <= RMSE_TOL
                                                         and parity["rgba8_max_abs_lsb"] <= 1)))

    # ---- end to end through the C ABI with host buffers: the volume goes host -> device and the float image comes back
    # device -> host inside the timed region, every step (next volume handed over as soon as a step is enqueued, image
    # copied out underneath the next step, as in bench.py's headline loop)
    Ke = c["e2e_steps"]
    pg_t, pg = _pinned(grid.nbytes, torch.uint8)
    pg[:] = grid
    outs = [_pinned((H, W, 4), torch.float32) for _ in range(2)]
    r.reset_accumulation()
    r.set_grid(pg)
    r.draw_frame(algo)
    r.read_accum_async(outs[0][1])
    r.read_wait()
    r.sync()
    r.reset_accumulation()
    e2e_ceff = 0
    t0 = time.perf_counter()
    r.set_grid(pg)
    for i in range(Ke):
        r.ubo = r.ubo
        r.draw_frame(algo)
        if i + 1 < Ke:
            r.set_grid(pg)
        r.read_accum_async(outs[i & 1][1])
        k = r.counters()
        e2e_ceff += k["S_eff"] * (k["light_count"] if algo != "path" else 1)
    r.read_wait()
    r.sync()
    e2e_s = time.perf_counter() - t0

    out = {
        "name": name, "workload": c["desc"], "metric": "light_pixel_contributions_per_s" if algo != "path" else "path_samples_per_s",
        "unit": "contributions/s" if algo != "path" else "non-zero-density camera samples/s",
        "value": ceff / (dev_ms * 1e-3), "ms_per_step": ms, "steps": K, "light_gen_groups": [c["gx"], 1, 1],
        "contributions_per_frame": ceff / K, "evals_per_frame": evals / K, "grid_bytes": int(grid.nbytes),
        "passes_ms_per_step": {k2: v / max(1, nfr) for k2, v in tsum.items()},
        "cpu_baseline": {"value": cpu_units / cpu_s, "unit": "contributions/s" if algo != "path" else "non-zero-density camera samples/s",
                         "cores": threads, "kind": "port",
                         "sample": f"oracle port (exact zero-density skip, {threads} threads over rows), every {s}th pixel in x and y of the "
                                   f"{W}x{H} frame: {npx} pixels, {check_frames} frame(s), S_eff {S_eff_sample} in the last, {cpu_s:.2f} s"},
        "e2e": {"value": e2e_ceff / e2e_s, "unit": "contributions/s" if algo != "path" else "non-zero-density camera samples/s",
                "ms_per_step": e2e_s * 1e3 / Ke, "steps": Ke, "h2d_bytes_per_step": int(grid.nbytes) + 96, "d2h_bytes_per_step": W * H * 16},
        "parity": parity,
    }
    out["speedup_vs_cpu_baseline"] = out["value"] / out["cpu_baseline"]["value"]
    out["e2e_speedup_vs_cpu_baseline"] = out["e2e"]["value"] / out["cpu_baseline"]["value"]
    eval_s = tsum["eval_ms"] * 1e-3
    if algo in FLOP_PER_EVAL and eval_s > 0:
        tf = evals * FLOP_PERFORMED[algo] / eval_s / 1e12
        rl = {"bound": "fp32", "kernel": KERNELS[algo], "achieved": tf, "peak": peaks["ffma_tflops"], "unit": "TFLOP/s",
              "frac": tf / peaks["ffma_tflops"], "flop_per_eval": FLOP_PERFORMED[algo], "evals_per_s": evals / eval_s,
              "shader_flop_per_eval": FLOP_PER_EVAL[algo], "shader_tflops": evals * FLOP_PER_EVAL[algo] / eval_s / 1e12,
              "peak_source": "FFMA microbenchmark measured in this run (vr_ubench kind 0)", "traffic": None}
        if algo in ("sphere", "beam"):  # these kernels sit on the MUFU pipe (1.5 MUFU per evaluation), not on FFMA issue
            mufu = evals / eval_s * 1.5
            rl.update({"bound": "mufu", "achieved": mufu / 1e12, "peak": peaks["mufu_rcp_tops"], "unit": "1e12 MUFU/s",
                       "frac": mufu / 1e12 / peaks["mufu_rcp_tops"], "mufu_per_eval": 1.5, "fp32_tflops": tf,
                       "fp32_frac": tf / peaks["ffma_tflops"],
                       "peak_source": "MUFU.RCP microbenchmark measured in this run (vr_ubench kind 1)"})
        out["roofline"] = rl
    else:
        steps_s = out["value"]
        out["roofline"] = {"bound": "latency", "kernel": "k_path", "achieved": steps_s, "peak": None, "unit": "camera samples/s", "frac": None,
                           "traffic": None,
                           "note": "one pixel per thread, divergent scattered walks through the grid: neither an FP32 nor an HBM roofline binds "
                                   "(DESIGN.md section 4); no meaningful peak to divide by"}
    return out


def run_all(vr, names, log, ffma_tflops=None, mufu_rcp_tops=None, hbm_gbs=None):
    """One entry per configuration; a failure becomes {"name", "error"} instead of taking the bench line down."""
    from oracle import pyoracle
    r = vr.VolumeRenderer(0)
    orc = pyoracle.Oracle()
    peaks = {"ffma_tflops": ffma_tflops or r.ubench("fma"), "mufu_rcp_tops": mufu_rcp_tops or r.ubench("rcp")}
    if hbm_gbs:
        peaks["hbm_gbs"] = hbm_gbs
    grids, out = {}, []
    for name in names:
        t0 = time.perf_counter()
        try:
            res = run_c1(vr, r, orc, peaks, log) if name == "C1" else run_gather(name, vr, r, orc, peaks, grids, log)
        except Exception as e:  # noqa: BLE001
            res = {"name": name, "error": f"{type(e).__name__}: {e}"}
        res["wall_s"] = round(time.perf_counter() - t0, 2)
        out.append(res)
        log(f"[configs] {name}: " + (res.get("error") or f"{res['ms_per_step']:.3f} ms/step, parity {res['parity']}") + f" ({res['wall_s']} s)")
    r.close()
    return out


# ------------------------------------------------------------------------------------------------------------------
# N > 1: the same configurations on all GPUs of the node through libvr's C++ multi-GPU host (vr_group_*: one process, one
# context and worker thread per GPU, peer access over NVLink). bench.py's rank 0 calls this after the headline; the other
# ranks wait at a barrier with their GPUs idle.
#   bands        C3 and C5 sphere as row bands balanced by measured feedback (vr_group_balance), timed on the devices
#                (vr_group_time_frames: CUDA events on every GPU's stream, maximum over the GPUs)
#   progressive  C5 path sample-parallel (vr_group_render_progressive): frame f on GPU f mod N, float sums, one reduce over
#                NVLink (peer reads on GPU 0), division
# Each entry also renders the same frames on ONE GPU in this run (device-timed), for the bit-equality check of the bands
# (or the reassociation bound of the sums) and as the in-run single-GPU time the multi-GPU time is divided into.
GROUP_ORDER = ["C3", "C5S", "C5P"]


def _configure(x, c, grid):
    x.set_grid(grid)
    x.resize(c["W"], c["H"])
    x.set_max_lights_capacity(max(87500, c["ml"]))
    x.ubo = x.default_ubo()
    x.update_ubo(maxLights=c["ml"])
    x.set_light_gen_groups(c["gx"], 1, 1)


def run_group_entry(name, vr, g, r1, ndev, grids, log):
    c = CONFIGS[name]
    n, W, H, algo, K = c["n"], c["W"], c["H"], c["algo"], c["steps"]
    if n not in grids:
        t0 = time.perf_counter()
        grids.clear()
        grids[n] = vr.build_cloud_grid(n)
        log(f"[{name}] cloud {n}^3: {grids[n].nbytes / 1e6:.1f} MB built in {time.perf_counter() - t0:.1f} s")
    grid = grids[n]
    _configure(g, c, grid)
    _configure(r1, c, grid)
    progressive = algo == "path"
    frames = 2 * ndev if progressive else K

    # ---- one GPU, the same frames (device-timed)
    for _ in range(2):
        r1.draw_frame(algo)
    r1.sync()
    r1.reset_accumulation()
    units = 0
    r1.timer_start()
    for _ in range(frames):
        r1.draw_frame(algo)
        k = r1.counters()
        units += k["S_eff"] * (k["light_count"] if algo != "path" else 1)
    one_ms = r1.timer_stop() / frames
    a1, i1 = r1.read_accum(), r1.read_image_rgba8()

    out = {"name": name, "workload": c["desc"], "n_gpus": ndev, "steps": frames,
           "metric": "light_pixel_contributions_per_s" if algo != "path" else "path_samples_per_s",
           "unit": "contributions/s" if algo != "path" else "non-zero-density camera samples/s",
           "host": "vr_group_* (csrc/vr_group.cu): one process, one context + worker thread per GPU, CUDA peer access",
           "single_gpu_ms_per_step_same_run": one_ms}
    if not progressive:
        bands = g.balance(algo, 4)
        g.time_frames(algo, 1)                       # first frames of the final split size buffers: not measured
        ms, per = g.time_frames(algo, K)
        g.reset_accumulation()
        for _ in range(frames):
            g.draw_frame(algo)
        g.sync()
        a, i = g.read_accum(), g.read_image_rgba8()
        equal = bool(np.array_equal(a, a1) and np.array_equal(i, i1))
        # end to end: volume host -> first GPU -> NVLink replication, frame, every GPU returns its rows to one host image
        Ke = c["e2e_steps"]
        g.reset_accumulation()
        g.set_grid(grid)
        g.draw_frame(algo)
        g.read_accum()
        t0 = time.perf_counter()
        for _ in range(Ke):
            g.set_grid(grid)
            g.draw_frame(algo)
            g.read_accum()
        e2e_s = (time.perf_counter() - t0) / Ke
        out.update({"partition": "row bands balanced by measured per-GPU time (vr_group_balance, 4 rounds); no exchange: every GPU keeps its band",
                    "bands": bands, "ms_per_step": ms, "per_gpu_ms_per_step": per, "value": units / frames / (ms * 1e-3),
                    "timing": "CUDA events on every GPU's stream, maximum over the GPUs (vr_group_time_frames)",
                    "speedup_vs_single_gpu_same_run": one_ms / ms, "scaling": "strong",
                    "e2e": {"value": units / frames / e2e_s, "ms_per_step": e2e_s * 1e3, "steps": Ke, "timing": "host wall clock, blocking read-back every step",
                            "h2d_bytes_per_step": int(grid.nbytes) + 96 * ndev, "d2h_bytes_per_step": W * H * 16,
                            "nvlink_replication_bytes_per_step": int(grid.nbytes) * (ndev - 1)},
                    "parity": {"multi_gpu_image_equals_single_gpu": equal, "frames": frames, "compared": "float accumulator and rgba8 image, bit for bit",
                               "ok": equal}})
    else:
        g.render_progressive(algo, ndev)             # warm-up: buffers on every GPU
        t0 = time.perf_counter()
        g.render_progressive(algo, frames)           # returns when GPU 0 holds the finished average
        wall_s = time.perf_counter() - t0
        a = g.read_accum()                           # (from GPU 0: the group routes read-backs there after a progressive run)
        rel, rmse = rel_err(a[..., 0], a1[..., 0])
        bound = frames * 2.0 ** -22
        g.reset_accumulation()
        out.update({"partition": f"sample-parallel: frame f on GPU f mod {ndev}, whole images, float sums; reduce = GPU 0 adds the other GPUs' "
                                 "accumulators in GPU order with peer reads over NVLink (k_add_accum), one division",
                    "ms_per_step": wall_s * 1e3 / frames, "value": units / wall_s,
                    "timing": "host wall clock around the blocking vr_group_render_progressive call (rendering on all GPUs + reduce + division); "
                              "the call spans N streams and a cross-GPU reduce, so no single stream's events cover it",
                    "speedup_vs_single_gpu_same_run": one_ms * frames / (wall_s * 1e3), "scaling": "strong (frames of one accumulation split over the GPUs)",
                    "nvlink_reduce_bytes": (ndev - 1) * W * H * 16,
                    "parity": {"max_rel_vs_sequential_average": rel, "rmse": rmse, "bound": bound, "frames": frames,
                               "compared": "float average against the one-GPU sequential running average of the same frames (equal up to float reassociation)",
                               "ok": bool(rel <= bound)}})
    return out


def run_group(vr, ndev, names, log):
    out = []
    g = r1 = None
    try:
        g = vr.RendererGroup(list(range(ndev)))
        r1 = vr.VolumeRenderer(0)
    except Exception as e:  # noqa: BLE001
        return [{"name": "group", "error": f"{type(e).__name__}: {e}"}]
    grids = {}
    for name in names:
        t0 = time.perf_counter()
        try:
            res = run_group_entry(name, vr, g, r1, ndev, grids, log)
        except Exception as e:  # noqa: BLE001
            res = {"name": name, "error": f"{type(e).__name__}: {e}"}
        res["wall_s"] = round(time.perf_counter() - t0, 2)
        out.append(res)
        log(f"[configs x{ndev}] {name}: " + (res.get("error") or f"{res['ms_per_step']:.3f} ms/step, {res['speedup_vs_single_gpu_same_run']:.2f}x of one GPU "
                                            f"({res['single_gpu_ms_per_step_same_run']:.3f} ms), parity {res['parity']}") + f" ({res['wall_s']} s)")
    r1.close()
    g.close()
    return out
