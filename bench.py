#!/usr/bin/env python3
"""bench.py -- BASELINE.json's metric on its configs[1] workload (C2 of SURVEY.md section 8d):

    Point VPLs, maxLights 1000, synthetic 256^3 procedural cloud, 1920x1080, all UBO defaults of the reference.

A step is one frame of the hot path, exactly what VolumeApp::drawFrame's compute section does (src/main.cpp:712-837):
frameCount++, counter reset, light generation (gx work-groups of 4x4 photons), camera march, light evaluation,
progressive average. metric = light-pixel contributions per second: C_eff = S_eff * lightCount per frame (camera
samples with non-zero density x lights) divided by the frame time; ms_per_step is ms/frame.

  python bench.py [--gpus N] [--steps K] [--warmup W]            our arm (N>1: launched by torchrun, NCCL)
  python bench.py --impl reference [...]                          the reference's shader source on the host cores

N>1 partitions the frame into row bands balanced by a measured per-row cost (grid and lights replicated, light
generation repeated on every rank). Pixels are independent, so the path has no exchange step: every rank keeps the
progressive accumulator of its band (--exchange none, the default) and the bands meet in host memory, each rank copying
its rows over its own PCIe link (the e2e leg). --exchange peer|nccl additionally leaves the whole image on every GPU after
every frame (fused peer stores from the resolve kernel, or one grouped NCCL send/recv). Strong scaling.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (grid N, W, H, algo, maxLights, light-gen groups gx, description)
    "C2": dict(n=256, W=1920, H=1080, algo="point", max_lights=1000, gx=32,
               desc="C2: Point VPLs, maxLights 1000, 256^3 procedural cloud (seed 1234, voxel 0.25, 14.2 MB NanoVDB), 1920x1080"),
    "C3": dict(n=256, W=1920, H=1080, algo="ray", max_lights=4096, gx=128,
               desc="C3: Ray lights, maxLights 4096, 256^3 procedural cloud, 1920x1080"),
}
FLOP_PER_EVAL = {"point": 11, "sphere": 29, "ray": 19, "beam": 37}  # SURVEY 8(d) algorithmic figures


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        # NVML in a thread (a sample every 5 ms: the timed region of the default run is ~0.3 s); nvidia-smi as the fallback
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            idx = self.gpu
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            ids = [v.strip() for v in vis.split(",") if v.strip()]
            if ids and all(v.isdigit() for v in ids) and self.gpu < len(ids):
                idx = int(ids[self.gpu])
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.nvml = pynvml
            self.samples = []
            self.running = True
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.QUERY}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        while self.running:
            try:
                sm = n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)
                mx = n.nvmlDeviceGetMaxClockInfo(self.handle, n.NVML_CLOCK_SM)
                get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
                self.samples.append((sm, mx, int(get(self.handle))))
            except Exception:
                pass
            time.sleep(0.005)

    def _stop_nvml(self):
        self.running = False
        self.thread.join(timeout=1)
        bits = {"hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40, "sw_power_cap": 0x4}
        sm = [a for a, _, _ in self.samples]
        mx = [b for _, b, _ in self.samples]
        reasons = sorted(k for k, bit in bits.items() if any(r & bit for _, _, r in self.samples))
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(max(mx)) if mx else None,
                "reasons": reasons, "samples": len(sm), "source": "nvml"}

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if getattr(self, "nvml", None):
            return self._stop_nvml()
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for row in self.rows:
            f = [x.strip() for x in row.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(max(mx)) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------ CPU arms
def cpu_workload(cfg, frame, ref=None):
    """Grid, UBO and the frame's light list for the CPU arms (neither is inside a timed region). With oracle/_ref present
    the grid comes from the reference's vendored NanoVDB builder (the same voxels; the reference arm then maps no product
    library at all), else from the product's host-side grid writer."""
    from oracle import pyoracle
    if ref is not None:
        grid = ref.make_cloud(cfg["n"])
    else:
        import volumerenderer_b200 as vr
        grid = vr.build_cloud_grid(cfg["n"])
    orc = pyoracle.Oracle()
    u = pyoracle.default_ubo(cfg["W"], cfg["H"], frameCount=frame, maxLights=cfg["max_lights"])
    lights, count, _, _ = orc.light_gen(grid, u, (cfg["gx"], 1, 1), 0)
    return grid, orc, u, lights


def sample_ceff(orc, grid, u, algo, n_lights, stride):
    """C_eff of the pixel-stride sample, counted exactly by a march without lights (outside any timed region)."""
    from oracle import pyoracle
    _, cnt = orc.gather(grid, u, algo, np.zeros(0, pyoracle.LIGHT_DTYPE), stride=stride, skip_zero=True)
    return cnt["S_eff"] * n_lights, cnt


def run_cpu(kind, orc, ref, grid, u, algo, lights, stride, threads):
    t0 = time.perf_counter()
    if kind == "reference":
        ref.gather(grid, u, algo, lights, stride=stride, nthreads=threads)
    else:
        orc.gather(grid, u, algo, lights, stride=stride, skip_zero=True, nthreads=threads)
    return time.perf_counter() - t0


def pick_stride(t_probe, probe_stride, target_s):
    s = probe_stride * (t_probe / target_s) ** 0.5
    for cand in (1, 2, 3, 4, 6, 8, 12, 16, 24, 32):
        if cand >= s:
            return cand
    return 32


def reference_arm(args, cfg):
    """--impl reference: the reference's own shader source (shaders/point_compute_color.comp etc. compiled as C++ in
    oracle/_ref) on all host threads; falls back to the oracle port when oracle/_ref was never built."""
    from oracle import pyoracle
    threads = os.cpu_count() or 1
    ref = pyoracle.Ref() if pyoracle.have_ref() else None
    grid, orc, u, lights = cpu_workload(cfg, 1, ref)
    kind = "reference" if ref is not None else "port"
    algo = cfg["algo"]
    t_probe = run_cpu(kind, orc, ref, grid, u, algo, lights, 32, threads)
    target = max(0.5, min(6.0, 150.0 / (args.steps + args.warmup)))
    stride = pick_stride(t_probe, 32, target)
    ceff, cnt = sample_ceff(orc, grid, u, algo, len(lights), stride)
    for _ in range(args.warmup):
        run_cpu(kind, orc, ref, grid, u, algo, lights, stride, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        run_cpu(kind, orc, ref, grid, u, algo, lights, stride, threads)
    dt = (time.perf_counter() - t0) / args.steps
    value = ceff / dt
    sample = (f"every {stride}th pixel in x and y of the {cfg['W']}x{cfg['H']} frame ({cnt['pixels']} pixels, "
              f"S_eff {cnt['S_eff']}, C_eff {ceff:.3e} per step); "
              + ("reference shader source evaluates every in-bbox sample (S_ref %d)" % cnt["S_ref"] if kind == "reference"
                 else "oracle port with the exact zero-density skip"))
    line = {
        "impl": "reference", "metric": "light_pixel_contributions_per_s", "value": value, "unit": "contributions/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": cfg["desc"]},
        "cpu_baseline": {"value": value, "unit": "contributions/s", "cores": threads, "kind": kind, "sample": sample,
                         "pixel_stride": stride},
        "e2e": {"value": value, "unit": "contributions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "full_frame_ms_equiv": dt * 1e3 * stride * stride,
    }
    print(json.dumps(line), flush=True)


class SharedHostImages:
    """N > 1: `count` (H, W, 4) float32 images in POSIX shared memory that every rank maps and page-locks
    (cudaHostRegister), so that each GPU copies the rows it rendered straight into ONE host image over its own PCIe link."""

    def __init__(self, count, H, W, rank, dist):
        import torch
        from multiprocessing import resource_tracker, shared_memory
        self.nbytes = count * H * W * 16
        box = [None]
        if rank == 0:
            self.shm = shared_memory.SharedMemory(create=True, size=self.nbytes)
            box[0] = self.shm.name
        dist.broadcast_object_list(box, src=0)
        if rank != 0:
            self.shm = shared_memory.SharedMemory(name=box[0])
            try:   # only the creating rank unlinks the segment
                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:
                pass
        self.owner = rank == 0
        self.images = np.ndarray((count, H, W, 4), np.float32, buffer=self.shm.buf)
        self.ptr = self.images.ctypes.data
        rc = torch.cuda.cudart().cudaHostRegister(self.ptr, self.nbytes, 0)
        if int(rc) != 0:
            raise RuntimeError(f"cudaHostRegister failed ({rc})")

    def close(self, dist):
        import torch
        torch.cuda.cudart().cudaHostUnregister(self.ptr)
        self.images = None
        dist.barrier()
        self.shm.close()
        if self.owner:
            self.shm.unlink()


# ------------------------------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=sorted(WORKLOADS))
    ap.add_argument("--exchange", default="none", choices=["none", "auto", "peer", "nccl"],
                    help="N>1: none = every rank keeps the accumulator of its band (default); peer/nccl/auto = all-gather of the "
                         "bands' frame colours every frame (whole image on every GPU)")
    ap.add_argument("--balance-rounds", type=int, default=6, help="N>1: rounds of measured feedback for the band split")
    ap.add_argument("--polish-moves", type=int, default=8, help="N>1: single-row moves tried after the feedback rounds")
    ap.add_argument("--eval-variant", type=int, default=8, choices=[0, 1, 2, 3, 8],
                    help="evaluation kernel for point/ray lists: 8 = ray-segment kernel (default), 2 = exact packed kernel")
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target duration of the cpu_baseline sample")
    ap.add_argument("--configs", default="all",
                    help="N=1: the other BASELINE.json configurations measured after the headline and reported in the line's `configs` "
                         "array (bench_configs.py): all | none | comma-separated subset of C1,C3,C4A,C5S,C5P")
    args = ap.parse_args()
    cfg = WORKLOADS[args.workload]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank == 0:
            reference_arm(args, cfg)
        return 0
    if args.warmup < 3:
        log("note: fewer than 3 warm-up steps requested; using 3")
        args.warmup = 3

    import torch
    import volumerenderer_b200 as vr
    from volumerenderer_b200 import tiles

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: volumerenderer_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local_rank}"))
    W, H, algo = cfg["W"], cfg["H"], cfg["algo"]

    grid = vr.build_cloud_grid(cfg["n"])
    r = vr.VolumeRenderer(local_rank)
    r.set_eval_variant(args.eval_variant)
    if world > 1:
        # one priority step above the default: libvr's side streams (slice units, odd segments, exact list) have the lowest
        # priority and must only take the SM slots this stream's kernels leave free (as with the context's own stream)
        shared = torch.cuda.Stream(priority=-1)
        torch.cuda.set_stream(shared)
        r.set_stream(shared.cuda_stream)  # one stream for the kernels and the NCCL band exchange
    r.set_grid(grid)
    r.resize(W, H)
    r.set_max_lights_capacity(max(87500, cfg["max_lights"]))
    r.ubo = r.default_ubo()
    r.update_ubo(maxLights=cfg["max_lights"])
    r.set_light_gen_groups(cfg["gx"], 1, 1)

    # row bands balanced by a per-row cost measured on one full frame (identical on every rank): the evaluation kernel's
    # time follows the row's non-zero-density samples, the camera march's the row count (tiles.row_costs)
    bands = [(0, H)]
    tile = None
    fc_t = None
    others = []
    def set_bands(new_bands):
        nonlocal bands, tile, others
        box = [new_bands]
        dist.broadcast_object_list(box, src=0)   # every rank uses rank 0's split
        bands = [tuple(int(v) for v in b) for b in box[0]]
        tile = tiles.band_tile(bands[rank], W)
        others = tiles.complement_tiles(bands[rank], W, H)

    if world > 1:
        r.draw_frame(algo)
        r.sync()
        tm = r.timings()
        cost = tiles.row_costs(r.row_sample_counts(), tm["march_ms"], tm["eval_ms"])
        set_bands(tiles.balanced_row_bands(cost, world))
        fc_t = tiles.frame_color_as_tensor(r)
        # measured feedback (untimed): each rank reports the device time of its band, the row costs are rescaled; the
        # per-band time is a staircase in CTA waves, so the split need not improve monotonically: the best one is kept
        def measure(frames=3):
            r.reset_accumulation()
            r.draw_frame(algo, tile)            # the first frame of a new split sizes buffers: not measured
            r.sync()
            r.timings_sum(reset=True)
            for _ in range(frames):
                r.draw_frame(algo, tile)
            ts, nf = r.timings_sum(reset=True)
            mine = torch.tensor([ts["total_ms"] / max(1, nf)], dtype=torch.float64, device="cuda")
            allms = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(allms, mine)
            return [float(x) for x in allms]

        best = (float("inf"), bands, None)
        for rnd in range(args.balance_rounds + 1):
            ms = measure()
            if max(ms) < best[0]:
                best = (max(ms), bands, ms)
            if rnd == args.balance_rounds:
                break
            cost = tiles.refine_row_costs(cost, bands, ms)
            set_bands(tiles.balanced_row_bands(cost, world))
        # polish: single-row moves away from the slowest rank, kept while the measured maximum goes down
        for _ in range(args.polish_moves):
            cand = tiles.polish_move(best[1], best[2])
            if cand is None:
                break
            set_bands(cand)
            ms = measure(5)
            if max(ms) >= best[0]:
                break
            best = (max(ms), bands, ms)
        set_bands(best[1])
        r.reset_accumulation()

    # N>1: the rank's band goes to every GPU straight from the resolve kernel (peer stores over NVLink into torch
    # symmetric memory) when that is available, else as one grouped NCCL send/recv of the 4 B/pixel frame colours
    peer_fc = None
    exchange = "none"
    if world > 1 and args.exchange == "none":
        exchange = "none (pixels are independent: every GPU keeps the progressive accumulator of its band; bands meet in host memory)"
    elif world > 1:
        exchange = "nccl grouped send/recv of frame colours"
        if args.exchange in ("auto", "peer"):
            try:
                peer_fc = tiles.PeerFrameColor(r, bands)
                exchange = "peer stores from the resolve kernel (symmetric memory) + device barrier"
            except Exception as e:  # noqa: BLE001
                if args.exchange == "peer":
                    raise
                log(f"rank {rank}: symmetric memory unavailable ({type(e).__name__}: {e}); using the NCCL exchange")
        flag = torch.tensor([1 if peer_fc is not None else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag) == 0 and peer_fc is not None:
            peer_fc.close()
            peer_fc = None
            exchange = "nccl grouped send/recv of frame colours"

    def step():
        if peer_fc is not None:
            peer_fc.draw_frame(algo)
            return
        # own band: light generation + gather + progressive average; the other ranks' rows arrive as frame colours
        # (4 B/pixel, one grouped NCCL send/recv on the kernels' stream) and are averaged in locally
        r.draw_frame(algo, tile)
        if world > 1 and args.exchange != "none":
            tiles.exchange_rows_p2p(fc_t, bands)
            for t in others:
                r.accumulate_frame_color(t)

    def barrier():
        r.sync()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # microbenchmarks for the roofline denominators (before the timed region)
    peak_fma = r.ubench("fma") if rank == 0 else 0.0
    peak_rcp = r.ubench("rcp") if rank == 0 else 0.0

    for _ in range(args.warmup):
        step()
    barrier()
    launches0 = r.launch_count()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ceff = 0
    evals = 0
    eval_ms = 0.0
    r.timings_sum(reset=True)
    r.timer_start()
    t_host0 = time.perf_counter()
    for _ in range(args.steps):
        step()
        c = r.counters()          # the frame's scalars (ready before its evaluation kernel starts): no frame-long wait
        ceff += c["S_eff"] * c["light_count"]
        evals += c["E"]
    dev_ms = r.timer_stop()
    barrier()
    host_ms = (time.perf_counter() - t_host0) * 1e3
    # per-pass CUDA-event times of exactly these frames, read after the timed region (lazy event ring in libvr)
    tsum, tframes = r.timings_sum(reset=True)
    assert tframes == args.steps, (tframes, args.steps)
    pass_ms = {k: tsum[k] for k in ("light_gen_ms", "march_ms", "prepare_ms", "eval_ms", "resolve_ms")}
    launches = r.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None

    per_rank = None
    if world > 1:
        mine = torch.tensor([pass_ms["march_ms"], pass_ms["eval_ms"], tsum["total_ms"], dev_ms], dtype=torch.float64, device="cuda") / args.steps
        allr = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        per_rank = [dict(zip(("march_ms", "eval_ms", "gather_ms", "step_ms"), (round(float(v), 4) for v in x))) for x in allr]
    t = torch.tensor([dev_ms, float(ceff), float(evals), pass_ms["eval_ms"]], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dev_ms_max, ceff_total, evals_total = float(tmax[0]), float(tsum[1]), float(tsum[2])
    else:
        dev_ms_max, ceff_total, evals_total = dev_ms, float(ceff), float(evals)
    ms_per_step = dev_ms_max / args.steps
    value = ceff_total / (dev_ms_max * 1e-3)

    # ---- e2e: the same frames through the C ABI with HOST buffers: grid bytes + UBO go host->device and the float
    # accumulator comes back device->host inside the timed region, every step
    pinned_grid = torch.empty(grid.nbytes, dtype=torch.uint8, pin_memory=True)
    pinned_grid.numpy()[:] = grid
    grid_host = pinned_grid.numpy()
    # two pinned result buffers: frame k's image is copied out (vr_read_accum_f32_async: device snapshot + copy stream)
    # underneath frame k+1's kernels; every frame's result reaches the host inside the timed region
    out_host_t = [torch.empty((H, W, 4), dtype=torch.float32, pin_memory=True) for _ in range(2)]
    out_host = [t.numpy() for t in out_host_t]
    e2e_frames = [0]

    # N > 1: the volume crosses PCIe once per node (rank 0), an NCCL broadcast replicates it over NVLink and every rank
    # installs it from device memory (tiles.ReplicatedGrid / vr_set_grid_device); every rank returns the rows it rendered
    # over its own PCIe link into one host image that all ranks map (shared, page-locked memory)
    replicated = tiles.ReplicatedGrid(r, grid.nbytes) if world > 1 else None
    grid_header = grid_host[:4096].copy()
    shared = SharedHostImages(2, H, W, rank, dist) if world > 1 else None

    def e2e_upload():
        if replicated is not None:
            replicated.set_grid(pinned_grid if rank == 0 else None, grid_header)   # H2D once + NVLink replication
        else:
            r.set_grid(grid_host)                   # H2D: the NanoVDB buffer (binding 2)

    # the volume is double-buffered on the device (vr_set_grid* do not drain the frames in flight), so a host that streams
    # volumes hands step k+1's bytes over as soon as step k is enqueued: the transfer, the replication and the next light
    # generation run underneath step k. Every timed step still uploads one volume and reads one image back; the first
    # timed step's upload is issued inside the timed region with nothing to hide under.
    def e2e_step(upload_next=True):
        r.ubo = r.ubo                               # H2D: the 96-byte UBO (kernel parameter)
        step()
        if upload_next:
            e2e_upload()
        if shared is not None:   # D2H: this rank's rows of the linear float image (waits for the previous frame's copy)
            r.read_accum_tile_async(tile, shared.images[e2e_frames[0] & 1])
        else:
            r.read_accum_async(out_host[e2e_frames[0] & 1])
        e2e_frames[0] += 1

    e2e_upload()
    for w in range(2):
        e2e_step(upload_next=(w == 0))
    r.read_wait()
    barrier()
    e2e_ceff = 0
    r.timings_sum(reset=True)
    t0 = time.perf_counter()
    e2e_upload()
    for k in range(args.steps):
        e2e_step(upload_next=(k + 1 < args.steps))
        c = r.counters()
        e2e_ceff += c["S_eff"] * c["light_count"]
    r.read_wait()                                   # the last frame's image has arrived
    barrier()
    e2e_s = time.perf_counter() - t0
    e2e_tsum, e2e_tframes = r.timings_sum(reset=True)   # this rank's per-pass device times of the end-to-end frames
    e2e_passes = {k: round(v / max(1, e2e_tframes), 4) for k, v in e2e_tsum.items()}
    te = torch.tensor([e2e_s, float(e2e_ceff)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = te.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = te.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        e2e_s, e2e_ceff = float(tmax[0]), float(tsum[1])
    e2e_value = e2e_ceff / e2e_s

    # ---- N>1: the bands, gathered in host memory by the ranks' own read-backs, must equal a whole-frame render of the
    # same frames bit for bit (untimed check); with an all-gather exchange every rank's device image must as well
    verified = None
    if world > 1:
        nfr = 3
        r.reset_accumulation()
        for _ in range(nfr):
            step()
        r.read_accum_tile_async(tile, shared.images[0])
        r.read_wait()
        band_i = r.read_image_rgba8()[bands[rank][0]:bands[rank][0] + bands[rank][1]]
        full_t = (r.read_accum(), r.read_image_rgba8()) if args.exchange != "none" else None
        barrier()
        if peer_fc is not None:
            r.set_frame_color_peers(None, [])
        r.reset_accumulation()
        for _ in range(nfr):
            r.draw_frame(algo)
        r.sync()
        a_f, i_f = r.read_accum(), r.read_image_rgba8()
        good = np.array_equal(i_f[bands[rank][0]:bands[rank][0] + bands[rank][1]], band_i)
        if rank == 0:
            good = good and np.array_equal(shared.images[0], a_f)     # the image the N read-backs assembled on the host
        if full_t is not None:
            good = good and np.array_equal(full_t[0], a_f) and np.array_equal(full_t[1], i_f)
        ok = torch.tensor([1 if good else 0], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        verified = bool(int(ok))
        barrier()
        shared.close(dist)

    if rank == 0:
        # ---- roofline of the dominant kernel (k_eval): FP32-issue bound, measured live with CUDA events
        flop = FLOP_PER_EVAL[algo]
        eval_s = pass_ms["eval_ms"] * 1e-3
        achieved = evals * flop / eval_s / 1e12 if eval_s > 0 else 0.0
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        samples_per_step = evals / max(1, c["eval_lights"]) / args.steps
        alg_bytes = samples_per_step * 20.0 + c["eval_lights"] * 16.0   # 16 B sample in + 4 B out, 16 B per light
        # measured DRAM traffic of one launch of the dominant kernel, from a committed ncu --set full capture. The entry is
        # stamped with the library version it was captured on: after a kernel change the stamp no longer matches the built
        # library and the figure is dropped (null + reason) instead of silently going stale.
        traffic, traffic_note, ncu_pipes = None, None, None
        try:
            with open(os.path.join(ROOT, "profiles", "eval_traffic.json")) as f:
                tj = json.load(f)
            ent = tj.get(args.workload if args.eval_variant == 8 else f"{args.workload}@{args.eval_variant}")
            built = vr.load().vr_version().decode()
            if isinstance(ent, dict) and ent.get("version") == built:
                traffic, traffic_note = ent["traffic"], f"{ent.get('source', '')} (captured on {ent.get('commit', '?')}, {built})"
                ncu_pipes = ent.get("ncu")   # pipe utilisation of the same capture: what actually bounds the kernel
            elif ent is not None:
                cap = ent.get("version") if isinstance(ent, dict) else "an unstamped round-1 build"
                traffic_note = f"dropped: captured on {cap}, the built library is {built}"
        except Exception as e:  # noqa: BLE001
            traffic_note = f"unavailable: {type(e).__name__}"
        point_kernels = {0: "k_eval<point,8>", 1: "k_eval_point2<8>", 2: "k_eval_point3<8,3,2>", 3: "k_eval_point3<6,4,2>",
                         8: "k_eval_seg<256,2,point,16,full> (+ its slice-unit instance on the last wave, k_eval_seg<256,2,point,8,*> on odd segments, k_eval_point3<8,3,2,list> on leftovers and near-light segments)"}
        kernel_name = point_kernels[args.eval_variant] if algo in ("point", "ray") else {0: "k_eval<sphere,8>", 8: "k_eval_seg<256,2,sphere,8,full> (+ its slice-unit instance on the last wave, k_eval_sphere2<8,list> on leftovers and near-light segments)"}.get(args.eval_variant, "k_eval_sphere2<8>")
        # MUFU.RCP per evaluation: the exact kernels need one (plus a rsqrt for sphere lists); the ray-segment kernel shares
        # one reciprocal between two lights
        mufu_per_eval = (0.5 if args.eval_variant == 8 else 1.0) if algo in ("point", "ray") else (1.5 if args.eval_variant == 8 else 2.0)
        roofline = {
            "bound": "fp32", "kernel": kernel_name, "achieved": achieved, "peak": peak_fma, "unit": "TFLOP/s",
            "frac": achieved / peak_fma if peak_fma else None,
            "peak_source": "FFMA microbenchmark measured in this run (vr_ubench kind 0); MEASURED_PEAKS.json has no non-tensor FP32 figure",
            "flop_per_eval": flop, "evals_per_s": evals / eval_s if eval_s > 0 else 0.0,
            "mufu_rcp_peak_per_s": peak_rcp * 1e12, "mufu_per_eval": mufu_per_eval,
            "mufu_frac": (evals / eval_s) * mufu_per_eval / (peak_rcp * 1e12) if peak_rcp and eval_s > 0 else None,
            "traffic": traffic, "traffic_source": traffic_note, "ncu_pipes": ncu_pipes,
            "hbm": {"bound": "hbm", "achieved": alg_bytes / (eval_s / args.steps) / 1e9 if eval_s > 0 else 0.0, "peak": hbm_peak,
                    "unit": "GB/s", "frac": (alg_bytes / (eval_s / args.steps) / 1e9) / hbm_peak if eval_s > 0 else None,
                    "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "fallback 6.65 TB/s (of fallback)",
                    "algorithmic_bytes_per_launch": alg_bytes},
        }
        cpu_baseline = None
        if world == 1 and not args.no_cpu_baseline:
            from oracle import pyoracle
            threads = os.cpu_count() or 1
            cgrid, orc, u, lights = cpu_workload(cfg, 1)
            t_probe = run_cpu("port", orc, None, cgrid, u, algo, lights, 32, threads)
            stride = pick_stride(t_probe, 32, args.cpu_seconds)
            cc, cnt = sample_ceff(orc, cgrid, u, algo, len(lights), stride)
            dt = run_cpu("port", orc, None, cgrid, u, algo, lights, stride, threads)
            cpu_baseline = {"value": cc / dt, "unit": "contributions/s", "cores": threads, "kind": "port",
                            "sample": f"oracle port (exact zero-density skip, {threads} threads over rows), every {stride}th pixel in x "
                                      f"and y of the {W}x{H} frame: {cnt['pixels']} pixels, C_eff {cc:.3e}, {dt:.2f} s"}
        line = {
            "metric": "light_pixel_contributions_per_s", "value": value, "unit": "contributions/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": cfg["desc"], "light_gen_groups": [cfg["gx"], 1, 1], "partition": f"row bands balanced by measured per-rank time; exchange: {exchange}" if world > 1 else "whole frame",
                       "l2": "per-step working set (1.3 GB of samples + 14 MB grid) exceeds the 126 MB L2; no flush needed",
                       "bands": bands if world > 1 else None, "eval_variant": args.eval_variant},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "contributions/s", "ms_per_step": e2e_s * 1e3 / args.steps,
                    "h2d_bytes_per_step": int(grid.nbytes) + 96 * world, "d2h_bytes_per_step": int(W * H * 16),
                    "d2h": "every rank copies the rows it rendered into one shared, page-locked host image" if world > 1 else "rank 0",
                    "nvlink_broadcast_bytes_per_step": int(grid.nbytes) * (world - 1) if world > 1 else 0,
                    "volume_replication": replicated.mode if replicated is not None else "single GPU: vr_set_grid from pinned host memory",
                    "passes_ms_per_step_rank0": e2e_passes},
            "gpu_launches": int(launches),
            "roofline": roofline,
            "cpu_baseline": cpu_baseline,
            "passes_ms_per_step": {k: v / args.steps for k, v in pass_ms.items()},
            "host_ms_per_step": host_ms / args.steps,
            "per_rank_ms_per_step": per_rank,
            "multi_gpu_image_equals_single_gpu": verified,
            "contributions_per_frame": ceff_total / args.steps,
        }
        if world == 1 and args.configs != "none" and args.workload == "C2":
            # the other BASELINE.json configurations (C1, C3, C4-A over all 64 frames, C5 sphere and path), each with its own
            # device time, roofline entry, CPU port on a pixel stride, end-to-end figure and parity against the oracle; the
            # headline keys above stay what they were (C2), so lines compare across rounds
            import bench_configs
            r.close()
            names = bench_configs.ORDER if args.configs == "all" else [n for n in args.configs.split(",") if n]
            line["configs"] = bench_configs.run_all(vr, names, log, ffma_tflops=peak_fma, mufu_rcp_tops=peak_rcp, hbm_gbs=hbm_peak)
        if world > 1 and args.configs != "none" and args.workload == "C2":
            # N > 1: C3 and C5 sphere as row bands, C5 path sample-parallel with its reduce over NVLink, on all GPUs of the node
            # through the C++ multi-GPU host (vr_group_*) from this process; the other ranks wait at the barrier below
            import bench_configs
            names = bench_configs.GROUP_ORDER if args.configs == "all" else [n for n in args.configs.split(",") if n in bench_configs.GROUP_ORDER]
            line["configs"] = bench_configs.run_group(vr, world, names, log)
        print(json.dumps(line), flush=True)
    if world > 1 and args.configs != "none" and args.workload == "C2":
        # the other ranks wait for rank 0's vr_group runs on the rendezvous store, not inside an NCCL barrier: a spinning NCCL
        # kernel would hold SMs of the very GPUs rank 0 is driving
        try:
            from datetime import timedelta
            store = dist.distributed_c10d._get_default_store()
            if rank == 0:
                store.set("vr_bench_configs_done", "1")
            else:
                store.wait(["vr_bench_configs_done"], timedelta(minutes=30))
        except Exception as e:  # noqa: BLE001
            log(f"rank {rank}: store wait unavailable ({type(e).__name__}: {e}); falling back to the NCCL barrier")
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    r.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
