"""SURVEY 8f rows on the GPU: the headless driver (N1) against the Python host layer, .nvdb input (N2), snapshot /
resume of the progressive state (N4), and path mode's shadow-term table against the oracle for several step sizes."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _read_ppm(path):
    data = open(path, "rb").read()
    parts = data.split(b"\n", 3)
    assert parts[0] == b"P6" and parts[2] == b"255"
    w, h = (int(x) for x in parts[1].split())
    return np.frombuffer(parts[3], np.uint8).reshape(h, w, 3)


def _read_pfm(path):
    data = open(path, "rb").read()
    parts = data.split(b"\n", 3)
    assert parts[0] == b"PF"
    w, h = (int(x) for x in parts[1].split())
    return np.frombuffer(parts[3], "<f4").reshape(h, w, 3)[::-1]


def test_headless_matches_host_layer(vr, cloud, tmp_path):
    """vr_headless (C++ over the C ABI, .nvdb input, JSON config) renders the same frames as VolumeRenderer.draw_frame"""
    grid = cloud(64)
    nv = str(tmp_path / "cloud64.nvdb")
    vr.save_nvdb_grid(nv, grid)
    cfg = tmp_path / "scene.json"
    cfg.write_text(json.dumps({"grid": nv, "algo": "ray", "frames": 3, "width": 80, "height": 60, "lightgen": [4, 1, 1],
                               "maxLights": 50, "photonInitialIntensity": 80.0, "cameraPos": [0.0, 22.0, -70.0]}))
    ppm, pfm = str(tmp_path / "o.ppm"), str(tmp_path / "o.pfm")
    run = subprocess.run([vr.HEADLESS_PATH, "--config", str(cfg), "--ppm", ppm, "--pfm", pfm], capture_output=True, text=True)
    assert run.returncode == 0, run.stderr
    info = json.loads(run.stdout.strip().splitlines()[-1])
    assert info["frames"] == 3 and info["frameCount"] == 3 and 0 < info["light_count"] <= 50
    with vr.VolumeRenderer(0) as r:
        r.set_grid(grid)
        r.resize(80, 60)
        r.ubo = r.default_ubo()
        r.update_ubo(maxLights=50, photonInitialIntensity=80.0, cameraPos=(0.0, 22.0, -70.0))
        r.set_light_gen_groups(4, 1, 1)
        for _ in range(3):
            r.draw_frame("ray")
        assert np.array_equal(_read_ppm(ppm), r.read_image_rgba8()[..., :3])
        assert np.array_equal(_read_pfm(pfm), r.read_accum()[..., :3])
        c = r.counters()
        assert (info["S_eff"], info["E"], info["light_count"]) == (c["S_eff"], c["E"], c["light_count"])


def test_headless_cpu_test_mode(vr, oracle, tmp_path):
    ppm = str(tmp_path / "d.ppm")
    run = subprocess.run([vr.HEADLESS_PATH, "--grid", "sphere:32", "--cpu-test", ppm, "--quiet"], capture_output=True, text=True)
    assert run.returncode == 0, run.stderr
    grid = vr.build_fog_sphere_grid(32.0, center=(0, 250, -200))
    _, grey = oracle.cpu_test(grid, 256, 256, nthreads=8)     # min(density*5, 255), CPU_test/main.cpp:118
    img = _read_ppm(ppm)
    assert np.array_equal(img[..., 0], grey) and np.array_equal(img[..., 1], grey)


def test_snapshot_resume_is_exact(vr, cloud, tmp_path):
    """4 frames straight == 2 frames, snapshot, fresh context, 2 more frames (float accumulator, rgba8 image, counter)"""
    grid = cloud(64)

    def fresh():
        r = vr.VolumeRenderer(0)
        r.set_grid(grid)
        r.set_light_gen_groups(4, 1, 1)
        return r

    a = fresh()
    a.resize(72, 40)
    a.ubo = a.default_ubo()
    a.update_ubo(maxLights=40, beamRadius=0.2)
    for _ in range(2):
        a.draw_frame("beam")
    snap = str(tmp_path / "state.vrs")
    a.save_snapshot(snap)
    for _ in range(2):
        a.draw_frame("beam")
    b = fresh()
    b.load_snapshot(snap)
    assert (b.width, b.height) == (72, 40)
    assert b.ubo.frameCount == 2 and b.ubo.maxLights == 40 and abs(b.ubo.beamRadius - 0.2) < 1e-7
    for _ in range(2):
        b.draw_frame("beam")
    assert b.ubo.frameCount == 4
    assert np.array_equal(a.read_accum(), b.read_accum())
    assert np.array_equal(a.read_image_rgba8(), b.read_image_rgba8())
    a.close(), b.close()
    # the driver resumes too
    out = subprocess.run([vr.HEADLESS_PATH, "--grid", "cloud:64", "--snapshot-in", snap, "--algo", "beam", "--frames", "2",
                          "--lightgen", "4", "--pfm", str(tmp_path / "r.pfm")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert json.loads(out.stdout.strip().splitlines()[-1])["frameCount"] == 4
    with pytest.raises(vr.VrError):
        fresh().load_snapshot(str(tmp_path / "r.pfm"))


@pytest.mark.parametrize("step,light", [(1.0, (-20.0, 15.0, -15.0)), (0.5, (3.0, 20.0, 4.0)), (0.25, (-40.0, 15.0, -15.0)),
                                        (2.0, (-20.0, 15.0, -15.0))])
def test_path_mode_shadow_table_bit_exact(vr, pyoracle, oracle, cloud, step, light):
    """step <= 1: the shadow term comes from the per-neighbour table; step 2: the per-sample loop. Light inside dense
    medium, inside thin medium and outside the volume."""
    grid = cloud(64)
    W, H = 64, 40
    kw = dict(frameCount=3, rayMarchingStepSize=step, lightSourceWorldPos=light, photonInitialIntensity=400.0)
    with vr.VolumeRenderer(0) as r:
        r.set_grid(grid)
        r.resize(W, H)
        r.ubo = r.default_ubo()
        r.update_ubo(**kw)
        r.clear_image()
        r.dispatch_compute_color("path")
        fc = r.read_frame_color()
        cnt = r.counters()
    ofc, ocnt = oracle.gather(grid, pyoracle.default_ubo(W, H, **kw), "path", np.zeros(0, pyoracle.LIGHT_DTYPE))
    assert (cnt["S_ref"], cnt["S_eff"]) == (ocnt["S_ref"], ocnt["S_eff"])
    assert np.array_equal(fc, ofc)
    assert ofc.max() > 0


@pytest.mark.parametrize("exchange", ["none", "allgather"])
def test_headless_multi_gpu_equals_single_context(vr, cloud, tmp_path, exchange):
    """vr_headless --devices ... (the C++ multi-GPU host: vr_group_*, band balancer, per-GPU read-back) writes the image a
    single context renders, bit for bit; with one GPU in the box the device is listed several times (bands on one GPU)."""
    import torch
    nd = max(1, torch.cuda.device_count())
    devs = ",".join(str(i % nd) for i in range(4))
    ppm, pfm = str(tmp_path / "g.ppm"), str(tmp_path / "g.pfm")
    env = dict(os.environ, CUDA_DEVICE_MAX_CONNECTIONS="32")
    run = subprocess.run([vr.HEADLESS_PATH, "--devices", devs, "--exchange", exchange, "--grid", "cloud:64", "--algo", "sphere",
                          "--size", "160x96", "--frames", "3", "--lightgen", "6", "--set", "maxLights=90", "--balance-rounds", "2",
                          "--bench", "2", "--ppm", ppm, "--pfm", pfm], capture_output=True, text=True, env=env, timeout=300)
    assert run.returncode == 0, run.stderr
    info = json.loads(run.stdout.strip().splitlines()[-1])
    assert info["gpus"] == 4 and info["frames"] == 3 and sum(h for _, h in info["bands"]) == 96 and info["ms_per_frame_device"] > 0
    with vr.VolumeRenderer(0) as r:
        r.set_grid(cloud(64))
        r.resize(160, 96)
        r.ubo = r.default_ubo()
        r.update_ubo(maxLights=90)
        r.set_light_gen_groups(6, 1, 1)
        for _ in range(3):
            r.draw_frame("sphere")
        assert np.array_equal(_read_ppm(ppm), r.read_image_rgba8()[..., :3])
        assert np.array_equal(_read_pfm(pfm), r.read_accum()[..., :3])
