"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json): light lists bit-exact in count and canonical order; radiance max relative error <= 1e-3 and
RMSE <= 1e-4 on the linear float frame colour (|ref| < 1e-6 -> absolute error); rgba8 accumulation within 1 LSB.
Work counters (S_ref, S_eff) and everything on the exact-arithmetic paths (march, path mode, CPU_test) bit-exact.
"""
import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu

REL_TOL = 1e-3   # max relative error on linear frame colour (north_star)
RMSE_TOL = 1e-4


@pytest.fixture(scope="module")
def renderer(vr):
    r = vr.VolumeRenderer(0)
    yield r
    r.close()


def _setup(r, grid, W, H, **ubo):
    r.set_grid(grid)
    r.resize(W, H)
    r.ubo = r.default_ubo()
    return r.update_ubo(**ubo)


@pytest.mark.parametrize("n,groups,max_lights,frame", [(64, (1, 1, 1), 1000, 1), (64, (8, 1, 1), 100, 1),
                                                       (64, (4, 2, 2), 50, 3), (128, (27, 1, 1), 1000, 1),
                                                       (128, (16, 1, 1), 37, 9)])
@pytest.mark.parametrize("variant", [0, 1])
def test_light_gen_bit_exact(renderer, pyoracle, oracle, cloud, n, groups, max_lights, frame, variant):
    grid = cloud(n)
    _setup(renderer, grid, 64, 64, maxLights=max_lights, frameCount=frame)
    renderer.reset_lights()
    renderer.dispatch_light_gen("point", groups, variant)
    lights, count, emitted = renderer.read_lights()
    u = pyoracle.default_ubo(64, 64, frameCount=frame, maxLights=max_lights)
    ol, oc, oem, _ = oracle.light_gen(grid, u, groups, variant)
    assert (count, emitted) == (oc, oem)
    assert lights.tobytes() == ol.tobytes()


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("max_lights", [5000, 300])
def test_light_gen_many_lights_per_photon(renderer, pyoracle, oracle, cloud, variant, max_lights):
    """scatteringProbability 0.9: photons emit far more lights than the kernel's private slab holds (the overflow
    walk), with and without the maxLights cap cutting into a photon's list."""
    grid = cloud(64)
    kw = dict(maxLights=max_lights, frameCount=2, scatteringProbability=0.9)
    _setup(renderer, grid, 64, 64, **kw)
    renderer.reset_lights()
    renderer.dispatch_light_gen("beam", (3, 1, 1), variant)
    lights, count, emitted = renderer.read_lights()
    ol, oc, oem, _ = oracle.light_gen(grid, pyoracle.default_ubo(64, 64, **kw), (3, 1, 1), variant)
    if variant == 0:
        assert emitted > 48 * 20, "the case must overflow the per-photon slab"
    assert (count, emitted) == (oc, oem)
    assert lights.tobytes() == ol.tobytes()


def test_path_light_gen_is_empty(renderer, cloud):
    _setup(renderer, cloud(64), 32, 32, frameCount=1)
    renderer.reset_lights()
    renderer.dispatch_light_gen("path", (4, 1, 1))
    _, count, emitted = renderer.read_lights()
    assert count == 0 and emitted == 0


@pytest.mark.parametrize("algo", ["point", "sphere", "ray", "beam"])
@pytest.mark.parametrize("n,W,H,max_lights,groups", [(64, 96, 64, 64, (4, 1, 1)), (128, 160, 90, 200, (8, 1, 1))])
def test_gather_parity(renderer, pyoracle, oracle, cloud, algo, n, W, H, max_lights, groups):
    grid = cloud(n)
    u = _setup(renderer, grid, W, H, maxLights=max_lights, frameCount=1)
    ou = pyoracle.default_ubo(W, H, frameCount=1, maxLights=max_lights)
    ol, oc, _, _ = oracle.light_gen(grid, ou, groups, 0)
    renderer.write_lights(ol)
    renderer.clear_image()
    renderer.dispatch_compute_color(algo)
    fc = renderer.read_frame_color()
    cnt = renderer.counters()
    ofc, ocnt = oracle.gather(grid, ou, algo, ol, skip_zero=True)
    assert cnt["S_ref"] == ocnt["S_ref"] and cnt["S_eff"] == ocnt["S_eff"]
    assert cnt["E"] == ocnt["E"]
    rel, rmse = rel_err(fc, ofc)
    assert rel <= REL_TOL and rmse <= RMSE_TOL, (rel, rmse)
    assert ofc.max() > 0.01  # the scene is not black


def test_path_mode_bit_exact(renderer, pyoracle, oracle, cloud):
    grid = cloud(64)
    W, H = 96, 64
    for frame in (1, 2, 5):
        _setup(renderer, grid, W, H, frameCount=frame)
        renderer.clear_image()
        renderer.dispatch_compute_color("path")
        fc = renderer.read_frame_color()
        ou = pyoracle.default_ubo(W, H, frameCount=frame)
        ofc, ocnt = oracle.gather(grid, ou, "path", np.zeros(0, pyoracle.LIGHT_DTYPE))
        cnt = renderer.counters()
        assert (cnt["S_ref"], cnt["S_eff"]) == (ocnt["S_ref"], ocnt["S_eff"])
        assert np.array_equal(fc, ofc)


def test_progressive_accumulation(renderer, pyoracle, oracle, cloud):
    """draw_frame x4 (new lights every frame): float accumulator and rgba8 image vs the oracle's running averages."""
    grid = cloud(64)
    W, H = 64, 48
    _setup(renderer, grid, W, H, maxLights=40)
    renderer.set_light_gen_groups(4, 1, 1)
    renderer.reset_accumulation()
    acc = np.zeros((H, W, 4), np.float32)
    acc[..., 3] = 1.0
    img8 = np.zeros((H, W, 4), np.uint8)
    img8[..., 3] = 255
    for frame in range(1, 5):
        renderer.draw_frame("point")
        ou = pyoracle.default_ubo(W, H, frameCount=frame, maxLights=40)
        ol, oc, _, _ = oracle.light_gen(grid, ou, (4, 1, 1), 0)
        gl, gc, _ = renderer.read_lights()
        assert gc == oc and gl.tobytes() == ol.tobytes()
        ofc, _ = oracle.gather(grid, ou, "point", ol)
        oracle.accumulate(ofc, frame, acc, img8)
        rel, rmse = rel_err(renderer.read_accum()[..., 0], acc[..., 0])
        assert rel <= REL_TOL and rmse <= RMSE_TOL, (frame, rel, rmse)
        d8 = np.abs(renderer.read_image_rgba8().astype(int) - img8.astype(int)).max()
        assert d8 <= 1, (frame, d8)
    assert renderer.ubo.frameCount == 4
    assert (renderer.read_accum()[..., 3] == 1.0).all() and (renderer.read_image_rgba8()[..., 3] == 255).all()


def test_light_prefetch_is_transparent(vr, cloud):
    """Frames rendered with the next-frame light prefetch (adopted, and discarded after a parameter change, a light-gen
    dispatch, or a reset) equal frames rendered without it, bit for bit: lights, counters, accumulator."""
    grid = cloud(64)

    def run(prefetch):
        out = []
        with vr.VolumeRenderer(0) as r:
            r.set_light_prefetch(prefetch)
            r.set_grid(grid)
            r.resize(64, 48)
            r.ubo = r.default_ubo()
            r.update_ubo(maxLights=60)
            r.set_light_gen_groups(4, 1, 1)
            for f in range(1, 12):
                if f == 5:
                    r.update_ubo(photonInitialIntensity=70.0)        # prefetched lights for frame 5 are stale
                if f == 8:
                    r.set_light_gen_groups(6, 1, 1)
                if f == 10:
                    r.reset_accumulation()                           # frameCount restarts at 1
                r.draw_frame("ray" if f % 2 else "point")
                lights, count, emitted = r.read_lights()
                out.append((lights.tobytes(), count, emitted, r.counters()["E"], r.read_accum().tobytes(), r.ubo.frameCount))
        return out

    assert run(True) == run(False)


def test_frame_color_peer_stores(vr, cloud):
    """vr_set_frame_color_peers: the resolve / path kernels store frame colours to the local buffer and to every peer
    address (here: further buffers on the same GPU standing in for NVLink-mapped peers), only inside the rendered tile;
    vr_accumulate_frame_color then completes the progressive average from such a buffer."""
    import torch
    grid = cloud(64)
    W, H = 64, 48
    with vr.VolumeRenderer(0) as r, vr.VolumeRenderer(0) as full:
        for q in (r, full):
            q.set_grid(grid)
            q.resize(W, H)
            q.ubo = q.default_ubo()
            q.update_ubo(maxLights=40)
            q.set_light_gen_groups(4, 1, 1)
        bufs = [torch.full((H, W), -1.0, device="cuda") for _ in range(3)]
        torch.cuda.synchronize()
        for algo in ("point", "path"):
            for q in (r, full):
                q.reset_accumulation()
            for frame in range(2):
                full.draw_frame(algo)
                r.set_frame_color_peers(bufs[0].data_ptr(), [b.data_ptr() for b in bufs])
                r.draw_frame(algo, (0, 10, W, 20))          # rows 10..29 rendered here
                r.sync()
                want = full.read_frame_color()
                for k, b in enumerate(bufs):
                    got = b.cpu().numpy()
                    assert np.array_equal(got[10:30], want[10:30])
                    if frame == 0 and algo == "point" and k > 0:   # (the local buffer was cleared with the image)
                        assert (got[:10] == -1.0).all() and (got[30:] == -1.0).all()
                # "receive" the other rows (as a peer's stores would have delivered them), then average them in
                bufs[0][:10] = torch.from_numpy(want[:10]).cuda()
                bufs[0][30:] = torch.from_numpy(want[30:]).cuda()
                torch.cuda.synchronize()
                r.accumulate_frame_color((0, 0, W, 10))
                r.accumulate_frame_color((0, 30, W, H - 30))
                assert np.array_equal(r.read_accum(), full.read_accum())
                assert np.array_equal(r.read_image_rgba8(), full.read_image_rgba8())
        r.set_frame_color_peers(None, [])
        r.draw_frame("point")
        assert r.read_frame_color().max() > 0


def test_tiles_equal_full_frame(renderer, pyoracle, oracle, cloud):
    grid = cloud(64)
    W, H = 100, 70  # not a multiple of the 16x8 block
    ou = pyoracle.default_ubo(W, H, frameCount=1, maxLights=30)
    ol, _, _, _ = oracle.light_gen(grid, ou, (4, 1, 1), 0)
    _setup(renderer, grid, W, H, maxLights=30, frameCount=1)
    renderer.write_lights(ol)
    renderer.clear_image()
    renderer.dispatch_compute_color("ray")
    full = renderer.read_frame_color()
    renderer.clear_image()
    for tile in [(0, 0, 37, 70), (37, 0, 63, 33), (37, 33, 63, 37)]:
        renderer.dispatch_compute_color("ray", tile)
    tiled = renderer.read_frame_color()
    assert np.array_equal(full, tiled)


def test_cpu_test_bit_exact(renderer, vr, oracle):
    """Config 1: CPU_test's density ray-march, fog sphere in front of its fixed camera."""
    grid = vr.build_fog_sphere_grid(32.0, center=(0, 250, -200))
    renderer.set_grid(grid)
    d = renderer.cpu_test_render(256, 256)
    od, _ = oracle.cpu_test(grid, 256, 256, nthreads=8)
    assert od.max() > 10
    assert np.array_equal(d, od)


def test_edge_cases(renderer, vr, pyoracle, oracle, cloud):
    grid = cloud(64)
    W, H = 48, 32
    # lightCount == 0 -> 0/0 -> frame colour defined as 0
    _setup(renderer, grid, W, H, frameCount=1)
    renderer.write_lights(np.zeros(0, pyoracle.LIGHT_DTYPE))
    renderer.clear_image()
    renderer.dispatch_compute_color("point")
    assert (renderer.read_frame_color() == 0).all()
    # camera looking away from the volume: no pixel hits the bbox
    _setup(renderer, grid, W, H, frameCount=1, cameraPos=(0.0, 20.0, 200.0))
    ou = pyoracle.default_ubo(W, H, frameCount=1, cameraPos=(0.0, 20.0, 200.0))
    ol, _, _, _ = oracle.light_gen(grid, ou, (2, 1, 1), 0)
    renderer.write_lights(ol)
    renderer.dispatch_compute_color("point")
    assert renderer.counters()["S_ref"] == 0 and (renderer.read_frame_color() == 0).all()
    # camera inside the volume (tmin clamps to 0)
    _setup(renderer, grid, W, H, frameCount=1, cameraPos=(0.0, 20.0, 0.0))
    ou = pyoracle.default_ubo(W, H, frameCount=1, cameraPos=(0.0, 20.0, 0.0))
    renderer.write_lights(ol)
    renderer.clear_image()
    renderer.dispatch_compute_color("sphere")
    ofc, ocnt = oracle.gather(grid, ou, "sphere", ol)
    rel, rmse = rel_err(renderer.read_frame_color(), ofc)
    assert renderer.counters()["S_eff"] == ocnt["S_eff"] and rel <= REL_TOL and rmse <= RMSE_TOL
    # a light closer than the 1e-4 guard to a sample contributes 0, not infinity
    _setup(renderer, grid, W, H, frameCount=1)
    near = np.zeros(1, pyoracle.LIGHT_DTYPE)
    near["to"][0] = (0.0, 20.0, 0.0)
    near["from"][0] = (0.0, 19.0, 0.0)
    near["intensity"][0] = 50.0
    renderer.write_lights(near)
    renderer.clear_image()
    renderer.dispatch_compute_color("point")
    ou = pyoracle.default_ubo(W, H, frameCount=1)
    ofc, _ = oracle.gather(grid, ou, "point", near)
    rel, rmse = rel_err(renderer.read_frame_color(), ofc)
    assert rel <= REL_TOL and rmse <= RMSE_TOL


@pytest.mark.parametrize("algo", ["point", "ray"])
def test_near_light_guard_all_eval_variants(renderer, pyoracle, oracle, cloud, algo):
    """Lights placed ON camera samples (d2 = 0), inside the guard radius (d2 < 1e-4, contributes 0) and just outside it
    (huge but finite): every evaluation kernel must agree with the oracle, and the speculative kernels (2, 3, 8) must
    take their guarded repeat path and still equal the guarded packed kernel (1) bit for bit."""
    grid = cloud(64)
    W, H = 96, 64
    ou = pyoracle.default_ubo(W, H, frameCount=1, maxLights=64)
    base, _, _, _ = oracle.light_gen(grid, ou, (4, 1, 1), 0)
    _setup(renderer, grid, W, H, maxLights=64, frameCount=1)
    renderer.write_lights(base)
    renderer.clear_image()
    renderer.dispatch_compute_color("point")
    smp = renderer.read_samples()
    assert len(smp) == renderer.counters()["S_eff"] > 1000
    picks = smp[[len(smp) // 7, len(smp) // 2, (len(smp) * 6) // 7], :3]
    near = np.zeros(4, pyoracle.LIGHT_DTYPE)
    near["intensity"] = 25.0
    pos = np.stack([picks[0],                                     # exactly on a sample: d2 = 0
                    picks[1] + np.float32([0.004, 0.0, 0.0]),     # d2 ~ 1.6e-5 < 1e-4: contributes 0
                    picks[2] + np.float32([0.0, 0.03, 0.0]),      # d2 ~ 9e-4: contributes I/(4 pi d2), large
                    picks[2] + np.float32([0.0, 0.0, 0.5])])
    if algo == "ray":   # the sub-points of a ray light start at `from` (s = 0, ray_compute_color.comp:16-20)
        near["from"], near["to"] = pos, pos + np.float32([0.7, 0.2, 0.1])
    else:               # point lights sit at `to` (point_compute_color.comp:12)
        near["to"], near["from"] = pos, pos - np.float32([0.7, 0.2, 0.1])
    lights = np.concatenate([base[:20], near, base[20:40]])
    ofc, _ = oracle.gather(grid, ou, algo, lights, skip_zero=True)
    out = {}
    for variant in (0, 1, 2, 3, 8):
        renderer.set_eval_variant(variant)
        redo0 = renderer.eval_redo_count()
        renderer.write_lights(lights)
        renderer.clear_image()
        renderer.dispatch_compute_color(algo)
        out[variant] = renderer.read_frame_color()
        rel, rmse = rel_err(out[variant], ofc)
        assert rel <= REL_TOL and rmse <= RMSE_TOL, (variant, rel, rmse)
        assert np.isfinite(out[variant]).all()
        if variant >= 2:
            assert renderer.eval_redo_count() > redo0, "the guarded repeat path was not exercised"
    renderer.set_eval_variant(8)
    assert np.array_equal(out[1], out[2]) and np.array_equal(out[1], out[3])
    # the ray-segment kernel leaves every near-light sample to the exact kernel: same pixels up to the far-field model
    rel, rmse = rel_err(out[8], out[2])
    assert rel <= 1e-4, (rel, rmse)


@pytest.mark.parametrize("algo,cam", [("point", None), ("ray", None), ("sphere", None), ("beam", None),
                                      ("point", dict(cameraPos=(14.0, 31.0, -58.0), fov=70.0)),
                                      ("sphere", dict(cameraPos=(-9.0, 8.0, -40.0), fov=28.0)),
                                      ("point", dict(cameraPos=(3.0, 22.0, -6.0), fov=60.0))])   # camera inside the cloud
def test_segment_kernel_near_lights(renderer, pyoracle, oracle, cloud, algo, cam):
    """The ray-segment kernel (variant 8) models a pixel's samples as points on the camera ray; whatever comes closer than
    the near distance to a light must reach the exact kernel. 300 lights are dropped 0.02 .. 0.6 world units away from
    randomly picked camera samples (all directions): the frame must match the exact kernel far inside the tolerance and
    the oracle within it, whole-frame and as two tiles, and a light next to the camera plane must switch every segment
    to the exact kernel (then the results are the exact kernel's, bit for bit)."""
    grid = cloud(64)
    W, H = 160, 96
    cam = cam or {}
    ou = pyoracle.default_ubo(W, H, frameCount=1, maxLights=400, **cam)
    base, _, _, _ = oracle.light_gen(grid, ou, (4, 1, 1), 0)
    _setup(renderer, grid, W, H, maxLights=400, frameCount=1, **cam)
    renderer.set_eval_variant(2)
    renderer.write_lights(base[:40])
    renderer.clear_image()
    renderer.dispatch_compute_color("point")
    smp = renderer.read_samples()
    rng = np.random.default_rng(7)
    picks = smp[rng.integers(0, len(smp), 300), :3]
    dirs = rng.normal(size=(300, 3))
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    pos = (picks + dirs * rng.uniform(0.02, 0.6, (300, 1))).astype(np.float32)
    near = np.zeros(300, pyoracle.LIGHT_DTYPE)
    near["intensity"] = 0.05
    if algo in ("ray", "beam", "sphere"):   # sub-points start at `from`; sphere lights sit at `from` (sphere_compute_color.comp:12)
        near["from"], near["to"] = pos, pos + np.float32([0.5, 0.3, 0.2])
    else:
        near["to"], near["from"] = pos, pos - np.float32([0.5, 0.3, 0.2])
    lights = np.concatenate([base[:40], near])
    res = {}
    for variant in (2, 8):
        renderer.set_eval_variant(variant)
        renderer.write_lights(lights)
        renderer.clear_image()
        renderer.dispatch_compute_color(algo)
        res[variant] = renderer.read_frame_color()
    seg = renderer.eval_segment_counts()
    assert seg["segments"] > 200 and seg["exact"] > seg["leftovers"], seg
    rel, rmse = rel_err(res[8], res[2])
    assert rel <= 1e-4 and rmse <= 1e-6, (rel, rmse)
    ofc, _ = oracle.gather(grid, ou, algo, lights, skip_zero=True)
    rel, rmse = rel_err(res[8], ofc)
    assert rel <= REL_TOL and rmse <= RMSE_TOL, (rel, rmse)
    renderer.clear_image()
    renderer.dispatch_compute_color(algo, (0, 0, W, 40))
    renderer.dispatch_compute_color(algo, (0, 40, W, H - 40))
    assert np.array_equal(renderer.read_frame_color(), res[8])
    # a light within the near distance of the camera plane has no bounded pixel footprint on that side: the flag kernel
    # clamps the near plane and tests the tile's pixels exactly, so only segments that really come near are handed over
    # (here none: the cloud is far from that light) and the frame still matches the exact kernel
    cam = np.zeros(1, pyoracle.LIGHT_DTYPE)
    cam["to"] = cam["from"] = np.float32(renderer.ubo.cameraPos[:]) + np.float32([0.3, -0.2, 0.1])
    cam["intensity"] = 1.0
    lights2 = np.concatenate([lights[:50], cam])
    for variant in (2, 8):
        renderer.set_eval_variant(variant)
        renderer.write_lights(lights2)
        renderer.clear_image()
        renderer.dispatch_compute_color("point")
        res[variant] = renderer.read_frame_color()
    seg = renderer.eval_segment_counts()
    assert seg["exact"] < seg["segments"] * 8 + seg["leftovers"], seg
    rel, rmse = rel_err(res[8], res[2])
    assert rel <= 1e-4 and rmse <= 1e-6, (rel, rmse)
    # an absurdly distant light (products of squared distances would leave FP32): everything goes the exact way, and the
    # results are the exact kernel's bit for bit
    far = np.zeros(1, pyoracle.LIGHT_DTYPE)
    far["to"] = far["from"] = np.float32([1.0e6, 2.0e5, 3.0e5])
    far["intensity"] = 1.0e9
    lights3 = np.concatenate([lights[:50], far])
    for variant in (2, 8):
        renderer.set_eval_variant(variant)
        renderer.write_lights(lights3)
        renderer.clear_image()
        renderer.dispatch_compute_color("point")
        res[variant] = renderer.read_frame_color()
    seg = renderer.eval_segment_counts()
    assert seg["exact"] == seg["segments"] * 8 + seg["leftovers"], seg
    assert np.array_equal(res[8], res[2])


def test_async_readback_equals_blocking(renderer, cloud):
    """vr_read_accum_f32_async snapshots the accumulator at the point of the call: the frames enqueued afterwards (which
    overwrite it) must not leak into the copy, and a second request waits for the first."""
    _setup(renderer, cloud(64), 96, 64, maxLights=64)
    renderer.set_light_gen_groups(4, 1, 1)
    renderer.reset_accumulation()
    renderer.draw_frame("point")
    want1 = renderer.read_accum()
    renderer.reset_accumulation()
    renderer.draw_frame("point")
    a, b = np.zeros((64, 96, 4), np.float32), np.zeros((64, 96, 4), np.float32)
    renderer.read_accum_async(a)
    for _ in range(3):
        renderer.draw_frame("point")        # frames 2..4 change the accumulator while the copy is in flight
    renderer.read_accum_async(b)
    renderer.read_wait()
    assert np.array_equal(a, want1)
    assert np.array_equal(b, renderer.read_accum()) and not np.array_equal(a, b)
    renderer.set_light_gen_groups(1, 1, 1)


def test_argument_errors(renderer, vr, cloud):
    grid = cloud(64)
    _setup(renderer, grid, 32, 32)
    with pytest.raises(vr.VrError) as e:
        renderer.update_ubo(rayMarchingStepSize=0.0)
    assert e.value.code == 1
    with pytest.raises(vr.VrError) as e:
        renderer.update_ubo(maxLights=1000000)  # the reference UI allows it and overruns its buffer
    assert e.value.code == 4
    renderer.update_ubo(lightRayStepSize=0.0)
    with pytest.raises(vr.VrError):
        renderer.dispatch_compute_color("ray")
    renderer.update_ubo(lightRayStepSize=0.3)
    renderer.update_ubo(rayMarchingStepSize=1.0e-6)  # t += step would stop advancing: refused instead of hanging the GPU
    with pytest.raises(vr.VrError) as e:
        renderer.dispatch_compute_color("point")
    assert e.value.code == 1 and "too small" in str(e.value)
    # ... and so would the photon walk of light generation: vr_dispatch_light_gen and vr_render_frame refuse it before
    # anything is enqueued, and the rejected frame leaves frameCount where it was
    with pytest.raises(vr.VrError) as e:
        renderer.dispatch_light_gen("point")
    assert e.value.code == 1 and "too small" in str(e.value)
    fc0 = renderer.ubo.frameCount
    with pytest.raises(vr.VrError) as e:
        renderer.draw_frame("point")
    assert e.value.code == 1 and renderer.ubo.frameCount == fc0
    with pytest.raises(vr.VrError):
        renderer.draw_frame("point", (0, 0, 64, 64))   # tile outside the image: rejected up front as well
    assert renderer.ubo.frameCount == fc0
    renderer.update_ubo(rayMarchingStepSize=1.0)
    renderer.draw_frame("point")                        # the context is still usable
    renderer.sync()
    assert renderer.ubo.frameCount == fc0 + 1
    with pytest.raises(vr.VrError) as e:
        renderer.set_grid(np.zeros(4096, np.uint8))
    assert e.value.code == 3
    with pytest.raises(vr.VrError):
        renderer.dispatch_compute_color("point", (0, 0, 64, 64))
    r2 = vr.VolumeRenderer(0)
    with pytest.raises(vr.VrError) as e:
        r2.dispatch_light_gen("point")
    assert e.value.code == 5
    r2.close()


@pytest.mark.gpu
def test_set_grid_device_equals_set_grid(vr, cloud):
    """vr_set_grid_device (grid taken from device memory in stream order, header from the host) renders what vr_set_grid
    renders, bit for bit; a header that stops before the root node is refused; tiles.ReplicatedGrid (one upload, staging
    buffers alternating) drives it the same way."""
    import torch
    from volumerenderer_b200 import tiles

    grid_a, grid_b = cloud(64), cloud(32)
    r = vr.VolumeRenderer(0)
    try:
        r.resize(96, 64)
        r.update_ubo(maxLights=64)

        def frames(setter):
            out = []
            for g in (grid_a, grid_b, grid_a):
                setter(g)
                r.reset_accumulation()
                r.draw_frame("point")
                r.draw_frame("point")
                out.append((r.read_accum(), r.read_lights()[0].tobytes()))
            return out

        want = frames(r.set_grid)

        def from_device(g):
            t = torch.from_numpy(g).cuda()
            torch.cuda.synchronize()  # the renderer runs on its own stream here: the source must be complete
            r.set_grid_device(t.data_ptr(), g.nbytes, g[:1024])
            r.sync()  # t goes out of scope: keep it alive until the copy has run

        got = frames(from_device)
        for (a, la), (b, lb) in zip(want, got):
            assert np.array_equal(a, b) and la == lb and len(la) > 0
        with pytest.raises(vr.VrError):
            r.set_grid_device(torch.from_numpy(grid_a).cuda().data_ptr(), grid_a.nbytes, grid_a[:700])
        # ReplicatedGrid on one rank: renderer on torch's stream, pinned source, no host synchronisation in between
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            r.set_stream(st.cuda_stream)
            pinned = {id(g): torch.from_numpy(g).pin_memory() for g in (grid_a, grid_b)}
            rg = {g.nbytes: tiles.ReplicatedGrid(r, g.nbytes) for g in (grid_a, grid_b)}
            got2 = frames(lambda g: rg[g.nbytes].set_grid(pinned[id(g)], g[:4096]))
            r.sync()
        r.set_stream(None)
        for (a, la), (b, lb) in zip(want, got2):
            assert np.array_equal(a, b) and la == lb and len(la) > 0
    finally:
        r.close()


@pytest.mark.gpu
def test_device_grid_installs_do_not_wait_on_the_host(vr, cloud):
    """A host that runs ahead: a new volume before every frame through vr_set_grid_device / ReplicatedGrid
    (vr_set_grid_device_from_stream), no host synchronisation anywhere in the sequence. The copy into the spare grid buffer
    waits ON THE DEVICE for that buffer's last readers (the frame and the light generation of two grids ago), so the
    accumulated image and the last light list must equal the ones of the same sequence driven through the blocking
    vr_set_grid."""
    import torch
    from volumerenderer_b200 import tiles

    grids = [cloud(64), cloud(64, seed=7), cloud(32)]
    order = [0, 1, 2, 1, 0, 2, 0, 1, 1, 2]
    r = vr.VolumeRenderer(0)
    try:
        r.resize(160, 96)
        r.update_ubo(maxLights=96)
        r.set_light_gen_groups(4)

        def run(install):
            for g in grids[:2] + grids[:2]:  # both grid buffers at their final capacity before the sequence starts
                r.set_grid(g)
            r.reset_accumulation()
            for i in order:
                install(i)
                r.draw_frame("point")
            r.sync()
            return r.read_accum(), r.read_lights()[0].tobytes(), r.counters()["S_eff"]

        want = run(lambda i: r.set_grid(grids[i]))
        assert want[0].any() and len(want[1]) > 0
        dev = [torch.from_numpy(g).cuda() for g in grids]
        torch.cuda.synchronize()
        got = run(lambda i: r.set_grid_device(dev[i].data_ptr(), grids[i].nbytes, grids[i][:4096]))
        assert np.array_equal(want[0], got[0]) and want[1] == got[1] and want[2] == got[2]
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            r.set_stream(st.cuda_stream)
            pinned = [torch.from_numpy(g).pin_memory() for g in grids]
            rg = {g.nbytes: tiles.ReplicatedGrid(r, g.nbytes) for g in grids}
            got2 = run(lambda i: rg[grids[i].nbytes].set_grid(pinned[i], grids[i][:4096]))
        r.set_stream(None)
        assert np.array_equal(want[0], got2[0]) and want[1] == got2[1] and want[2] == got2[2]
    finally:
        r.close()


@pytest.mark.gpu
@pytest.mark.parametrize("algo", ["point", "beam"])
def test_single_pass_march_equals_two_pass(vr, cloud, algo):
    """The single-pass march (slots from the lookup-free step bound, sample buffer with gaps) and the two-pass march
    (count, scan, fill: compact buffer) give the same samples, counters and images bit for bit, on the whole frame, on
    tiles, with a new grid every frame and with the march cache."""
    from volumerenderer_b200.tiles import band_tile

    grid, grid2 = cloud(64), cloud(32)
    r = vr.VolumeRenderer(0)
    try:
        W, H = 160, 96
        r.resize(W, H)
        r.update_ubo(maxLights=48, beamRadius=0.4)

        def run(single):
            r.set_march_single_pass(single)
            out = []
            for g in (grid, grid2, grid):
                r.set_grid(g)
                r.reset_accumulation()
                for _ in range(3):
                    r.draw_frame(algo)
                out.append((r.read_accum(), r.read_image_rgba8(), r.read_samples(), r.counters(), r.row_sample_counts()))
            # two tiles of one frame, then the march cache on repeated frames
            r.clear_image()
            r.dispatch_compute_color(algo, band_tile((0, 40), W))
            s0 = r.read_samples()
            r.dispatch_compute_color(algo, band_tile((40, H - 40), W))
            out.append((r.read_frame_color(), s0, r.read_samples()))
            r.set_march_cache(True)
            r.reset_accumulation()
            for _ in range(3):
                r.draw_frame(algo)
            r.set_march_cache(False)
            out.append((r.read_accum(),))
            return out

        two, one = run(False), run(True)
        for a, b in zip(two, one):
            for x, y in zip(a, b):
                if isinstance(x, dict):
                    assert x == y
                else:
                    assert np.array_equal(x, y)
        assert two[0][3]["S_eff"] > 10000 and len(two[0][2]) == two[0][3]["S_eff"]
    finally:
        r.close()


@pytest.mark.parametrize("algo", ["point", "beam"])
def test_eval_side_stream_is_transparent(vr, cloud, algo):
    """vr_set_eval_side_stream: the small evaluation kernels on a second stream or behind the big one: same bits."""
    grid = cloud(64)
    out = {}
    for side in (1, 0):
        with vr.VolumeRenderer(0) as r:
            _setup(r, grid, 160, 96, maxLights=120)
            r.set_eval_side_stream(side)
            r.set_light_gen_groups(8, 1, 1)
            for _ in range(3):
                r.draw_frame(algo)
            out[side] = (r.read_accum(), r.read_image_rgba8(), r.eval_segment_counts())
    assert np.array_equal(out[0][0], out[1][0]) and np.array_equal(out[0][1], out[1][1]) and out[0][2] == out[1][2]


def test_tile_readback_async(renderer, cloud):
    """vr_read_accum_f32_tile_async: the rows (or rectangle) a GPU rendered, placed into the caller's full image."""
    W, H = 96, 64
    _setup(renderer, cloud(64), W, H, maxLights=64)
    renderer.set_light_gen_groups(4, 1, 1)
    renderer.reset_accumulation()
    renderer.draw_frame("point")
    want = renderer.read_accum()
    img = np.full((H, W, 4), -1.0, np.float32)
    renderer.read_accum_tile_async((0, 16, W, 24), img)
    renderer.draw_frame("point")                       # overwrites the accumulator while the copy is in flight
    renderer.read_wait()
    assert np.array_equal(img[16:40], want[16:40]) and (img[:16] == -1.0).all() and (img[40:] == -1.0).all()
    img[:] = -1.0
    want = renderer.read_accum()
    renderer.read_accum_tile_async((8, 4, 40, 10), img)
    renderer.read_wait()
    assert np.array_equal(img[4:14, 8:48], want[4:14, 8:48])
    img[4:14, 8:48] = -1.0
    assert (img == -1.0).all()
    renderer.set_light_gen_groups(1, 1, 1)


def test_sample_parallel_progressive_sum(vr, cloud):
    """SURVEY 8e, sample-parallel mode: G contexts render every G-th frame of a progressive run (vr_set_frame_step) into
    float SUM accumulators (vr_set_accumulate_mode); sum of the sums / frames equals the sequential running average of
    shaders/point_compute_color.comp:99-105 up to float reassociation (bound: frames * 2^-24 relative), with exactly
    the light lists of the sequential run."""
    grid = cloud(64)
    W, H, G, frames = 96, 64, 3, 12
    with vr.VolumeRenderer(0) as seq:
        _setup(seq, grid, W, H, maxLights=80)
        seq.set_light_gen_groups(6, 1, 1)
        lists = []
        for f in range(frames):
            seq.draw_frame("point")
            lists.append(seq.read_lights()[0].tobytes())
        want = seq.read_accum()
    total = np.zeros((H, W, 4), np.float64)
    for rank in range(G):
        with vr.VolumeRenderer(0) as r:
            _setup(r, grid, W, H, maxLights=80)
            r.set_light_gen_groups(6, 1, 1)
            r.set_frame_step(rank + 1, G)
            r.set_accumulate_mode("sum")
            r.clear_image()
            for k in range(frames // G):
                r.draw_frame("point")
                assert r.ubo.frameCount == rank + 1 + k * G
                assert r.read_lights()[0].tobytes() == lists[rank + k * G]
            part = r.read_accum()
            total += part
            if rank == 0:   # finish on one context: sum / frames, rgba8 from the average
                r.finish_sum(frames // G)
                avg = r.read_accum()
                assert np.allclose(avg[..., :3], part[..., :3] / np.float32(frames // G), rtol=1e-6, atol=0)
                assert np.array_equal(r.read_image_rgba8()[..., 0], np.rint(np.clip(avg[..., 0], 0, 1) * 255).astype(np.uint8))
                assert r.ubo.frameCount == frames // G
    got = (total[..., 0] / frames).astype(np.float32)
    rel, rmse = rel_err(got, want[..., 0])
    assert rel <= frames * 2.0 ** -22 and want[..., 0].max() > 0.01, (rel, rmse)
