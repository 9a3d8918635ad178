"""BASELINE.json configurations at full size (1920x1080, 256^3 cloud): bit-exact light lists and work counters, a
pixel-stride comparison against the oracle, and size-independent properties (tiles, linearity, determinism)."""
import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu
W, H = 1920, 1080
REL_TOL, RMSE_TOL = 1e-3, 1e-4
DEFAULT_VARIANT = 8  # the library default: ray-segment kernel for point/ray lists


@pytest.fixture(scope="module")
def scene(vr, pyoracle, oracle, cloud):
    grid = cloud(256)
    r = vr.VolumeRenderer(0)
    r.set_grid(grid)
    r.resize(W, H)
    yield r, grid
    r.close()


@pytest.mark.parametrize("algo,max_lights,gx,stride", [("point", 1000, 32, 16), ("sphere", 1000, 32, 24), ("ray", 4096, 128, 40),
                                                       ("beam", 512, 16, 40)])
def test_config_against_oracle_on_a_pixel_stride(scene, pyoracle, oracle, algo, max_lights, gx, stride):
    r, grid = scene
    r.ubo = r.default_ubo()
    r.update_ubo(maxLights=max_lights, frameCount=0)
    r.set_light_gen_groups(gx, 1, 1)
    r.reset_accumulation()
    r.draw_frame(algo)
    lights, count, emitted = r.read_lights()
    ou = pyoracle.default_ubo(W, H, frameCount=1, maxLights=max_lights)
    ol, oc, oem, _ = oracle.light_gen(grid, ou, (gx, 1, 1), 0)
    assert (count, emitted) == (oc, oem) and lights.tobytes() == ol.tobytes()
    fc = r.read_frame_color()
    ofc, ocnt = oracle.gather(grid, ou, algo, ol, stride=stride)
    rel, rmse = rel_err(fc[::stride, ::stride], ofc[::stride, ::stride])
    assert rel <= REL_TOL and rmse <= RMSE_TOL, (algo, rel, rmse)
    assert ofc[::stride, ::stride].mean() > 0.01
    r.set_light_gen_groups(1, 1, 1)


def test_full_frame_counters_match_oracle_march(scene, pyoracle, oracle):
    r, grid = scene
    r.ubo = r.default_ubo()
    r.update_ubo(frameCount=1)
    r.write_lights(np.zeros(0, pyoracle.LIGHT_DTYPE))
    r.dispatch_compute_color("point")
    cnt = r.counters()
    ou = pyoracle.default_ubo(W, H, frameCount=1)
    _, ocnt = oracle.gather(grid, ou, "point", np.zeros(0, pyoracle.LIGHT_DTYPE))
    assert (cnt["S_ref"], cnt["S_eff"]) == (ocnt["S_ref"], ocnt["S_eff"])
    rows = r.row_sample_counts()
    assert int(rows.sum()) == cnt["S_eff"]


def test_tiles_linearity_determinism(scene, pyoracle, oracle):
    r, grid = scene
    ou = pyoracle.default_ubo(W, H, frameCount=1, maxLights=200, photonInitialIntensity=10.0)
    ol, _, _, _ = oracle.light_gen(grid, ou, (8, 1, 1), 0)
    r.ubo = r.default_ubo()
    r.update_ubo(frameCount=1, maxLights=200)
    r.write_lights(ol)
    r.clear_image()
    r.dispatch_compute_color("point")
    full = r.read_frame_color()
    r.clear_image()
    r.dispatch_compute_color("point")
    assert np.array_equal(full, r.read_frame_color()), "the gather must be deterministic"
    # image tiles (row bands like the multi-GPU partition) reproduce the full frame bit for bit
    from volumerenderer_b200.tiles import balanced_row_bands, band_tile
    bands = balanced_row_bands(r.row_sample_counts(), 3)
    r.clear_image()
    for b in bands:
        r.dispatch_compute_color("point", band_tile(b, W))
    assert np.array_equal(full, r.read_frame_color())
    # linear in the light intensities wherever the clamp is inactive: doubling I doubles the colour (exact in FP32)
    ol2 = ol.copy()
    ol2["intensity"] *= 2.0
    r.write_lights(ol2)
    r.clear_image()
    r.dispatch_compute_color("point")
    twice = r.read_frame_color()
    m = (twice < 1.0) & (full > 0)
    assert m.sum() > 100000 and np.array_equal(twice[m], 2.0 * full[m])
    # the exact evaluation kernels agree with the default (ray-segment, variant 8) far inside the tolerance, with each
    # other to rounding, and the guarded and speculative packed kernels bit for bit
    res = {}
    for variant in (0, 1, 2, 3):
        r.set_eval_variant(variant)
        r.write_lights(ol)
        r.clear_image()
        r.dispatch_compute_color("point")
        res[variant] = r.read_frame_color()
        rel, rmse = rel_err(full, res[variant])
        assert rel <= 1e-4 and rmse <= 1e-6, (variant, rel, rmse)
    r.set_eval_variant(DEFAULT_VARIANT)
    rel, rmse = rel_err(res[0], res[2])
    assert rel <= 1e-5 and rmse <= 1e-7, (rel, rmse)
    assert np.array_equal(res[1], res[2]) and np.array_equal(res[1], res[3])
    seg = r.eval_segment_counts()
    assert seg["segments"] * 8 + seg["leftovers"] == r.counters()["S_eff"] and seg["exact"] >= seg["leftovers"]


def test_sphere_kernels_agree(scene, pyoracle, oracle):
    """packed sphere kernel (merged roundings) against the scalar kernel that keeps every rounding of the shader"""
    r, grid = scene
    ou = pyoracle.default_ubo(W, H, frameCount=1, maxLights=300)
    ol, _, _, _ = oracle.light_gen(grid, ou, (8, 1, 1), 0)
    r.ubo = r.default_ubo()
    r.update_ubo(frameCount=1, maxLights=300)
    res = {}
    for variant in (0, 2):
        r.set_eval_variant(variant)
        r.write_lights(ol)
        r.clear_image()
        r.dispatch_compute_color("sphere")
        res[variant] = r.read_frame_color()
    r.set_eval_variant(DEFAULT_VARIANT)
    rel, rmse = rel_err(res[2], res[0])
    assert rel <= REL_TOL and rmse <= 1e-6, (rel, rmse)
