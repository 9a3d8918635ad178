"""The CUDA path against the committed golden vectors (outputs of the reference's own shader source and of its vendored
NanoVDB, tools/make_golden.py). Runs on the GPU box, where /root/reference does not exist."""
import gzip
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
MAN = json.load(open(os.path.join(GOLD, "manifest.json")))
W, H = MAN["image"]
REL_TOL, RMSE_TOL = 1e-3, 1e-4  # north_star tolerance on linear radiance


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def load_gz(name):
    with gzip.open(os.path.join(GOLD, name), "rb") as f:
        return np.frombuffer(f.read(), np.uint8).copy()


@pytest.fixture(scope="module")
def r(vr, cloud):
    rr = vr.VolumeRenderer(0)
    rr.set_grid(cloud(64))
    rr.resize(W, H)
    yield rr
    rr.close()


@pytest.mark.parametrize("name", ["nanovdb_cloud32", "nanovdb_ball80"])
def test_device_traversal_on_nanovdb_built_grids(vr, name):
    g = load_gz(name + ".nvdb.raw.gz")
    ijk = np.load(os.path.join(GOLD, name + "_ijk.npy"))
    want = np.load(os.path.join(GOLD, name + "_values.npy"))
    with vr.VolumeRenderer(0) as rr:
        rr.set_grid(g)
        assert np.array_equal(rr.lookup_values(ijk), want)
        # sorted order makes consecutive lookups hit the accessor's cached leaf / lower / upper nodes
        order = np.lexsort((ijk[:, 2], ijk[:, 1], ijk[:, 0]))
        assert np.array_equal(rr.lookup_values(ijk[order]), want[order])


@pytest.mark.parametrize("name", [k for k in MAN["cases"]["light_gen"] if k != "path_light_gen_is_empty"])
def test_light_gen_golden(r, name):
    c = MAN["cases"]["light_gen"][name]
    r.ubo = r.default_ubo()
    r.update_ubo(frameCount=c["frame"], maxLights=c["max_lights"])
    r.reset_lights()
    algo = {"light_gen": "point"}.get(c["shader"], c["shader"])
    r.dispatch_light_gen(algo, tuple(c["groups"]), c["variant"])
    lights, count, _ = r.read_lights()
    assert count == c["count"] and sha(lights) == c["sha256"]


@pytest.mark.parametrize("algo", ["point", "sphere", "ray", "beam"])
@pytest.mark.parametrize("variant", [0, 1, 8])
def test_gather_golden(r, algo, variant):
    r.ubo = r.default_ubo()
    r.update_ubo(frameCount=1, maxLights=100)
    r.set_eval_variant(variant)
    r.write_lights(np.load(os.path.join(GOLD, "lights_f1_8x1x1_cap100_reset.npy")))
    r.clear_image()
    r.dispatch_compute_color(algo)
    want = np.load(os.path.join(GOLD, f"frame_{algo}.npy"))
    rel, rmse = rel_err(r.read_frame_color(), want)
    r.set_eval_variant(8)
    assert rel <= REL_TOL and rmse <= RMSE_TOL, (rel, rmse)


@pytest.mark.parametrize("frame", [1, 3])
def test_path_golden_bit_exact(r, frame):
    r.ubo = r.default_ubo()
    r.update_ubo(frameCount=frame)
    r.clear_image()
    r.dispatch_compute_color("path")
    want = np.load(os.path.join(GOLD, f"frame_path_f{frame}.npy"))  # storage image: (0*(n-1) + frameColor)/n
    assert np.array_equal(r.read_accum()[..., 0], want)


def test_accumulation_golden(r):
    c = MAN["cases"]["accumulate"]
    r.ubo = r.default_ubo()
    r.update_ubo(maxLights=c["max_lights"])
    r.set_light_gen_groups(*c["groups"])
    r.reset_accumulation()
    for _ in range(c["frames"]):
        r.draw_frame(c["algo"])
    rel, rmse = rel_err(r.read_accum()[..., 0], np.load(os.path.join(GOLD, "accum3_point_f32.npy")))
    assert rel <= REL_TOL and rmse <= RMSE_TOL, (rel, rmse)
    d8 = np.abs(r.read_image_rgba8().astype(int) - np.load(os.path.join(GOLD, "accum3_point_rgba8.npy")).astype(int))
    assert d8.max() <= 1  # rgba8-faithful mode: +-1 LSB
    r.set_light_gen_groups(1, 1, 1)


def test_cpu_test_golden_bit_exact(vr):
    sphere = vr.build_fog_sphere_grid(32.0, center=(0, 250, -200))
    assert sha(sphere) == MAN["sphere_sha256"]
    with vr.VolumeRenderer(0) as rr:
        rr.set_grid(sphere)
        assert np.array_equal(rr.cpu_test_render(64, 64), np.load(os.path.join(GOLD, "cpu_test_density_64.npy")))
