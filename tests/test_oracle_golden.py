"""The oracle (CPU restatement) against the committed golden vectors, which are OUTPUTS OF THE REFERENCE'S OWN CODE:
its shader sources compiled as C++ and its vendored NanoVDB (tools/make_golden.py). Bit-exact everywhere."""
import gzip
import hashlib
import json
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(__file__), "golden")
MAN = json.load(open(os.path.join(GOLD, "manifest.json")))
W, H = MAN["image"]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def load_gz(name):
    with gzip.open(os.path.join(GOLD, name), "rb") as f:
        return np.frombuffer(f.read(), np.uint8).copy()


@pytest.fixture(scope="module")
def cloud64(cloud):
    g = cloud(64)
    assert sha(g) == MAN["cloud64_sha256"], "the host-side grid writer must reproduce the fixture grid bit for bit"
    return g


def test_grid_facts(oracle, cloud64):
    info = oracle.grid_info(cloud64)
    f = MAN["cloud64_facts"]
    assert info["bbox_min"].tolist() == f["bbox_min"] and info["bbox_max"].tolist() == f["bbox_max"]
    assert info["node_counts"].tolist() == f["node_counts"] and info["background"] == f["background"]


@pytest.mark.parametrize("name", ["nanovdb_cloud32", "nanovdb_ball80"])
def test_lookup_on_nanovdb_built_grids(oracle, name):
    """Grids built by the vendored NanoVDB (createNanoGrid); values from NanoGrid::tree().getValue. ball80 is the
    known-answer grid of SURVEY Appendix D: interior leaves pruned to lower-level tiles, root bbox [11,69]^3."""
    g = load_gz(name + ".nvdb.raw.gz")
    assert sha(g) == MAN[name]["sha256"]
    info = oracle.grid_info(g)
    assert info["bbox_min"].tolist() == MAN[name]["facts"]["bbox_min"]
    assert info["bbox_max"].tolist() == MAN[name]["facts"]["bbox_max"]
    assert info["node_counts"].tolist() == MAN[name]["facts"]["node_counts"]
    ijk = np.load(os.path.join(GOLD, name + "_ijk.npy"))
    want = np.load(os.path.join(GOLD, name + "_values.npy"))
    assert np.array_equal(oracle.lookup(g, ijk), want)
    if name == "nanovdb_ball80":
        assert g.nbytes == 772480 and info["node_counts"].tolist() == [218, 1, 1]
        assert oracle.lookup(g, [[40, 40, 40]])[0] == 0.5  # inside a pruned (tile) region
        assert oracle.index_to_world(g, [0, 0, 0]).tolist() == [-32.0, -12.0, -32.0]
        assert oracle.world_to_index(g, [-32.0, -12.0, -32.0]).tolist() == [0.0, 0.0, 0.0]


@pytest.mark.parametrize("name", [k for k in MAN["cases"]["light_gen"] if k != "path_light_gen_is_empty"])
def test_light_gen_golden(pyoracle, oracle, cloud64, name):
    c = MAN["cases"]["light_gen"][name]
    u = pyoracle.default_ubo(W, H, frameCount=c["frame"], maxLights=c["max_lights"])
    lights, count, emitted, _ = oracle.light_gen(cloud64, u, tuple(c["groups"]), c["variant"])
    want = np.load(os.path.join(GOLD, f"lights_{name}.npy"))
    assert count == c["count"] and emitted >= count
    assert lights.tobytes() == want.tobytes() and sha(lights) == c["sha256"]


@pytest.mark.parametrize("algo", ["point", "sphere", "ray", "beam"])
def test_gather_golden(pyoracle, oracle, cloud64, algo):
    u = pyoracle.default_ubo(W, H, frameCount=1, maxLights=100)
    lights = np.load(os.path.join(GOLD, "lights_f1_8x1x1_cap100_reset.npy"))
    want = np.load(os.path.join(GOLD, f"frame_{algo}.npy"))
    faithful, c0 = oracle.gather(cloud64, u, algo, lights, skip_zero=False)
    fast, c1 = oracle.gather(cloud64, u, algo, lights, skip_zero=True)
    assert np.array_equal(faithful, want)
    assert np.array_equal(fast, want), "skipping zero-density samples must be exact"
    assert c0["S_ref"] == c1["S_ref"] and c0["S_eff"] == c1["S_eff"] and c0["E"] > c1["E"]


@pytest.mark.parametrize("frame", [1, 3])
def test_path_golden(pyoracle, oracle, cloud64, frame):
    u = pyoracle.default_ubo(W, H, frameCount=frame)
    fc, _ = oracle.gather(cloud64, u, "path", np.zeros(0, pyoracle.LIGHT_DTYPE))
    want = np.load(os.path.join(GOLD, f"frame_path_f{frame}.npy"))  # (0*(n-1) + frameColor) / n from a cleared image
    acc = np.zeros((H, W, 4), np.float32)
    oracle.accumulate(fc, frame, acc, None, path_mode=True)
    assert np.array_equal(acc[..., 0], want)


def test_accumulation_golden(pyoracle, oracle, cloud64):
    c = MAN["cases"]["accumulate"]
    acc = np.zeros((H, W, 4), np.float32)
    img8 = np.zeros((H, W, 4), np.uint8)
    for frame in range(1, c["frames"] + 1):
        u = pyoracle.default_ubo(W, H, frameCount=frame, maxLights=c["max_lights"])
        lights, _, _, _ = oracle.light_gen(cloud64, u, tuple(c["groups"]), 0)
        fc, _ = oracle.gather(cloud64, u, c["algo"], lights)
        oracle.accumulate(fc, frame, acc, img8)
    assert np.array_equal(acc[..., 0], np.load(os.path.join(GOLD, "accum3_point_f32.npy")))
    assert np.array_equal(img8, np.load(os.path.join(GOLD, "accum3_point_rgba8.npy")))


def test_cpu_test_golden(vr, oracle):
    sphere = vr.build_fog_sphere_grid(32.0, center=(0, 250, -200))
    assert sha(sphere) == MAN["sphere_sha256"]
    d, grey = oracle.cpu_test(sphere, 64, 64, nthreads=4)
    assert np.array_equal(d, np.load(os.path.join(GOLD, "cpu_test_density_64.npy")))
    assert d.max() > 10 and grey.max() == min(255, int(d.max() * 5.0))
