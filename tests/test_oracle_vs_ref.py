"""The oracle against oracle/_ref (the reference's shader sources compiled as C++), live, on more cases than the
committed golden set. Skipped where oracle/_ref was never built (no reference checkout)."""
import numpy as np
import pytest


@pytest.mark.parametrize("frame,groups,max_lights", [(1, (1, 1, 1), 1000), (4, (6, 1, 1), 64), (11, (3, 2, 2), 25)])
def test_light_gen_all_shader_bodies(pyoracle, oracle, ref, cloud, frame, groups, max_lights):
    grid = cloud(64)
    u = pyoracle.default_ubo(64, 64, frameCount=frame, maxLights=max_lights)
    for shader, variant in [("light_gen", 0), ("beam", 0), ("point", 1), ("ray", 1), ("sphere", 1)]:
        a, ca, _, _ = oracle.light_gen(grid, u, groups, variant)
        b, cb = ref.light_gen(grid, u, groups, shader)
        assert ca == cb and a.tobytes() == b.tobytes(), shader
    _, c = ref.light_gen(grid, u, groups, "path")
    assert c == 0


@pytest.mark.parametrize("algo", ["point", "sphere", "ray", "beam", "path"])
def test_gather_bit_exact(pyoracle, oracle, ref, cloud, algo):
    grid = cloud(64)
    W, H = 40, 30
    for frame, cam in [(1, (0.0, 20.0, -75.0)), (2, (10.0, 25.0, -60.0)), (5, (0.0, 20.0, 0.0))]:
        u = pyoracle.default_ubo(W, H, frameCount=frame, maxLights=48, cameraPos=cam)
        lights, _, _, _ = oracle.light_gen(grid, u, (4, 1, 1), 0)
        fc, _ = oracle.gather(grid, u, algo, lights, skip_zero=False)
        acc = np.zeros((H, W, 4), np.float32)
        oracle.accumulate(fc, frame, acc, None, path_mode=(algo == "path"))
        img = ref.gather(grid, u, algo, lights)
        assert np.array_equal(acc[..., 0], img[..., 0]), (algo, frame)


def test_non_uniform_step_and_absorption(pyoracle, oracle, ref, cloud):
    grid = cloud(64)
    u = pyoracle.default_ubo(32, 24, frameCount=3, maxLights=80, rayMarchingStepSize=0.7, absorptionCoefficient=0.11,
                             scatteringProbability=0.2, lightRayStepSize=0.45, beamRadius=0.3)
    a, ca, _, _ = oracle.light_gen(grid, u, (4, 1, 1), 0)
    b, cb = ref.light_gen(grid, u, (4, 1, 1), "light_gen")
    assert ca == cb and a.tobytes() == b.tobytes()
    for algo in ("point", "beam"):
        fc, _ = oracle.gather(grid, u, algo, a)
        img = ref.gather(grid, u, algo, a)
        acc = np.zeros((24, 32, 4), np.float32)
        oracle.accumulate(fc, 3, acc, None)
        assert np.array_equal(acc[..., 0], img[..., 0])


def test_rgba8_image_matches_reference_quantisation(pyoracle, oracle, ref, cloud):
    grid = cloud(64)
    W, H = 32, 24
    img = np.zeros((H, W, 4), np.float32)
    img8 = np.zeros((H, W, 4), np.uint8)
    for frame in (1, 2, 3, 4):
        u = pyoracle.default_ubo(W, H, frameCount=frame, maxLights=30)
        lights, _, _, _ = oracle.light_gen(grid, u, (2, 1, 1), 0)
        ref.gather(grid, u, "sphere", lights, image_rgba=img, unorm8=True)
        fc, _ = oracle.gather(grid, u, "sphere", lights)
        oracle.accumulate(fc, frame, None, img8)
    assert np.array_equal(np.rint(img[..., :3] * 255).astype(np.uint8), img8[..., :3])


def test_lookup_matches_nanovdb_accessors(oracle, ref):
    """PNanoVDB-layout restatement vs nanovdb::ReadAccessor and tree().getValue on a NanoVDB-built fog sphere
    (mirrors external/openvdb/nanovdb/nanovdb/unittest/TestNanoVDB.cc:4531 createFogVolumeSphere, :5672 accessor)."""
    g = ref.make_fog_sphere(30.0, voxel=1.0)
    facts = ref.grid_facts(g)
    assert facts["bbox_min"].tolist() == [-30, -30, -30] and facts["bbox_max"].tolist() == [30, 30, 30]
    ax = np.arange(-34, 35, 3)
    ijk = np.stack(np.meshgrid(ax, ax, ax, indexing="ij"), -1).reshape(-1, 3).astype(np.int32)
    far = np.array([[5000, 0, 0], [-4097, 3, 9], [0, 0, 0], [2 ** 20, -2 ** 20, 7]], np.int32)
    ijk = np.concatenate([ijk, far])
    a = oracle.lookup(g, ijk)
    assert np.array_equal(a, ref.get_values(g, ijk)) and np.array_equal(a, ref.get_values(g, ijk, accessor=True))
    assert a[-4] == facts["background"] == 3.0  # levelSetToFog leaves the root background at the SDF half-width
    assert oracle.lookup(g, [[0, 0, 0]])[0] == 1.0


def test_cpu_test_matches_nanovdb_getvalue(oracle, ref):
    g = ref.make_fog_sphere(32.0, center=(0.0, 250.0, -200.0))
    a, _ = oracle.cpu_test(g, 48, 48, nthreads=4)
    assert np.array_equal(a, ref.cpu_test(g, 48, 48, nthreads=4))
    assert a.max() > 10
