#!/usr/bin/env python3
"""Generate tests/golden/ from the reference's own code (oracle/_ref): run in a container that has /root/reference.

Every vector below is an OUTPUT OF THE REFERENCE'S SHADER SOURCE (shaders/*.comp compiled as C++ through
oracle/ref/glsl_shim.h) or of its vendored NanoVDB (getValue, createNanoGrid), not of our restatement. The committed
fixtures let the oracle and the CUDA path be checked on machines without the reference checkout (the GPU box).

  python tools/make_golden.py        # rewrites tests/golden/*
"""
import gzip
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import volumerenderer_b200 as vr  # noqa: E402  (host-side grid writer only)
from oracle import pyoracle  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
W, H = 48, 32


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = pyoracle.Ref()
    man = {"generator": "tools/make_golden.py via oracle/_ref (reference shaders compiled as C++, vendored NanoVDB 32.7.0)",
           "image": [W, H], "cases": {}}

    # ---- grids. cloud64: our writer (deterministic); the reference side only consumes the bytes.
    cloud = vr.build_cloud_grid(64)
    man["cloud64_sha256"] = sha(cloud)
    man["cloud64_facts"] = {k: (v.tolist() if hasattr(v, "tolist") else v) for k, v in ref.grid_facts(cloud).items()}

    # ---- a grid BUILT BY the vendored NanoVDB (functor route), committed gzipped, with getValue known answers
    nv = ref.make_cloud(32, seed=7, voxel=2.0, translation=(-32.0, -12.0, -32.0))
    with gzip.GzipFile(os.path.join(OUT, "nanovdb_cloud32.nvdb.raw.gz"), "wb", mtime=0) as f:
        f.write(nv.tobytes())
    rng = np.random.default_rng(1)
    ijk = np.concatenate([rng.integers(-8, 40, (4000, 3)), rng.integers(-5000, 5000, (200, 3)),
                          np.array([[0, 0, 0], [31, 31, 31], [-1, -1, -1], [4096, 0, 0], [16, 16, 16]])]).astype(np.int32)
    np.save(os.path.join(OUT, "nanovdb_cloud32_ijk.npy"), ijk)
    np.save(os.path.join(OUT, "nanovdb_cloud32_values.npy"), ref.get_values(nv, ijk))
    man["nanovdb_cloud32"] = {"sha256": sha(nv), "facts": {k: (v.tolist() if hasattr(v, "tolist") else v) for k, v in ref.grid_facts(nv).items()}}
    # the known-answer grid of SURVEY Appendix D: 80^3 box, value 0.5 inside a radius-30 ball (pruned interior tiles)
    i = np.arange(80)
    d2 = (i[:, None, None] - 40.0) ** 2 + (i[None, :, None] - 40.0) ** 2 + (i[None, None, :] - 40.0) ** 2
    ball = np.where(d2 < 900.0, np.float32(0.5), np.float32(0.0)).astype(np.float32)
    nb = ref.make_dense(ball, voxel=0.25, translation=(-32.0, -12.0, -32.0))
    with gzip.GzipFile(os.path.join(OUT, "nanovdb_ball80.nvdb.raw.gz"), "wb", mtime=0) as f:
        f.write(nb.tobytes())
    ijk2 = np.concatenate([rng.integers(-4, 84, (4000, 3)), np.array([[40, 40, 40], [11, 40, 40], [10, 40, 40], [69, 40, 40]])]).astype(np.int32)
    np.save(os.path.join(OUT, "nanovdb_ball80_ijk.npy"), ijk2)
    np.save(os.path.join(OUT, "nanovdb_ball80_values.npy"), ref.get_values(nb, ijk2))
    man["nanovdb_ball80"] = {"sha256": sha(nb), "size": int(nb.nbytes),
                             "facts": {k: (v.tolist() if hasattr(v, "tolist") else v) for k, v in ref.grid_facts(nb).items()}}

    # ---- RNG known answers (shaders/common_functions.h:2-25)
    man["rng"] = {"hash_111": ref.hash(1, 1, 1), "randf_111": ref.randf(1, 1, 1), "randf_222": ref.randf(2, 2, 2),
                  "random_dir_0.3_0.6": ref.random_dir(0.3, 0.6).tolist()}

    # ---- light generation (light_gen.comp and the no-reset bodies)
    lg = {}
    for name, frame, groups, ml, shader, variant in [
            ("f1_8x1x1_cap100_reset", 1, (8, 1, 1), 100, "light_gen", 0), ("f2_8x1x1_cap100_reset", 2, (8, 1, 1), 100, "light_gen", 0),
            ("f7_4x2x2_cap37_reset", 7, (4, 2, 2), 37, "light_gen", 0), ("f1_8x2x2_cap100_keep", 1, (8, 2, 2), 100, "point", 1),
            ("f3_2x1x1_cap1000_keep", 3, (2, 1, 1), 1000, "ray", 1), ("f1_1x1x1_cap1000_beam", 1, (1, 1, 1), 1000, "beam", 0)]:
        u = pyoracle.default_ubo(W, H, frameCount=frame, maxLights=ml)
        lights, count = ref.light_gen(cloud, u, groups, shader)
        np.save(os.path.join(OUT, f"lights_{name}.npy"), lights)
        lg[name] = dict(frame=frame, groups=groups, max_lights=ml, shader=shader, variant=variant, count=count, sha256=sha(lights))
    u = pyoracle.default_ubo(W, H, frameCount=1, maxLights=100)
    _, count = ref.light_gen(cloud, u, (4, 1, 1), "path")
    lg["path_light_gen_is_empty"] = dict(count=count)
    man["cases"]["light_gen"] = lg

    # ---- gathers: storage image after one frame from a cleared image (= frameColor for frameCount 1)
    u1 = pyoracle.default_ubo(W, H, frameCount=1, maxLights=100)
    lights = np.load(os.path.join(OUT, "lights_f1_8x1x1_cap100_reset.npy"))
    g = {}
    for algo in ("point", "sphere", "ray", "beam"):
        img = ref.gather(cloud, u1, algo, lights, nthreads=8)
        np.save(os.path.join(OUT, f"frame_{algo}.npy"), img[..., 0].copy())
        assert np.array_equal(img[..., 0], img[..., 1]) and (img[..., 3] == 1).all()
        g[algo] = dict(mean=float(img[..., 0].mean()), sha256=sha(img[..., 0]))
    for frame in (1, 3):
        u = pyoracle.default_ubo(W, H, frameCount=frame)
        img = ref.gather(cloud, u, "path", np.zeros(0, pyoracle.LIGHT_DTYPE), nthreads=8)
        np.save(os.path.join(OUT, f"frame_path_f{frame}.npy"), img[..., 0].copy())  # (0*(n-1) + c)/n
        g[f"path_f{frame}"] = dict(mean=float(img[..., 0].mean()), sha256=sha(img[..., 0]))
    man["cases"]["gather"] = g

    # ---- progressive accumulation over 3 frames (new lights each frame), float image and rgba8 image
    imgf = np.zeros((H, W, 4), np.float32)
    img8 = np.zeros((H, W, 4), np.float32)
    for frame in (1, 2, 3):
        u = pyoracle.default_ubo(W, H, frameCount=frame, maxLights=60)
        lights, _ = ref.light_gen(cloud, u, (4, 1, 1), "light_gen")
        ref.gather(cloud, u, "point", lights, image_rgba=imgf, nthreads=8)
        ref.gather(cloud, u, "point", lights, image_rgba=img8, nthreads=8, unorm8=True)
    np.save(os.path.join(OUT, "accum3_point_f32.npy"), imgf[..., 0].copy())
    np.save(os.path.join(OUT, "accum3_point_rgba8.npy"), np.rint(img8 * 255.0).astype(np.uint8))
    man["cases"]["accumulate"] = dict(frames=3, groups=(4, 1, 1), max_lights=60, algo="point")

    # ---- CPU_test loop around nanovdb's tree().getValue on a fog sphere in front of its camera (64x64 image)
    sphere = vr.build_fog_sphere_grid(32.0, center=(0, 250, -200))
    man["sphere_sha256"] = sha(sphere)
    np.save(os.path.join(OUT, "cpu_test_density_64.npy"), ref.cpu_test(sphere, 64, 64, nthreads=8))

    with open(os.path.join(OUT, "manifest.json"), "w") as f:
        json.dump(man, f, indent=1, sort_keys=True)
    total = sum(os.path.getsize(os.path.join(OUT, x)) for x in os.listdir(OUT))
    print("golden fixtures written:", len(os.listdir(OUT)), "files,", total, "bytes")


if __name__ == "__main__":
    main()
