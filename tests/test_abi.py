"""The C-ABI library: loads, exports every symbol include/vr.h declares, struct layouts match the reference's
std140/std430 blocks, and it fails loudly (no CPU fallback) when there is no CUDA device. No compute calls here."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "vr.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(vr_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(vr):
    lib = vr.load()
    syms = declared_symbols()
    assert len(syms) >= 40
    for s in syms:
        assert hasattr(lib, s), f"libvr.so does not export {s}"
    from volumerenderer_b200 import _lib
    assert sorted(_lib.EXPORTS) == syms, "python binding list and include/vr.h disagree"


def test_ubo_layout_is_std140_block(vr):
    # offsets of shaders/common_bindings.h:19-34 under std140
    want = dict(frameCount=0, framebufferDim=8, cameraPos=16, fov=28, photonInitialIntensity=32, scatteringProbability=36,
                absorptionCoefficient=40, maxLights=44, rayMaxDistance=48, rayMarchingStepSize=52,
                lightSourceWorldPos=64, beamRadius=76, lightRayStepSize=80, radiusFalloff=84)
    for name, off in want.items():
        assert getattr(vr.VrUbo, name).offset == off, name
    assert ctypes.sizeof(vr.VrUbo) == 96
    assert vr.LIGHT_DTYPE.itemsize == 32 and vr.LIGHT_DTYPE.fields["to"][1] == 16 and vr.LIGHT_DTYPE.fields["intensity"][1] == 28


def test_default_params_are_the_reference_defaults(vr):
    u = vr.VrUbo()
    vr.load().vr_default_params(ctypes.byref(u), 1024, 1024)
    # src/main.cpp:546-559
    assert list(u.cameraPos) == [0.0, 20.0, -75.0] and u.fov == 45.0
    assert u.photonInitialIntensity == 100.0 and u.scatteringProbability == np.float32(0.05)
    assert u.absorptionCoefficient == np.float32(0.05) and u.maxLights == 1000
    assert u.rayMaxDistance == 2500.0 and u.rayMarchingStepSize == 1.0
    assert list(u.lightSourceWorldPos) == [-20.0, 15.0, -15.0]
    assert u.beamRadius == np.float32(0.1) and u.lightRayStepSize == np.float32(0.3)
    assert list(u.framebufferDim) == [1024, 1024] and u.frameCount == 0


def test_algorithm_enum_order(vr):
    # enum class Algorithm { Beam, Ray, Point, Sphere, Path } (src/main.cpp:65-68)
    assert vr.ALGO == {"beam": 0, "ray": 1, "point": 2, "sphere": 3, "path": 4}


def test_no_cpu_fallback(vr, has_gpu):
    if has_gpu:
        pytest.skip("a CUDA device is present")
    with pytest.raises(vr.VrError) as e:
        vr.VolumeRenderer(0)
    assert e.value.code == 2 and "no CPU fallback" in str(e.value)


def test_product_does_not_reference_the_oracle():
    """The oracle is test infrastructure: nothing under the package or include/ may import, link or open it."""
    pkg = os.path.join(ROOT, "volumerenderer_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(base, f), errors="replace").read()
                assert "libvr_oracle" not in text and "pyoracle" not in text and "vr_oracle.h" not in text, f
    out = os.popen(f"ldd {os.path.join(pkg, 'libvr.so')}").read()
    assert "oracle" not in out and "libvr_ref" not in out


def test_grid_fixture_entry_points_run_without_gpu(vr):
    g = vr.build_cloud_grid(16)
    assert isinstance(g, np.ndarray) and g.nbytes > 672


def test_nvtx_build_flag_compiles(tmp_path):
    """The profiling build (-DVR_NVTX: one NVTX range per pass) must keep compiling."""
    import shutil
    import subprocess
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("no nvcc")
    obj = str(tmp_path / "vr_api_nvtx.o")
    src = os.path.join(ROOT, "volumerenderer_b200", "csrc", "vr_api.cu")
    subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-Xcompiler", "-fPIC,-ffp-contract=off",
                           "-DVR_NVTX", "-c", src, "-o", obj])
    syms = subprocess.run(["nm", obj], capture_output=True, text=True).stdout
    assert "nvtx" in syms.lower()


def test_traffic_stamp_matches_the_built_library(vr):
    """bench.py reports roofline.traffic only when profiles/eval_traffic.json was captured on the library version that is
    built; a kernel change must come with a version bump and a new capture, so the committed stamp equals vr_version()."""
    import json
    import os

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(os.path.join(root, "profiles", "eval_traffic.json")) as f:
        ent = json.load(f)["C2"]
    assert ent["version"] == vr.load().vr_version().decode()
    assert ent["traffic"] > 1.0e9 and "ncu" in ent and 0.5 < ent["ncu"]["fma_pipe_busy"] < 1.0
