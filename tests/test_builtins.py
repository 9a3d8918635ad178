"""The pinned GLSL built-in layer (vr_glsl_math.h) and the shaders' RNG: known answers and accuracy against libm."""
import json
import os

import numpy as np

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def test_hash_known_answers(oracle):
    # SURVEY Appendix A.2 (computed from the definition in shaders/common_functions.h:2-14)
    assert oracle.hash(1, 1, 1) == 1273441199
    assert oracle.hash(2, 2, 2) == 2620179431
    assert oracle.randf(1, 1, 1) == np.float32(0.29649612307548523)
    assert oracle.randf(2, 2, 2) == np.float32(0.610058069229126)
    # float(0xFFFFFFFF) * 2^-32 == 1.0 exactly: randf can return 1.0
    assert np.float32(np.float32(0xFFFFFFFF) * np.float32(2.0 ** -32)) == np.float32(1.0)


def test_rng_matches_reference_golden(oracle):
    man = json.load(open(os.path.join(GOLD, "manifest.json")))["rng"]
    assert oracle.hash(1, 1, 1) == man["hash_111"]
    assert oracle.randf(1, 1, 1) == np.float32(man["randf_111"])
    assert oracle.randf(2, 2, 2) == np.float32(man["randf_222"])
    assert oracle.random_dir(0.3, 0.6).tolist() == [np.float32(x) for x in man["random_dir_0.3_0.6"]]


def test_rng_matches_reference_shader(oracle, ref):
    rng = np.random.default_rng(0)
    for x, y, z in rng.integers(0, 2 ** 32, (200, 3), dtype=np.uint64):
        assert oracle.hash(int(x), int(y), int(z)) == ref.hash(int(x), int(y), int(z))
        assert oracle.randf(int(x), int(y), int(z)) == ref.randf(int(x), int(y), int(z))
    for r1, r2 in rng.random((200, 2)).astype(np.float32):
        assert np.array_equal(oracle.random_dir(float(r1), float(r2)), ref.random_dir(float(r1), float(r2)))


def _ulps(a, b):
    a = np.asarray(a, np.float32)
    b64 = np.asarray(b, np.float64)
    spacing = np.spacing(np.abs(b64).astype(np.float32)).astype(np.float64)
    return np.abs(a.astype(np.float64) - b64) / spacing


def test_exp_accuracy(oracle):
    xs = np.concatenate([np.linspace(-100, 88, 40001), np.linspace(-1, 0, 20001)]).astype(np.float32)
    got = oracle.builtin("exp", xs)
    assert _ulps(got, np.exp(xs.astype(np.float64))).max() <= 2.0
    assert oracle.builtin("exp", [0.0])[0] == 1.0 and oracle.builtin("exp", [-0.0])[0] == 1.0
    assert oracle.builtin("exp", [-200.0])[0] == 0.0 and np.isinf(oracle.builtin("exp", [100.0])[0])
    assert np.isnan(oracle.builtin("exp", [np.nan])[0])


def test_trig_accuracy(oracle):
    xs = np.linspace(-0.1, 6.4, 40001).astype(np.float32)
    assert np.abs(oracle.builtin("sin", xs) - np.sin(xs.astype(np.float64))).max() <= 1.5e-7
    assert np.abs(oracle.builtin("cos", xs) - np.cos(xs.astype(np.float64))).max() <= 1.5e-7
    xs = np.linspace(-1, 1, 40001).astype(np.float32)
    assert np.abs(oracle.builtin("acos", xs) - np.arccos(xs.astype(np.float64))).max() <= 5e-7
    assert oracle.builtin("acos", [1.0])[0] == 0.0
    assert abs(oracle.builtin("tan", [np.float32(45 * 0.5 * np.pi / 180)])[0] - np.tan(np.deg2rad(22.5))) < 1e-7


def test_random_dir_is_unit(oracle):
    rng = np.random.default_rng(3)
    for r1, r2 in rng.random((500, 2)).astype(np.float32):
        d = oracle.random_dir(float(r1), float(r2)).astype(np.float64)
        assert abs(np.linalg.norm(d) - 1.0) < 1e-6
    # r1 = 1.0 is a legal roll (randf can return 1.0): theta = acos(-1) = pi
    assert np.isfinite(oracle.random_dir(1.0, 1.0)).all()


# ---- exhaustive sweeps where the hot path lives. The kernels, the oracle and the oracle/_ref shim share vr_glsl_math.h, so
# the bit-exact parity tests cannot see an error in it; these sweeps are its own guard: EVERY float of the range, against
# double-precision libm.
def _floats_between(lo, hi, step=1):
    """every step-th positive float32 from lo to hi inclusive, in chunks of 2^24 bit patterns"""
    a = int(np.float32(lo).view(np.uint32))
    b = int(np.float32(hi).view(np.uint32))
    for start in range(a, b + 1, (1 << 24) * step):
        yield np.arange(start, min(b + 1, start + (1 << 24) * step), step, dtype=np.uint32).view(np.float32)


def test_exp_every_float_of_the_attenuation_range(oracle):
    """exp(-val * sigma * step) (light_gen.comp:66, point_compute_color.comp:80): arguments in [-0.06, 0] for the UBO
    defaults (sigma 0.05, step 1, densities up to ~1). Every float in [-0.06, -2^-30], then every 64th down to -2^-60
    (there exp rounds to 1 or 1 - 2^-24 and nothing can go wrong slowly)."""
    worst = 0.0
    for xs in _floats_between(2.0 ** -30, 0.06):
        x = -xs
        worst = max(worst, float(_ulps(oracle.builtin("exp", x), np.exp(x.astype(np.float64))).max()))
    for xs in _floats_between(2.0 ** -60, 2.0 ** -30, 64):
        x = -xs
        worst = max(worst, float(_ulps(oracle.builtin("exp", x), np.exp(x.astype(np.float64))).max()))
    assert worst <= 1.0, worst


def test_acos_every_float_near_one_and_sin_cos_on_the_circle(oracle):
    """randomDir (common_functions.h:24-33): theta = acos(1 - 2 r1), phi = 2 pi r2, then sin/cos of both. acos is
    ill-conditioned near +-1: every float with 0.5 <= |x| <= 1, every 64th below. sin/cos: every float in [2^-12, 2 pi],
    every 256th below. Absolute bounds, the ones the sparse sweeps above use."""
    worst = 0.0
    for sign in (1.0, -1.0):
        for xs in list(_floats_between(0.5, 1.0)) + list(_floats_between(2.0 ** -30, 0.5, 64)):
            x = (xs * np.float32(sign)).astype(np.float32)
            worst = max(worst, float(np.abs(oracle.builtin("acos", x) - np.arccos(x.astype(np.float64))).max()))
    assert worst <= 5e-7, worst
    ws = wc = 0.0
    for xs in list(_floats_between(2.0 ** -12, 2.0 * np.pi)) + list(_floats_between(2.0 ** -40, 2.0 ** -12, 256)):
        x64 = xs.astype(np.float64)
        ws = max(ws, float(np.abs(oracle.builtin("sin", xs) - np.sin(x64)).max()))
        wc = max(wc, float(np.abs(oracle.builtin("cos", xs) - np.cos(x64)).max()))
    assert ws <= 1.5e-7 and wc <= 1.5e-7, (ws, wc)
