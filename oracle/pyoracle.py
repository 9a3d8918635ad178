"""ctypes bindings of the CPU checker (oracle/libvr_oracle.so) and, when present, of oracle/_ref/libvr_ref.so.

TEST INFRASTRUCTURE: import this only from tests/, __graft_entry__.smoke() and bench.py's CPU legs.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libvr_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libvr_ref.so")

ALGO = {"beam": 0, "ray": 1, "point": 2, "sphere": 3, "path": 4}


class Ubo(C.Structure):
    """shaders/common_bindings.h:19-34 (std140, 96 bytes)."""
    _fields_ = [
        ("frameCount", C.c_uint32), ("_pad0", C.c_uint32), ("framebufferDim", C.c_uint32 * 2),
        ("cameraPos", C.c_float * 3), ("fov", C.c_float),
        ("photonInitialIntensity", C.c_float), ("scatteringProbability", C.c_float),
        ("absorptionCoefficient", C.c_float), ("maxLights", C.c_uint32),
        ("rayMaxDistance", C.c_float), ("rayMarchingStepSize", C.c_float), ("_pad1", C.c_uint32 * 2),
        ("lightSourceWorldPos", C.c_float * 3), ("beamRadius", C.c_float),
        ("lightRayStepSize", C.c_float), ("radiusFalloff", C.c_float), ("_pad2", C.c_uint32 * 2),
    ]


assert C.sizeof(Ubo) == 96

LIGHT_DTYPE = np.dtype([("from", "<f4", 3), ("_pad0", "<f4"), ("to", "<f4", 3), ("intensity", "<f4")])
assert LIGHT_DTYPE.itemsize == 32


class Options(C.Structure):
    """vro_options: opt-in physics the reference's shaders lack; all zero = the reference."""
    _fields_ = [("trilinear", C.c_uint32), ("phase_hg", C.c_uint32), ("hg_g", C.c_float), ("light_transmittance", C.c_uint32)]


class Counters(C.Structure):
    _fields_ = [("S_ref", C.c_uint64), ("S_eff", C.c_uint64), ("E", C.c_uint64), ("pixels", C.c_uint64),
                ("hit_pixels", C.c_uint64)]


def default_ubo(width, height, **kw):
    """UBO defaults of src/main.cpp:546-559."""
    u = Ubo()
    u.frameCount = 0
    u.framebufferDim[0], u.framebufferDim[1] = width, height
    u.cameraPos[:] = (0.0, 20.0, -75.0)
    u.fov = 45.0
    u.photonInitialIntensity = 100.0
    u.scatteringProbability = 0.05
    u.absorptionCoefficient = 0.05
    u.maxLights = 1000
    u.rayMaxDistance = 2500.0
    u.rayMarchingStepSize = 1.0
    u.lightSourceWorldPos[:] = (-20.0, 15.0, -15.0)
    u.beamRadius = 0.1
    u.lightRayStepSize = 0.3
    u.radiusFalloff = 0.0
    for k, v in kw.items():
        if isinstance(v, (tuple, list)):
            getattr(u, k)[:] = v
        else:
            setattr(u, k, v)
    return u


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self, path=ORACLE_SO):
        if not os.path.exists(path):
            raise FileNotFoundError(f"{path} missing: run `make -C oracle oracle` (or __graft_entry__.build())")
        L = self.lib = C.CDLL(path)
        L.vro_hash.restype = C.c_uint32
        L.vro_hash.argtypes = [C.c_uint32] * 3
        L.vro_randf.restype = C.c_float
        L.vro_randf.argtypes = [C.c_uint32] * 3
        for f in ("vro_exp", "vro_sin", "vro_cos", "vro_acos", "vro_tan"):
            getattr(L, f).restype = C.c_float
            getattr(L, f).argtypes = [C.c_float]
        L.vro_random_dir.argtypes = [C.c_float, C.c_float, C.c_void_p]
        L.vro_grid_info.argtypes = [C.c_void_p, C.c_uint64] + [C.c_void_p] * 5
        L.vro_lookup.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.vro_world_to_index.argtypes = [C.c_void_p] * 3
        L.vro_index_to_world.argtypes = [C.c_void_p] * 3
        L.vro_light_gen.argtypes = [C.c_void_p, C.POINTER(Ubo), C.c_int, C.c_uint32, C.c_uint32, C.c_uint32,
                                    C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.vro_gather.argtypes = [C.c_void_p, C.POINTER(Ubo), C.c_int, C.c_void_p, C.c_uint32] + [C.c_uint32] * 5 + \
                                [C.c_int, C.c_int, C.c_void_p, C.POINTER(Counters)]
        L.vro_gather_ex.argtypes = [C.c_void_p, C.POINTER(Ubo), C.c_int, C.c_void_p, C.c_uint32] + [C.c_uint32] * 5 + \
            [C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(Counters)]
        L.vro_trilinear.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.vro_accumulate.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_int, C.c_void_p, C.c_void_p]
        L.vro_cpu_test.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]

    def hash(self, x, y, z):
        return self.lib.vro_hash(x, y, z)

    def randf(self, x, y, z):
        return self.lib.vro_randf(x, y, z)

    def random_dir(self, r1, r2):
        out = np.zeros(3, np.float32)
        self.lib.vro_random_dir(r1, r2, _ptr(out))
        return out

    def builtin(self, name, xs):
        xs = np.ascontiguousarray(xs, np.float32)
        out = np.empty(xs.shape, np.float32)
        self.lib.vro_builtin_array.argtypes = [C.c_int, C.c_void_p, C.c_uint64, C.c_void_p]
        rc = self.lib.vro_builtin_array(("exp", "sin", "cos", "acos", "tan").index(name), _ptr(xs), xs.size, _ptr(out))
        assert rc == 0
        return out

    def grid_info(self, grid):
        bmin = np.zeros(3, np.int32)
        bmax = np.zeros(3, np.int32)
        bg = C.c_float()
        counts = np.zeros(3, np.uint32)
        tiles = C.c_uint32()
        rc = self.lib.vro_grid_info(_ptr(grid), grid.nbytes, _ptr(bmin), _ptr(bmax), C.byref(bg), _ptr(counts), C.byref(tiles))
        if rc:
            raise ValueError(f"not a NanoVDB float grid (code {rc})")
        return dict(bbox_min=bmin, bbox_max=bmax, background=bg.value, node_counts=counts, root_tiles=tiles.value)

    def lookup(self, grid, ijk):
        ijk = np.ascontiguousarray(ijk, np.int32).reshape(-1, 3)
        out = np.zeros(len(ijk), np.float32)
        self.lib.vro_lookup(_ptr(grid), _ptr(ijk), len(ijk), _ptr(out))
        return out

    def world_to_index(self, grid, w):
        w = np.asarray(w, np.float32)
        o = np.zeros(3, np.float32)
        self.lib.vro_world_to_index(_ptr(grid), _ptr(w), _ptr(o))
        return o

    def index_to_world(self, grid, i):
        i = np.asarray(i, np.float32)
        o = np.zeros(3, np.float32)
        self.lib.vro_index_to_world(_ptr(grid), _ptr(i), _ptr(o))
        return o

    def light_gen(self, grid, ubo, groups=(1, 1, 1), variant=0):
        lights = np.zeros(max(int(ubo.maxLights), 1), LIGHT_DTYPE)
        count = C.c_uint32()
        emitted = C.c_uint64()
        steps = C.c_uint64()
        self.lib.vro_light_gen(_ptr(grid), C.byref(ubo), variant, groups[0], groups[1], groups[2], _ptr(lights),
                               C.byref(count), C.byref(emitted), C.byref(steps))
        return lights[:min(count.value, int(ubo.maxLights))].copy(), count.value, emitted.value, steps.value

    def trilinear(self, grid, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        self.lib.vro_trilinear(_ptr(grid), _ptr(xyz), len(xyz), _ptr(out))
        return out

    def gather(self, grid, ubo, algo, lights, rect=None, stride=1, skip_zero=True, nthreads=0, out=None, options=None):
        W, H = ubo.framebufferDim[0], ubo.framebufferDim[1]
        if rect is None:
            rect = (0, 0, W, H)
        if nthreads <= 0:
            nthreads = os.cpu_count() or 1
        fc = out if out is not None else np.zeros((H, W), np.float32)
        lights = np.ascontiguousarray(lights, LIGHT_DTYPE)
        cnt = Counters()
        self.lib.vro_gather_ex(_ptr(grid), C.byref(ubo), ALGO[algo] if isinstance(algo, str) else algo, _ptr(lights),
                               len(lights), rect[0], rect[1], rect[2], rect[3], stride, int(skip_zero), nthreads,
                               C.byref(options) if options is not None else None, _ptr(fc), C.byref(cnt))
        return fc, {k: getattr(cnt, k) for k, _ in Counters._fields_}

    def accumulate(self, frame_color, frame_count, accum=None, rgba8=None, path_mode=False):
        fc = np.ascontiguousarray(frame_color, np.float32)
        self.lib.vro_accumulate(_ptr(fc), fc.size, frame_count, int(path_mode),
                                _ptr(accum) if accum is not None else None, _ptr(rgba8) if rgba8 is not None else None)

    def cpu_test(self, grid, width=256, height=256, nthreads=1):
        d = np.zeros((height, width), np.float32)
        g = np.zeros((height, width), np.uint8)
        self.lib.vro_cpu_test(_ptr(grid), width, height, nthreads, _ptr(d), _ptr(g))
        return d, g


class Ref:
    """The reference's own shader sources + vendored NanoVDB, compiled here into oracle/_ref (absent on the GPU box
    only if the build never ran in a container that has /root/reference)."""

    LG_SHADER = {"light_gen": 0, "point": 1, "ray": 2, "sphere": 3, "beam": 4, "path": 5}

    def __init__(self, path=REF_SO):
        if not os.path.exists(path):
            raise FileNotFoundError(path)
        L = self.lib = C.CDLL(path)
        L.vrr_set_grid.argtypes = [C.c_void_p]
        L.vrr_light_gen.argtypes = [C.POINTER(Ubo), C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint32)]
        L.vrr_gather.argtypes = [C.POINTER(Ubo), C.c_int, C.c_void_p, C.c_uint32] + [C.c_uint32] * 5 + [C.c_int, C.c_void_p, C.c_int]
        L.vrr_hash.restype = C.c_uint32
        L.vrr_hash.argtypes = [C.c_uint32] * 3
        L.vrr_randf.restype = C.c_float
        L.vrr_randf.argtypes = [C.c_uint32] * 3
        L.vrr_random_dir.argtypes = [C.c_float, C.c_float, C.c_void_p]
        L.vrr_free.argtypes = [C.c_void_p]
        L.vrr_make_fog_sphere.argtypes = [C.c_double, C.c_void_p, C.c_double, C.c_double, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        L.vrr_make_cloud.argtypes = [C.c_int32, C.c_uint32, C.c_double, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        L.vrr_make_dense.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_double, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        L.vrr_get_values.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.c_void_p]
        if hasattr(L, "vrr_trilinear"):
            L.vrr_trilinear.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.vrr_grid_facts.argtypes = [C.c_void_p] * 5
        L.vrr_cpu_test.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        L.vrr_write_nvdb.argtypes = [C.c_void_p, C.c_uint64, C.c_char_p]
        L.vrr_read_nvdb.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
        self._grid = None

    def _take(self, p, n):
        a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,)).copy()
        self.lib.vrr_free(p)
        return a

    def make_fog_sphere(self, radius, center=(0.0, 0.0, 0.0), voxel=1.0, half_width=3.0):
        c = np.asarray(center, np.float64)
        p, n = C.c_void_p(), C.c_uint64()
        assert self.lib.vrr_make_fog_sphere(radius, _ptr(c), voxel, half_width, C.byref(p), C.byref(n)) == 0
        return self._take(p, n)

    def make_cloud(self, n, seed=1234, voxel=None, translation=(-32.0, -12.0, -32.0)):
        t = np.asarray(translation, np.float64)
        p, sz = C.c_void_p(), C.c_uint64()
        assert self.lib.vrr_make_cloud(n, seed, voxel if voxel else 64.0 / n, _ptr(t), C.byref(p), C.byref(sz)) == 0
        return self._take(p, sz)

    def make_dense(self, dense, origin=(0, 0, 0), background=0.0, voxel=1.0, translation=(0.0, 0.0, 0.0)):
        d = np.ascontiguousarray(dense, np.float32)
        dims = np.asarray(d.shape, np.int32)
        o = np.asarray(origin, np.int32)
        t = np.asarray(translation, np.float64)
        p, sz = C.c_void_p(), C.c_uint64()
        assert self.lib.vrr_make_dense(_ptr(d), _ptr(dims), _ptr(o), background, voxel, _ptr(t), C.byref(p), C.byref(sz)) == 0
        return self._take(p, sz)

    def trilinear(self, grid, xyz):
        """nanovdb::math::createSampler<1>(accessor)(xyz) of the vendored NanoVDB at index-space positions."""
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        assert self.lib.vrr_trilinear(_ptr(grid), _ptr(xyz), len(xyz), _ptr(out)) == 0
        return out

    def get_values(self, grid, ijk, accessor=False):
        ijk = np.ascontiguousarray(ijk, np.int32).reshape(-1, 3)
        out = np.zeros(len(ijk), np.float32)
        assert self.lib.vrr_get_values(_ptr(grid), _ptr(ijk), len(ijk), int(accessor), _ptr(out)) == 0
        return out

    def grid_facts(self, grid):
        bbox = np.zeros(6, np.int32)
        active = C.c_uint64()
        counts = np.zeros(3, np.uint32)
        bg = C.c_float()
        assert self.lib.vrr_grid_facts(_ptr(grid), _ptr(bbox), C.byref(active), _ptr(counts), C.byref(bg)) == 0
        return dict(bbox_min=bbox[:3].copy(), bbox_max=bbox[3:].copy(), active=active.value, node_counts=counts, background=bg.value)

    def write_nvdb(self, grid, path):
        """nanovdb::io::writeGrid(path, handle, Codec::NONE)"""
        assert self.lib.vrr_write_nvdb(_ptr(grid), grid.nbytes, os.fsencode(path)) == 0

    def read_nvdb(self, path):
        """nanovdb::io::readGrid(path, 0) -> grid bytes"""
        p, n = C.c_void_p(), C.c_uint64()
        rc = self.lib.vrr_read_nvdb(os.fsencode(path), C.byref(p), C.byref(n))
        if rc:
            raise ValueError(f"nanovdb::io::readGrid failed on {path} (code {rc})")
        return self._take(p, n)

    def set_grid(self, grid):
        self._grid = grid  # keep alive
        self.lib.vrr_set_grid(_ptr(grid))

    def hash(self, x, y, z):
        return self.lib.vrr_hash(x, y, z)

    def randf(self, x, y, z):
        return self.lib.vrr_randf(x, y, z)

    def random_dir(self, r1, r2):
        out = np.zeros(3, np.float32)
        self.lib.vrr_random_dir(r1, r2, _ptr(out))
        return out

    def light_gen(self, grid, ubo, groups=(1, 1, 1), shader="light_gen"):
        self.set_grid(grid)
        lights = np.zeros(max(int(ubo.maxLights), 1), LIGHT_DTYPE)
        count = C.c_uint32()
        self.lib.vrr_light_gen(C.byref(ubo), self.LG_SHADER[shader], groups[0], groups[1], groups[2], _ptr(lights), C.byref(count))
        return lights[:min(count.value, int(ubo.maxLights))].copy(), count.value

    def gather(self, grid, ubo, algo, lights, image_rgba=None, rect=None, stride=1, nthreads=0, unorm8=False):
        """Runs *_compute_color.comp; returns the storage image (float RGBA) after the shader's running average."""
        self.set_grid(grid)
        W, H = ubo.framebufferDim[0], ubo.framebufferDim[1]
        if rect is None:
            rect = (0, 0, W, H)
        if nthreads <= 0:
            nthreads = os.cpu_count() or 1
        img = image_rgba if image_rgba is not None else np.zeros((H, W, 4), np.float32)
        lights = np.ascontiguousarray(lights, LIGHT_DTYPE)
        self.lib.vrr_gather(C.byref(ubo), ALGO[algo] if isinstance(algo, str) else algo, _ptr(lights), len(lights),
                            rect[0], rect[1], rect[2], rect[3], stride, nthreads, _ptr(img), int(unorm8))
        return img

    def cpu_test(self, grid, width=256, height=256, nthreads=1):
        d = np.zeros((height, width), np.float32)
        assert self.lib.vrr_cpu_test(_ptr(grid), width, height, nthreads, _ptr(d)) == 0
        return d


def have_ref():
    return os.path.exists(REF_SO)
