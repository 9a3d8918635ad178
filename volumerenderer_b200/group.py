"""ctypes mirror of the vr_group_* entry points (include/vr.h): several GPUs of one node driven from one process by
libvr's own C++ host layer (csrc/vr_group.cu) -- worker threads, CUDA peer access, device-side barrier, band balancer.
bench.py's N > 1 arm uses one process per GPU instead (torchrun); this class is what tests and tools use to drive the
C++ path, and what `vr_headless --gpus N` does from C++."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import ALGO, VrUbo
from .renderer import VrError

EXCHANGE = {"none": 0, "allgather": 1}


class RendererGroup:
    def __init__(self, devices):
        self._L = _lib.load()
        self._g = C.c_void_p()
        devs = (C.c_int * len(devices))(*devices)
        rc = self._L.vr_group_create(devs, len(devices), C.byref(self._g))
        if rc != 0:
            raise VrError(rc, self._L.vr_group_last_error(None).decode())
        self.n = len(devices)
        self.width = self.height = 0

    def _check(self, rc):
        if rc != 0:
            raise VrError(rc, self._L.vr_group_last_error(self._g).decode())

    def close(self):
        if getattr(self, "_g", None):
            self._L.vr_group_destroy(self._g)
            self._g = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def ctx(self, i):
        """Raw vr_ctx* of GPU i (for the vr_* calls the group does not wrap)."""
        return C.c_void_p(self._L.vr_group_ctx(self._g, i))

    def set_grid(self, grid):
        g = np.ascontiguousarray(grid, np.uint8)
        self._check(self._L.vr_group_set_grid(self._g, g.ctypes.data_as(C.c_void_p), g.nbytes))

    def resize(self, w, h):
        self._check(self._L.vr_group_resize(self._g, w, h))
        self.width, self.height = w, h

    @property
    def ubo(self):
        u = VrUbo()
        self._check(self._L.vr_group_get_params(self._g, C.byref(u)))
        return u

    @ubo.setter
    def ubo(self, u):
        self._check(self._L.vr_group_set_params(self._g, C.byref(u)))

    def update_ubo(self, **kw):
        u = self.ubo
        for k, v in kw.items():
            if isinstance(v, (tuple, list, np.ndarray)):
                getattr(u, k)[:] = [float(x) for x in v]
            else:
                setattr(u, k, v)
        self.ubo = u
        return u

    def default_ubo(self):
        u = VrUbo()
        self._L.vr_default_params(C.byref(u), self.width, self.height)
        return u

    def set_max_lights_capacity(self, n):
        self._check(self._L.vr_group_set_max_lights_capacity(self._g, n))

    def set_light_gen_groups(self, gx=1, gy=1, gz=1, variant=0):
        self._check(self._L.vr_group_set_light_gen_groups(self._g, gx, gy, gz, variant))

    def set_exchange(self, mode):
        self._check(self._L.vr_group_set_exchange(self._g, EXCHANGE.get(mode, mode)))

    def reset_accumulation(self):
        self._check(self._L.vr_group_reset_accumulation(self._g))

    def set_bands(self, bands):
        y0 = (C.c_uint32 * self.n)(*[int(b[0]) for b in bands])
        h = (C.c_uint32 * self.n)(*[int(b[1]) for b in bands])
        self._check(self._L.vr_group_set_bands(self._g, y0, h))

    def bands(self):
        y0, h = (C.c_uint32 * self.n)(), (C.c_uint32 * self.n)()
        self._check(self._L.vr_group_get_bands(self._g, y0, h))
        return [(int(a), int(b)) for a, b in zip(y0, h)]

    def balance(self, algo, rounds=6):
        self._check(self._L.vr_group_balance(self._g, ALGO[algo] if isinstance(algo, str) else algo, rounds))
        return self.bands()

    def draw_frame(self, algo):
        self._check(self._L.vr_group_render_frame(self._g, ALGO[algo] if isinstance(algo, str) else algo))

    def sync(self):
        self._check(self._L.vr_group_sync(self._g))

    def read_accum(self):
        out = np.zeros((self.height, self.width, 4), np.float32)
        self._check(self._L.vr_group_read_accum_f32(self._g, out.ctypes.data_as(C.c_void_p)))
        return out

    def read_image_rgba8(self):
        out = np.zeros((self.height, self.width, 4), np.uint8)
        self._check(self._L.vr_group_read_image_rgba8(self._g, out.ctypes.data_as(C.c_void_p)))
        return out

    def render_progressive(self, algo, frames):
        self._check(self._L.vr_group_render_progressive(self._g, ALGO[algo] if isinstance(algo, str) else algo, frames))

    def time_frames(self, algo, frames):
        worst = C.c_float()
        per = (C.c_float * self.n)()
        self._check(self._L.vr_group_time_frames(self._g, ALGO[algo] if isinstance(algo, str) else algo, frames, C.byref(worst), per))
        return worst.value, [float(x) for x in per]
