"""Host-side mirror of the compute section of the reference's VolumeApp (src/main.cpp), on top of the libvr C ABI.

`VolumeRenderer.draw_frame()` is `VolumeApp::drawFrame` (src/main.cpp:620-837) without the window: frameCount++,
UBO update, light-counter reset, image clear on the first frame, light-generation dispatch, gather dispatch. The
`ubo` attribute is the 96-byte `uboData` block (src/main.cpp:83-98) with the defaults of :546-559; algorithm names
follow the `Algorithm` enum (:65-68). All compute runs in libvr's sm_100a kernels; there is no CPU fallback.
"""
import ctypes as C
import os

import numpy as np

from . import _lib
from ._lib import ALGO, LG_KEEP_ORIGIN, LG_RESET_ORIGIN, LIGHT_DTYPE, VrCounters, VrOptions, VrTile, VrTimings, VrUbo


class VrError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{_lib.STATUS_NAMES.get(code, code)}: {msg}")
        self.code = code


def _algo(a):
    return ALGO[a] if isinstance(a, str) else int(a)


class VolumeRenderer:
    """One CUDA device, one stream, one set of bindings (grid, lights, counter, image, UBO)."""

    def __init__(self, device=0):
        self._L = _lib.load()
        self._ctx = C.c_void_p()
        rc = self._L.vr_create(device, C.byref(self._ctx))
        if rc != 0:
            raise VrError(rc, self._L.vr_last_error(None).decode())
        self.device = device
        self.width = self.height = 0
        self._grid_keepalive = None

    # ------------------------------------------------------------------ plumbing
    def _check(self, rc):
        if rc != 0:
            raise VrError(rc, self._L.vr_last_error(self._ctx).decode())

    def close(self):
        if getattr(self, "_ctx", None):
            self._L.vr_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ------------------------------------------------------------------ bindings
    def set_grid(self, grid_bytes):
        """binding 2: raw nanovdb::GridHandle bytes (src/main.cpp:1191-1212)."""
        g = np.ascontiguousarray(np.frombuffer(grid_bytes, np.uint8) if not isinstance(grid_bytes, np.ndarray) else grid_bytes)
        self._check(self._L.vr_set_grid(self._ctx, g.ctypes.data_as(C.c_void_p), g.nbytes))

    def set_grid_device(self, device_ptr, nbytes, header):
        """vr_set_grid_device: the grid from `nbytes` bytes of device memory at `device_ptr` (int), copied in stream order
        on the renderer's stream; `header` = host bytes of the grid's first >= 1024 bytes (or the whole grid)."""
        h = np.ascontiguousarray(np.frombuffer(header, np.uint8) if not isinstance(header, np.ndarray) else header)
        self._check(self._L.vr_set_grid_device(self._ctx, C.c_void_p(int(device_ptr)), int(nbytes), h.ctypes.data_as(C.c_void_p), h.nbytes))

    def set_grid_device_from_stream(self, device_ptr, nbytes, header, ready_stream):
        """vr_set_grid_device_from_stream: as set_grid_device, for bytes produced in the order of the caller's CUDA stream
        `ready_stream` (int handle); the copy runs next to the frames in flight on libvr's upload stream."""
        h = np.ascontiguousarray(np.frombuffer(header, np.uint8) if not isinstance(header, np.ndarray) else header)
        self._check(self._L.vr_set_grid_device_from_stream(self._ctx, C.c_void_p(int(device_ptr)), int(nbytes),
                                                           h.ctypes.data_as(C.c_void_p), h.nbytes, C.c_void_p(int(ready_stream))))

    def resize(self, width, height):
        """binding 0: storage image sized to the framebuffer; resets the progressive state (src/main.cpp:936-982)."""
        self._check(self._L.vr_resize(self._ctx, width, height))
        self.width, self.height = width, height

    @property
    def ubo(self):
        u = VrUbo()
        self._check(self._L.vr_get_params(self._ctx, C.byref(u)))
        return u

    @ubo.setter
    def ubo(self, u):
        self._check(self._L.vr_set_params(self._ctx, C.byref(u)))

    def update_ubo(self, **kw):
        """uniformBuffer->updateBuffer(&uboData) with some fields changed (the ImGui sliders, src/main.cpp:662-691)."""
        u = self.ubo
        for k, v in kw.items():
            if isinstance(v, (tuple, list, np.ndarray)):
                getattr(u, k)[:] = [float(x) for x in v] if k != "framebufferDim" else [int(x) for x in v]
            else:
                setattr(u, k, v)
        self.ubo = u
        return u

    def default_ubo(self):
        u = VrUbo()
        self._L.vr_default_params(C.byref(u), self.width, self.height)
        return u

    def set_max_lights_capacity(self, n):
        self._check(self._L.vr_set_max_lights_capacity(self._ctx, n))

    def set_light_gen_groups(self, gx=1, gy=1, gz=1, variant=LG_RESET_ORIGIN):
        self._check(self._L.vr_set_light_gen_groups(self._ctx, gx, gy, gz, variant))

    def set_options(self, trilinear=False, phase_hg_g=None, light_transmittance=False):
        """Opt-in physics the reference's shaders lack (include/vr.h, vr_options): trilinear density at the camera
        samples, Henyey-Greenstein phase factor with asymmetry `phase_hg_g` (None = off), light->sample transmittance.
        All off (the default) is the reference's arithmetic."""
        o = VrOptions(int(bool(trilinear)), int(phase_hg_g is not None), float(phase_hg_g or 0.0), int(bool(light_transmittance)))
        self._check(self._L.vr_set_options(self._ctx, C.byref(o)))

    def options(self):
        o = VrOptions()
        self._check(self._L.vr_get_options(self._ctx, C.byref(o)))
        return dict(trilinear=bool(o.trilinear), phase_hg_g=o.hg_g if o.phase_hg else None,
                    light_transmittance=bool(o.light_transmittance))

    def trilinear_values(self, xyz):
        """Device trilinear sampler (nanovdb TrilinearSampler semantics) at index-space positions; test hook."""
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        self._check(self._L.vr_trilinear_values(self._ctx, xyz.ctypes.data_as(C.c_void_p), len(xyz), out.ctypes.data_as(C.c_void_p)))
        return out

    def update_beam_radius(self, alpha):
        """ubo.beamRadius = calculateRadius(ubo.beamRadius, alpha) (src/main.cpp:609-618, the call commented out at :713)."""
        return self.update_ubo(beamRadius=float(self._L.vr_progressive_radius(float(self.ubo.beamRadius), float(alpha))))

    def set_eval_variant(self, variant):
        """0: scalar FP32 evaluation kernel, 1: packed FP32x2 kernel with the guard in the loop, 2: packed speculative
        kernel (unguarded loop, guarded repeat when needed; same results as 1; default), 3: as 2, 6 samples per thread,
        8: ray-segment kernel (include/vr.h)."""
        self._check(self._L.vr_set_eval_variant(self._ctx, int(variant)))

    def set_eval_near_distance(self, distance):
        """Variant 8: lights closer than this to any sample of a thread send its samples to the exact kernel."""
        self._check(self._L.vr_set_eval_near_distance(self._ctx, float(distance)))

    def eval_segment_counts(self):
        """Variant 8 bookkeeping: dict(segments, leftovers, exact) of the last evaluation."""
        a = (C.c_uint64 * 3)()
        self._check(self._L.vr_get_eval_segment_counts(self._ctx, a))
        return dict(segments=int(a[0]), leftovers=int(a[1]), exact=int(a[2]))

    def read_samples(self):
        """(n, 4) float32: world x, y, z, weight of the last gather's camera samples (pixel-major, march order)."""
        n = C.c_uint64(0)
        self._check(self._L.vr_read_samples(self._ctx, None, 0, C.byref(n)))
        out = np.zeros((int(n.value), 4), np.float32)
        if n.value:
            self._check(self._L.vr_read_samples(self._ctx, out.ctypes.data_as(C.c_void_p), n.value, C.byref(n)))
        return out

    def eval_redo_count(self):
        n = C.c_uint64(0)
        self._check(self._L.vr_get_eval_redo_count(self._ctx, C.byref(n)))
        return int(n.value)

    def set_light_prefetch(self, enabled):
        """Next-frame light generation underneath the current frame's evaluation kernel (default on); see include/vr.h."""
        self._check(self._L.vr_set_light_prefetch(self._ctx, int(bool(enabled))))

    def set_march_cache(self, enabled):
        self._check(self._L.vr_set_march_cache(self._ctx, int(bool(enabled))))

    # ------------------------------------------------------------------ commands
    def reset_lights(self):
        self._check(self._L.vr_reset_lights(self._ctx))

    def clear_image(self):
        self._check(self._L.vr_clear_image(self._ctx))

    def write_lights(self, lights):
        a = np.ascontiguousarray(lights, LIGHT_DTYPE)
        self._check(self._L.vr_write_lights(self._ctx, a.ctypes.data_as(C.c_void_p), len(a)))

    def dispatch_light_gen(self, algo, groups=(1, 1, 1), variant=LG_RESET_ORIGIN):
        self._check(self._L.vr_dispatch_light_gen(self._ctx, _algo(algo), groups[0], groups[1], groups[2], variant))

    def dispatch_compute_color(self, algo, tile=None):
        t = C.byref(VrTile(*tile)) if tile is not None else None
        self._check(self._L.vr_dispatch_compute_color(self._ctx, _algo(algo), t))

    def draw_frame(self, algo, tile=None):
        """The compute section of VolumeApp::drawFrame (src/main.cpp:712-837)."""
        t = C.byref(VrTile(*tile)) if tile is not None else None
        self._check(self._L.vr_render_frame(self._ctx, _algo(algo), t))

    def accumulate_frame_color(self, tile=None):
        """Progressive average of `tile` from frame colours already in the frame-colour buffer (rows received from
        another GPU); see include/vr.h."""
        t = C.byref(VrTile(*tile)) if tile is not None else None
        self._check(self._L.vr_accumulate_frame_color(self._ctx, t))

    def set_frame_color_peers(self, local_ptr, peer_ptrs):
        """Fused tile all-gather: resolve stores frame colours to `local_ptr` and every address in `peer_ptrs` (the same
        buffer on the other GPUs, peer-mapped); (None, []) restores the private buffer. See include/vr.h."""
        arr = (C.c_void_p * max(1, len(peer_ptrs)))(*[C.c_void_p(int(p)) for p in peer_ptrs])
        self._check(self._L.vr_set_frame_color_peers(self._ctx, C.c_void_p(int(local_ptr)) if local_ptr else None, arr, len(peer_ptrs)))

    def reset_accumulation(self):
        self._check(self._L.vr_reset_accumulation(self._ctx))

    def sync(self):
        self._check(self._L.vr_sync(self._ctx))

    # ------------------------------------------------------------------ read-back
    def read_lights(self):
        cnt = (C.c_uint32 * 2)()
        total = C.c_uint64()
        self._check(self._L.vr_read_light_counter(self._ctx, cnt, C.byref(total)))
        n = C.c_uint32()
        out = np.zeros(max(cnt[0], 1), LIGHT_DTYPE)
        self._check(self._L.vr_read_lights(self._ctx, out.ctypes.data_as(C.c_void_p), len(out), C.byref(n)))
        return out[:min(n.value, len(out))], n.value, total.value

    def read_frame_color(self):
        out = np.zeros((self.height, self.width), np.float32)
        self._check(self._L.vr_read_frame_color_f32(self._ctx, out.ctypes.data_as(C.c_void_p)))
        return out

    def read_accum(self):
        out = np.zeros((self.height, self.width, 4), np.float32)
        self._check(self._L.vr_read_accum_f32(self._ctx, out.ctypes.data_as(C.c_void_p)))
        return out

    def read_accum_async(self, out):
        """vr_read_accum_f32_async into `out` ((H, W, 4) float32, ideally pinned): valid after read_wait()."""
        assert out.dtype == np.float32 and out.size == self.width * self.height * 4 and out.flags["C_CONTIGUOUS"]
        self._check(self._L.vr_read_accum_f32_async(self._ctx, out.ctypes.data_as(C.c_void_p)))

    def read_wait(self):
        self._check(self._L.vr_read_wait(self._ctx))

    def read_accum_tile_async(self, tile, out_image):
        """vr_read_accum_f32_tile_async: the accumulator's pixels of `tile` into the same pixels of `out_image`, an
        (H, W, 4) float32 array of the full image (ideally pinned; may be shared between processes)."""
        assert out_image.dtype == np.float32 and out_image.shape == (self.height, self.width, 4) and out_image.flags["C_CONTIGUOUS"]
        x0, y0 = int(tile[0]), int(tile[1])
        ptr = out_image.ctypes.data + (y0 * self.width + x0) * 16
        self._check(self._L.vr_read_accum_f32_tile_async(self._ctx, C.byref(VrTile(*tile)), C.c_void_p(ptr), self.width * 16))

    def set_eval_side_stream(self, enabled):
        """vr_set_eval_side_stream: second stream for the small evaluation kernels (default on); results identical."""
        self._check(self._L.vr_set_eval_side_stream(self._ctx, int(bool(enabled))))

    # ------------------------------------------------------------------ sample-parallel progressive mode
    def set_frame_step(self, first, step):
        """vr_set_frame_step: the next draw_frame is frame `first` (0: as usual), later ones advance frameCount by `step`
        (rank r of G: first = r + 1, step = G)."""
        self._check(self._L.vr_set_frame_step(self._ctx, int(first), int(step)))

    def set_accumulate_mode(self, mode):
        """'average' (the reference's running average) or 'sum' (float sum of the frame colours; include/vr.h)."""
        self._check(self._L.vr_set_accumulate_mode(self._ctx, {"average": 0, "sum": 1}.get(mode, mode)))

    def finish_sum(self, total_frames):
        """vr_finish_sum: accumulator = sum / total_frames, rgba8 from that average, frameCount = total_frames."""
        self._check(self._L.vr_finish_sum(self._ctx, int(total_frames)))

    def read_image_rgba8(self):
        out = np.zeros((self.height, self.width, 4), np.uint8)
        self._check(self._L.vr_read_image_rgba8(self._ctx, out.ctypes.data_as(C.c_void_p)))
        return out

    def write_accum(self, accum, frame_count):
        a = np.ascontiguousarray(accum, np.float32)
        assert a.shape == (self.height, self.width, 4)
        self._check(self._L.vr_write_accum_f32(self._ctx, a.ctypes.data_as(C.c_void_p), frame_count))

    def write_image_rgba8(self, img):
        a = np.ascontiguousarray(img, np.uint8)
        assert a.shape == (self.height, self.width, 4)
        self._check(self._L.vr_write_image_rgba8(self._ctx, a.ctypes.data_as(C.c_void_p)))

    def counters(self):
        c = VrCounters()
        self._check(self._L.vr_get_counters(self._ctx, C.byref(c)))
        return {k: getattr(c, k) for k, _ in VrCounters._fields_}

    def timings(self):
        t = VrTimings()
        self._check(self._L.vr_get_timings(self._ctx, C.byref(t)))
        return {k: getattr(t, k) for k, _ in VrTimings._fields_}

    def timings_sum(self, reset=False):
        """Per-pass times summed over the frames finished since the last reset, and the frame count (no per-frame sync)."""
        t, n = VrTimings(), C.c_uint64()
        self._check(self._L.vr_get_timings_sum(self._ctx, C.byref(t), C.byref(n), int(bool(reset))))
        return {k: getattr(t, k) for k, _ in VrTimings._fields_}, int(n.value)

    def row_sample_counts(self):
        out = np.zeros(self.height, np.uint64)
        self._check(self._L.vr_row_sample_counts(self._ctx, out.ctypes.data_as(C.c_void_p)))
        return out

    def device_buffers(self):
        """Raw device pointers (accum float4, rgba8, frame colour) and the CUDA stream, for zero-copy interop."""
        a, i, f, s = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        self._check(self._L.vr_get_device_buffers(self._ctx, C.byref(a), C.byref(i), C.byref(f), C.byref(s)))
        return dict(accum=a.value, rgba8=i.value, frame_color=f.value, stream=s.value)

    def march_info(self):
        """Layout of the last gather's sample buffer: dict(single_pass, slots, samples)."""
        a = (C.c_uint64 * 3)()
        self._check(self._L.vr_get_march_info(self._ctx, a))
        return dict(single_pass=bool(a[0]), slots=int(a[1]), samples=int(a[2]))

    def set_march_single_pass(self, enabled):
        """vr_set_march_single_pass: 1 (default) = one march per frame into a sample buffer with gaps; 0 = count + fill."""
        self._check(self._L.vr_set_march_single_pass(self._ctx, int(bool(enabled))))

    def set_stream(self, cuda_stream):
        """Run on a caller-owned CUDA stream (int handle, e.g. torch.cuda.current_stream().cuda_stream); None = own."""
        self._check(self._L.vr_set_stream(self._ctx, C.c_void_p(cuda_stream) if cuda_stream else None))

    def launch_count(self):
        n = C.c_uint64()
        self._check(self._L.vr_get_launch_count(self._ctx, C.byref(n)))
        return n.value

    def lookup_values(self, ijk):
        """Grid values at integer index coordinates through the device traversal."""
        a = np.ascontiguousarray(ijk, np.int32).reshape(-1, 3)
        out = np.zeros(len(a), np.float32)
        self._check(self._L.vr_lookup_values(self._ctx, a.ctypes.data_as(C.c_void_p), len(a), out.ctypes.data_as(C.c_void_p)))
        return out

    def timer_start(self):
        self._check(self._L.vr_timer_start(self._ctx))

    def timer_stop(self):
        ms = C.c_float()
        self._check(self._L.vr_timer_stop(self._ctx, C.byref(ms)))
        return ms.value

    def ubench(self, kind):
        """0: FP32 FMA TFLOP/s, 1: MUFU.RCP 1e12 op/s, 2: point-light inner loop 1e12 eval/s (registers only)."""
        out = C.c_double()
        self._check(self._L.vr_ubench(self._ctx, {"fma": 0, "rcp": 1, "point": 2}.get(kind, kind), C.byref(out)))
        return out.value

    def cpu_test_render(self, width=256, height=256):
        """CPU_test/main.cpp's density ray-march (fixed camera, 1200 unit steps) as a kernel."""
        out = np.zeros((height, width), np.float32)
        self._check(self._L.vr_cpu_test_render(self._ctx, width, height, out.ctypes.data_as(C.c_void_p)))
        return out


    # ------------------------------------------------------------------ snapshots (progressive state)
    def save_snapshot(self, path):
        rc = self._L.vr_snapshot_save(self._ctx, os.fsencode(path))
        if rc:
            raise VrError(rc, (self._L.vr_io_last_error() or b"").decode() or self._L.vr_last_error(self._ctx).decode())

    def load_snapshot(self, path):
        rc = self._L.vr_snapshot_load(self._ctx, os.fsencode(path))
        if rc:
            raise VrError(rc, (self._L.vr_io_last_error() or b"").decode() or self._L.vr_last_error(self._ctx).decode())
        u = VrUbo()
        self._check(self._L.vr_get_params(self._ctx, C.byref(u)))
        self.width, self.height = int(u.framebufferDim[0]), int(u.framebufferDim[1])


# ---------------------------------------------------------------------- files (host side, no GPU needed)
def _io_check(L, rc, what):
    if rc:
        raise VrError(rc, f"{what}: {(L.vr_io_last_error() or b'').decode()}")


def load_nvdb_grid(path, grid_name=None):
    """First float grid (or the one called grid_name) of an uncompressed .nvdb file / raw grid buffer, as uint8 array."""
    L = _lib.load()
    p, sz = C.c_void_p(), C.c_uint64()
    _io_check(L, L.vr_grid_load_nvdb(os.fsencode(path), grid_name.encode() if grid_name else None, C.byref(p), C.byref(sz)),
              "vr_grid_load_nvdb")
    return _take(L, p, sz)


def save_nvdb_grid(path, grid, raw=False):
    L = _lib.load()
    g = np.ascontiguousarray(grid, np.uint8)
    _io_check(L, L.vr_grid_save_nvdb(os.fsencode(path), g.ctypes.data_as(C.c_void_p), g.nbytes, int(bool(raw))), "vr_grid_save_nvdb")


def write_ppm(path, rgba8):
    L = _lib.load()
    a = np.ascontiguousarray(rgba8, np.uint8)
    assert a.ndim == 3 and a.shape[2] == 4
    _io_check(L, L.vr_image_write_ppm(os.fsencode(path), a.ctypes.data_as(C.c_void_p), a.shape[1], a.shape[0]), "vr_image_write_ppm")


def write_pfm(path, image):
    L = _lib.load()
    a = np.ascontiguousarray(image, np.float32)
    if a.ndim == 2:
        a = a[..., None]
    _io_check(L, L.vr_image_write_pfm(os.fsencode(path), a.ctypes.data_as(C.c_void_p), a.shape[1], a.shape[0], a.shape[2]), "vr_image_write_pfm")


HEADLESS_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "vr_headless")


# ---------------------------------------------------------------------- grid fixtures (host side, no GPU needed)
def _take(L, p, n):
    a = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,)).copy()
    L.vr_grid_free(p)
    return a


def build_cloud_grid(n, seed=1234, voxel_size=None, translation=(-32.0, -12.0, -32.0), nthreads=0):
    """Procedural cloud of SURVEY 8(d): n^3 voxels, background 0, uniform map voxel 64/n + translation."""
    L = _lib.load()
    t = np.asarray(translation, np.float64)
    p, sz = C.c_void_p(), C.c_uint64()
    rc = L.vr_grid_build_cloud(n, seed, voxel_size if voxel_size else 64.0 / n, t.ctypes.data_as(C.c_void_p), nthreads,
                               C.byref(p), C.byref(sz))
    if rc:
        raise VrError(rc, "vr_grid_build_cloud")
    return _take(L, p, sz)


def build_fog_sphere_grid(radius, center=(0, 0, 0), voxel_size=1.0, translation=(0.0, 0.0, 0.0), nthreads=0):
    L = _lib.load()
    c = np.asarray(center, np.int32)
    t = np.asarray(translation, np.float64)
    p, sz = C.c_void_p(), C.c_uint64()
    rc = L.vr_grid_build_fog_sphere(radius, c.ctypes.data_as(C.c_void_p), voxel_size, t.ctypes.data_as(C.c_void_p), nthreads,
                                    C.byref(p), C.byref(sz))
    if rc:
        raise VrError(rc, "vr_grid_build_fog_sphere")
    return _take(L, p, sz)


def build_dense_grid(dense, origin=(0, 0, 0), background=0.0, voxel_size=1.0, translation=(0.0, 0.0, 0.0), nthreads=0):
    L = _lib.load()
    d = np.ascontiguousarray(dense, np.float32)
    dims = np.asarray(d.shape, np.int32)
    o = np.asarray(origin, np.int32)
    t = np.asarray(translation, np.float64)
    p, sz = C.c_void_p(), C.c_uint64()
    rc = L.vr_grid_build_dense(d.ctypes.data_as(C.c_void_p), dims.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p),
                               background, voxel_size, t.ctypes.data_as(C.c_void_p), nthreads, C.byref(p), C.byref(sz))
    if rc:
        raise VrError(rc, "vr_grid_build_dense")
    return _take(L, p, sz)
