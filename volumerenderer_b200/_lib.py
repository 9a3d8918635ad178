"""ctypes binding of libvr.so (the C ABI declared in include/vr.h). No torch types cross this boundary."""
import ctypes as C
import os

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libvr.so")

VR_OK, VR_E_ARG, VR_E_CUDA, VR_E_BADGRID, VR_E_CAPACITY, VR_E_STATE = range(6)
STATUS_NAMES = {0: "VR_OK", 1: "VR_E_ARG", 2: "VR_E_CUDA", 3: "VR_E_BADGRID", 4: "VR_E_CAPACITY", 5: "VR_E_STATE"}

# enum order of src/main.cpp:65-68
ALGO = {"beam": 0, "ray": 1, "point": 2, "sphere": 3, "path": 4}
LG_RESET_ORIGIN, LG_KEEP_ORIGIN = 0, 1


class VrUbo(C.Structure):
    """struct vr_ubo: the 96-byte std140 UBO of shaders/common_bindings.h:19-34."""
    _fields_ = [
        ("frameCount", C.c_uint32), ("_pad0", C.c_uint32), ("framebufferDim", C.c_uint32 * 2),
        ("cameraPos", C.c_float * 3), ("fov", C.c_float),
        ("photonInitialIntensity", C.c_float), ("scatteringProbability", C.c_float),
        ("absorptionCoefficient", C.c_float), ("maxLights", C.c_uint32),
        ("rayMaxDistance", C.c_float), ("rayMarchingStepSize", C.c_float), ("_pad1", C.c_uint32 * 2),
        ("lightSourceWorldPos", C.c_float * 3), ("beamRadius", C.c_float),
        ("lightRayStepSize", C.c_float), ("radiusFalloff", C.c_float), ("_pad2", C.c_uint32 * 2),
    ]


assert C.sizeof(VrUbo) == 96


class VrTile(C.Structure):
    _fields_ = [("x0", C.c_uint32), ("y0", C.c_uint32), ("w", C.c_uint32), ("h", C.c_uint32)]


class VrOptions(C.Structure):
    """struct vr_options: opt-in physics; all zero = the reference's arithmetic."""
    _fields_ = [("trilinear", C.c_uint32), ("phase_hg", C.c_uint32), ("hg_g", C.c_float), ("light_transmittance", C.c_uint32)]


class VrCounters(C.Structure):
    _fields_ = [("S_ref", C.c_uint64), ("S_eff", C.c_uint64), ("E", C.c_uint64), ("eval_lights", C.c_uint64),
                ("light_count", C.c_uint64)]


class VrTimings(C.Structure):
    _fields_ = [("light_gen_ms", C.c_float), ("march_ms", C.c_float), ("prepare_ms", C.c_float),
                ("eval_ms", C.c_float), ("resolve_ms", C.c_float), ("total_ms", C.c_float)]


# struct vr_ray_light (std430 stride 32)
LIGHT_DTYPE = np.dtype([("from", "<f4", 3), ("_pad0", "<f4"), ("to", "<f4", 3), ("intensity", "<f4")])
assert LIGHT_DTYPE.itemsize == 32

# every symbol include/vr.h declares (tests check that the library exports all of them)
EXPORTS = [
    "vr_create", "vr_destroy", "vr_last_error", "vr_set_grid", "vr_resize", "vr_set_params", "vr_get_params",
    "vr_default_params", "vr_set_max_lights_capacity", "vr_reset_lights", "vr_write_lights", "vr_clear_image",
    "vr_dispatch_light_gen", "vr_dispatch_compute_color", "vr_set_light_gen_groups", "vr_render_frame",
    "vr_reset_accumulation", "vr_sync", "vr_read_lights", "vr_read_light_counter", "vr_read_image_rgba8",
    "vr_read_accum_f32", "vr_read_frame_color_f32", "vr_write_accum_f32", "vr_write_image_rgba8", "vr_get_counters",
    "vr_get_timings", "vr_get_timings_sum", "vr_get_device_buffers", "vr_row_sample_counts", "vr_accumulate_frame_color", "vr_set_frame_color_peers", "vr_set_march_cache", "vr_set_march_single_pass", "vr_set_light_prefetch", "vr_read_accum_f32_async", "vr_read_wait", "vr_set_options", "vr_get_options", "vr_trilinear_values", "vr_progressive_radius", "vr_set_eval_variant", "vr_set_eval_near_distance", "vr_get_eval_segment_counts", "vr_get_eval_redo_count", "vr_read_samples", "vr_lookup_values", "vr_cpu_test_render",
    "vr_set_stream", "vr_set_grid_device", "vr_set_grid_device_from_stream", "vr_get_launch_count", "vr_timer_start", "vr_timer_stop", "vr_ubench", "vr_grid_build_cloud", "vr_grid_build_fog_sphere", "vr_grid_build_dense", "vr_grid_free", "vr_version",
    "vr_grid_load_nvdb", "vr_grid_save_nvdb", "vr_image_write_ppm", "vr_image_write_pfm", "vr_snapshot_save", "vr_snapshot_load",
    "vr_io_last_error",
    "vr_read_accum_f32_tile_async", "vr_set_eval_side_stream", "vr_set_frame_step", "vr_set_accumulate_mode", "vr_finish_sum", "vr_get_march_info", "vr_read_image_rgba8_tile",
    "vr_group_create", "vr_group_destroy", "vr_group_last_error", "vr_group_size", "vr_group_ctx", "vr_group_set_grid", "vr_group_resize",
    "vr_group_set_params", "vr_group_get_params", "vr_group_set_max_lights_capacity", "vr_group_set_light_gen_groups",
    "vr_group_set_exchange", "vr_group_reset_accumulation", "vr_group_set_bands", "vr_group_get_bands", "vr_group_balance",
    "vr_group_render_frame", "vr_group_sync", "vr_group_read_accum_f32", "vr_group_read_accum_f32_async", "vr_group_read_wait",
    "vr_group_read_image_rgba8", "vr_group_render_progressive", "vr_group_time_frames",
    "vr_light_gen_static",
]

_lib = None


class VrLibraryMissing(RuntimeError):
    pass


def load():
    """Load libvr.so. There is no fallback: a missing library is an error."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise VrLibraryMissing(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(nvcc, sm_100a). volumerenderer_b200 has no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    P, U32, U64, I = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
    L.vr_version.restype = C.c_char_p
    L.vr_last_error.restype = C.c_char_p
    L.vr_last_error.argtypes = [P]
    L.vr_create.argtypes = [I, C.POINTER(P)]
    L.vr_destroy.argtypes = [P]
    L.vr_set_grid.argtypes = [P, P, U64]
    L.vr_resize.argtypes = [P, U32, U32]
    L.vr_set_params.argtypes = [P, C.POINTER(VrUbo)]
    L.vr_get_params.argtypes = [P, C.POINTER(VrUbo)]
    L.vr_default_params.argtypes = [C.POINTER(VrUbo), U32, U32]
    L.vr_default_params.restype = None
    L.vr_set_max_lights_capacity.argtypes = [P, U32]
    L.vr_reset_lights.argtypes = [P]
    L.vr_write_lights.argtypes = [P, P, U32]
    L.vr_clear_image.argtypes = [P]
    L.vr_dispatch_light_gen.argtypes = [P, I, U32, U32, U32, I]
    L.vr_dispatch_compute_color.argtypes = [P, I, C.POINTER(VrTile)]
    L.vr_set_light_gen_groups.argtypes = [P, U32, U32, U32, I]
    L.vr_render_frame.argtypes = [P, I, C.POINTER(VrTile)]
    L.vr_reset_accumulation.argtypes = [P]
    L.vr_sync.argtypes = [P]
    L.vr_read_lights.argtypes = [P, P, U32, C.POINTER(U32)]
    L.vr_read_light_counter.argtypes = [P, P, C.POINTER(U64)]
    L.vr_read_image_rgba8.argtypes = [P, P]
    L.vr_read_accum_f32.argtypes = [P, P]
    L.vr_read_accum_f32_async.argtypes = [P, P]
    L.vr_read_wait.argtypes = [P]
    L.vr_read_frame_color_f32.argtypes = [P, P]
    L.vr_write_accum_f32.argtypes = [P, P, U32]
    L.vr_write_image_rgba8.argtypes = [P, P]
    L.vr_get_counters.argtypes = [P, C.POINTER(VrCounters)]
    L.vr_get_timings.argtypes = [P, C.POINTER(VrTimings)]
    L.vr_get_timings_sum.argtypes = [P, C.POINTER(VrTimings), C.POINTER(U64), I]
    L.vr_get_device_buffers.argtypes = [P, C.POINTER(P), C.POINTER(P), C.POINTER(P), C.POINTER(P)]
    L.vr_row_sample_counts.argtypes = [P, P]
    L.vr_accumulate_frame_color.argtypes = [P, C.POINTER(VrTile)]
    L.vr_set_frame_color_peers.argtypes = [P, P, C.POINTER(P), U32]
    L.vr_set_march_cache.argtypes = [P, I]
    L.vr_set_light_prefetch.argtypes = [P, I]
    L.vr_set_march_single_pass.argtypes = [P, I]
    L.vr_cpu_test_render.argtypes = [P, U32, U32, P]
    L.vr_set_options.argtypes = [P, C.POINTER(VrOptions)]
    L.vr_get_options.argtypes = [P, C.POINTER(VrOptions)]
    L.vr_trilinear_values.argtypes = [P, P, U64, P]
    L.vr_progressive_radius.argtypes = [C.c_double, C.c_double]
    L.vr_progressive_radius.restype = C.c_double
    L.vr_set_eval_variant.argtypes = [P, I]
    L.vr_get_eval_redo_count.argtypes = [P, C.POINTER(U64)]
    L.vr_set_eval_near_distance.argtypes = [P, C.c_float]
    L.vr_get_eval_segment_counts.argtypes = [P, C.POINTER(U64)]
    L.vr_read_samples.argtypes = [P, P, U64, C.POINTER(U64)]
    L.vr_lookup_values.argtypes = [P, P, U64, P]
    L.vr_set_stream.argtypes = [P, P]
    L.vr_set_grid_device.argtypes = [P, P, C.c_uint64, P, C.c_uint64]
    L.vr_set_grid_device_from_stream.argtypes = [P, P, C.c_uint64, P, C.c_uint64, P]
    L.vr_get_launch_count.argtypes = [P, C.POINTER(U64)]
    L.vr_timer_start.argtypes = [P]
    L.vr_timer_stop.argtypes = [P, C.POINTER(C.c_float)]
    L.vr_ubench.argtypes = [P, I, C.POINTER(C.c_double)]
    L.vr_grid_build_cloud.argtypes = [C.c_int32, U32, C.c_double, P, I, C.POINTER(P), C.POINTER(U64)]
    L.vr_grid_build_fog_sphere.argtypes = [C.c_double, P, C.c_double, P, I, C.POINTER(P), C.POINTER(U64)]
    L.vr_grid_build_dense.argtypes = [P, P, P, C.c_float, C.c_double, P, I, C.POINTER(P), C.POINTER(U64)]
    L.vr_grid_free.argtypes = [P]
    L.vr_grid_free.restype = None
    L.vr_grid_load_nvdb.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(P), C.POINTER(U64)]
    L.vr_grid_save_nvdb.argtypes = [C.c_char_p, P, U64, I]
    L.vr_image_write_ppm.argtypes = [C.c_char_p, P, U32, U32]
    L.vr_image_write_pfm.argtypes = [C.c_char_p, P, U32, U32, U32]
    L.vr_snapshot_save.argtypes = [P, C.c_char_p]
    L.vr_snapshot_load.argtypes = [P, C.c_char_p]
    L.vr_io_last_error.restype = C.c_char_p
    L.vr_read_accum_f32_tile_async.argtypes = [P, C.POINTER(VrTile), P, U64]
    L.vr_set_eval_side_stream.argtypes = [P, I]
    L.vr_set_frame_step.argtypes = [P, U32, U32]
    L.vr_set_accumulate_mode.argtypes = [P, I]
    L.vr_finish_sum.argtypes = [P, U32]
    L.vr_get_march_info.argtypes = [P, C.POINTER(U64)]
    L.vr_read_image_rgba8_tile.argtypes = [P, C.POINTER(VrTile), P, U64]
    L.vr_group_create.argtypes = [C.POINTER(I), U32, C.POINTER(P)]
    L.vr_group_destroy.argtypes = [P]
    L.vr_group_last_error.restype = C.c_char_p
    L.vr_group_last_error.argtypes = [P]
    L.vr_group_size.argtypes = [P]
    L.vr_group_size.restype = U32
    L.vr_group_ctx.argtypes = [P, U32]
    L.vr_group_ctx.restype = P
    L.vr_group_set_grid.argtypes = [P, P, U64]
    L.vr_group_resize.argtypes = [P, U32, U32]
    L.vr_group_set_params.argtypes = [P, C.POINTER(VrUbo)]
    L.vr_group_get_params.argtypes = [P, C.POINTER(VrUbo)]
    L.vr_group_set_max_lights_capacity.argtypes = [P, U32]
    L.vr_group_set_light_gen_groups.argtypes = [P, U32, U32, U32, I]
    L.vr_group_set_exchange.argtypes = [P, I]
    L.vr_group_reset_accumulation.argtypes = [P]
    L.vr_group_set_bands.argtypes = [P, C.POINTER(U32), C.POINTER(U32)]
    L.vr_group_get_bands.argtypes = [P, C.POINTER(U32), C.POINTER(U32)]
    L.vr_group_balance.argtypes = [P, I, U32]
    L.vr_group_render_frame.argtypes = [P, I]
    L.vr_group_sync.argtypes = [P]
    L.vr_group_read_accum_f32.argtypes = [P, P]
    L.vr_group_read_accum_f32_async.argtypes = [P, P]
    L.vr_group_read_wait.argtypes = [P]
    L.vr_group_read_image_rgba8.argtypes = [P, P]
    L.vr_group_render_progressive.argtypes = [P, I, U32]
    L.vr_group_time_frames.argtypes = [P, I, U32, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L.vr_light_gen_static.argtypes = [I, P, P, U64, P, U64, C.POINTER(VrUbo), U32, U32, U32, I, P, P, C.c_char_p, U64]
    _lib = L
    return L
