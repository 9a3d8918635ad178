"""vr_light_gen_static: light generation for hosts that may not allocate device memory (include/vr.h).

The host this stands for owns all GPU memory through its own pool, or calls in the middle of a frame where a cudaMalloc or
cudaFree inside a plug-in would stall the device, so the library must get by on memory it already has: the working
buffers of the pass live in one fixed array inside libvr. Here torch plays that host: it owns the volume, the light list
and the counter (tensors out of its caching allocator) and hands over raw device pointers and its stream.

What is checked: the light list, the counter and the scatter-event total are byte-identical with what the allocating
entry points (vr_reset_lights + vr_dispatch_light_gen of a context) produce for the same volume and parameters -- light
generation is integer/bit-exact work, so the tolerance is zero, as in tests/test_gpu_parity.py. Several calls follow each
other on one stream with different photon counts and seeds, so a call that trusted what its predecessor left in the shared
array would be caught. Free device memory is compared around the later calls: it must not move.

The file name sorts last on purpose: pytest -x stops at the first failure, and a problem in this add-on must not hide
the results of the parity tests.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

VR_OK, VR_E_ARG, VR_E_CAPACITY = 0, 1, 4
MAX_PHOTONS = 4096  # VR_STATIC_MAX_PHOTONS in include/vr.h


def _context_lights(r, groups, variant, **ubo):
    """The existing path: counter reset + vr_dispatch_light_gen on a context, read back through the C ABI."""
    r.update_ubo(**ubo)
    r.reset_lights()
    r.dispatch_light_gen("point", groups, variant)
    lights, count, emitted = r.read_lights()
    return lights[:count].tobytes(), count, emitted


def _static_lights(L, torch, grid, grid_dev, u, groups, variant, lights_dev, counter_dev):
    """The same pass through vr_light_gen_static on torch's current stream; outputs are the caller's tensors."""
    err = C.create_string_buffer(512)
    rc = L.vr_light_gen_static(0, C.c_void_p(torch.cuda.current_stream().cuda_stream), C.c_void_p(grid_dev.data_ptr()),
                               grid.nbytes, grid[:1024].ctypes.data_as(C.c_void_p), 1024, C.byref(u), groups[0], groups[1],
                               groups[2], variant, C.c_void_p(lights_dev.data_ptr()), C.c_void_p(counter_dev.data_ptr()),
                               err, len(err))
    assert rc == VR_OK, (rc, err.value)
    torch.cuda.synchronize()
    counter = counter_dev.cpu().numpy()
    count = int(counter[:4].view(np.uint32)[0])
    emitted = int(counter[8:16].view(np.uint64)[0])
    return lights_dev.cpu().numpy()[:count * 32].tobytes(), count, emitted


def test_static_light_gen_equals_context_light_gen(vr, cloud):
    import torch
    from volumerenderer_b200 import _lib

    L = _lib.load()
    grid = cloud(64)
    torch.cuda.set_device(0)
    grid_dev = torch.from_numpy(grid).cuda()
    max_lights = 1000
    lights_dev = torch.empty(max_lights * 32, dtype=torch.uint8, device="cuda")
    counter_dev = torch.empty(16, dtype=torch.uint8, device="cuda")
    # (work-groups, variant, frameCount, maxLights): a large dispatch first, then a small one that must not see its slabs,
    # a capped list (more scatter events than maxLights), the no-reset shader body, and the largest dispatch allowed
    cases = [((8, 4, 1), 0, 1, 1000), ((1, 1, 1), 0, 2, 1000), ((4, 2, 2), 0, 3, 40), ((2, 2, 1), 1, 7, 1000),
             ((16, 16, 1), 0, 5, 1000), ((2, 1, 1), 0, 5, 1000)]
    with vr.VolumeRenderer(0) as r:
        r.set_grid(grid)
        r.resize(64, 64)
        r.set_max_lights_capacity(max_lights)
        want = [_context_lights(r, g, v, frameCount=f, maxLights=m) for g, v, f, m in cases]
        assert want[0][1] > 100 and want[2][2] > want[2][1] == 40 and want[0][0] != want[1][0]  # the cases do differ
        free_before = None
        for i, (g, v, f, m) in enumerate(cases):
            u = r.ubo
            u.frameCount, u.maxLights = f, m
            lights_dev.fill_(0xAB)
            counter_dev.fill_(0xCD)  # the call resets the counter itself
            torch.cuda.synchronize()
            if i == 1:  # the first call may load device code; from the second on free memory must stay where it is
                free_before = torch.cuda.mem_get_info()[0]
            got = _static_lights(L, torch, grid, grid_dev, u, g, v, lights_dev, counter_dev)
            assert got[1:] == want[i][1:], (i, got[1:], want[i][1:])
            assert got[0] == want[i][0], f"case {i}: light list differs from vr_dispatch_light_gen"
        assert torch.cuda.mem_get_info()[0] == free_before, "device memory was allocated or freed by vr_light_gen_static"
        # two calls back to back on the stream without a host wait in between: the second overwrites the first's outputs
        u = r.ubo
        err = C.create_string_buffer(512)
        for g, v, f, m in (cases[0], cases[3]):
            u.frameCount, u.maxLights = f, m
            rc = L.vr_light_gen_static(0, C.c_void_p(torch.cuda.current_stream().cuda_stream), C.c_void_p(grid_dev.data_ptr()),
                                       grid.nbytes, grid[:1024].ctypes.data_as(C.c_void_p), 1024, C.byref(u), g[0], g[1], g[2], v,
                                       C.c_void_p(lights_dev.data_ptr()), C.c_void_p(counter_dev.data_ptr()), err, len(err))
            assert rc == VR_OK, err.value
        torch.cuda.synchronize()
        count = int(counter_dev.cpu().numpy()[:4].view(np.uint32)[0])
        assert count == want[3][1] and lights_dev.cpu().numpy()[:count * 32].tobytes() == want[3][0]


def test_static_light_gen_rejects_what_does_not_fit(vr, cloud):
    """More photons than the fixed working memory holds, and malformed arguments, are refused with a message and without
    touching the outputs."""
    import torch
    from volumerenderer_b200 import _lib

    L = _lib.load()
    grid = cloud(64)
    grid_dev = torch.from_numpy(grid).cuda()
    lights_dev = torch.full((64 * 32,), 0x5A, dtype=torch.uint8, device="cuda")
    counter_dev = torch.full((16,), 0x5A, dtype=torch.uint8, device="cuda")
    u = _lib.VrUbo()
    L.vr_default_params(C.byref(u), 64, 64)
    u.frameCount, u.maxLights = 1, 64
    err = C.create_string_buffer(512)

    def call(groups, header_bytes=1024, grid_ptr=None, step=None):
        uu = _lib.VrUbo.from_buffer_copy(u)
        if step is not None:
            uu.rayMarchingStepSize = step
        err.value = b""
        return L.vr_light_gen_static(0, None, C.c_void_p(grid_dev.data_ptr() if grid_ptr is None else grid_ptr), grid.nbytes,
                                     grid[:1024].ctypes.data_as(C.c_void_p), header_bytes, C.byref(uu), groups[0], groups[1],
                                     groups[2], 0, C.c_void_p(lights_dev.data_ptr()), C.c_void_p(counter_dev.data_ptr()), err,
                                     len(err))

    assert MAX_PHOTONS == 16 * 16 * 16
    assert call((16, 16, 2)) == VR_E_CAPACITY and str(MAX_PHOTONS).encode() in err.value
    assert call((1 << 31, 1 << 31, 4)) == VR_E_CAPACITY  # products that wrap 64 bits are caught per factor
    assert call((0, 1, 1)) == VR_E_ARG and err.value
    assert call((1, 1, 1), header_bytes=100) == VR_E_ARG and b"header" in err.value
    assert call((1, 1, 1), grid_ptr=grid_dev.data_ptr() + 4) == VR_E_ARG and b"aligned" in err.value
    assert call((1, 1, 1), step=0.0) == VR_E_ARG and b"rayMarchingStepSize" in err.value
    torch.cuda.synchronize()
    assert bool((lights_dev == 0x5A).all()) and bool((counter_dev == 0x5A).all())
    assert call((1, 1, 1)) == VR_OK
    torch.cuda.synchronize()
    assert int(counter_dev.cpu().numpy()[:4].view(np.uint32)[0]) <= 64
