"""vr_group_* (csrc/vr_group.cu): the C++ multi-GPU host path -- one process, one context and one worker thread per GPU,
peer access, device-side barrier, band balancer, per-GPU read-back, sample-parallel progressive runs. On a box with one
GPU the group lists that device several times (several bands on one GPU): same code path, same kernels, peer stores to
the device's own memory. Everything must equal the one-context render bit for bit."""
import ctypes as C

import numpy as np
import pytest

from conftest import rel_err

pytestmark = pytest.mark.gpu


def _ndev():
    import torch
    return torch.cuda.device_count()


def _devices(n):
    nd = max(1, _ndev())
    return [i % nd for i in range(n)]


def _single(vr, grid, W, H, algo, frames, **ubo):
    with vr.VolumeRenderer(0) as r:
        r.set_grid(grid)
        r.resize(W, H)
        r.ubo = r.default_ubo()
        r.update_ubo(**ubo)
        r.set_light_gen_groups(6, 1, 1)
        for _ in range(frames):
            r.draw_frame(algo)
        return r.read_accum(), r.read_image_rgba8()


@pytest.mark.parametrize("exchange", ["none", "allgather"])
@pytest.mark.parametrize("algo,n", [("point", 4), ("beam", 3), ("path", 2)])
def test_group_tiled_frames_equal_single_context(vr, cloud, algo, n, exchange):
    grid = cloud(64)
    W, H, frames = 160, 96, 4
    want_a, want_i = _single(vr, grid, W, H, algo, frames, maxLights=90)
    with vr.RendererGroup(_devices(n)) as g:
        g.set_grid(grid)
        g.resize(W, H)
        g.ubo = g.default_ubo()
        g.update_ubo(maxLights=90)
        g.set_light_gen_groups(6, 1, 1)
        g.set_exchange(exchange)
        bands = g.balance(algo, rounds=2)
        assert len(bands) == n and sum(h for _, h in bands) == H and all(h > 0 for _, h in bands)
        assert g.ubo.frameCount == 0
        for _ in range(frames):
            g.draw_frame(algo)
        g.sync()
        assert np.array_equal(g.read_accum(), want_a)
        assert np.array_equal(g.read_image_rgba8(), want_i)
        if exchange == "allgather":   # every GPU holds the whole image
            L = vr.load()
            out = np.zeros((H, W, 4), np.float32)
            for i in range(n):
                assert L.vr_read_accum_f32(g.ctx(i), out.ctypes.data_as(C.c_void_p)) == 0
                assert np.array_equal(out, want_a), i
        # explicit bands, a new grid through the group (one upload, GPU-to-GPU replication), and more frames
        g.set_bands([(0, 10)] + [(10 + k * ((H - 10) // (n - 1)), (H - 10) // (n - 1) if k < n - 2 else H - 10 - (n - 2) * ((H - 10) // (n - 1))) for k in range(n - 1)])
        g.set_grid(grid)
        g.reset_accumulation()
        for _ in range(frames):
            g.draw_frame(algo)
        assert np.array_equal(g.read_accum(), want_a)
        ms, per = g.time_frames(algo, 2)
        assert ms > 0 and len(per) == n and max(per) == pytest.approx(ms)


def test_group_sample_parallel_progressive(vr, cloud):
    """frames 1..12 round-robin over 3 contexts, float sums, one reduce + division: the sequential running average up to
    float reassociation (bound 12 * 2^-22 relative), rgba8 = UNORM8 of the average"""
    grid = cloud(64)
    W, H, frames = 128, 80, 12
    want_a, _ = _single(vr, grid, W, H, "beam", frames, maxLights=60)
    with vr.RendererGroup(_devices(3)) as g:
        g.set_grid(grid)
        g.resize(W, H)
        g.ubo = g.default_ubo()
        g.update_ubo(maxLights=60)
        g.set_light_gen_groups(6, 1, 1)
        g.render_progressive("beam", frames)
        L = vr.load()
        got = np.zeros((H, W, 4), np.float32)
        assert L.vr_read_accum_f32(g.ctx(0), got.ctypes.data_as(C.c_void_p)) == 0
        img = np.zeros((H, W, 4), np.uint8)
        assert L.vr_read_image_rgba8(g.ctx(0), img.ctypes.data_as(C.c_void_p)) == 0
        rel, rmse = rel_err(got[..., 0], want_a[..., 0])
        assert rel <= frames * 2.0 ** -22 and want_a[..., 0].max() > 0.01, (rel, rmse)
        assert np.array_equal(img[..., 0], np.rint(np.clip(got[..., 0], 0, 1) * 255).astype(np.uint8))
        assert g.ubo.frameCount == frames
        # the result lives on the first GPU only (the others keep raw partial sums): the group's own read-backs return it
        # whole, blocking and asynchronous, and tiled frames are refused until the accumulation is reset
        assert np.array_equal(g.read_accum(), got)
        assert np.array_equal(g.read_image_rgba8(), img)
        got2 = np.full((H, W, 4), -1.0, np.float32)
        assert L.vr_group_read_accum_f32_async(g._g, got2.ctypes.data_as(C.c_void_p)) == 0
        assert L.vr_group_read_wait(g._g) == 0
        assert np.array_equal(got2, got)
        with pytest.raises(vr.VrError) as e:
            g.draw_frame("beam")
        assert e.value.code == 5 and "vr_group_reset_accumulation" in str(e.value)
        with pytest.raises(vr.VrError):
            g.time_frames("beam", 1)
        # the group is back in its default mode: a tiled frame sequence after the progressive run works as before
        g.reset_accumulation()
        g.draw_frame("beam")
        g.sync()
        one, _ = _single(vr, grid, W, H, "beam", 1, maxLights=60)
        assert np.array_equal(g.read_accum(), one)


@pytest.mark.parametrize("exchange", ["none", "allgather"])
def test_group_on_distinct_gpus(vr, cloud, exchange):
    """The same checks with every context on its OWN physical GPU (skipped on a one-GPU box, where the tests above list the
    device several times): CUDA peer access, k_peer_barrier's flags and the resolve kernel's peer stores cross NVLink, the
    volume is replicated GPU to GPU, the progressive reduce reads the peers' accumulators."""
    nd = _ndev()
    if nd < 2:
        pytest.skip("needs at least two GPUs")
    n = min(nd, 4)
    grid = cloud(64)
    W, H, frames = 160, 96, 4
    want_a, want_i = _single(vr, grid, W, H, "point", frames, maxLights=90)
    with vr.RendererGroup(list(range(n))) as g:
        g.set_grid(grid)
        g.resize(W, H)
        g.ubo = g.default_ubo()
        g.update_ubo(maxLights=90)
        g.set_light_gen_groups(6, 1, 1)
        g.set_exchange(exchange)
        g.balance("point", rounds=2)
        for _ in range(frames):
            g.draw_frame("point")
        g.sync()
        assert np.array_equal(g.read_accum(), want_a)
        assert np.array_equal(g.read_image_rgba8(), want_i)
        if exchange == "allgather":   # every GPU holds the whole image after the barrier
            L = vr.load()
            out = np.zeros((H, W, 4), np.float32)
            for i in range(n):
                assert L.vr_read_accum_f32(g.ctx(i), out.ctypes.data_as(C.c_void_p)) == 0
                assert np.array_equal(out, want_a), i
        # sample-parallel progressive run: reduce over NVLink on GPU 0, read back through the group
        pframes = 3 * n
        seq_a, _ = _single(vr, grid, W, H, "point", pframes, maxLights=90)
        g.render_progressive("point", pframes)
        rel, _ = rel_err(g.read_accum()[..., 0], seq_a[..., 0])
        assert rel <= pframes * 2.0 ** -22, rel


def test_group_argument_errors(vr, cloud):
    with pytest.raises(vr.VrError):
        vr.RendererGroup([99])
    with vr.RendererGroup(_devices(2)) as g:
        with pytest.raises(vr.VrError) as e:
            g.draw_frame("point")
        assert e.value.code == 5
        g.set_grid(cloud(64))
        g.resize(64, 48)
        with pytest.raises(vr.VrError):
            g.set_bands([(0, 20), (24, 24)])     # gap
        with pytest.raises(vr.VrError):
            g.set_bands([(0, 48), (48, 0)])      # empty band
        g.ubo = g.default_ubo()
        with pytest.raises(vr.VrError) as e:
            g.update_ubo(maxLights=10 ** 6)
        assert e.value.code == 4
