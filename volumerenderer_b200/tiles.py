"""Image-tile partitioning across GPUs: pixels are independent given (grid, lights, UBO), so
each rank renders a band of rows with the grid and the light list replicated (light generation is deterministic and
is simply repeated on every rank), and the bands are exchanged with one broadcast per band over NCCL.

The reference is single-GPU (one queue, src/main.cpp:265-270); this module is new capability, not a mirror.
Only the band exchange uses torch.distributed; everything else is plain Python/numpy and is tested on CPU with gloo.
"""
import numpy as np


def balanced_row_bands(row_cost, n_parts, min_rows=1):
    """Split rows [0, H) into n_parts contiguous bands of near-equal total cost.

    row_cost: per-row work estimate (e.g. VolumeRenderer.row_sample_counts(): non-zero-density camera samples per row;
    the evaluation kernel's time is proportional to it). Returns [(y0, h), ...] covering every row exactly once.
    """
    cost = np.asarray(row_cost, np.float64)
    H = len(cost)
    if n_parts < 1:
        raise ValueError("n_parts must be >= 1")
    if H < n_parts * min_rows:
        raise ValueError(f"{H} rows cannot be split into {n_parts} bands of >= {min_rows} rows")
    total = float(cost.sum())
    if total <= 0.0:
        cost = np.ones(H)
        total = float(H)
    prefix = np.concatenate([[0.0], np.cumsum(cost)])
    cuts = [0]
    for p in range(1, n_parts):
        target = total * p / n_parts
        y = int(np.searchsorted(prefix, target, side="left"))
        # pick the closer of the two neighbouring cut positions
        if y > 0 and abs(prefix[y - 1] - target) <= abs(prefix[min(y, H)] - target):
            y -= 1
        y = max(y, cuts[-1] + min_rows)
        y = min(y, H - (n_parts - p) * min_rows)
        cuts.append(y)
    cuts.append(H)
    return [(cuts[i], cuts[i + 1] - cuts[i]) for i in range(n_parts)]


def band_tile(band, width):
    """(y0, h) -> vr_tile tuple (x0, y0, w, h)."""
    return (0, int(band[0]), int(width), int(band[1]))


def exchange_bands(image, bands, group=None):
    """In-place all-gather of row bands: rank r owns rows bands[r] of `image` (a torch tensor whose first dimension is
    the image row, identical shape on every rank); afterwards every rank holds the full image.

    One broadcast per band (bands have unequal heights, which rules out a plain all_gather); NCCL runs them back to
    back on the communicator's stream. Works with CPU tensors over gloo (tests) and CUDA tensors over NCCL.
    """
    import torch.distributed as dist

    world = dist.get_world_size(group)
    if len(bands) != world:
        raise ValueError(f"{len(bands)} bands for world size {world}")
    works = []
    for r, (y0, h) in enumerate(bands):
        if h == 0:
            continue
        src = dist.get_global_rank(group, r) if group is not None else r
        works.append(dist.broadcast(image[y0:y0 + h], src=src, group=group, async_op=True))
    for w in works:
        w.wait()
    return image


def exchange_rows_p2p(image, bands, group=None):
    """Same result as exchange_bands, as ONE grouped NCCL operation: every rank sends its band to every peer and
    receives theirs (ncclGroupStart / Send / Recv / GroupEnd through batch_isend_irecv), so an N-way exchange costs one
    launch instead of N broadcasts. Meant for the 4 B/pixel frame-colour image; the receiver rebuilds its accumulator
    with VolumeRenderer.accumulate_frame_color."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if len(bands) != world:
        raise ValueError(f"{len(bands)} bands for world size {world}")
    ops = []
    y0, h = bands[rank]
    for r, (py0, ph) in enumerate(bands):
        if r == rank:
            continue
        peer = dist.get_global_rank(group, r) if group is not None else r
        if h:
            ops.append(dist.P2POp(dist.isend, image[y0:y0 + h], peer, group))
        if ph:
            ops.append(dist.P2POp(dist.irecv, image[py0:py0 + ph], peer, group))
    if ops:
        for w in dist.batch_isend_irecv(ops):
            w.wait()
    return image


def complement_tiles(band, width, height):
    """The (up to two) row rectangles of a width x height image outside `band` = (y0, h), as vr_tile tuples."""
    y0, h = int(band[0]), int(band[1])
    out = []
    if y0 > 0:
        out.append((0, 0, int(width), y0))
    if y0 + h < height:
        out.append((0, y0 + h, int(width), int(height) - (y0 + h)))
    return out


def row_costs(row_samples, march_ms, eval_ms):
    """Per-row cost for balanced_row_bands from one measured full frame: the evaluation kernel's time is proportional
    to the row's non-zero-density samples, the camera march's to the row itself (every pixel walks the bbox chord
    whether or not it finds density). Cost unit = one sample; a row adds march_ms/H worth of samples."""
    s = np.asarray(row_samples, np.float64)
    total = float(s.sum())
    if total <= 0.0 or eval_ms <= 0.0:
        return np.ones(len(s))
    per_row = (march_ms / len(s)) / (eval_ms / total)
    return s + per_row


def refine_row_costs(row_cost, bands, rank_ms):
    """One step of measured feedback: rank r spent rank_ms[r] on the rows of bands[r], so the cost of those rows is
    rescaled to make the model's band totals proportional to the measurements. Feed the result to balanced_row_bands;
    two or three rounds (a couple of untimed frames each) settle the split to within a row or two."""
    cost = np.asarray(row_cost, np.float64).copy()
    ms = np.asarray(rank_ms, np.float64)
    if len(ms) != len(bands) or not np.all(ms > 0):
        return cost
    total_model = float(cost.sum())
    for (y0, h), t in zip(bands, ms):
        if h <= 0:
            continue
        model = float(cost[y0:y0 + h].sum())
        if model > 0:
            cost[y0:y0 + h] *= (t / ms.sum()) / (model / total_model)
    return cost


def polish_move(bands, rank_ms):
    """One local move after the feedback rounds: the slowest band gives one boundary row to the faster of its two
    neighbours. Row costs differ by two orders of magnitude between the image's empty top and its centre, so where the
    model's residual is a fraction of a band's time a single measured row move is finer than another rescaling of the
    model. Returns the new bands, or None when the slowest band cannot give a row away."""
    ms = [float(x) for x in rank_ms]
    n = len(bands)
    if n < 2 or len(ms) != n:
        return None
    w = max(range(n), key=lambda i: ms[i])
    if bands[w][1] <= 1:
        return None
    nb = [i for i in (w - 1, w + 1) if 0 <= i < n]
    t = min(nb, key=lambda i: ms[i])
    out = [list(b) for b in bands]
    out[w][1] -= 1
    out[t][1] += 1
    if t < w:          # the upper neighbour takes the slowest band's first row
        out[w][0] += 1
    else:              # the lower neighbour takes its last row
        out[t][0] -= 1
    return [tuple(b) for b in out]


class PeerFrameColor:
    """Fused tile all-gather over NVLink peer memory (the "compute + collective in one kernel" form).

    Two (H, W) float32 frame-colour buffers are allocated as torch symmetric memory, i.e. every rank maps every other
    rank's buffers into its own address space. libvr's resolve kernel then stores each pixel of the rank's band to all
    ranks' buffers directly (vr_set_frame_color_peers), a device-side barrier on the renderer's stream orders the
    stores, and each rank averages the rows it did not render from its local copy (vr_accumulate_frame_color). No NCCL
    call per frame. The two buffers alternate by frame so that a fast rank's stores for the next frame cannot land in
    memory a slow rank is still reading.

    Raises whatever torch raises if symmetric memory is unavailable; callers fall back to exchange_rows_p2p.
    """

    def __init__(self, renderer, bands, group=None):
        import torch
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem

        self.r = renderer
        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world = dist.get_world_size(self.group)
        dev = torch.device(f"cuda:{renderer.device}")
        self.bufs, self.hdls, self.ptrs = [], [], []
        for _ in range(2):
            t = symm_mem.empty((renderer.height, renderer.width), dtype=torch.float32, device=dev)
            t.zero_()
            h = symm_mem.rendezvous(t, self.group)
            self.bufs.append(t)
            self.hdls.append(h)
            self.ptrs.append([int(p) for p in h.buffer_ptrs])
        self.set_bands(bands)
        self.parity = 0
        torch.cuda.synchronize()
        dist.barrier(self.group)

    def set_bands(self, bands):
        self.bands = [tuple(b) for b in bands]
        self.tile = band_tile(self.bands[self.rank], self.r.width)
        self.others = complement_tiles(self.bands[self.rank], self.r.width, self.r.height)

    def draw_frame(self, algo):
        """One frame: own band rendered with its frame colours stored to every rank, barrier, rest of the image averaged."""
        p = self.parity
        self.r.set_frame_color_peers(self.ptrs[p][self.rank], self.ptrs[p])
        self.r.draw_frame(algo, self.tile)
        self.hdls[p].barrier(channel=0)          # device-side, on the current (= renderer's) stream
        for t in self.others:
            self.r.accumulate_frame_color(t)
        self.parity ^= 1

    def close(self):
        self.r.set_frame_color_peers(None, [])


def reduce_progressive_sums(accum_sum, dst=0, group=None):
    """Sample-parallel progressive mode: every rank accumulates a float SUM of the frames it rendered;
    one reduce(sum) to `dst` combines them (divide by the total frame count afterwards). Not available for the
    rgba8-faithful running average, which is sequential by construction."""
    import torch.distributed as dist

    dist.reduce(accum_sum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum_sum


class DeviceArray:
    """Zero-copy view of libvr device memory for torch.as_tensor (CUDA array interface v2)."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def frame_color_as_tensor(renderer):
    """The renderer's frame-colour buffer as a (H, W) float32 CUDA tensor sharing memory with libvr."""
    import torch

    bufs = renderer.device_buffers()
    return torch.as_tensor(DeviceArray(bufs["frame_color"], (renderer.height, renderer.width), "<f4"),
                           device=f"cuda:{renderer.device}")


def accum_as_tensor(renderer):
    """The renderer's float4 accumulator as a (H, W, 4) float32 CUDA tensor sharing memory with libvr."""
    import torch

    bufs = renderer.device_buffers()
    return torch.as_tensor(DeviceArray(bufs["accum"], (renderer.height, renderer.width, 4), "<f4"),
                           device=f"cuda:{renderer.device}")


class ReplicatedGrid:
    """A new volume every frame on N GPUs with ONE host->device transfer per node: rank `src` uploads the NanoVDB bytes
    into a device staging buffer (a side stream, two buffers: the upload of frame k+1 runs under frame k) and replicates
    them over NVLink; every rank then installs them with vr_set_grid_device in stream order. Nothing here blocks the host.
    The renderer must run on torch's current stream (renderer.set_stream).

    Replication, preferred: the staging buffers are torch symmetric memory, rank `src` pushes its buffer into every peer's
    with copy-engine peer copies on the side stream (no SMs involved), and a device-side barrier on the side streams tells
    the peers that the bytes have landed -- the whole exchange runs underneath the previous frame's kernels. Fallback
    (symmetric memory unavailable): one NCCL broadcast on the render stream. (An NCCL broadcast on a side stream was
    measured and rejected: its CTAs become resident one by one as evaluation CTAs retire and then hold whole SMs while
    they wait for the rest of their grid -- C3 at two GPUs went from 287.5 to 306.9 ms per frame end to end.)"""

    def __init__(self, renderer, nbytes, src=0, group=None, symmetric=True):
        import torch
        import torch.distributed as dist

        self.r, self.src, self.nbytes = renderer, src, int(nbytes)
        self.group = group if group is not None else (dist.group.WORLD if dist.is_initialized() else None)
        self.world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        dev = torch.device(f"cuda:{renderer.device}")
        self.hdls = None
        ok = 0
        if symmetric and self.world > 1:
            try:
                import torch.distributed._symmetric_memory as symm_mem

                self.staging, hdls = [], []
                for _ in range(2):
                    t = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=dev)
                    hdls.append(symm_mem.rendezvous(t, self.group))
                    self.staging.append(t)
                self.hdls = hdls
                ok = 1
            except Exception:  # noqa: BLE001
                self.hdls = None
        if self.world > 1:  # every rank must take the same path
            flag = torch.tensor([ok], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
            if int(flag) == 0:
                self.hdls = None
        if self.hdls is None:
            self.staging = [torch.empty(self.nbytes, dtype=torch.uint8, device=dev) for _ in range(2)]
        else:
            self.peer_views = [[h.get_buffer(p, (self.nbytes,), torch.uint8) for p in range(self.world)] for h in self.hdls]
        self.mode = "symmetric memory: copy-engine peer copies + device barrier on a side stream" if self.hdls is not None else \
            ("nccl broadcast on the render stream" if self.world > 1 else "single GPU")
        self.upload = torch.cuda.Stream(device=dev)
        self.uploaded = [torch.cuda.Event() for _ in range(2)]
        self.consumed = [None, None]
        self.k = 0

    def set_grid(self, pinned_grid, header):
        """pinned_grid: uint8 CPU tensor (pinned) with the grid on rank `src`, ignored elsewhere; header: host bytes of the
        grid's first >= 1024 bytes, on every rank (they carry the map, the bounding box and the root offset)."""
        import torch
        import torch.distributed as dist

        i = self.k & 1
        main = torch.cuda.current_stream()
        if self.consumed[i] is not None:
            self.upload.wait_event(self.consumed[i])  # this rank's frame of two calls ago has installed the buffer's contents
        with torch.cuda.stream(self.upload):
            if self.hdls is not None:
                self.hdls[i].barrier(channel=0)      # ... and so has every peer's (src is about to overwrite their buffers)
            if self.rank == self.src:
                self.staging[i].copy_(pinned_grid, non_blocking=True)
                if self.hdls is not None:
                    for p in range(self.world):
                        if p != self.rank:
                            self.peer_views[i][p].copy_(self.staging[i], non_blocking=True)
            if self.hdls is not None:
                self.hdls[i].barrier(channel=1)      # the bytes have landed everywhere
            self.uploaded[i].record(self.upload)
        if self.hdls is not None or self.world == 1:
            # the bytes are complete in the side stream's order: libvr copies them into its free grid buffer next to the
            # frames in flight, and the next frame's light generation can start underneath the previous frame
            self.r.set_grid_device_from_stream(self.staging[i].data_ptr(), self.nbytes, header, self.upload.cuda_stream)
        else:
            main.wait_event(self.uploaded[i])
            dist.broadcast(self.staging[i], src=self.src, group=self.group)
            self.r.set_grid_device(self.staging[i].data_ptr(), self.nbytes, header)
        ev = torch.cuda.Event()
        ev.record(main)
        self.consumed[i] = ev
        self.k += 1
