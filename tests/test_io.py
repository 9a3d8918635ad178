"""Host-side file formats: .nvdb container, PPM/PFM writers, the headless driver's CLI surface.
No GPU needed. The .nvdb reader/writer is checked against nanovdb::io::writeGrid / readGrid of the vendored NanoVDB
(oracle/_ref, where the reference checkout exists) and against a committed file written by it (tests/golden)."""
import gzip
import os
import struct
import subprocess

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _same_grid(a, b):
    """equal up to gridIndex/gridCount (bytes 24..32), which readers rewrite for a single-grid buffer"""
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and np.array_equal(a[:24], b[:24]) and np.array_equal(a[32:], b[32:])


def test_nvdb_roundtrip_file_and_raw(vr, cloud, tmp_path):
    grid = cloud(32)
    for raw in (False, True):
        p = str(tmp_path / f"g{int(raw)}.nvdb")
        vr.save_nvdb_grid(p, grid, raw=raw)
        assert os.path.getsize(p) >= grid.nbytes
        back = vr.load_nvdb_grid(p)
        assert _same_grid(back, grid)
    # container layout: FileHeader 16 B, FileMetaData 176 B, name, grid (NanoVDB.h:6286-6325)
    data = open(str(tmp_path / "g0.nvdb"), "rb").read()
    magic, version, count, codec = struct.unpack_from("<QIHH", data, 0)
    assert magic == 0x304244566F6E614E and count == 1 and codec == 0
    grid_size, file_size = struct.unpack_from("<QQ", data, 16)
    name_size = struct.unpack_from("<I", data, 16 + 136)[0]
    assert grid_size == file_size == grid.nbytes and len(data) == 16 + 176 + name_size + grid.nbytes
    assert struct.unpack_from("<I", data, 16 + 32)[0] == 1  # GridType::Float


def test_nvdb_golden_written_by_nanovdb(vr, oracle):
    """tests/golden/nanovdb_ball80.nvdb.gz was written by nanovdb::io::writeGrid (tools/make_golden.py)"""
    path = os.path.join(GOLD, "nanovdb_ball80.nvdb.gz")
    raw = np.frombuffer(gzip.open(os.path.join(GOLD, "nanovdb_ball80.nvdb.raw.gz")).read(), np.uint8)
    import tempfile
    with tempfile.NamedTemporaryFile(suffix=".nvdb") as f:
        f.write(gzip.open(path).read())
        f.flush()
        g = vr.load_nvdb_grid(f.name)
    assert _same_grid(g, raw)
    ijk = np.load(os.path.join(GOLD, "nanovdb_ball80_ijk.npy"))
    vals = np.load(os.path.join(GOLD, "nanovdb_ball80_values.npy"))
    assert np.array_equal(oracle.lookup(g, ijk), vals)


def test_nvdb_interop_with_nanovdb_io(vr, ref, cloud, tmp_path):
    grid = cloud(32)
    theirs = str(tmp_path / "theirs.nvdb")
    ref.write_nvdb(grid, theirs)                      # nanovdb::io::writeGrid
    assert _same_grid(vr.load_nvdb_grid(theirs), grid)
    ours = str(tmp_path / "ours.nvdb")
    vr.save_nvdb_grid(ours, grid)
    assert _same_grid(ref.read_nvdb(ours), grid)      # nanovdb::io::readGrid
    # a NanoVDB-built grid (its own name, stats, checksum) through both directions
    sphere = ref.make_fog_sphere(20.0)
    vr.save_nvdb_grid(ours, sphere)
    assert _same_grid(ref.read_nvdb(ours), sphere)
    assert _same_grid(vr.load_nvdb_grid(ours, "sphere_fog"), sphere)
    with pytest.raises(vr.VrError):
        vr.load_nvdb_grid(ours, "no_such_grid")


def test_nvdb_rejects_compressed_and_garbage(vr, cloud, tmp_path):
    grid = cloud(32)
    p = str(tmp_path / "c.nvdb")
    vr.save_nvdb_grid(p, grid)
    data = bytearray(open(p, "rb").read())
    struct.pack_into("<Q", data, 0, 0x324244566F6E614E)   # "NanoVDB2" file magic (unambiguous)
    struct.pack_into("<H", data, 14, 1)                   # Codec::ZIP
    open(p, "wb").write(data)
    with pytest.raises(vr.VrError) as e:
        vr.load_nvdb_grid(p)
    assert e.value.code == 3 and "compressed" in str(e.value)
    open(p, "wb").write(b"not a grid at all" * 100)
    with pytest.raises(vr.VrError) as e:
        vr.load_nvdb_grid(p)
    assert e.value.code == 3
    with pytest.raises(vr.VrError):
        vr.load_nvdb_grid(str(tmp_path / "missing.nvdb"))
    with pytest.raises(vr.VrError):
        vr.save_nvdb_grid(p, np.zeros(4096, np.uint8))
    open(p, "wb").write(bytes(data[:3000]))               # truncated
    with pytest.raises(vr.VrError):
        vr.load_nvdb_grid(p)


@pytest.mark.timeout(120)
def test_nvdb_loader_survives_corrupt_headers(vr, cloud, tmp_path):
    """Seeded fuzz of vr_grid_load_nvdb (file container and raw buffer): special 64-bit values (0, 2^32, 2^63, 2^64-1, ...) and
    random words written over the headers, byte flips, truncation. Every case must return -- VR_E_BADGRID or a grid
    that lies inside the file -- never hang, over-read or grow without bound (sizes near 2^64 used to wrap the bounds check)."""
    grid = cloud(32)
    p = str(tmp_path / "g.nvdb")
    vr.save_nvdb_grid(p, grid)
    file_bytes = open(p, "rb").read()
    vr.save_nvdb_grid(p, grid, raw=True)
    raw_bytes = open(p, "rb").read()
    rng = np.random.default_rng(7)
    specials = [0, 1, 31, 32, 2**31, 2**32 - 1, 2**32, 2**63, 2**64 - 1, 2**64 - 32, len(file_bytes), len(file_bytes) + 1]
    loaded = rejected = 0
    for it in range(600):
        data = bytearray(file_bytes if it % 2 == 0 else raw_bytes)
        mode = it % 5
        if mode == 0:
            struct.pack_into("<Q", data, int(rng.integers(0, 128)) * 8, specials[int(rng.integers(0, len(specials)))])
        elif mode == 1:
            for _ in range(int(rng.integers(1, 8))):
                data[int(rng.integers(0, min(len(data), 1400)))] = int(rng.integers(0, 256))
        elif mode == 2:
            data = data[:int(rng.integers(0, len(data)))]
        elif mode == 3:
            struct.pack_into("<I", data, int(rng.integers(0, 256)) * 4, int(rng.integers(0, 2**32)))
        else:
            struct.pack_into("<Q", data, int(rng.integers(0, 90)) * 8, specials[int(rng.integers(0, len(specials)))])
            data = data[:int(rng.integers(700, len(data)))]
        open(p, "wb").write(bytes(data))
        try:
            g = vr.load_nvdb_grid(p)
            assert g.nbytes <= len(data) and g.nbytes % 32 == 0, (it, g.nbytes, len(data))
            loaded += 1
        except vr.VrError as e:
            assert e.code in (1, 3), (it, e.code, str(e))   # VR_E_BADGRID; VR_E_ARG for an unreadable file
            rejected += 1
    assert loaded > 50 and rejected > 50


def test_ppm_and_pfm_writers(vr, tmp_path):
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (7, 5, 4), dtype=np.uint8)
    p = str(tmp_path / "a.ppm")
    vr.write_ppm(p, img)
    data = open(p, "rb").read()
    assert data.startswith(b"P6\n5 7\n255\n")            # CPU_test/main.cpp:130
    assert np.array_equal(np.frombuffer(data[len(b"P6\n5 7\n255\n"):], np.uint8).reshape(7, 5, 3), img[..., :3])
    acc = rng.random((7, 5, 4), dtype=np.float32)
    p = str(tmp_path / "a.pfm")
    vr.write_pfm(p, acc)
    data = open(p, "rb").read()
    head = b"PF\n5 7\n-1.0\n"
    assert data.startswith(head)
    assert np.array_equal(np.frombuffer(data[len(head):], "<f4").reshape(7, 5, 3)[::-1], acc[..., :3])
    vr.write_pfm(p, acc[..., 0])
    data = open(p, "rb").read()
    assert data.startswith(b"Pf\n5 7\n-1.0\n")


def test_headless_cli_without_gpu(vr, tmp_path, has_gpu):
    exe = vr.HEADLESS_PATH
    assert os.path.exists(exe), "vr_headless not built (run __graft_entry__.build())"
    out = subprocess.run([exe, "--help"], capture_output=True, text=True)
    assert out.returncode == 0 and "--lightgen" in out.stdout and "--snapshot-in" in out.stdout
    bad = subprocess.run([exe, "--algo"], capture_output=True, text=True)
    assert bad.returncode == 2
    # host-only leg: build a volume and write it as .nvdb happens before any CUDA call
    nv = str(tmp_path / "s.nvdb")
    run = subprocess.run([exe, "--grid", "cloud:16", "--save-nvdb", nv, "--frames", "1", "--size", "16x16", "--quiet"],
                         capture_output=True, text=True)
    assert os.path.exists(nv) and vr.load_nvdb_grid(nv).nbytes > 1000
    if not has_gpu:
        assert run.returncode == 1 and "no CUDA device" in run.stderr   # fails loudly: there is no CPU fallback
    cfg = tmp_path / "c.json"
    cfg.write_text('{"grid": "cloud:16", "algo": "nonsense", "frames": 1}')
    bad = subprocess.run([exe, "--config", str(cfg)], capture_output=True, text=True)
    assert bad.returncode != 0
