#!/usr/bin/env python3
"""Golden vectors for the opt-in trilinear density: nanovdb::math::createSampler<1>(accessor)(xyz) of
the VENDORED NanoVDB (oracle/_ref, built where /root/reference exists) on the two committed NanoVDB-built grids.
Writes tests/golden/trilinear_<grid>_{xyz,values}.npy. Run in the container that has the reference checkout."""
import gzip
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    ref = pyoracle.Ref()
    rng = np.random.default_rng(20261104)
    for name, lo, hi in (("nanovdb_cloud32", -3.0, 35.0), ("nanovdb_ball80", -3.0, 83.0)):
        with gzip.open(os.path.join(GOLD, name + ".nvdb.raw.gz"), "rb") as f:
            grid = np.frombuffer(f.read(), np.uint8).copy()
        xyz = np.concatenate([rng.uniform(lo, hi, (6000, 3)), rng.integers(int(lo), int(hi), (500, 3)).astype(np.float64),
                              rng.uniform(-300.0, 300.0, (200, 3))]).astype(np.float32)
        np.save(os.path.join(GOLD, f"trilinear_{name}_xyz.npy"), xyz)
        vals = ref.trilinear(grid, xyz)
        np.save(os.path.join(GOLD, f"trilinear_{name}_values.npy"), vals)
        print(name, len(xyz), "positions,", int((vals != 0).sum()), "non-zero, max", float(vals.max()))


if __name__ == "__main__":
    main()
