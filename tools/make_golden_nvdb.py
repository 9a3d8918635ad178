"""tests/golden/nanovdb_ball80.nvdb.gz: the known-answer grid of SURVEY Appendix D written as an .nvdb FILE by the
vendored NanoVDB (nanovdb::io::writeGrid, Codec::NONE, through oracle/_ref). Needs the reference checkout; run here,
the fixture is committed. tools/make_golden.py makes the raw buffer this starts from."""
import gzip
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from oracle import pyoracle

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    ref = pyoracle.Ref()
    raw = np.frombuffer(gzip.open(os.path.join(OUT, "nanovdb_ball80.nvdb.raw.gz")).read(), np.uint8).copy()
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "ball80.nvdb")
        ref.write_nvdb(raw, p)
        data = open(p, "rb").read()
    with gzip.GzipFile(os.path.join(OUT, "nanovdb_ball80.nvdb.gz"), "wb", mtime=0) as f:
        f.write(data)
    print("wrote nanovdb_ball80.nvdb.gz:", len(data), "bytes uncompressed; header", data[:8])


if __name__ == "__main__":
    main()
