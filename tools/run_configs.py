#!/usr/bin/env python3
"""Command-line front end of bench_configs.py: the BASELINE.json configurations next to the CPU port, one JSON line each.

    python tools/run_configs.py [C1 C3 C4A C5S C5P] > gpurun_out/configs.jsonl

bench.py reports the same entries in its `configs` array; this prints them without the C2 headline run.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench_configs
import volumerenderer_b200 as vr


def main():
    names = sys.argv[1:] or bench_configs.ORDER
    for res in bench_configs.run_all(vr, names, lambda *a: print(*a, file=sys.stderr, flush=True)):
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
