#!/usr/bin/env python
"""Headless counterpart of python_src/play_simulation.py: parses a `.output` file
with the reader logic of play_simulation.py:45-101 (no pygame / OpenGL) and
reports what a replay would show.

  python tools/check_output.py -i test.output [--bytes 8] [--compare other.output]
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from particle_simulation_b200 import formats  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("-i", "--input", required=True, help="path to input particle file")
ap.add_argument("--bytes", type=int, default=4, choices=(4, 8), help="record length (4 = real32 as shipped)")
ap.add_argument("--compare", help="second .output file: report the largest position difference per frame")
args = ap.parse_args()
area, declared, frames = formats.read_output(args.input, args.bytes)
print("simulation area:", area)
print("Number of blocks: %d (header says %d)" % (len(frames), declared))
ok = declared == len(frames)
for k, (ids, pos, vel) in enumerate(frames, 1):
    inside = ((pos >= 0) & (pos <= np.array(area))).all(1).mean() if len(ids) else 1.0
    print("%d. block: %d records, %d particles, %.1f%% inside the box, |v| max %.4g" % (
        k, 7 * len(ids), len(ids), 100 * inside, np.abs(vel).max() if len(ids) else 0.0))
    ok = ok and np.isfinite(pos).all() and np.isfinite(vel).all()
if args.compare:
    _, _, other = formats.read_output(args.compare, args.bytes)
    for k, (a, b) in enumerate(zip(frames, other), 1):
        if len(a[0]) != len(b[0]):
            print("%d. block: particle counts differ (%d vs %d)" % (k, len(a[0]), len(b[0])))
            ok = False
        else:
            print("%d. block: max |dx| = %.3g, max |dv| = %.3g" % (k, np.abs(a[1] - b[1]).max(), np.abs(a[2] - b[2]).max()))
sys.exit(0 if ok else 1)
