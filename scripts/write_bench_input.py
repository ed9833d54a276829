"""Input file of the C++ drivers (badger_amcl_b200/host/test/test_host_api bench, bench_dist): header, cells, scan, poses of
the bench workload (4000^2 map, 1080 beams, P global-uniform particles). usage: write_bench_input.py <out.bin> <P>"""
import os
import struct
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from badger_amcl_b200 import synth  # noqa: E402

out, P = sys.argv[1], int(sys.argv[2])
cells = synth.make_grid(4000, 4000, seed=7)
truth = synth.pick_truth_pose(cells, 0.05, seed=5)
ranges, angles = synth.make_scan(cells, 0.05, truth, n_beams=1080)
poses = synth.uniform_free_poses(cells, 0.05, P, seed=1)
with open(out, "wb") as f:
    f.write(struct.pack("<9i", 0x414D434C, cells.shape[1], cells.shape[0], P, len(ranges), P // 10, 1, 1, 12345))
    f.write(struct.pack("<3d", 0.05, 2.0, 30.0))
    f.write(np.ascontiguousarray(cells, np.int8).tobytes())
    f.write(np.ascontiguousarray(ranges, np.float64).tobytes())
    f.write(np.ascontiguousarray(angles, np.float64).tobytes())
    f.write(np.ascontiguousarray(poses[:, :3], np.float64).tobytes())
