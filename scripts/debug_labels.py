import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from badger_amcl_b200 import cabi, synth
from oracle import oracle as O
RES = 0.05
ctx = cabi.Context(0)
cells = synth.make_grid(2000, 2000, seed=7)
truth = synth.pick_truth_pose(cells, RES, seed=5)
n = 4000
poses = synth.gaussian_poses(truth, n, seed=3)
rnd = synth.uniform_free_poses(cells, RES, n, seed=9)
rng = np.random.default_rng(5)
w = rng.random(n); w /= w.sum()
port = O.PortPF(500, n)
port.init_with_poses(poses)
aos = np.concatenate([poses, w[:, None]], axis=1)
port.set_samples(aos); port.set_averages(0.010, 0.007); port.set_random_poses(rnd); port.srand48(99)
pf = cabi.ParticleFilter(ctx, 500, n)
pf.init_with_poses(poses); pf.upload_samples(aos)
pf.w_slow, pf.w_fast = 0.010, 0.007
pf.set_random_poses(rnd); pf.srand48(99)
port.update_resample(); r = pf.update_resample()
g, wn = pf.sample_clusters(), port.sample_clusters()
bad = np.nonzero(g != wn)[0]
print("n", len(g), "bad", len(bad), "C", r.cluster_count, port.cluster_count())
s = pf.samples()
cs = np.array([0.5, 0.5, 10 * np.pi / 180])
keys = np.floor(s[:, :3] / cs).astype(int)
for i in bad[:40]:
    print(i, g[i], wn[i], keys[i], s[i, :3])
np.savez("gpurun_out/labels.npz", g=g, w=wn, s=s)
