"""The distributed filter over REAL NCCL ranks (one process per GPU, torch.distributed.run), against the oracle's single
set: bit-identical indices / poses / labels / KLD counts, statistics to 1e-6. Skipped on a box with fewer than two GPUs
(tests/test_gpu_sharded.py runs the same library path there over the in-process communicator)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from badger_amcl_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RES = 0.05


def _gpu_count():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=30).stdout
        return sum(1 for ln in out.splitlines() if ln.startswith("GPU "))
    except Exception:
        return 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("n,cloud", [(200000, "global"), (200000, "tracking")])
def test_nccl_ranks_reproduce_the_oracles_single_set(oracle_built, tmp_path, n, cloud):
    world = min(_gpu_count(), 8)
    if world < 2:
        pytest.skip("needs at least two GPUs")
    O = oracle_built
    cells = synth.make_grid(1000, 1000, seed=7)
    truth = synth.pick_truth_pose(cells, RES, seed=5)
    poses = (synth.uniform_free_poses(cells, RES, n, seed=1) if cloud == "global" else synth.gaussian_poses(truth, n, seed=2))
    rng = np.random.default_rng(n)
    port = O.PortPF(n // 10, n)
    port.init_with_poses(poses)
    case = dict(poses=poses, steps=2, init=np.array([port.leaf_count(), port.node_count(), port.cluster_count()]))
    port.srand48(7)
    for s in range(2):
        m = port.sample_count()
        w = np.exp(rng.normal(0.0, 2.0, m))
        w /= w.sum()
        smp = port.samples()
        smp[:, 3] = w
        port.set_samples(smp)
        port.set_averages(1.0, 1.0)
        port.update_resample()
        case.update({"w%d" % s: w, "count%d" % s: port.sample_count(), "src%d" % s: port.resample_sources(),
                     "samples%d" % s: port.samples(), "labels%d" % s: port.sample_clusters(),
                     "tree%d" % s: np.array([port.leaf_count(), port.node_count(), port.cluster_count()]),
                     "stats%d" % s: port.set_stats(), "conv%d" % s: port.is_converged()})
    path = str(tmp_path / "case.npz")
    np.savez(path, **case)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(ROOT, "tests", "nccl_worker.py"), path]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count(" ok") == world
