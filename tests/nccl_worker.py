"""One rank of the real-NCCL parity run (launched by tests/test_gpu_nccl.py through torch.distributed.run): the
distributed update + resample of the C library over NCCL ranks, one GPU each. Inputs and the oracle's expected outputs
come from the .npz the parent wrote; every rank checks its own window and exits non-zero on a mismatch."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from badger_amcl_b200 import cabi, sharded
    case = np.load(sys.argv[1])
    rank, local_rank, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local_rank)
    dist.init_process_group("gloo")  # host channel for the NCCL id only; the data path is the library's own NCCL
    ctx = cabi.Context(local_rank)
    poses = case["poses"]
    n = len(poses)
    pf = cabi.ParticleFilter(ctx, n // 10, n)
    comm = sharded.Comm.nccl(ctx, rank, world, dist)
    f = sharded.ShardedFilter(pf, comm)
    f.set_global_samples(np.concatenate([poses, np.full((n, 1), 1.0 / n)], axis=1))
    init = f.init_stats()
    assert (init.leaf_count, init.node_count, init.cluster_count) == tuple(int(v) for v in case["init"]), "init tree"
    f.srand48(7)
    steps = int(case["steps"])
    for s in range(steps):
        w = case["w%d" % s]
        off, cnt = f.window()
        pf.upload_weights(w[off:off + cnt])
        f.w_slow = f.w_fast = 1.0
        res = f.update_resample()
        off, cnt = f.window()
        sl = slice(off, off + cnt)
        assert res.sample_count == int(case["count%d" % s])
        assert np.array_equal(f.window_sources(), case["src%d" % s][sl]), "sources"
        assert np.array_equal(f.download_window(), case["samples%d" % s][sl]), "samples"
        assert np.array_equal(f.window_clusters(), case["labels%d" % s][sl]), "labels"
        assert (res.leaf_count, res.node_count, res.cluster_count) == tuple(int(v) for v in case["tree%d" % s]), "tree"
        assert np.allclose(np.array(list(res.mean) + list(res.cov)), case["stats%d" % s], rtol=1e-6, atol=1e-9)
        assert bool(res.converged) == bool(case["conv%d" % s])
    pf.close()
    comm.close()
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()
    print("rank %d ok" % rank, flush=True)


if __name__ == "__main__":
    main()
