"""Test-only helpers for the distributed filter: G "ranks" as threads of one process on one GPU (each with its own
context, filter and an in-process communicator, bamcl_comm_create_local), so a single-GPU box runs the whole multi-rank
path of the C library. Nothing here ships."""
import threading


def run_ranks(world, fn):
    """fn(rank, group) on `world` threads (group = the local communicator group); returns the per-rank results,
    re-raising the first failure"""
    from badger_amcl_b200 import sharded
    group = sharded.Comm.local_group(world)
    out, err = [None] * world, [None] * world

    def body(r):
        try:
            out[r] = fn(r, group)
        except BaseException as e:  # noqa: BLE001
            err[r] = e

    ts = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    sharded.Comm.free_local_group(group)
    first = [e for e in err if e is not None and "peer rank failed" not in str(e)] or [e for e in err if e is not None]
    if first:
        raise first[0]
    return out
