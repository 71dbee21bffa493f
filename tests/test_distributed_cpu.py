"""CPU suite, world_size 2 over gloo: the host-side glue of the row-sharded path
(label statistics, row offsets, shard arithmetic).  The kernels themselves need a GPU."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.distributed.init_process_group("gloo", rank=rank, world_size=world)
    from cobalt_smart_lender_ai_b200 import parallel
    from cobalt_smart_lender_ai_b200.synth import make_lending
    n = 1001
    b, e = parallel.shard_rows(n, rank, world)
    X, y, _ = make_lending(e - b, 12, seed=9, row_offset=b)
    pos = float((y == 1).sum()) * 3.0
    neg = float((y == 0).sum())
    gp, gn, wmax, off = parallel.global_stats(pos, neg, 1.0 + rank, e - b)
    Xf, yf, _ = make_lending(n, 12, seed=9)
    ok = (off == b and wmax == float(world) and gp == float((yf == 1).sum()) * 3.0 and gn == float((yf == 0).sum())
          and torch.equal(torch.nan_to_num(Xf[b:e]), torch.nan_to_num(X)) and torch.equal(yf[b:e], y))
    # an int64 histogram-like buffer summed over ranks is identical whatever the shard count (order-free)
    h = torch.arange(64, dtype=torch.int64) * (rank + 1) - 7
    torch.distributed.all_reduce(h)
    ok = ok and torch.equal(h, torch.arange(64, dtype=torch.int64) * sum(range(1, world + 1)) - 7 * world)
    out[rank] = bool(ok)
    torch.distributed.destroy_process_group()


def test_global_stats_and_sharding_world2():
    world, port = 2, _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}


def _sketch_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.distributed.init_process_group("gloo", rank=rank, world_size=world)
    from cobalt_smart_lender_ai_b200 import parallel
    from cobalt_smart_lender_ai_b200.synth import make_lending
    from oracle import oracle as O
    ok = True
    for n, f in ((1501, 12), (400, 2)):                       # second case: fewer features than ranks at world 3
        Xf, _, _ = make_lending(n, f, seed=4)
        b, e = parallel.shard_rows(n, rank, world)
        seen = []

        def cpu_sketch(X, max_bin):                           # stands in for K1 (needs a GPU); records what it was given
            seen.append(tuple(X.shape))
            return tuple(torch.from_numpy(a) for a in O.compute_cuts(X.numpy(), max_bin))
        cuts, n_cuts, min_vals = parallel.sketch_cuts_global(Xf[b:e].clone(), 256, sketch=cpu_sketch)
        want = O.compute_cuts(Xf.numpy(), 256)
        ok &= np.array_equal(cuts.numpy().view(np.uint32), want[0].view(np.uint32))
        ok &= np.array_equal(n_cuts.numpy(), want[1]) and np.array_equal(min_vals.numpy().view(np.uint32), want[2].view(np.uint32))
        f0, f1 = parallel.feature_slices(f, world)[rank]
        ok &= seen == ([(n, f1 - f0)] if f1 > f0 else [])       # all rows, own features only
    out[rank] = bool(ok)
    torch.distributed.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_sketch_moves_columns_not_the_matrix(world):
    """Multi-GPU K1 host logic: rows sharded in, each rank sorts ALL rows of ITS feature slice only, cuts bit-equal to
    the single-process result (the oracle's R1 stands in for the GPU sketch here)."""
    port = _free_port()
    out = mp.Manager().dict()
    mp.spawn(_sketch_worker, args=(world, port, out), nprocs=world, join=True)
    assert dict(out) == {r: True for r in range(world)}
