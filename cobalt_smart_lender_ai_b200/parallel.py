"""Row-sharded data parallelism: the host-side glue around the one collective of the path.

Rows are split contiguously over ranks (one process per GPU).  Replicated: cuts, parameters,
the tree being grown.  Exchanged: label statistics (4 scalars, once), the feature columns for
the exact quantile sketch (once, re-sharded by feature: no rank ever holds other ranks' rows of features it does not sort), and per tree level the int64 histogram buffer
(``GpuTrainer.boost_one``).  Integer sums make every exchange order independent, so split
decisions are bit-identical for any world size."""
from __future__ import annotations

import torch

from . import ops


_LOCAL_ONLY = 0


class local_only:
    """Inside this context a process ignores its process group: fits use this GPU's rows alone and exchange nothing.
    It is how a job made of many independent fits (the tuning step's fold x candidate grid, reference
    src/model_train_test/model_tree_train_test.py:148-159) puts ONE whole fit on each GPU instead of sharding every
    fit over all of them."""

    def __enter__(self):
        global _LOCAL_ONLY
        _LOCAL_ONLY += 1
        return self

    def __exit__(self, *a):
        global _LOCAL_ONLY
        _LOCAL_ONLY -= 1


def _dist():
    d = torch.distributed
    if _LOCAL_ONLY:
        return None
    return d if d.is_available() and d.is_initialized() and d.get_world_size() > 1 else None


def world_size() -> int:
    d = _dist()
    return d.get_world_size() if d is not None else 1


def shard_rows(n_total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous shard [begin, end) of rank; the first n_total % world ranks get one extra row."""
    base, rem = divmod(n_total, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def feature_slices(n_features: int, world: int) -> list[tuple[int, int]]:
    return [shard_rows(n_features, r, world) for r in range(world)]


def global_stats(pos_w: float, neg_w: float, wmax: float, n_local: int, device=None):
    """Sums the weighted label counts over ranks, takes the max weight, and returns this rank's
    global row offset.  Works on any backend (gloo on CPU tensors, nccl on CUDA tensors)."""
    d = _dist()
    if d is None:
        return pos_w, neg_w, wmax, 0
    dev = device if device is not None else "cpu"
    s = torch.tensor([pos_w, neg_w], dtype=torch.float64, device=dev)
    d.all_reduce(s)
    m = torch.tensor([wmax], dtype=torch.float64, device=dev)
    d.all_reduce(m, op=d.ReduceOp.MAX)
    counts = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(d.get_world_size())]
    d.all_gather(counts, torch.tensor([n_local], dtype=torch.int64, device=dev))
    offset = int(sum(int(c.item()) for c in counts[: d.get_rank()]))
    return float(s[0].item()), float(s[1].item()), float(m[0].item()), offset


def exchange_columns(Xd: torch.Tensor, slices, sizes) -> torch.Tensor | None:
    """Row shards in, column shards out: every rank ends up with ALL rows (in global row order) of ITS feature slice,
    f32 [N, f1 - f0].  Point-to-point: rank k sends columns [f0_r, f1_r) of its rows to owner r and receives the other
    ranks' rows of its own slice straight into place; each rank sends its shard exactly once in total and holds
    N x F / world values afterwards -- rows of features it does not sort never arrive.  Works on gloo and nccl."""
    d = torch.distributed
    world, rank = d.get_world_size(), d.get_rank()
    f0, f1 = slices[rank]
    offs = [0]
    for k in sizes:
        offs.append(offs[-1] + k)
    mine = None
    ops_, keep = [], []
    if f1 > f0:
        mine = torch.empty((offs[-1], f1 - f0), dtype=Xd.dtype, device=Xd.device)
        mine[offs[rank]:offs[rank + 1]] = Xd[:, f0:f1]
        ops_ += [d.P2POp(d.irecv, mine[offs[k]:offs[k + 1]], k) for k in range(world) if k != rank and sizes[k] > 0]
    if Xd.shape[0] > 0:
        for r in range(world):
            g0, g1 = slices[r]
            if r != rank and g1 > g0:                 # more ranks than features: such an owner has nothing to sort
                keep.append(Xd[:, g0:g1].contiguous())
                ops_.append(d.P2POp(d.isend, keep[-1], r))
    if ops_:
        for w in d.batch_isend_irecv(ops_):
            w.wait()
    return mine


def sketch_cuts_global(Xd: torch.Tensor, max_bin: int, sketch=None):
    """Exact R1 cuts over ALL ranks' rows, identical on every rank.  The exact sketch sorts whole columns, so the
    columns are re-sharded by feature (exchange_columns: each rank receives every row of its F / world features and
    nothing else), each rank runs K1 on its slice, and the small cut arrays are summed into place on every rank."""
    d = _dist()
    sketch = sketch or ops.sketch_cuts
    if d is None:
        return sketch(Xd, max_bin)
    world, rank = d.get_world_size(), d.get_rank()
    n_loc = torch.tensor([Xd.shape[0]], dtype=torch.int64, device=Xd.device)
    ns = [torch.zeros_like(n_loc) for _ in range(world)]
    d.all_gather(ns, n_loc)
    ns = [int(t.item()) for t in ns]
    f = Xd.shape[1]
    sl = feature_slices(f, world)
    f0, f1 = sl[rank]
    full = exchange_columns(Xd, sl, ns)
    cuts = torch.full((f, 256), float("inf"), dtype=torch.float32, device=Xd.device)
    n_cuts = torch.zeros(f, dtype=torch.int32, device=Xd.device)
    min_vals = torch.zeros(f, dtype=torch.float32, device=Xd.device)
    if f1 > f0:
        c, k, m = sketch(full, max_bin)
        cuts[f0:f1], n_cuts[f0:f1], min_vals[f0:f1] = c, k, m
    del full
    # every rank owns a disjoint feature slice; a SUM/MAX over zero-initialised slices assembles the whole
    cuts_bits = torch.where(torch.arange(f, device=Xd.device)[:, None].ge(f0) & torch.arange(f, device=Xd.device)[:, None].lt(f1),
                            cuts.view(torch.int32), torch.zeros_like(cuts.view(torch.int32)))
    d.all_reduce(cuts_bits)
    d.all_reduce(n_cuts)
    mv_bits = torch.where((torch.arange(f, device=Xd.device) >= f0) & (torch.arange(f, device=Xd.device) < f1),
                          min_vals.view(torch.int32), torch.zeros(f, dtype=torch.int32, device=Xd.device))
    d.all_reduce(mv_bits)
    return cuts_bits.view(torch.float32), n_cuts, mv_bits.view(torch.float32)
