"""Property tests (hypothesis) of the invariants the row-sharded design rests on:
fixed-point histograms are additive over ANY split of the rows -- parent = left + right, and the sum over 1, 2, 4 or 8
row shards is the same bytes -- the partition is a stable permutation, and the bin mapping is monotone.  The CPU
versions run against the oracle; the GPU versions push the same properties through the C ABI."""
import numpy as np
import pytest
from hypothesis import given, settings
from hypothesis import strategies as st

from oracle import oracle as O


def _case(seed, n, f):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, f)).astype(np.float32)
    X[:, 0] = rng.integers(0, 3, n)                          # a low-cardinality column
    X[rng.random((n, f)) < 0.05] = np.nan
    cuts, nc, mv = O.compute_cuts(X, 256)
    bins = O.bin_matrix(X, cuts, nc)
    q = np.stack([rng.integers(-(1 << 17), 1 << 17, n), rng.integers(0, 1 << 17, n)], 1).astype(np.int32)
    return X, cuts, nc, bins, O.bin_ptr_of(nc), q


@settings(max_examples=25, deadline=None)
@given(seed=st.integers(0, 2 ** 31 - 1), n=st.integers(1, 3000), f=st.integers(1, 9), world=st.sampled_from([1, 2, 4, 8]))
def test_histogram_is_additive_over_row_shards_and_children(seed, n, f, world):
    X, cuts, nc, bins, bp, q = _case(seed, n, f)
    whole = O.hist(bins, f, q, None, bp)
    # contiguous row shards, as parallel.shard_rows deals them: the int64 sum over shards is the same bytes
    from cobalt_smart_lender_ai_b200.parallel import shard_rows
    acc = np.zeros_like(whole)
    for r in range(world):
        b, e = shard_rows(n, r, world)
        if e > b:
            acc += O.hist(bins, f, q, np.arange(b, e, dtype=np.uint32), bp)
    assert np.array_equal(acc, whole)
    # any split of a node's rows into two children: parent = left + right, so sibling = parent - built is exact
    rng = np.random.default_rng(seed ^ 0x5bd1e995)
    rows = rng.permutation(n)[: rng.integers(1, n + 1)].astype(np.uint32)
    go_left = rng.random(len(rows)) < 0.3
    hp = O.hist(bins, f, q, rows, bp)
    hl = O.hist(bins, f, q, rows[go_left], bp) if go_left.any() else np.zeros_like(hp)
    hr = O.hist(bins, f, q, rows[~go_left], bp) if (~go_left).any() else np.zeros_like(hp)
    assert np.array_equal(hl + hr, hp)


@settings(max_examples=25, deadline=None)
@given(seed=st.integers(0, 2 ** 31 - 1), n=st.integers(1, 3000), f=st.integers(1, 6))
def test_partition_is_a_stable_permutation_and_bins_are_monotone(seed, n, f):
    X, cuts, nc, bins, bp, q = _case(seed, n, f)
    rng = np.random.default_rng(seed + 1)
    seg = np.sort(rng.permutation(n)[: rng.integers(1, n + 1)]).astype(np.uint32)
    feat = int(rng.integers(0, f))
    sb = int(rng.integers(-1, nc[feat]))
    dl = int(rng.integers(0, 2))
    out, nl = O.partition(bins, seg, feat, sb, dl)
    assert sorted(out.tolist()) == seg.tolist()                                  # a permutation of the segment
    assert np.all(np.diff(out[:nl].astype(np.int64)) > 0) and np.all(np.diff(out[nl:].astype(np.int64)) > 0)   # stable
    code = bins[out.astype(np.int64), feat]
    left = np.where(code == 255, bool(dl), code <= sb)
    assert left[:nl].all() and not left[nl:].any()
    # R2: x <= y  =>  bin(x) <= bin(y), and bin(x) <= k  <=>  x < cuts[k]
    col = X[:, feat]
    ok = ~np.isnan(col)
    order = np.argsort(col[ok], kind="stable")
    assert np.all(np.diff(bins[ok, feat][order].astype(np.int64)) >= 0)
    k = int(rng.integers(0, nc[feat]))
    assert np.array_equal(bins[ok, feat] <= k, (col[ok] < cuts[feat, k]) | (k == nc[feat] - 1))


@pytest.mark.gpu
@settings(max_examples=12, deadline=None)
@given(seed=st.integers(0, 2 ** 31 - 1), n=st.integers(1, 40000), f=st.integers(1, 40), world=st.sampled_from([1, 2, 4, 8]))
def test_gpu_histogram_shard_invariance(seed, n, f, world):
    """cb_hist_build on every row shard, summed as int64, equals the oracle's histogram of all rows (and therefore the
    all-reduced level buffer of a sharded run does not depend on the world size)."""
    import torch

    from cobalt_smart_lender_ai_b200 import ops
    from cobalt_smart_lender_ai_b200.parallel import shard_rows
    X, cuts, nc, bins, bp, q = _case(seed, n, f)
    ldb = (f + 15) // 16 * 16
    bins = O.bin_matrix(X, cuts, nc, ldb=ldb)
    bd, qd, bpd = (torch.from_numpy(np.ascontiguousarray(a)).cuda() for a in (bins, q, bp))
    acc = torch.zeros((int(bp[-1]), 2), dtype=torch.int64, device="cuda")
    for r in range(world):
        b, e = shard_rows(n, r, world)
        if e > b:
            rows = torch.arange(b, e, dtype=torch.int32, device="cuda")
            acc += ops.hist_build(bd, f, qd, rows, bpd, int(bp[-1]))
    assert np.array_equal(acc.cpu().numpy(), O.hist(bins, f, q, None, bp))
