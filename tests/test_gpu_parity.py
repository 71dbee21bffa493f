"""GPU parity suite (-m gpu): every CUDA entry point is called through the C ABI
(cobalt_smart_lender_ai_b200.ops / trainer -> ctypes -> libcobalt_gbdt.so) and compared with
the CPU oracle on the same seeded inputs.  Integer, byte and index results must be
bit-exact; probabilities within 1e-5 (they are in fact bit-equal); AUC within 1e-4."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from cobalt_smart_lender_ai_b200 import ops as _ops
    return _ops


@pytest.fixture(scope="module")
def data():
    from cobalt_smart_lender_ai_b200.synth import make_lending
    X, y, names = make_lending(60000, 60, seed=7)
    return X.numpy(), y.numpy(), names


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _edge_matrix(n=5003, seed=3):
    rng = np.random.default_rng(seed)
    X = np.zeros((n, 11), np.float32)
    X[:, 0] = rng.normal(size=n)
    X[:, 1] = np.nan                                   # empty column
    X[:, 2] = 3.25                                     # constant
    X[:, 3] = rng.integers(0, 2, n)                    # binary
    X[:, 4] = rng.integers(-5, 12, n)                  # few distinct, negatives
    X[:, 5] = np.where(rng.random(n) < 0.5, -0.0, 0.0) # signed zeros
    X[:, 6] = rng.lognormal(3, 2, n)
    X[:, 7] = np.round(rng.normal(size=n) * 40)        # ~250 distinct: around the 255 limit
    X[:, 8] = np.round(rng.normal(size=n) * 60)        # > 255 distinct with heavy ties
    X[:, 9] = rng.normal(size=n) * 1e30
    X[:, 10] = rng.integers(0, 300, n)                 # exactly 300 distinct
    X[rng.random(X.shape) < 0.07] = np.nan
    X[0, 0], X[1, 0], X[2, 6] = np.inf, -np.inf, np.inf
    return X


@pytest.mark.parametrize("max_bin", [256, 16, 2])
def test_sketch_cuts_bit_exact(ops, max_bin):
    X = _edge_matrix()
    cuts, nc, mv = O.compute_cuts(X, max_bin)
    c, k, m = ops.sketch_cuts(_dev(X), max_bin)
    assert np.array_equal(k.cpu().numpy(), nc)
    assert np.array_equal(m.cpu().numpy().view(np.uint32), mv.view(np.uint32))
    assert np.array_equal(c.cpu().numpy().view(np.uint32), cuts.view(np.uint32))


def test_sketch_cuts_schema(ops, data):
    X = data[0]
    cuts, nc, mv = O.compute_cuts(X, 256)
    c, k, m = ops.sketch_cuts(_dev(X), 256)
    assert np.array_equal(k.cpu().numpy(), nc)
    assert np.array_equal(c.cpu().numpy().view(np.uint32), cuts.view(np.uint32))
    assert np.array_equal(m.cpu().numpy().view(np.uint32), mv.view(np.uint32))


@pytest.mark.parametrize("n", [1, 31, 4097])
def test_sketch_ragged_sizes(ops, n):
    X = _edge_matrix(n=max(n, 8))[:n]
    cuts, nc, mv = O.compute_cuts(X, 256)
    c, k, m = ops.sketch_cuts(_dev(X), 256)
    assert np.array_equal(k.cpu().numpy(), nc)
    assert np.array_equal(c.cpu().numpy().view(np.uint32), cuts.view(np.uint32))


def test_bin_pack_bit_exact(ops, data):
    for X in (_edge_matrix(), data[0][:20011]):
        cuts, nc, _ = O.compute_cuts(X, 256)
        ldb = ops.padded_ldb(X.shape[1])
        want = O.bin_matrix(X, cuts, nc, ldb=ldb)
        got = ops.bin_pack(_dev(X), _dev(cuts), _dev(nc), ldb).cpu().numpy()
        assert np.array_equal(got, want)
        pos, _ = ops.plan_layout(nc, X.shape[1], ldb)          # permuted column placement
        got = ops.bin_pack(_dev(X), _dev(cuts), _dev(nc), ldb, _dev(pos)).cpu().numpy()
        placed = pos >= 0
        assert sorted(pos[placed].tolist()) == list(range(X.shape[1]))
        assert np.array_equal(got[:, placed], want[:, pos[placed]]) and np.all(got[:, ~placed] == 255)
        got_b = ops.bin_pack(_dev(X), _dev(cuts), _dev(nc), ldb, _dev(pos), blocked=True).cpu().numpy()
        assert np.array_equal(got_b, _block_major(got))


def test_sigmoid_and_grad_bit_exact(ops):
    rng = np.random.default_rng(0)
    m = np.concatenate([rng.normal(size=100000) * 6, [0, -0.0, 100, -100, 88.7, -87.1, 1e-30]]).astype(np.float32)
    assert np.array_equal(ops.sigmoid(_dev(m)).cpu().numpy().view(np.uint32), O.sigmoid(m).view(np.uint32))
    y = (rng.random(len(m)) < 0.13).astype(np.uint8)
    w = rng.random(len(m)).astype(np.float32) + 0.5
    for weight, sub in ((None, 1.0), (w, 1.0), (None, 0.8)):
        wmax = 1.0 if weight is None else float(weight.max())
        sg, sh = O.quant_shifts(6.75 * wmax)
        assert (sg, sh) == ops.quant_shifts(6.75 * wmax)
        want = O.grad(m, y, weight, 6.75, sg, sh, subsample=sub, seed=78, tree=3, row_offset=1000)
        gp, tot = ops.grad_logistic(_dev(m), _dev(y), None if weight is None else _dev(weight), 6.75, sg, sh,
                                    subsample=sub, seed=78, tree=3, row_offset=1000)
        assert np.array_equal(gp.cpu().numpy(), want)
        assert tot.cpu().numpy().tolist() == want.astype(np.int64).sum(0).tolist()
        assert np.abs(want).max() <= 2 ** 17


def _binned(ops, X):
    cuts, nc, mv = O.compute_cuts(X, 256)
    ldb = ops.padded_ldb(X.shape[1])
    bins = O.bin_matrix(X, cuts, nc, ldb=ldb)
    return cuts, nc, mv, bins, O.bin_ptr_of(nc)


def test_hist_build_bit_exact(ops, data):
    X, y, _ = data
    cuts, nc, mv, bins, bp = _binned(ops, X)
    rng = np.random.default_rng(1)
    f = X.shape[1]
    q = O.grad(rng.normal(size=len(y)).astype(np.float32), y, None, 6.75, 15, 17)
    bd, qd, bpd = _dev(bins), _dev(q), _dev(bp)
    S = int(bp[-1])
    # root (identity rows), then gathered subsets of ragged sizes incl. empty
    assert np.array_equal(ops.hist_build(bd, f, qd, None, bpd, S).cpu().numpy(), O.hist(bins, f, q, None, bp))
    for cnt in (0, 1, 33, 4096, 4097, 25001):
        rows = np.sort(rng.choice(len(y), cnt, replace=False)).astype(np.uint32)
        got = ops.hist_build(bd, f, qd, _dev(rows.view(np.int32)) if cnt else torch.zeros(0, dtype=torch.int32, device="cuda"),
                             bpd, S, count=cnt).cpu().numpy()
        assert np.array_equal(got, O.hist(bins, f, q, rows, bp)), cnt
    # additivity (K9): parent - built == sibling
    rows = rng.permutation(len(y)).astype(np.uint32)
    a, b = np.sort(rows[:20000]), np.sort(rows[20000:])
    ha = ops.hist_build(bd, f, qd, _dev(a.view(np.int32)), bpd, S)
    root = ops.hist_build(bd, f, qd, None, bpd, S)
    assert np.array_equal(ops.hist_subtract(root, ha).cpu().numpy(), O.hist(bins, f, q, b, bp))


def test_hist_many_bins_uses_feature_tiles(ops):
    """150 continuous features x 255 bins = 38 250 slots: more than one shared-memory tile."""
    rng = np.random.default_rng(2)
    X = rng.normal(size=(9000, 150)).astype(np.float32)
    X[rng.random(X.shape) < 0.02] = np.nan
    cuts, nc, mv, bins, bp = _binned(ops, X)
    assert bp[-1] > 30000
    y = (rng.random(9000) < 0.2).astype(np.uint8)
    q = O.grad(np.zeros(9000, np.float32), y, None, 4.0, 16, 18)
    got = ops.hist_build(_dev(bins), 150, _dev(q), None, _dev(bp), int(bp[-1])).cpu().numpy()
    assert np.array_equal(got, O.hist(bins, 150, q, None, bp))


def _tp(ops, **kw):
    from cobalt_smart_lender_ai_b200._lib import TreeParams
    d = dict(eta=0.3, reg_lambda=1.0, reg_alpha=0.0, gamma=0.0, min_child_weight=1.0, max_depth=6, sg=15, sh=17)
    d.update(kw)
    return TreeParams(np.float32(d["eta"]), d["reg_lambda"], d["reg_alpha"], d["gamma"], d["min_child_weight"],
                      d["max_depth"], d["sg"], d["sh"]), O.TreeParams(**d)


def _assert_split_equal(got, want):
    assert got["feature"] == want["feature"]
    if want["feature"] < 0:
        return
    for k in ("split_bin", "default_left", "GL", "HL", "GR", "HR"):
        assert got[k] == want[k], k
    assert np.float32(got["loss_chg"]).view(np.uint32) == np.float32(want["loss_chg"]).view(np.uint32)
    assert np.float32(got["threshold"]).view(np.uint32) == np.float32(want["threshold"]).view(np.uint32)


def test_split_scan_bit_exact(ops, data):
    X, y, _ = data
    cuts, nc, mv, bins, bp = _binned(ops, X)
    rng = np.random.default_rng(4)
    f = X.shape[1]
    q = O.grad(rng.normal(size=len(y)).astype(np.float32), y, None, 6.75, 15, 17)
    for trial in range(6):
        cnt = [len(y), 30000, 5000, 700, 60, 5][trial]
        rows = None if trial == 0 else np.sort(rng.choice(len(y), cnt, replace=False)).astype(np.uint32)
        h = O.hist(bins, f, q, rows, bp)
        tot = q.astype(np.int64).sum(0) if rows is None else q[rows].astype(np.int64).sum(0)
        for kw in (dict(), dict(reg_alpha=0.5, min_child_weight=3.0), dict(reg_lambda=0.0, min_child_weight=0.0)):
            tpc, tpo = _tp(ops, **kw)
            mask = None if trial % 2 == 0 else (rng.random(f) < 0.6).astype(np.uint8)
            want = O.scan(h, bp, cuts, mv, tot[0], tot[1], tpo, mask)
            got = ops.split_scan(_dev(h), _dev(bp), _dev(cuts), _dev(mv), _dev(tot), tpc,
                                 None if mask is None else _dev(mask))
            _assert_split_equal(got, want)


def test_split_scan_ties_and_missing(ops):
    bp = np.array([0, 3, 6], np.int32)
    cuts = np.full((2, 256), np.inf, np.float32)
    cuts[:, :3] = [1, 2, 3]
    mv = np.array([-1, -1], np.float32)
    tpc, tpo = _tp(ops, eta=1.0, min_child_weight=0.0, max_depth=3, sg=4, sh=4)
    cases = [
        (np.array([[-160, 160], [0, 160], [160, 160]] * 2, np.int64), (0, 480)),
        (np.array([[-160, 160], [160, 160], [0, 0], [0, 16], [0, 16], [0, 16]], np.int64), (-480, 640)),
        (np.zeros((6, 2), np.int64), (0, 0)),
    ]
    for h, tot in cases:
        t = np.array(tot, np.int64)
        _assert_split_equal(ops.split_scan(_dev(h), _dev(bp), _dev(cuts), _dev(mv), _dev(t), tpc),
                            O.scan(h, bp, cuts, mv, t[0], t[1], tpo))


@pytest.mark.parametrize("cnt", [0, 1, 31, 32, 33, 2047, 2048, 2049, 50021])
def test_partition_stable_bit_exact(ops, data, cnt):
    X = data[0]
    cuts, nc, mv, bins, bp = _binned(ops, X)
    rng = np.random.default_rng(cnt)
    seg = np.sort(rng.choice(len(X), cnt, replace=False)).astype(np.uint32)
    bd = _dev(bins)
    for feat, sb, dl in ((0, 100, 0), (1, 30, 1), (40, 0, 0), (33, -1, 1), (59, 0, 1)):
        want, nl = O.partition(bins, seg, feat, sb, dl)
        got, gl = ops.partition_stable(bd, _dev(seg.view(np.int32)) if cnt else torch.zeros(0, dtype=torch.int32, device="cuda"),
                                       feat, sb, dl)
        assert gl == nl
        assert np.array_equal(got.cpu().numpy().view(np.uint32), want)


def test_leaf_update_standalone(ops):
    """K7 as its own entry point: rows of a leaf's segment take the leaf value, one f32 add each (R10)."""
    rng = np.random.default_rng(4)
    m = rng.normal(size=5000).astype(np.float32)
    rows = rng.permutation(5000)[:1777].astype(np.uint32)
    leaf = np.float32(-0.123456789)
    md = _dev(m.copy())
    ops.leaf_update(md, _dev(rows.view(np.int32)), float(leaf))
    want = m.copy()
    want[rows] = want[rows] + leaf
    assert np.array_equal(md.cpu().numpy().view(np.uint32), want.view(np.uint32))
    ops.leaf_update(md, None, 0.5, count=100)                     # identity segment: rows 0..99
    want[:100] = want[:100] + np.float32(0.5)
    assert np.array_equal(md.cpu().numpy().view(np.uint32), want.view(np.uint32))
    ops.leaf_update(md, None, 1.0, count=0)                       # empty segment is a no-op
    assert np.array_equal(md.cpu().numpy().view(np.uint32), want.view(np.uint32))


def test_predict_q8_refuses_a_foreign_plane_base(ops, fixture_forest):
    """A forest laid out for another shared-memory window is refused with a status (no trap, the context survives)."""
    import ctypes as C

    from cobalt_smart_lender_ai_b200 import _lib
    chunk, base = ops.predict_q8_plan(20, 7)
    assert chunk >= 8
    lib = _lib.load()
    ldc = ops.q8_ldc(20)
    codes = torch.zeros((128, ldc), dtype=torch.uint8, device="cuda")
    forest = torch.zeros((chunk, 2 << 7), dtype=torch.int32, device="cuda")
    out = torch.zeros(128, dtype=torch.float32, device="cuda")
    rc = lib.cb_predict_q8(C.c_void_p(codes.data_ptr()), 128, 20, ldc, C.c_void_p(forest.data_ptr()), chunk, 7, 0.0,
                           base + 128, C.c_void_p(out.data_ptr()), None, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == -1 and b"plane_base" in lib.cb_last_error()
    torch.cuda.synchronize()                                      # still a healthy context
    assert float(out.sum()) == 0.0


def _block_major(b):
    """[n][ldb] row-major -> the library's column-block-major storage (same shape, different byte order)."""
    n, ldb = b.shape
    parts = [np.ascontiguousarray(b[:, k:k + 128]).reshape(-1) for k in range(0, ldb, 128)]
    return np.concatenate(parts).reshape(n, ldb)


def _train_gpu(ops, ref, X, y, n_trees, spw, subsample=1.0, seed=0, masks=None, weight=None, natural=False,
               with_cm=True, **kw):
    from cobalt_smart_lender_ai_b200.trainer import GpuTrainer
    tp = ref["tp"]
    pos, _cf = None, False
    if natural:
        bins = _dev(ref["bins"])
    else:  # the product's column placement applied to the oracle's bins
        pos_np, _cf = ops.plan_layout(ref["n_cuts"], X.shape[1], ref["bins"].shape[1])
        b = np.full_like(ref["bins"], 255)
        placed = pos_np >= 0
        b[:, placed] = ref["bins"][:, pos_np[placed]]
        bins, pos = _dev(_block_major(b) if _cf else b), _dev(pos_np)
    tr = GpuTrainer(bins, _dev(y), None if weight is None else _dev(weight), X.shape[1], _dev(ref["bin_ptr"]),
                    _dev(ref["cuts"]), _dev(ref["min_vals"]), max_depth=tp.max_depth, eta=tp.eta,
                    reg_lambda=tp.reg_lambda, reg_alpha=tp.reg_alpha, gamma=tp.gamma,
                    min_child_weight=tp.min_child_weight, scale_pos_weight=spw, sg=tp.sg, sh=tp.sh,
                    subsample=subsample, seed=seed, max_trees=n_trees, base_margin=O.base_margin_of(ref["base_score"]),
                    group=False, pos_feature=pos, blocked=_cf,
                    bins_cm=ops.bin_transpose(bins, blocked=_cf) if with_cm else None)
    tr.boost(n_trees, None if masks is None else _dev(masks))
    return tr


def _assert_forest_equal(got, want_flat):
    assert np.array_equal(got["tree_offsets"], want_flat["tree_offsets"])
    for k in ("left", "right", "parent", "split_index", "split_bin", "default_left"):
        assert np.array_equal(got[k], want_flat[k]), k
    for k in ("split_cond", "base_weight", "loss_chg", "sum_hess"):
        assert np.array_equal(got[k].view(np.uint32), want_flat[k].view(np.uint32)), k


def _final_rows(ref, tr, bins):
    """GPU rows are not moved at the last level (children are leaves): finish the stable
    partition of the last level on the host before comparing with the oracle's row order."""
    rows = tr.row_order()
    rows = np.arange(tr.n, dtype=np.uint32) if rows is None else rows.cpu().numpy().view(np.uint32).copy()
    fo = tr.forest_numpy()
    t = tr.n_trees - 1
    a = int(fo["tree_offsets"][t])
    D = tr.max_depth
    # walk the tree to find depth-(D-1) internal nodes and their segments
    L, R = fo["left"][a:], fo["right"][a:]
    segs = {0: (0, tr.n)}
    depth = {0: 0}
    order = [0]
    for nid in order:
        if L[nid] < 0:
            continue
        b, c = segs[nid]
        seg = rows[b:b + c]
        if depth[nid] == D - 1:
            out, nl = O.partition(bins, seg, int(fo["split_index"][a + nid]), int(fo["split_bin"][a + nid]),
                                  int(fo["default_left"][a + nid]))
            rows[b:b + c] = out
            continue
        code = bins[seg.astype(np.int64), int(fo["split_index"][a + nid])]
        left = np.where(code == 255, bool(fo["default_left"][a + nid]), code <= fo["split_bin"][a + nid])
        nl = int(left.sum())
        segs[L[nid]], segs[R[nid]] = (b, nl), (b + nl, c - nl)
        depth[L[nid]] = depth[R[nid]] = depth[nid] + 1
        order += [L[nid], R[nid]]
    return rows


def test_train_matches_committed_golden(ops, oracle_small):
    g = oracle_small
    X, y = g["X"], g["y"]
    ref = O.fit_reference(X, y, n_estimators=5, max_depth=4, scale_pos_weight=float(g["spw"]), base_score=0.5,
                          ldb=32, want_rows=True, want_hist=True)
    assert np.array_equal(ref["forest"]["split_index"], g["forest_split_index"])
    tr = _train_gpu(ops, ref, X, y, 5, float(g["spw"]))
    _assert_forest_equal(tr.forest_numpy(), ref["flat"])
    assert np.array_equal(tr.margin.cpu().numpy().view(np.uint32), g["margin"].view(np.uint32))
    assert np.array_equal(tr.node_hist(0).cpu().numpy(), g["forest_root_hist"])
    assert np.array_equal(_final_rows(ref, tr, g["bins"]), g["forest_row_order"])


@pytest.mark.parametrize("cfg", [
    dict(n_estimators=12, max_depth=6),
    dict(n_estimators=6, max_depth=3, learning_rate=0.05, gamma=5.0),
    dict(n_estimators=5, max_depth=8, min_child_weight=5.0, reg_lambda=0.5, reg_alpha=0.1),
    dict(n_estimators=6, max_depth=5, subsample=0.8, seed=78),
    dict(n_estimators=3, max_depth=1),
])
def test_train_bit_exact(ops, data, cfg):
    X, y, _ = data
    spw = float((y == 0).sum() / (y == 1).sum())
    ref = O.fit_reference(X, y, scale_pos_weight=spw, want_rows=True, want_hist=True, **cfg)
    tr = _train_gpu(ops, ref, X, y, cfg["n_estimators"], spw, subsample=cfg.get("subsample", 1.0), seed=cfg.get("seed", 0))
    _assert_forest_equal(tr.forest_numpy(), ref["flat"])
    assert np.array_equal(tr.margin.cpu().numpy().view(np.uint32), ref["margin"].view(np.uint32))
    assert np.array_equal(tr.node_hist(0).cpu().numpy(), ref["forest"]["root_hist"])
    assert np.array_equal(_final_rows(ref, tr, ref["bins"]), ref["forest"]["row_order"])


def test_train_natural_order_and_tiled_fallback(ops, data):
    """Natural column order (pos_feature = NULL) and a bin count that does not fit one
    shared-memory image (150 x 255 bins): both run the tiled generic histogram kernel."""
    X, y, _ = data
    X, y = X[:15000], y[:15000]
    ref = O.fit_reference(X, y, n_estimators=3, max_depth=4, scale_pos_weight=2.0)
    tr = _train_gpu(ops, ref, X, y, 3, 2.0, natural=True)
    _assert_forest_equal(tr.forest_numpy(), ref["flat"])
    rng = np.random.default_rng(2)
    Xw = rng.normal(size=(9000, 150)).astype(np.float32)
    Xw[rng.random(Xw.shape) < 0.02] = np.nan
    yw = (rng.random(9000) < 0.2).astype(np.uint8)
    ref = O.fit_reference(Xw, yw, n_estimators=2, max_depth=3, scale_pos_weight=4.0)
    assert ref["bin_ptr"][-1] > 30000
    assert not ops.plan_layout(ref["n_cuts"], 150)[1]
    tr = _train_gpu(ops, ref, Xw, yw, 2, 4.0)
    _assert_forest_equal(tr.forest_numpy(), ref["flat"])
    assert np.array_equal(tr.margin.cpu().numpy().view(np.uint32), ref["margin"].view(np.uint32))


def test_train_150_features_two_column_blocks(ops):
    """The BASELINE schema (150 columns -> two column blocks, row-slot copies folded, block-major storage)."""
    from cobalt_smart_lender_ai_b200.synth import make_lending
    X, y, _ = make_lending(30000, 150, seed=5)
    X, y = X.numpy(), y.numpy()
    spw = float((y == 0).sum() / (y == 1).sum())
    ref = O.fit_reference(X, y, n_estimators=4, max_depth=6, scale_pos_weight=spw, want_rows=True, want_hist=True)
    assert ops.plan_layout(ref["n_cuts"], 150)[1]
    tr = _train_gpu(ops, ref, X, y, 4, spw)
    _assert_forest_equal(tr.forest_numpy(), ref["flat"])
    assert np.array_equal(tr.margin.cpu().numpy().view(np.uint32), ref["margin"].view(np.uint32))
    assert np.array_equal(tr.node_hist(0).cpu().numpy(), ref["forest"]["root_hist"])
    tr2 = _train_gpu(ops, ref, X, y, 4, spw, with_cm=False)        # without the position-major copy
    _assert_forest_equal(tr2.forest_numpy(), ref["flat"])
    assert torch.equal(tr2.margin, tr.margin)
    assert np.array_equal(_final_rows(ref, tr, ref["bins"]), ref["forest"]["row_order"])


@pytest.mark.parametrize("n,f,depth", [(1, 1, 2), (2, 3, 3), (33, 5, 4), (257, 33, 5), (1000, 129, 9), (5000, 2, 6)])
def test_train_tiny_and_odd_shapes(ops, n, f, depth):
    """Ragged sizes: single rows, one feature, more than one column block with few rows, maximum depth."""
    rng = np.random.default_rng(n * 131 + f)
    X = rng.normal(size=(n, f)).astype(np.float32)
    X[:, ::2] = np.round(X[:, ::2] * 2)                       # low-cardinality columns
    X[rng.random(X.shape) < 0.1] = np.nan
    y = (rng.random(n) < 0.4).astype(np.uint8)
    if n > 1:
        y[0], y[1] = 0, 1
    ref = O.fit_reference(X, y, n_estimators=5, max_depth=depth, scale_pos_weight=1.5, min_child_weight=0.0, base_score=0.5)
    tr = _train_gpu(ops, ref, X, y, 5, 1.5)
    _assert_forest_equal(tr.forest_numpy(), ref["flat"])
    assert np.array_equal(tr.margin.cpu().numpy().view(np.uint32), ref["margin"].view(np.uint32))
    from cobalt_smart_lender_ai_b200.booster import Booster
    m = Booster(tr.forest_numpy(), f, None, 0.5).predict_device(_dev(X), output_margin=True).cpu().numpy()
    assert np.array_equal(m.view(np.uint32), ref["margin"].view(np.uint32))


def test_train_colsample_and_weights(ops, data):
    from cobalt_smart_lender_ai_b200.sklearn import colsample_masks
    X, y, _ = data
    X, y = X[:20000], y[:20000]
    rng = np.random.default_rng(9)
    w = (rng.random(len(y)) + 0.25).astype(np.float32)
    masks = colsample_masks(4, X.shape[1], 0.5, 78)
    ref = O.fit_reference(X, y, n_estimators=4, max_depth=5, scale_pos_weight=3.0, feature_masks=masks, sample_weight=w)
    tr = _train_gpu(ops, ref, X, y, 4, 3.0, masks=masks, weight=w)
    _assert_forest_equal(tr.forest_numpy(), ref["flat"])
    assert np.array_equal(tr.margin.cpu().numpy().view(np.uint32), ref["margin"].view(np.uint32))


def test_predict_fixture_forest(ops, fixture_forest):
    """The reference's shipped 300-tree model scored on the GPU vs the oracle (R11)."""
    from cobalt_smart_lender_ai_b200.booster import Booster
    ff = fixture_forest
    rng = np.random.default_rng(5)
    internal = ff["left"] != -1
    X = np.zeros((4099, 20), np.float32)
    for j in range(20):
        tj = ff["split_cond"][internal & (ff["split_index"] == j)]
        X[:, j] = rng.choice(tj, size=len(X)) + rng.choice([-1e-3, 0.0, 1e-3], size=len(X))
    X[rng.random(X.shape) < 0.1] = np.nan
    want = O.predict_margin(X, ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"], ff["split_cond"],
                            ff["default_left"], 0.0)
    b = Booster(ff, 20, [str(s) for s in ff["feature_names"]], float(ff["base_score"]))
    got = b.predict_device(_dev(X), output_margin=True).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    p = b.predict_device(_dev(X)).cpu().numpy()
    assert np.max(np.abs(p - O.sigmoid(want))) <= 1e-5
    # the pointer-chasing kernel (used for very deep trees) gives the same bits as the complete-tree kernel
    nodes, toff = b._device_forest(torch.device("cuda"))
    m2, _ = ops.predict(_dev(X), nodes, toff, 0.0, want_margin=True, want_proba=False)
    assert np.array_equal(m2.cpu().numpy().view(np.uint32), want.view(np.uint32))
    # ragged: 0 rows and 1 row
    assert b.predict_device(_dev(X[:1])).shape == (1,)
    assert b.predict_device(_dev(X[:0])).shape == (0,)


from _forests import random_forest as _random_forest  # noqa: E402


@pytest.mark.parametrize("n_rows,n_features,n_trees,depth", [(1, 3, 1, 1), (127, 7, 5, 3), (128, 150, 37, 6), (1000, 250, 19, 6),
                                                            (4133, 150, 70, 8), (513, 20, 300, 6)])
def test_predict_quantised_matches_float_kernels(ops, n_rows, n_features, n_trees, depth, monkeypatch):
    """K8 on quantised inputs (codes against the forest's own thresholds) = the f32 kernels = the oracle, bit for bit:
    values exactly on, just below and just above thresholds, +-inf, -0.0, NaN with either default direction."""
    from cobalt_smart_lender_ai_b200.booster import Booster
    rng = np.random.default_rng(n_rows + n_trees)
    fo, pools = _random_forest(rng, n_trees, n_features, depth, 40)
    X = np.zeros((n_rows, n_features), np.float32)
    for j in range(n_features):
        base = rng.choice(pools[j], size=n_rows)
        kind = rng.integers(0, 6, size=n_rows)
        X[:, j] = np.where(kind == 0, base, np.where(kind == 1, np.nextafter(base, np.float32(-np.inf)),
                           np.where(kind == 2, np.nextafter(base, np.float32(np.inf)), rng.normal(size=n_rows)))).astype(np.float32)
    X[rng.random(X.shape) < 0.08] = np.nan
    X[rng.random(X.shape) < 0.01] = np.inf
    X[rng.random(X.shape) < 0.01] = -np.inf
    X[rng.random(X.shape) < 0.01] = -0.0
    want = O.predict_margin(X, fo["tree_offsets"], fo["left"], fo["right"], fo["split_index"], fo["split_cond"],
                            fo["default_left"], 0.0)
    for mode in ("q8", "f32"):
        monkeypatch.setenv("COBALT_PREDICT", mode)
        b = Booster(fo, n_features, None, 0.5)
        got = b.predict_device(_dev(X), output_margin=True).cpu().numpy()
        assert (b._device_q8(torch.device("cuda")) is not None) == (mode == "q8")
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), mode


def test_predict_quantised_falls_back(ops):
    """More than 254 distinct thresholds on one feature, or too many features for the 16-bit offset field (about
    250): the f32 kernel scores instead."""
    from cobalt_smart_lender_ai_b200.booster import Booster
    rng = np.random.default_rng(9)
    fo, _ = _random_forest(rng, 400, 2, 6, 4000)
    X = rng.normal(size=(300, 2)).astype(np.float32)
    b = Booster(fo, 2, None, 0.5)
    assert b._device_q8(torch.device("cuda")) is None
    want = O.predict_margin(X, fo["tree_offsets"], fo["left"], fo["right"], fo["split_index"], fo["split_cond"],
                            fo["default_left"], 0.0)
    assert np.array_equal(b.predict_device(_dev(X), output_margin=True).cpu().numpy().view(np.uint32), want.view(np.uint32))
    fo2, _ = _random_forest(rng, 5, 300, 4, 10)
    X2 = rng.normal(size=(50, 300)).astype(np.float32)
    b2 = Booster(fo2, 300, None, 0.5)
    assert b2._device_q8(torch.device("cuda")) is None
    want2 = O.predict_margin(X2, fo2["tree_offsets"], fo2["left"], fo2["right"], fo2["split_index"], fo2["split_cond"],
                             fo2["default_left"], 0.0)
    assert np.array_equal(b2.predict_device(_dev(X2), output_margin=True).cpu().numpy().view(np.uint32), want2.view(np.uint32))


@pytest.mark.parametrize("n_rows,n_features,n_trees,depth", [(1, 3, 1, 1), (37, 6, 9, 4), (300, 40, 50, 6), (64, 150, 20, 9)])
def test_tree_shap_bit_exact(ops, n_rows, n_features, n_trees, depth):
    """K10 = the oracle's R13 bit for bit (ragged trees, repeated features on a path, NaN either way), and every
    row's attributions add up to its margin."""
    from cobalt_smart_lender_ai_b200.booster import Booster
    from cobalt_smart_lender_ai_b200.explain import TreeExplainer
    rng = np.random.default_rng(n_rows)
    fo, pools = _random_forest(rng, n_trees, n_features, depth, 6)
    X = rng.normal(size=(n_rows, n_features)).astype(np.float32)
    X[rng.random(X.shape) < 0.1] = np.nan
    want = O.tree_shap(X, fo["tree_offsets"], fo["left"], fo["right"], fo["split_index"], fo["split_cond"],
                       fo["default_left"], fo["sum_hess"], 0.0)
    b = Booster(fo, n_features, None, 0.5)
    got = b.predict_contribs_device(_dev(X)).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    margin = b.predict_device(_dev(X), output_margin=True).cpu().numpy()
    assert np.max(np.abs(got.astype(np.float64).sum(1) - margin)) <= 1e-4 * max(1.0, np.abs(margin).max())
    ex = TreeExplainer(b)
    assert np.array_equal(ex.shap_values(X[0]), want[:1, :-1]) and ex.expected_value == float(want[0, -1])


def test_tree_shap_fixture_forest(ops, fixture_forest):
    """The reference's shipped 300-tree model explained on the GPU: oracle bits, additivity to the margin."""
    from cobalt_smart_lender_ai_b200.booster import Booster
    ff = fixture_forest
    rng = np.random.default_rng(11)
    X = (rng.normal(size=(200, 20)) * 200 + 400).astype(np.float32)
    X[:, 12:] = rng.integers(0, 2, (200, 8))
    X[rng.random(X.shape) < 0.05] = np.nan
    b = Booster(ff, 20, [str(s) for s in ff["feature_names"]], float(ff["base_score"]))
    got = b.predict_contribs_device(_dev(X)).cpu().numpy()
    want = O.tree_shap(X, ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"], ff["split_cond"],
                       ff["default_left"], ff["sum_hess"], 0.0)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    margin = O.predict_margin(X, ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"], ff["split_cond"],
                              ff["default_left"], 0.0)
    assert np.max(np.abs(got.astype(np.float64).sum(1) - margin)) <= 2e-4


def test_predict_proba_streams_large_host_matrices(ops, fixture_forest, monkeypatch):
    """Host matrices above the streaming threshold are scored chunk by chunk behind their own H2D copies (pinned
    source: direct; pageable source: through pinned staging buffers): same bits as the one-shot path."""
    import cobalt_smart_lender_ai_b200.sklearn as sk
    ff = fixture_forest
    clf = sk.XGBClassifier().load_fixture_forest(ff, [str(s) for s in ff["feature_names"]], float(ff["base_score"]))
    rng = np.random.default_rng(2)
    X = (rng.normal(size=(10007, 20)) * 200 + 400).astype(np.float32)
    X[rng.random(X.shape) < 0.05] = np.nan
    want = clf.predict_proba(X)
    monkeypatch.setattr(sk, "_STREAM_BYTES", 20 * 4 * 1000)          # 1000-row chunks: 11 chunks, the last one ragged
    assert np.array_equal(clf.predict_proba(X), want)                  # pageable source
    Xp = torch.from_numpy(X).pin_memory()
    assert np.array_equal(clf.predict_proba(Xp.numpy()), want)         # pinned source


def test_estimator_end_to_end(ops, data):
    """fit/predict_proba/predict/feature_importances_/get_score/pickle through the drop-in class."""
    import pickle

    import pandas as pd
    from sklearn.metrics import roc_auc_score

    from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier
    X, y, names = data
    df = pd.DataFrame(X[:40000], columns=names)
    for c in names[-5:]:
        df[c] = df[c].astype(bool)               # get_dummies emits bool columns
    yt = pd.Series(y[:40000])
    spw = float((yt == 0).sum() / (yt == 1).sum())
    clf = XGBClassifier(eval_metric="logloss", use_label_encoder=False, scale_pos_weight=spw, random_state=42,
                        n_estimators=20)
    clf.fit(df, yt)
    ref = O.fit_reference(X[:40000], y[:40000], n_estimators=20, max_depth=6, learning_rate=0.3, scale_pos_weight=spw)
    _assert_forest_equal(clf.get_booster().forest | {"split_bin": ref["flat"]["split_bin"]}, ref["flat"])
    Xt, ytest = X[40000:], y[40000:]
    proba = clf.predict_proba(pd.DataFrame(Xt, columns=names))
    want = O.predict_proba_reference(ref, Xt)
    assert proba.shape == (len(Xt), 2) and proba.dtype == np.float32
    assert np.max(np.abs(proba[:, 1] - want)) <= 1e-5
    assert abs(roc_auc_score(ytest, proba[:, 1]) - roc_auc_score(ytest, want)) <= 1e-4
    assert np.array_equal(clf.predict(Xt), (want > 0.5).astype(np.int64))
    imp = clf.feature_importances_
    assert imp.shape == (60,) and abs(imp.sum() - 1.0) < 1e-5
    gain = clf.get_booster().get_score(importance_type="gain")
    fo = ref["flat"]
    internal = fo["left"] >= 0
    j = int(fo["split_index"][0])
    sel = internal & (fo["split_index"] == j)
    assert gain[names[j]] == pytest.approx(float(fo["loss_chg"][sel].astype(np.float64).mean()))
    clf2 = pickle.loads(pickle.dumps(clf))
    assert np.array_equal(clf2.predict_proba(Xt[:100]), proba[:100])
    with pytest.raises(ValueError):
        clf.predict_proba(pd.DataFrame(Xt[:5], columns=names[::-1]))


def _train_grouped(ref, X, y, n_trees, spw, masks=None, **kw):
    """The product's own preparation: K2 on the GPU with one-hot groups packed as one code byte each."""
    from cobalt_smart_lender_ai_b200.binned import BinnedMatrix
    from cobalt_smart_lender_ai_b200.trainer import GpuTrainer
    tp = ref["tp"]
    bm = BinnedMatrix.from_dense(_dev(X), _dev(ref["cuts"]), _dev(ref["n_cuts"]), **kw)
    tr = GpuTrainer(bm, _dev(y), None, X.shape[1], _dev(ref["bin_ptr"]), _dev(ref["cuts"]), _dev(ref["min_vals"]),
                    max_depth=tp.max_depth, eta=tp.eta, gamma=tp.gamma, scale_pos_weight=spw, sg=tp.sg, sh=tp.sh,
                    max_trees=n_trees, base_margin=O.base_margin_of(ref["base_score"]), group=False)
    tr.boost(n_trees, None if masks is None else _dev(masks))
    return tr, bm


def test_onehot_groups_one_code_byte_bit_exact(ops):
    """SURVEY 8f-2 (the part on the hot path): the engineered schema's one-hot dummies are found on the data, a group
    shares one code byte, 150 columns fit one 128-byte line -- and codes, root histogram, every split, the row order
    and the margins still equal the oracle's, which reads one byte per feature."""
    from cobalt_smart_lender_ai_b200.synth import make_lending
    X, y, names = make_lending(30000, 150, seed=5)
    X, y = X.numpy(), y.numpy()
    spw = float((y == 0).sum() / (y == 1).sum())
    ref = O.fit_reference(X, y, n_estimators=4, max_depth=6, scale_pos_weight=spw, want_rows=True, want_hist=True)
    tr, bm = _train_grouped(ref, X, y, 4, spw)
    assert bm.n_groups >= 6 and bm.ldb == 128 and not bm.blocked and bm.cf
    grouped = {ft for g in bm.groups for ft in g}
    assert all(ref["n_cuts"][ft] == 2 for ft in grouped) and len(grouped) >= 30
    assert np.array_equal(bm.natural_codes(), ref["bins"][:, :150])
    _assert_forest_equal(tr.forest_numpy(), ref["flat"])
    assert np.array_equal(tr.margin.cpu().numpy().view(np.uint32), ref["margin"].view(np.uint32))
    assert np.array_equal(tr.node_hist(0).cpu().numpy(), ref["forest"]["root_hist"])
    assert np.array_equal(_final_rows(ref, tr, ref["bins"]), ref["forest"]["row_order"])
    # the same fit with the groups switched off gives the same forest (and uses two column blocks)
    tr2, bm2 = _train_grouped(ref, X, y, 4, spw, group_onehot=False)
    assert bm2.n_groups == 0 and bm2.ldb == 160
    _assert_forest_equal(tr2.forest_numpy(), ref["flat"])
    # per-tree column masks (colsample_bytree / RFE) address member features one by one
    masks = (np.random.default_rng(3).random((3, 150)) < 0.6).astype(np.uint8)
    ref3 = O.fit_reference(X, y, n_estimators=3, max_depth=5, scale_pos_weight=spw, feature_masks=masks)
    tr3, _ = _train_grouped(ref3, X, y, 3, spw, masks=masks)
    _assert_forest_equal(tr3.forest_numpy(), ref3["flat"])


@pytest.mark.parametrize("group_onehot", [True, False])
def test_conflict_free_histogram_kernel_standalone(ops, group_onehot):
    """The kernel that is 60 % of a boosting round, called on its own through the C ABI with the row lists a trainer
    would hand it -- all rows (root form), a gathered ragged subset, a tiny subset, one row, a subset larger than one
    int32 chunk per CTA -- against the oracle's histogram, with and without one-hot group columns (one and two column
    blocks)."""
    from cobalt_smart_lender_ai_b200.binned import BinnedMatrix
    from cobalt_smart_lender_ai_b200.synth import make_lending
    n, f = 60000, 150
    X, y, _ = make_lending(n, f, seed=17)
    X, y = X.numpy(), y.numpy()
    cuts, nc, mv = O.compute_cuts(X, 256)
    bins = O.bin_matrix(X, cuts, nc)
    bp = O.bin_ptr_of(nc)
    rng = np.random.default_rng(3)
    q = O.grad(rng.normal(size=n).astype(np.float32), y, None, 6.75, 14, 16)
    bm = BinnedMatrix.from_dense(_dev(X), _dev(cuts), _dev(nc), group_onehot=group_onehot)
    assert bm.cf and (bm.n_groups > 0) == group_onehot
    qd, bpd, S = _dev(q), _dev(bp), int(bp[-1])
    for rows in (None, np.sort(rng.permutation(n)[:23456]), np.sort(rng.permutation(n)[:300]), np.array([n - 1]),
                 rng.permutation(n)[:40001]):                    # the last one unsorted: order must not matter
        rd = None if rows is None else _dev(rows.astype(np.uint32).view(np.int32))
        got = ops.hist_build_cf(bm, qd, rd, bpd, S).cpu().numpy()
        want = O.hist(bins, f, q, None if rows is None else rows.astype(np.uint32), bp)
        assert np.array_equal(got, want), None if rows is None else len(rows)
    # more rows per CTA than one int32 partial may hold (16 352): 148 CTAs x 2 chunks needs > 2.4 M rows -- repeat rows
    big = np.tile(np.arange(n, dtype=np.uint32), 45)[:2_500_000]
    got = ops.hist_build_cf(bm, qd, _dev(big.view(np.int32)), bpd, S).cpu().numpy()
    assert np.array_equal(got, O.hist(bins, f, q, big, bp))


def test_onehot_detection_is_exact():
    """Grouping is decided on the data: one row where two indicators are both set, or one missing value, keeps a
    column out of every group; splits on grouped members (forced by the data) send exactly the member's rows right."""
    from cobalt_smart_lender_ai_b200 import ops as _ops
    from cobalt_smart_lender_ai_b200.binned import BinnedMatrix, detect_onehot_groups
    rng = np.random.default_rng(11)
    n = 20000
    cat = rng.integers(0, 5, n)                                   # 5 levels -> 4 dummies (drop_first)
    cat[11] = 1
    X = np.zeros((n, 9), np.float32)
    for m in range(4):
        X[:, m] = cat == m + 1
    X[:, 4] = rng.random(n) < 0.3                                 # independent indicator: co-occurs with the dummies
    X[:, 5] = (cat == 0) & (rng.random(n) < 0.5)                  # exclusive with the dummies by construction
    X[:, 6] = X[:, 5]; X[7, 6] = np.nan                           # ... but has ONE missing value
    X[:, 7] = X[:, 5]; X[11, 7] = 1.0                             # ... but is set together with dummy 0 ONCE (row 11)
    X[:, 8] = rng.normal(size=n)
    y = ((cat == 2) | (X[:, 8] > 1.0)).astype(np.uint8)
    cuts, nc, mv = O.compute_cuts(X, 256)
    groups = detect_onehot_groups(_dev(X), _dev(cuts), _dev(nc))
    assert groups == [[0, 1, 2, 3, 5]], groups
    ref = O.fit_reference(X, y, n_estimators=3, max_depth=4, scale_pos_weight=1.0, want_rows=True)
    tr, bm = _train_grouped(ref, X, y, 3, 1.0)
    assert bm.groups == groups and np.array_equal(bm.natural_codes(), ref["bins"][:, :9])
    assert 1 in ref["flat"]["split_index"][ref["flat"]["left"] >= 0]          # a member of the group is split on
    _assert_forest_equal(tr.forest_numpy(), ref["flat"])
    assert np.array_equal(tr.margin.cpu().numpy().view(np.uint32), ref["margin"].view(np.uint32))
    assert np.array_equal(_final_rows(ref, tr, ref["bins"]), ref["forest"]["row_order"])


def test_full_size_properties(ops):
    """BASELINE size (2.2M x 150): properties that need no oracle run -- histogram additivity
    (parent = left + right, exact), node totals, margin replay through the scoring kernel."""
    from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier
    from cobalt_smart_lender_ai_b200.synth import make_lending
    from cobalt_smart_lender_ai_b200.trainer import GpuTrainer
    n, f = 2_200_000, 150
    X, y, _ = make_lending(n, f, device="cuda")
    from cobalt_smart_lender_ai_b200.binned import BinnedMatrix
    cuts, n_cuts, min_vals = ops.sketch_cuts(X, 256)
    bm = BinnedMatrix.from_dense(X, cuts, n_cuts)
    assert bm.cf and bm.n_groups > 0 and bm.ldb == 128
    bp = ops.bin_ptr_of(n_cuts)
    spw = float(((y == 0).sum() / (y == 1).sum()).item())
    sg, sh = ops.quant_shifts(spw)
    tr = GpuTrainer(bm, y, None, f, bp, cuts, min_vals, max_depth=6, eta=0.3, scale_pos_weight=spw, sg=sg, sh=sh,
                    max_trees=3, group=False)
    tr.boost(3)
    fo = tr.forest_numpy()
    a = int(fo["tree_offsets"][2])
    L, R = fo["left"][a:], fo["right"][a:]
    checked = 0
    for nid in range(len(L)):
        if L[nid] >= 0 and R[nid] < 63:          # children with a histogram slot (depth < 6)
            hp, hl, hr = tr.node_hist(nid), tr.node_hist(int(L[nid])), tr.node_hist(int(R[nid]))
            assert torch.equal(hp, hl + hr)
            checked += 1
    assert checked >= 20
    # every feature's bins + its missing rows sum to the node total: check the dense last one-hot column
    root = tr.node_hist(0)
    b0, b1 = int(bp[f - 1]), int(bp[f])
    gp, tot = ops.grad_logistic(torch.zeros(1, device="cuda"), y[:1], None, spw, sg, sh)  # smoke of the standalone call
    assert root[b0:b1].sum(0).tolist() == root[int(bp[f - 2]):b0].sum(0).tolist()
    # margin replay: scoring kernel on raw floats == training-time margins on bins (bit-exact)
    from cobalt_smart_lender_ai_b200.booster import Booster
    m = Booster(fo, f, None, 0.5).predict_device(X, output_margin=True)
    assert torch.equal(m.view(torch.int32), tr.margin.view(torch.int32))
    # the sorted-ness property of the stable partition: every leaf segment is ascending in row id
    rows = tr.row_order().cpu().numpy().view(np.uint32)
    assert np.array_equal(np.sort(rows), np.arange(n, dtype=np.uint32))


def test_c3_size_properties(ops):
    """BASELINE configs[3] size on one GPU's share and beyond (20M x 150, depth 8): the oracle cannot run this in
    seconds, so the size-independent properties stand in -- exact histogram additivity down to depth 8, margin replay
    through the scoring kernel bit for bit, the row order a permutation with ascending leaf segments."""
    from cobalt_smart_lender_ai_b200.binned import BinnedMatrix
    from cobalt_smart_lender_ai_b200.booster import Booster
    from cobalt_smart_lender_ai_b200.synth import make_lending
    from cobalt_smart_lender_ai_b200.trainer import GpuTrainer
    free, _total = torch.cuda.mem_get_info()
    if free < 90e9:
        pytest.skip("needs ~90 GB of free device memory")
    n, f, D = 20_000_000, 150, 8
    X = torch.empty((n, f), dtype=torch.float32, device="cuda")
    y = torch.empty(n, dtype=torch.uint8, device="cuda")
    for b in range(0, n, 2_000_000):
        Xc, yc, _ = make_lending(2_000_000, f, row_offset=b, device="cuda")
        X[b:b + 2_000_000], y[b:b + 2_000_000] = Xc, yc
    del Xc, yc
    cuts, n_cuts, min_vals = ops.sketch_cuts(X, 256)
    bm = BinnedMatrix.from_dense(X, cuts, n_cuts)
    bp = ops.bin_ptr_of(n_cuts)
    spw = float(((y == 0).sum() / (y == 1).sum()).item())
    sg, sh = ops.quant_shifts(spw)
    tr = GpuTrainer(bm, y, None, f, bp, cuts, min_vals, max_depth=D, eta=0.3, scale_pos_weight=spw, sg=sg, sh=sh,
                    max_trees=2, group=False)
    tr.boost(2)
    fo = tr.forest_numpy()
    a = int(fo["tree_offsets"][1])
    L, R = fo["left"][a:], fo["right"][a:]
    checked = 0
    for nid in range(len(L)):
        if L[nid] >= 0 and R[nid] < (1 << D) - 1:                 # children with a histogram slot (depth < 8)
            assert torch.equal(tr.node_hist(nid), tr.node_hist(int(L[nid])) + tr.node_hist(int(R[nid])))
            checked += 1
    assert checked >= 100
    m = Booster(fo, f, None, 0.5).predict_device(X, output_margin=True)
    assert torch.equal(m.view(torch.int32), tr.margin.view(torch.int32))
    rows = tr.row_order().cpu().numpy().view(np.uint32)
    assert np.array_equal(np.sort(rows), np.arange(n, dtype=np.uint32))
    tr.close()


def test_scoring_service_on_gpu(ops, fixture_forest, tmp_path):
    """src/api path end to end: the reference's shipped forest served through the FastAPI mirror
    (/predict_bulk_csv -> predict_proba_df -> cb_predict_heap), model written and re-read as a .ubj file."""
    import pandas as pd
    from fastapi.testclient import TestClient

    from cobalt_smart_lender_ai_b200 import ubjson
    from cobalt_smart_lender_ai_b200.api import cobalt_fast_api as api
    from cobalt_smart_lender_ai_b200.booster import Booster
    ff = fixture_forest
    names = [str(s) for s in ff["feature_names"]]
    path = tmp_path / "model.ubj"
    ubjson.save_model(Booster(ff, 20, names, float(ff["base_score"])), str(path))
    api.xgb_model = api.load_model(path)
    api.explainer = api.TreeExplainer(api.xgb_model)       # what lifespan() does at start-up (reference :45-46)
    rng = np.random.default_rng(3)
    X = (rng.normal(size=(50, 20)) * 200 + 400).astype(np.float32)
    X[:, 12:] = rng.integers(0, 2, (50, 8))
    df = pd.DataFrame(X.astype(np.float64), columns=names)
    want = O.sigmoid(O.predict_margin(X, ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"],
                                      ff["split_cond"], ff["default_left"], 0.0))
    assert np.max(np.abs(api.predict_proba_df(df) - want)) <= 1e-5
    client = TestClient(api.app)
    r = client.post("/predict_bulk_csv", files={"file": ("x.csv", df.to_csv(index=False), "text/csv")})
    assert r.status_code == 200
    got = np.array([rec["prob_default"] for rec in r.json()["predictions"]])
    assert np.max(np.abs(got - want)) <= 1e-5
    row = {n: float(df.iloc[0][n]) for n in names}
    r = client.post("/predict", json=row)
    assert r.status_code == 200 and abs(r.json()["prob_default"] - want[0]) <= 1e-5
    phi = O.tree_shap(X[:1], ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"], ff["split_cond"],
                      ff["default_left"], ff["sum_hess"], 0.0)
    assert np.array_equal(np.array(r.json()["shap_values"], np.float32).view(np.uint32), phi[0, :20].view(np.uint32))
    assert np.float32(r.json()["base_value"]) == phi[0, 20]
    r = client.post("/feature_importance_bulk", json={"data": [row]})
    assert r.json()["top_features"][0]["feature"] in names


def test_rfe_bins_once_and_matches_sklearn_rfe(ops, data):
    """the elimination loop on resident bins (one sketch + binning, a feature mask per step) selects and
    ranks exactly like sklearn's RFE refitting the same estimator on every column subset."""
    import time

    import pandas as pd
    from sklearn.feature_selection import RFE as SkRFE

    from cobalt_smart_lender_ai_b200.feature_selection import RFE
    from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier
    X, y, names = data
    df = pd.DataFrame(X[:20000, :40], columns=names[:40])
    yt = pd.Series(y[:20000])
    base = XGBClassifier(eval_metric="logloss", use_label_encoder=False, scale_pos_weight=3.0, random_state=42,
                         n_estimators=12, max_depth=4)
    t0 = time.perf_counter()
    fast = RFE(estimator=base, n_features_to_select=20, step=1).fit(df, yt)
    t1 = time.perf_counter()
    slow = SkRFE(estimator=base, n_features_to_select=20, step=1).fit(df, yt)
    t2 = time.perf_counter()
    assert fast.binned_once_ and fast.fits_ == 20
    assert np.array_equal(fast.support_, slow.support_) and np.array_equal(fast.ranking_, slow.ranking_)
    assert fast.n_features_ == 20
    Xs = df.iloc[:64]
    assert np.array_equal(fast.predict_proba(Xs), slow.predict_proba(Xs))
    assert list(fast.get_feature_names_out()) == list(slow.get_feature_names_out())
    print(f"RFE 40 -> 20 on 20000 rows: resident bins {t1 - t0:.2f} s, sklearn refits {t2 - t1:.2f} s")
    # step > 1 and a fractional target go through the same bookkeeping
    f2 = RFE(estimator=base, n_features_to_select=0.25, step=3).fit(df, yt)
    s2 = SkRFE(estimator=base, n_features_to_select=0.25, step=3).fit(df, yt)
    assert np.array_equal(f2.ranking_, s2.ranking_)
    # colsample_bytree < 1 is delegated to sklearn's RFE
    f3 = RFE(estimator=XGBClassifier(n_estimators=3, max_depth=2, colsample_bytree=0.5), n_features_to_select=35).fit(df, yt)
    assert f3.binned_once_ is False and f3.support_.sum() == 35


def test_randomized_search_on_prepared_folds_matches_sklearn(ops, data):
    """the search with every fold uploaded, sketched and binned once gives sklearn's numbers exactly --
    same candidates, same per-fold AUCs, same winner."""
    import time

    import pandas as pd
    from sklearn.model_selection import RandomizedSearchCV as SkSearch
    from sklearn.model_selection import StratifiedKFold

    from cobalt_smart_lender_ai_b200.model_selection import RandomizedSearchCV
    from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier
    X, y, names = data
    df = pd.DataFrame(X[:30000, :25], columns=names[:25])
    yt = pd.Series(y[:30000])
    base = XGBClassifier(eval_metric="logloss", use_label_encoder=False, scale_pos_weight=4.0, random_state=78)
    dist = {"n_estimators": [10, 20], "max_depth": [3, 5], "learning_rate": [0.05, 0.1], "subsample": [0.8, 1.0],
            "colsample_bytree": [0.5, 1.0], "gamma": [0, 1]}
    kw = dict(param_distributions=dist, n_iter=6, scoring="roc_auc", cv=StratifiedKFold(3), random_state=22)
    t0 = time.perf_counter()
    fast = RandomizedSearchCV(estimator=base, **kw).fit(df, yt)
    t1 = time.perf_counter()
    slow = SkSearch(estimator=base, n_jobs=1, **kw).fit(df, yt)
    t2 = time.perf_counter()
    assert fast.prepared_folds_ == 3
    assert fast.cv_results_["params"] == slow.cv_results_["params"]
    for k in ("mean_test_score", "split0_test_score", "split1_test_score", "split2_test_score", "rank_test_score"):
        assert np.array_equal(np.asarray(fast.cv_results_[k]), np.asarray(slow.cv_results_[k])), k
    assert fast.best_params_ == slow.best_params_ and fast.best_score_ == slow.best_score_
    assert np.array_equal(fast.best_estimator_.predict_proba(df.iloc[:64]), slow.best_estimator_.predict_proba(df.iloc[:64]))
    print(f"search 6 x 3 on 30000 rows: prepared folds {t1 - t0:.2f} s, sklearn {t2 - t1:.2f} s")
    # a candidate that changes what the preparation depends on gets its own prepared fold
    f2 = RandomizedSearchCV(estimator=base, param_distributions={"max_bin": [16, 64]}, n_iter=2, scoring="roc_auc",
                            cv=StratifiedKFold(2), random_state=1).fit(df, yt)
    s2 = SkSearch(estimator=base, param_distributions={"max_bin": [16, 64]}, n_iter=2, scoring="roc_auc",
                  cv=StratifiedKFold(2), random_state=1, n_jobs=1).fit(df, yt)
    assert f2.prepared_folds_ == 4 and np.array_equal(f2.cv_results_["mean_test_score"], s2.cv_results_["mean_test_score"])


def test_reference_training_job_flow(ops):
    """The reference's training script, steps 2-7 (leakage drop, split, RFE to 20 features,
    RandomizedSearchCV, test metrics, joblib dump/load), on synthetic loans -- with this package's
    XGBClassifier as the only change (model_train_test mirror)."""
    import io

    import joblib
    import pandas as pd

    from cobalt_smart_lender_ai_b200.model_train_test.model_tree_train_test import train_pipeline
    from cobalt_smart_lender_ai_b200.synth import make_lending
    X, y, names = make_lending(12000, 26, seed=21)
    df = pd.DataFrame(X.numpy(), columns=names)
    for c in names[-4:]:
        df[c] = df[c].astype(bool)
    df["loan_default"] = y.numpy().astype(int)
    df["total_pymnt"] = 1.0                                       # a leakage column that must be dropped
    out = train_pipeline(df, n_features_to_select=20, n_iter=3, cv=3,
                         param_distributions={"n_estimators": [20, 30], "max_depth": [3, 5], "learning_rate": [0.1],
                                              "subsample": [0.8, 1.0], "colsample_bytree": [0.5, 1.0], "gamma": [0, 1]})
    assert len(out["selected_features"]) == 20 and "total_pymnt" not in out["selected_features"]
    assert 0.6 < out["metrics"]["auc"] <= 1.0 and 0.6 < out["cv_best_score"] <= 1.0
    assert abs(sum(out["feature_importances"].values()) - 1.0) < 1e-4
    buf = io.BytesIO()
    joblib.dump(out["model"], buf)
    buf.seek(0)
    again = joblib.load(buf)
    Xs = df[out["selected_features"]].iloc[:50]
    assert np.array_equal(again.predict_proba(Xs), out["model"].predict_proba(Xs))
