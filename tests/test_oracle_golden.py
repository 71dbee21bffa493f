"""CPU suite: pins the oracle (oracle/gbdt_oracle.c) to the only golden artifact the
reference ships -- src/api/models/xgb_model_tree.pkl, decoded into
tests/golden/fixture_forest.npz by tests/golden/make_golden.py -- and to its own
committed outputs (oracle_small.npz)."""
import numpy as np
import pytest

from oracle import oracle as O


def _per_tree(ff):
    offs = ff["tree_offsets"]
    for t in range(len(offs) - 1):
        yield t, int(offs[t]), int(offs[t + 1])


def _unscaled_G(ff):
    """G = -w (H + lambda) with w the un-scaled weight: base_weight for internal nodes,
    leaf / eta for leaves (the file stores eta-scaled leaves)."""
    lam, eta = float(ff["reg_lambda"]), float(ff["eta"])
    is_leaf = ff["left"] == -1
    w = ff["base_weight"].astype(np.float64)
    w = np.where(is_leaf, w / eta, w)
    H = ff["sum_hess"].astype(np.float64)
    return -w * (H + lam), H


def test_fixture_shape(fixture_forest):
    ff = fixture_forest
    assert list(ff["version"]) == [3, 0, 0]
    assert len(ff["tree_offsets"]) == 301 and ff["tree_offsets"][-1] == 59216
    assert int(ff["max_depth"]) == 7 and int(ff["max_bin"]) == 256
    assert len(ff["feature_names"]) == 20
    assert (ff["left"] == -1).sum() == 29758


def test_gain_formula_known_answers(fixture_forest):
    """R6: loss_chg = G_L^2/(H_L+l) + G_R^2/(H_R+l) - G^2/(H+l), no 1/2 factor (29 458 tuples)."""
    ff = fixture_forest
    G, H = _unscaled_G(ff)
    lam = float(ff["reg_lambda"])
    rel = []
    for _t, a, b in _per_tree(ff):
        L, R = ff["left"][a:b], ff["right"][a:b]
        for k in np.nonzero(L != -1)[0]:
            l, r = a + L[k], a + R[k]
            got = (O.calc_gain(G[l], H[l], lam) + O.calc_gain(G[r], H[r], lam) - O.calc_gain(G[a + k], H[a + k], lam))
            rel.append(abs(got - ff["loss_chg"][a + k]) / max(1e-12, abs(ff["loss_chg"][a + k])))
    rel = np.array(rel)
    assert len(rel) == 29458
    assert np.median(rel) < 2e-6 and np.quantile(rel, 0.99) < 1e-3


def test_leaf_formula_known_answers(fixture_forest):
    """R10: leaf = f32(eta * -(G_parent - G_sibling)/(H + lambda)) on leaves whose sibling is internal."""
    ff = fixture_forest
    G, H = _unscaled_G(ff)
    lam, eta = float(ff["reg_lambda"]), float(ff["eta"])
    rel = []
    for _t, a, b in _per_tree(ff):
        L, R, P = ff["left"][a:b], ff["right"][a:b], ff["parent"][a:b]
        for k in range(1, b - a):
            if L[k] != -1:
                continue
            p = P[k]
            sib = R[p] if L[p] == k else L[p]
            if L[sib] == -1:
                continue
            g = G[a + p] - G[a + sib]
            got = eta * O.calc_weight(g, H[a + k], lam)
            want = ff["split_cond"][a + k]
            if abs(want) > 1e-4:
                rel.append(abs(got - want) / abs(want))
    rel = np.array(rel)
    assert len(rel) > 1000
    assert np.median(rel) < 1e-5 and np.quantile(rel, 0.99) < 2e-3
    leaves = ff["left"] == -1
    assert np.array_equal(ff["base_weight"][leaves].view(np.uint32), ff["split_cond"][leaves].view(np.uint32))


def test_hessian_additivity_and_validity(fixture_forest):
    """R5 additivity parent = L + R; R7: loss_chg >= gamma, child hessian >= min_child_weight."""
    ff = fixture_forest
    internal = np.nonzero(ff["left"] != -1)[0]
    tree_of = np.searchsorted(ff["tree_offsets"], internal, side="right") - 1
    base = ff["tree_offsets"][tree_of]
    l, r = base + ff["left"][internal], base + ff["right"][internal]
    H = ff["sum_hess"].astype(np.float64)
    assert np.max(np.abs(H[l] + H[r] - H[internal]) / H[internal]) < 1e-6
    assert ff["loss_chg"][internal].min() >= float(ff["gamma"])
    assert min(H[l].min(), H[r].min()) >= float(ff["min_child_weight"])


def test_gain_formula_every_node_within_input_rounding(fixture_forest):
    """R6 on ALL 29 458 internal nodes, no quantile: the stored loss_chg is reproduced from the stored (f32) weights and
    hessians to within the rounding noise of those inputs, 3.5e-7 of the three gain terms' magnitude.  A factor 1/2, a
    lambda left out of a denominator or gamma subtracted from the gain would be off by O(1) of that scale."""
    ff = fixture_forest
    G, H = _unscaled_G(ff)
    lam = float(ff["reg_lambda"])
    internal = np.nonzero(ff["left"] != -1)[0]
    base = ff["tree_offsets"][np.searchsorted(ff["tree_offsets"], internal, side="right") - 1]
    l, r = base + ff["left"][internal], base + ff["right"][internal]
    gain = lambda g, h: g * g / (h + lam)  # noqa: E731
    gl, gr, gp = gain(G[l], H[l]), gain(G[r], H[r]), gain(G[internal], H[internal])
    err = np.abs(gl + gr - gp - ff["loss_chg"][internal].astype(np.float64))
    assert len(err) == 29458 and np.max(err / (gl + gr + gp)) < 3.5e-7
    for wrong in (0.5 * (gl + gr - gp), gain(G[l], H[l] - lam) + gr - gp):
        assert np.median(np.abs(wrong - ff["loss_chg"][internal]) / (gl + gr + gp)) > 1e-4


def test_validity_boundaries_are_inclusive(fixture_forest):
    """R7 strictness as far as the shipped model shows it: a child whose hessian sum EQUALS min_child_weight (1.0) was
    accepted, and every accepted loss_chg is >= gamma = 5 (smallest stored: 5.0035); the oracle's scan accepts
    H == min_child_weight and rejects anything below."""
    ff = fixture_forest
    internal = np.nonzero(ff["left"] != -1)[0]
    base = ff["tree_offsets"][np.searchsorted(ff["tree_offsets"], internal, side="right") - 1]
    kids = np.concatenate([base + ff["left"][internal], base + ff["right"][internal]])
    assert ff["sum_hess"][kids].min() == np.float32(1.0) == np.float32(ff["min_child_weight"])
    assert ff["loss_chg"][internal].min() >= float(ff["gamma"]) > 0
    bp = np.array([0, 2], np.int32)
    cuts = np.full((1, 256), np.inf, np.float32)
    cuts[0, :2] = [1, 2]
    mv = np.array([-1], np.float32)
    tp = O.TreeParams(eta=1.0, reg_lambda=1.0, min_child_weight=1.0, max_depth=3, sg=4, sh=4)
    h = np.array([[-64, 16], [640, 1600]], np.int64)             # left child: H = 16 * 2^-4 = 1.0 exactly
    s = O.scan(h, bp, cuts, mv, 576, 1616, tp)
    assert s["feature"] == 0 and s["split_bin"] == 0 and s["HL"] == 16
    h[0, 1] = 15                                                   # H = 0.9375 < min_child_weight: no valid split
    assert O.scan(h, bp, cuts, mv, 576, 1615, tp)["feature"] == -1


def test_bfs_numbering(fixture_forest):
    """R9: children are allocated consecutively (right = left + 1) in node-id order."""
    ff = fixture_forest
    for _t, a, b in _per_tree(ff):
        L, R = ff["left"][a:b], ff["right"][a:b]
        nxt = 1
        for k in range(b - a):
            if L[k] != -1:
                assert L[k] == nxt and R[k] == nxt + 1
                nxt += 2
        assert nxt == b - a


def _py_traverse(ff, x):
    m = np.float32(0.0)
    for _t, a, _b in _per_tree(ff):
        k = 0
        while ff["left"][a + k] != -1:
            v = x[ff["split_index"][a + k]]
            if np.isnan(v):
                left = bool(ff["default_left"][a + k])
            else:
                left = bool(np.float32(v) < ff["split_cond"][a + k])
            k = ff["left"][a + k] if left else ff["right"][a + k]
        m = np.float32(m + ff["split_cond"][a + k])
    return m


def test_traversal_matches_plain_python(fixture_forest):
    """R11 on the shipped forest: x < threshold goes left, NaN follows default_left."""
    ff = fixture_forest
    rng = np.random.default_rng(5)
    thr = ff["split_cond"][ff["left"] != -1]
    feat = ff["split_index"][ff["left"] != -1]
    X = np.zeros((24, 20), np.float32)
    for j in range(20):
        tj = thr[feat == j]
        X[:, j] = rng.choice(tj, size=24) + rng.choice([-1e-3, 0.0, 1e-3], size=24)  # on and around thresholds
    X[rng.random(X.shape) < 0.15] = np.nan
    got = O.predict_margin(X, ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"], ff["split_cond"],
                           ff["default_left"], 0.0)
    want = np.array([_py_traverse(ff, X[i]) for i in range(len(X))], np.float32)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_cut_scheme_evidence(fixture_forest):
    """R1 evidence (v): small-cardinality features split on observed values and on the
    last-cut sentinel max + (|max| + 1e-5); dummies split only at 1.0; <= 254 thresholds."""
    ff = fixture_forest
    internal = ff["left"] != -1
    names = [str(s) for s in ff["feature_names"]]
    thr = {n: np.unique(ff["split_cond"][internal & (ff["split_index"] == j)]) for j, n in enumerate(names)}
    emp = thr["emp_length_num"]
    assert np.array_equal(emp[:-1], np.arange(1, 11, dtype=np.float32))
    assert emp[-1] == np.float32(10.0) + (np.abs(np.float32(10.0)) + np.float32(1e-5))
    for n in names[12:]:
        assert list(thr[n]) == [1.0]
    assert max(len(v) for v in thr.values()) <= 254


def test_sigmoid_spec_accuracy():
    x = np.linspace(-30, 30, 20001).astype(np.float32)
    got = O.sigmoid(x)
    want = 1.0 / (1.0 + np.exp(-x.astype(np.float64)))
    assert np.max(np.abs(got - want)) < 2e-7
    assert O.sigmoid(np.array([200.0, -200.0], np.float32)).tolist() == [1.0, pytest.approx(0.0, abs=1e-37)]


def test_oracle_regression_small(oracle_small):
    g = oracle_small
    X, y = g["X"], g["y"]
    cuts, nc, mv = O.compute_cuts(X, 256)
    assert np.array_equal(nc, g["n_cuts"]) and np.array_equal(mv, g["min_vals"])
    assert np.array_equal(cuts.view(np.uint32), g["cuts"].view(np.uint32))
    bins = O.bin_matrix(X, cuts, nc, ldb=32)
    assert np.array_equal(bins, g["bins"])
    tp = O.TreeParams(eta=0.3, max_depth=4, sg=int(g["sg"]), sh=int(g["sh"]))
    margin = np.zeros(len(y), np.float32)
    fo = O.train(bins, 24, O.bin_ptr_of(nc), cuts, mv, y, None, float(g["spw"]), tp, 5, margin,
                 want_rows=True, want_hist=True)
    for k in ("left", "right", "split_index", "split_bin", "default_left", "n_nodes", "row_order", "root_hist"):
        assert np.array_equal(fo[k], g[f"forest_{k}"]), k
    for k in ("split_cond", "base_weight", "loss_chg", "sum_hess"):
        assert np.array_equal(fo[k].view(np.uint32), g[f"forest_{k}"].view(np.uint32)), k
    assert np.array_equal(margin.view(np.uint32), g["margin"].view(np.uint32))


def test_oracle_cut_rules():
    """R1/R2 on hand-made columns."""
    n = 1000
    X = np.zeros((n, 4), np.float32)
    X[:, 0] = np.arange(n) % 3                      # 3 distinct -> cuts {1, 2, sentinel}
    X[:, 1] = np.nan                                # empty column
    X[:, 2] = np.arange(n, dtype=np.float32)        # 1000 distinct -> rank quantiles
    X[:, 3] = np.where(np.arange(n) % 2 == 0, -0.0, 5.0)
    cuts, nc, mv = O.compute_cuts(X, 256)
    assert nc[0] == 3 and list(cuts[0, :3]) == [1.0, 2.0, np.float32(2.0) + (np.float32(2.0) + np.float32(1e-5))]
    assert nc[1] == 1 and mv[1] == np.float32(1e-5) - (np.float32(1e-5) + np.float32(1e-5))
    assert nc[2] == 255 and np.all(np.diff(cuts[2, :255]) > 0)
    assert list(cuts[2, :3]) == [float(1 * n // 255), float(2 * n // 255), float(3 * n // 255)]
    assert nc[3] == 2 and cuts[3, 0] == 5.0
    bins = O.bin_matrix(X, cuts, nc)
    assert set(bins[:, 0]) == {0, 1, 2} and np.all(bins[:, 1] == 255)
    assert bins[:, 2].max() == 254 and bins[0, 2] == 0
    b2 = np.searchsorted(cuts[2, :255], X[:, 2], side="right")
    assert np.array_equal(bins[:, 2], np.minimum(b2, 254))


def test_oracle_split_rules():
    """R6/R7 on a hand-made node: forward/backward scans, missing direction and tie-breaks."""
    bp = np.array([0, 3, 6], np.int32)
    cuts = np.full((2, 256), np.inf, np.float32)
    cuts[0, :3] = [1, 2, 3]
    cuts[1, :3] = [1, 2, 3]
    mv = np.array([-1, -1], np.float32)
    tp = O.TreeParams(eta=1.0, reg_lambda=1.0, min_child_weight=0.0, max_depth=3, sg=4, sh=4)
    # identical histograms on both features -> lower feature index must win
    h = np.array([[-160, 160], [0, 160], [160, 160]] * 2, np.int64)
    s = O.scan(h, bp, cuts, mv, 0, 480, tp)
    assert s["feature"] == 0 and s["default_left"] == 0 and s["split_bin"] == 0 and s["threshold"] == 1.0
    # symmetric gains at bins 0 and 1 within one feature -> first in scan order wins
    assert s["GL"] == -160 and s["HL"] == 160
    # missing rows (totals exceed the bins' sum): backward scan may send missing left
    h2 = np.array([[-160, 160], [160, 160], [0, 0], [0, 16], [0, 16], [0, 16]], np.int64)
    s2 = O.scan(h2, bp, cuts, mv, -160 - 320, 320 + 320, tp)  # missing carries G=-480,H=320 on feature 0
    assert s2["feature"] == 0 and s2["default_left"] == 1
    assert s2["GL"] + s2["GR"] == -480 and s2["HL"] + s2["HR"] == 640


def test_tree_shap_is_the_shapley_value_of_the_path_dependent_expectation():
    """R13 pinned to its DEFINITION (no shap / xgboost here to compare with): enumerate all feature subsets of small
    random forests -- ragged trees, features repeated along a path, missing values -- and compare (f64 enumeration
    vs the f32 algorithm: 1e-5)."""
    from _forests import random_forest
    rng = np.random.default_rng(0)
    for T, F, D in [(3, 4, 3), (5, 6, 4), (4, 5, 5), (2, 3, 6)]:
        fo, _ = random_forest(rng, T, F, D, 3)
        X = rng.normal(size=(5, F)).astype(np.float32)
        X[rng.random(X.shape) < 0.2] = np.nan
        args = (fo["tree_offsets"], fo["left"], fo["right"], fo["split_index"], fo["split_cond"], fo["default_left"])
        got = O.tree_shap(X, *args, fo["sum_hess"], 0.5)
        margin = O.predict_margin(X, *args, 0.5)
        for r in range(len(X)):
            phi, base = O.shapley_brute_force(X[r], *args, fo["sum_hess"])
            assert np.max(np.abs(got[r, :F] - phi)) <= 1e-5
            assert abs(got[r, F] - 0.5 - base) <= 1e-5
            assert abs(float(got[r].astype(np.float64).sum()) - margin[r]) <= 1e-5


def test_tree_shap_additivity_on_the_shipped_model(fixture_forest):
    """Local accuracy on the reference's own 300-tree model: attributions + expected value = margin."""
    ff = fixture_forest
    rng = np.random.default_rng(4)
    X = (rng.normal(size=(64, 20)) * 200 + 400).astype(np.float32)
    X[rng.random(X.shape) < 0.05] = np.nan
    args = (ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"], ff["split_cond"], ff["default_left"])
    got = O.tree_shap(X, *args, ff["sum_hess"], 0.0)
    margin = O.predict_margin(X, *args, 0.0)
    assert np.max(np.abs(got.astype(np.float64).sum(1) - margin)) <= 2e-4
    unused = np.setdiff1d(np.arange(20), np.unique(ff["split_index"][ff["left"] != -1]))
    assert np.all(got[:, unused] == 0.0)                 # a feature no tree splits on gets nothing (dummy axiom)


def test_training_rules_learn_like_an_independent_histogram_gbdt():
    """Training end to end is unpinned against xgboost (not installable); as an independent cross-check
    the restated rules R1-R10 must reach the held-out AUC of sklearn's HistGradientBoostingClassifier -- another
    histogram GBDT with different binning and growth details -- within a sanity band."""
    from sklearn.ensemble import HistGradientBoostingClassifier
    from sklearn.metrics import roc_auc_score

    from cobalt_smart_lender_ai_b200.synth import make_lending
    X, y, _ = make_lending(24000, 30, seed=5)
    X, y = X.numpy(), y.numpy().astype(int)
    Xtr, ytr, Xte, yte = X[:16000], y[:16000], X[16000:], y[16000:]
    ref = O.fit_reference(Xtr, ytr, n_estimators=60, max_depth=6, learning_rate=0.1, scale_pos_weight=1.0)
    auc_oracle = roc_auc_score(yte, O.predict_proba_reference(ref, Xte))
    hgb = HistGradientBoostingClassifier(max_iter=60, max_depth=6, max_leaf_nodes=None, learning_rate=0.1, l2_regularization=1.0,
                                         min_samples_leaf=1, max_bins=255, early_stopping=False).fit(Xtr, ytr)
    auc_hgb = roc_auc_score(yte, hgb.predict_proba(Xte)[:, 1])
    assert auc_oracle > 0.7 and abs(auc_oracle - auc_hgb) <= 0.01, (auc_oracle, auc_hgb)
