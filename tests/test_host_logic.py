"""CPU suite: host-side logic of the product (no compute calls, no GPU)."""
import ctypes as C
import os
import pickle
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    """The C-ABI library builds/loads without a GPU and exports everything include/*.h declares."""
    from cobalt_smart_lender_ai_b200 import _lib
    lib = _lib.load()
    assert lib.cb_version() == 200
    hdr = open(os.path.join(ROOT, "include", "cobalt_gbdt.h")).read()
    declared = set(re.findall(r"^(?:int|void|int64_t|const char\*)\s+(cb_[a-z0-9_]+)\s*\(", hdr, flags=re.M))
    assert declared, "no declarations found"
    raw = C.CDLL(_lib.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(raw, name), f"{name} declared in the header but not exported"
    assert declared <= set(_lib.EXPORTS) | {"cb_version"}, declared - set(_lib.EXPORTS)


def test_no_cpu_fallback_fails_loudly():
    import torch

    from cobalt_smart_lender_ai_b200 import _lib
    from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.CobaltError):
        XGBClassifier().fit(np.zeros((4, 2), np.float32), np.array([0, 1, 0, 1]))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "cobalt_smart_lender_ai_b200")
    for dp, _, fns in os.walk(pkg):
        for fn in fns:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, fn)).read()
                assert "oracle" not in src.replace("the oracle", "").replace("oracle's", "").replace("oracle.QBITS", "") \
                    or "import oracle" not in src and "from oracle" not in src, fn
                assert "from oracle" not in src and "import oracle" not in src, fn


def test_estimator_contract_clone_params_pickle():
    from sklearn.base import clone

    from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier
    m = XGBClassifier(eval_metric="logloss", use_label_encoder=False, scale_pos_weight=6.7, random_state=42)
    c = clone(m).set_params(n_estimators=300, max_depth=3, learning_rate=0.05, subsample=0.8, colsample_bytree=0.5, gamma=0)
    assert c.get_params()["max_depth"] == 3 and c.get_params()["scale_pos_weight"] == 6.7
    assert m.get_params()["n_estimators"] == 100 and m.get_params()["max_depth"] == 6 and m.get_params()["learning_rate"] == 0.3
    pickle.loads(pickle.dumps(c))            # an unfitted estimator travels to loky workers
    from sklearn.exceptions import NotFittedError
    with pytest.raises(NotFittedError):
        c.get_booster()


def test_quant_shifts_match_oracle_and_bound():
    from cobalt_smart_lender_ai_b200 import ops
    from oracle import oracle as O
    for w in (0.3, 1.0, 1.5, 2.0, 6.746, 8.0, 100.0):
        sg, sh = ops.quant_shifts(w)
        assert (sg, sh) == O.quant_shifts(w)
        assert w * 2.0 ** sg <= 2 ** 17 < w * 2.0 ** (sg + 1) * 1.0000001 * 2 or w * 2.0 ** sg <= 2 ** 17


def test_colsample_masks_and_hash_match_oracle_hash():
    from cobalt_smart_lender_ai_b200.sklearn import _hash32, colsample_masks
    from oracle import oracle as O
    for s, a, b in ((78, 0, 0), (78, 3, 1234567), (42, 299, 2 ** 33 + 5)):
        assert _hash32(s, a, b) == O.lib().orc_hash(s, a, b)
    m = colsample_masks(5, 20, 0.5, 78)
    assert m.shape == (5, 20) and (m.sum(1) == 10).all() and len({tuple(r) for r in m}) > 1
    assert colsample_masks(5, 20, 1.0, 78) is None


def test_base_score_estimate():
    from cobalt_smart_lender_ai_b200.sklearn import estimate_base_score
    from oracle import oracle as O
    y = np.array([0] * 870 + [1] * 130)
    assert estimate_base_score(130 * (870 / 130), 870) == 0.5          # spw = neg/pos -> exactly 0.5 (fixture)
    assert estimate_base_score(130 * 2.0, 870) == O.estimate_base_score(y, 2.0)


def test_shard_rows_and_feature_slices():
    from cobalt_smart_lender_ai_b200 import parallel
    for n, w in ((2_200_000, 8), (10, 3), (7, 8)):
        spans = [parallel.shard_rows(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(spans[i][1] == spans[i + 1][0] for i in range(w - 1))
        assert max(e - b for b, e in spans) - min(e - b for b, e in spans) <= 1
    assert parallel.feature_slices(150, 8)[-1][1] == 150


def test_heap_forest_packing_matches_oracle_traversal(fixture_forest):
    """Complete-tree (heap) repacking of the shipped forest walks to the same leaves as the wire format."""
    from cobalt_smart_lender_ai_b200.booster import heap_forest, pack_forest
    from oracle import oracle as O
    ff = fixture_forest
    thr, meta, leaf, depth = heap_forest(ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"],
                                         ff["split_cond"], ff["default_left"])
    assert depth == 7 and thr.shape == (300, 127) and leaf.shape == (300, 128)
    rng = np.random.default_rng(0)
    X = (rng.normal(size=(64, 20)) * 300 + 500).astype(np.float32)
    X[rng.random(X.shape) < 0.1] = np.nan
    want = O.predict_margin(X, ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"], ff["split_cond"],
                            ff["default_left"], 0.0)
    NI, T = thr.shape[1], thr.shape[0]
    ar = np.arange(T)
    for r in range(len(X)):
        idx = np.zeros(T, np.int64)
        for _ in range(depth):
            q = meta[ar, idx].astype(np.int64)
            v = X[r, q & 0x7FFF]
            left = np.where(np.isnan(v), (q >> 15) == 1, v < thr[ar, idx])
            idx = 2 * idx + np.where(left, 1, 2)
        m = np.float32(0)
        for x in leaf[ar, idx - NI]:
            m = np.float32(m + x)
        assert m.view(np.uint32) == want[r].view(np.uint32)
    nodes, toff = pack_forest(ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"], ff["split_cond"],
                              ff["default_left"])
    assert nodes.shape == (59216,) and toff[-1] == 59216


def test_ubjson_roundtrip_and_reference_pickle(fixture_forest):
    from cobalt_smart_lender_ai_b200 import ubjson
    ff = fixture_forest
    names = [str(s) for s in ff["feature_names"]]
    sub = {k: ff[k][: int(ff["tree_offsets"][5])] for k in ("left", "right", "parent", "split_index", "split_cond",
                                                           "default_left", "base_weight", "loss_chg", "sum_hess")}
    sub["tree_offsets"] = ff["tree_offsets"][:6]
    doc = ubjson.decode(ubjson.encode(ubjson.model_document(sub, names, 0.5)))
    fo, nm, base = ubjson.forest_from_model(doc)
    assert nm == names and base == 0.5
    for k, v in sub.items():
        assert np.array_equal(fo[k], v), k
    ref = "/root/reference/src/api/models/xgb_model_tree.pkl"
    if os.path.exists(ref):                  # only in the build container; the GPU box has the golden npz
        fo, nm, base = ubjson.forest_from_model(ubjson.decode(ubjson.booster_bytes(ref)))
        assert nm == names and base == float(ff["base_score"])
        for k in ("left", "right", "split_index", "default_left", "tree_offsets"):
            assert np.array_equal(fo[k], ff[k]), k
        assert np.array_equal(fo["split_cond"].view(np.uint32), ff["split_cond"].view(np.uint32))


def test_booster_get_score_semantics(fixture_forest):
    from cobalt_smart_lender_ai_b200.booster import Booster
    ff = fixture_forest
    names = [str(s) for s in ff["feature_names"]]
    b = Booster(ff, 20, names, 0.5)
    w = b.get_score(importance_type="weight")
    assert [int(w[n]) for n in names] == [1926, 690, 3188, 2365, 4789, 657, 1442, 3152, 2299, 423, 1689, 3159, 337, 433,
                                          320, 522, 520, 205, 428, 914]
    g, tg = b.get_score(importance_type="gain"), b.get_score(importance_type="total_gain")
    assert all(abs(g[n] * w[n] - tg[n]) <= 1e-6 * tg[n] for n in names)
    with pytest.raises(ValueError):
        b.get_score(importance_type="nope")
    b2 = pickle.loads(pickle.dumps(b))
    assert b2.get_score(importance_type="weight") == w


def test_api_routes_with_stub_model():
    """Route logic and JSON shapes of the scoring service mirror (reference src/api/cobalt_fast_api.py:96-143)."""
    from fastapi.testclient import TestClient

    from cobalt_smart_lender_ai_b200.api import cobalt_fast_api as api

    class _Booster:
        def get_score(self, importance_type="weight"):
            return {f"f{i}": float(i) for i in range(15)}

    class _Stub:
        def predict_proba(self, df):
            p = np.full(len(df), 0.25, np.float32)
            return np.stack([1 - p, p], 1)

        def get_booster(self):
            return _Booster()

    class _Explainer:                       # shap.TreeExplainer's surface as /predict uses it (reference :100-105)
        expected_value = -1.5

        def shap_values(self, df):
            return np.tile(np.arange(df.shape[1], dtype=np.float32), (len(df), 1))

    api.xgb_model = _Stub()
    api.explainer = _Explainer()
    client = TestClient(api.app)            # no context manager: lifespan (model load) is not run
    fields = {n: 1.0 for n in ("loan_amnt", "term", "installment", "fico_range_low", "last_fico_range_high",
                               "open_il_12m", "open_il_24m", "max_bal_bc", "num_rev_accts", "pub_rec_bankruptcies",
                               "emp_length_num", "earliest_cr_line_days")}
    fields.update({n: 0 for n in ("grade_E", "home_ownership_MORTGAGE", "verification_status_Verified",
                                  "application_type_Joint App", "hardship_status_BROKEN", "hardship_status_COMPLETE",
                                  "hardship_status_COMPLETED", "hardship_status_No Hardship")})
    r = client.post("/predict", json=fields)
    assert r.status_code == 200 and r.json()["prob_default"] == 0.25
    assert r.json()["features"][15] == "application_type_Joint App" and len(r.json()["features"]) == 20
    assert r.json()["shap_values"] == [float(i) for i in range(20)] and r.json()["base_value"] == -1.5
    csv = "a,b\n1,2\n3,\n"
    r = client.post("/predict_bulk_csv", files={"file": ("x.csv", csv, "text/csv")})
    assert r.status_code == 200
    recs = r.json()["predictions"]
    assert len(recs) == 2 and recs[0]["prob_default"] == 0.25 and recs[1]["b"] == "null"
    # +-inf and empty cells all come back as the string "null" (reference :121 replace([inf, -inf], nan).fillna("null"))
    csv = "a,b\ninf,2\n-inf,\n5,7\n"
    r = client.post("/predict_bulk_csv", files={"file": ("x.csv", csv, "text/csv")})
    assert r.status_code == 200, r.text
    recs = r.json()["predictions"]
    assert recs[0]["a"] == "null" and recs[1]["a"] == "null" and recs[1]["b"] == "null" and recs[2]["a"] == 5.0
    r = client.post("/feature_importance_bulk", json={"data": [{"x": 1}]})
    top = r.json()["top_features"]
    assert len(top) == 10 and top[0] == {"feature": "f14", "importance": 14.0}
    assert client.post("/feature_importance_bulk", json={"data": []}).status_code == 400


def test_bench_reference_arm_contract():
    """`bench.py --impl reference` (the oracle port on the host cores) prints ONE JSON line with the contract's keys."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                          "--ref-rows", "20000"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["metric"] == "boosting rounds/s" and d["value"] > 0
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["cpu_baseline"]["kind"] == "port" and "workload" in d["config"]


def test_q8_forest_encoding_decides_like_the_float_compare():
    """Host side of the quantised scorer (booster.q8_forest): threshold tables, node words (span << 16 | plane/feature
    offset) and the two code planes, walked here in numpy exactly as predict_q8_kernel does, give the oracle's margins
    bit for bit -- values on / next to thresholds, +-inf, -0.0 and NaN with either default direction."""
    from _forests import random_forest

    from cobalt_smart_lender_ai_b200.booster import heap_forest, q8_forest
    from oracle import oracle as O
    rng = np.random.default_rng(1)
    for n, f, T, depth, base in [(200, 7, 9, 3, 0), (300, 150, 20, 6, 1024), (100, 250, 5, 6, 1024)]:
        fo, pools = random_forest(rng, T, f, depth, 40)
        X = np.zeros((n, f), np.float32)
        for j in range(f):
            b = rng.choice(pools[j], size=n)
            kind = rng.integers(0, 6, size=n)
            X[:, j] = np.where(kind == 0, b, np.where(kind == 1, np.nextafter(b, np.float32(-np.inf)),
                               np.where(kind == 2, np.nextafter(b, np.float32(np.inf)), rng.normal(size=n)))).astype(np.float32)
        for val, p in ((np.nan, 0.08), (np.inf, 0.01), (-np.inf, 0.01), (-0.0, 0.01)):
            X[rng.random(X.shape) < p] = val
        args = (fo["tree_offsets"], fo["left"], fo["right"], fo["split_index"], fo["split_cond"], fo["default_left"])
        want = O.predict_margin(X, *args, 0.0)
        thr, meta, leaf, d = heap_forest(*args)
        cuts, n_cuts, words = q8_forest(thr, meta, leaf, d, f, base)
        codes = np.full((n, f), 255, np.int64)                      # K2: number of thresholds <= x, clamped; NaN -> 255
        for j in range(f):
            c = cuts[j, :n_cuts[j]]
            lo = np.searchsorted(c, X[:, j], side="right")
            codes[:, j] = np.where(np.isnan(X[:, j]), 255, np.minimum(lo, n_cuts[j] - 1))
        planes = np.stack([codes, (codes + 1) & 255])               # plane 1: code + 1, missing wraps to 0
        m = np.zeros(n, np.float32)
        rows = np.arange(n)
        for t in range(words.shape[0]):
            i = np.ones(n, np.int64)
            for _ in range(d):
                q = words[t, i].astype(np.int64)
                off = ((q & 0xFFFF) - base) // 128                  # (plane * f + feature)
                x = planes[off // f, rows, off % f]
                i = 2 * i + np.where((x << 16) <= q, 0, 1)
            m = (m + words[t, i].view(np.float32)).astype(np.float32)
        assert np.array_equal(m.view(np.uint32), want.view(np.uint32))
    # more thresholds on a feature than a byte can index: no quantised form
    fo, _ = random_forest(rng, 400, 2, 6, 4000)
    thr, meta, leaf, d = heap_forest(fo["tree_offsets"], fo["left"], fo["right"], fo["split_index"], fo["split_cond"],
                                     fo["default_left"])
    assert q8_forest(thr, meta, leaf, d, 2, 1024) is None


def test_orchestration_classes_delegate_what_they_do_not_cover():
    """feature_selection.RFE / model_selection.RandomizedSearchCV take the prepared-once path only for this package's
    estimator (and roc_auc for the search); anything else is handed to sklearn's own classes unchanged."""
    from sklearn.feature_selection import RFE as SkRFE
    from sklearn.linear_model import LogisticRegression
    from sklearn.model_selection import RandomizedSearchCV as SkSearch

    from cobalt_smart_lender_ai_b200.feature_selection import RFE
    from cobalt_smart_lender_ai_b200.model_selection import RandomizedSearchCV
    from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier
    rng = np.random.default_rng(0)
    X = rng.normal(size=(300, 8))
    y = (X[:, 0] + 0.5 * X[:, 3] + 0.2 * rng.normal(size=300) > 0).astype(int)
    lr = LogisticRegression(max_iter=200)
    a = RFE(estimator=lr, n_features_to_select=3).fit(X, y)
    b = SkRFE(estimator=lr, n_features_to_select=3).fit(X, y)
    assert a.binned_once_ is False and np.array_equal(a.support_, b.support_) and np.array_equal(a.ranking_, b.ranking_)
    assert a.transform(X).shape == (300, 3)
    kw = dict(param_distributions={"C": [0.1, 1.0, 10.0]}, n_iter=2, scoring="roc_auc", cv=3, random_state=0)
    s1 = RandomizedSearchCV(estimator=lr, **kw).fit(X, y)
    s2 = SkSearch(estimator=lr, **kw).fit(X, y)
    assert s1.prepared_folds_ is None and s1.best_params_ == s2.best_params_ and s1.best_score_ == s2.best_score_
    assert np.array_equal(s1.predict(X), s2.predict(X))
    # which settings take the fast path
    assert RFE(estimator=XGBClassifier(), n_features_to_select=3)._fast_path()
    assert not RFE(estimator=XGBClassifier(colsample_bytree=0.5), n_features_to_select=3)._fast_path()
    assert RandomizedSearchCV(XGBClassifier(), {}, scoring="roc_auc")._fast_path()
    assert not RandomizedSearchCV(XGBClassifier(), {}, scoring="accuracy")._fast_path()


def test_load_model_never_unpickles_a_foreign_file(tmp_path, fixture_forest):
    """load_model scans for an xgboost booster payload first (no code of the file runs) and only unpickles files that
    hold none -- this package's own estimator, saved with joblib or pickle."""
    import pickle

    import joblib

    from cobalt_smart_lender_ai_b200.api.cobalt_fast_api import load_model
    from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier
    from cobalt_smart_lender_ai_b200 import ubjson
    keys = ["tree_offsets", "left", "right", "parent", "split_index", "split_cond", "default_left", "base_weight",
            "loss_chg", "sum_hess"]
    m = XGBClassifier().load_fixture_forest({k: fixture_forest[k] for k in keys},
                                            [str(x) for x in fixture_forest["feature_names"]], 0.5)
    a, b = tmp_path / "a.pkl", tmp_path / "b.pkl"
    joblib.dump(m, a)
    b.write_bytes(pickle.dumps(m))
    for p in (a, b):
        with pytest.raises(ubjson.NoBoosterPayload):
            ubjson.booster_bytes(str(p))
        got = load_model(p)
        assert np.array_equal(got.get_booster().forest["split_cond"], fixture_forest["split_cond"])

    class Evil:
        def __reduce__(self):
            return (os.system, ("exit 1",))
    c = tmp_path / "c.pkl"
    c.write_bytes(pickle.dumps({"x": Evil()}))
    with pytest.raises(Exception):
        load_model(c)          # joblib.load is reached only for payload-free files; a foreign object is refused


def test_row_stride_group_arrays_and_code_unpacking():
    """Host logic of the binned-matrix layer: row strides never let a column block's row straddle a 128-byte line;
    group arrays are CSR; natural_codes() undoes placement, column blocks and one-hot packing."""
    import torch

    from cobalt_smart_lender_ai_b200 import ops
    from cobalt_smart_lender_ai_b200.binned import BinnedMatrix
    for n_cols in range(1, 400):
        ldb = ops.padded_ldb(n_cols)
        last = ldb - 128 * ((ldb - 1) // 128)
        assert ldb >= n_cols and ldb % 4 == 0 and last & (last - 1) == 0 and last >= 16, (n_cols, ldb)
    assert [ops.padded_ldb(k) for k in (112, 128, 150)] == [128, 128, 160]
    ptr, feat = ops.group_arrays([[3, 5, 9], [1, 2]])
    assert ptr.tolist() == [0, 3, 5] and feat.tolist() == [3, 5, 9, 1, 2] and ops.group_arrays([]) == (None, None)
    # 6 features: 0 and 4 plain, group A = [1, 2, 3] in one byte, feature 5 plain; placement scrambled, ldb 16
    rng = np.random.default_rng(0)
    n = 50
    nat = np.zeros((n, 6), np.uint8)
    nat[:, 0] = rng.integers(0, 200, n)
    nat[:, 4] = rng.integers(0, 9, n)
    nat[:, 5] = np.where(rng.random(n) < 0.2, 255, rng.integers(0, 3, n))
    which = rng.integers(0, 4, n)                                  # 0: none of the group set
    for m in range(3):
        nat[:, 1 + m] = which == m + 1
    pos = np.full(16, -1, np.int32)
    pos[[7, 2, 9, 12]] = [0, 4, -2, 5]                            # byte 9 holds group 0
    b = np.full((n, 16), 255, np.uint8)
    b[:, 7], b[:, 2], b[:, 9], b[:, 12] = nat[:, 0], nat[:, 4], which.astype(np.uint8), nat[:, 5]
    bm = BinnedMatrix(torch.from_numpy(b), None, None, pos, None, None, [[1, 2, 3]], False, True, n, 6, 16)
    assert np.array_equal(bm.natural_codes(), nat)
    # block-major storage of a 160-byte row: block 0 = [n][128] then block 1 = [n][32]
    wide = rng.integers(0, 255, (n, 160)).astype(np.uint8)
    stored = np.concatenate([wide[:, :128].reshape(-1), wide[:, 128:].reshape(-1)]).reshape(n, 160)
    bm2 = BinnedMatrix(torch.from_numpy(stored), None, None, np.arange(160, dtype=np.int32), None, None, [], True, True, n, 160, 160)
    assert np.array_equal(bm2.natural_codes(), wide)


def test_local_only_hides_the_process_group():
    """parallel.local_only(): inside it the path behaves as single-process whatever torch.distributed says (nesting ok)."""
    from cobalt_smart_lender_ai_b200 import parallel
    assert parallel.world_size() == 1 and parallel._dist() is None
    with parallel.local_only():
        with parallel.local_only():
            assert parallel._LOCAL_ONLY == 2 and parallel.world_size() == 1
        assert parallel._LOCAL_ONLY == 1
    assert parallel._LOCAL_ONLY == 0
    assert parallel.global_stats(3.0, 5.0, 2.0, 10) == (3.0, 5.0, 2.0, 0)
