"""CPU suite: the preprocessing mirrors against outputs of the REFERENCE's own modules
(src/data_preprocessing/{clean_data,feature_engineering}.py, imported in the build container by
tests/golden/make_preprocessing_golden.py; the frames are committed as tests/golden/prep_*.pkl)."""
import os
from datetime import datetime

import numpy as np
import pandas as pd
import pytest

from cobalt_smart_lender_ai_b200.data_preprocessing import clean_data as cd
from cobalt_smart_lender_ai_b200.data_preprocessing import feature_engineering as fe

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TODAY = datetime(2025, 1, 15)


def _load(name):
    return pd.read_pickle(os.path.join(G, f"prep_{name}.pkl"))


def test_clean_data_flow_matches_reference():
    raw, want = _load("raw"), _load("cleaned")
    before = raw.copy()
    got = cd.clean_data_flow(raw)
    pd.testing.assert_frame_equal(raw, before)                 # input not mutated
    pd.testing.assert_frame_equal(got, want)
    assert "mostly_missing" not in got.columns and got["term"].dtype.kind == "i"
    assert got[["inq_last_12m", "open_acc_6m"]].isnull().sum().sum() == 0      # the intended (notebook) fill


def test_drop_columns_with_missing_values():
    df = pd.DataFrame({"a": [1, None, None, None], "b": [1, 2, None, 4], "c": [None] * 4})
    assert list(cd.drop_columns_with_missing_values(df, 70.0).columns) == ["b"]
    assert list(cd.drop_columns_with_missing_values(df, 80.0).columns) == ["a", "b"]


def test_clean_lending_data_matches_reference():
    got = fe.clean_lending_data(_load("cleaned"), today=TODAY)
    pd.testing.assert_frame_equal(got, _load("cl2"))
    assert got["loan_default"].isnull().sum() > 0              # unmapped statuses stay NaN, like the reference


def test_feature_engineer_matches_reference():
    tree, nn = fe.feature_engineer_lending_data(_load("cl2"))
    want_tree, want_nn = _load("tree"), _load("nn")
    assert list(tree.columns) == list(want_tree.columns) and list(nn.columns) == list(want_nn.columns)
    for got, want in ((tree, want_tree), (nn, want_nn)):
        for c in want.columns:
            if want[c].dtype.kind == "f":
                np.testing.assert_allclose(got[c].to_numpy(float), want[c].to_numpy(float), rtol=0, atol=1e-12, equal_nan=True)
            else:
                assert got[c].astype(str).tolist() == want[c].astype(str).tolist(), c
    dummies = [c for c in tree.columns if c.startswith(("grade_", "purpose_", "hardship_status_"))]
    assert dummies and all(tree[c].dtype == bool for c in dummies)             # what XGBClassifier.fit must accept


def test_log1p_rule():
    df = pd.DataFrame({"loan_amnt": [10.0, 0.0, -1.0, np.nan], "dti": [-1.0, 0.0, -3.0, np.nan], "other": [5.0] * 4})
    out = fe.log1p_positive(df)
    np.testing.assert_allclose(out["loan_amnt"], [np.log1p(10.0), 0.0, -1.0, np.nan], equal_nan=True)
    assert out["dti"].equals(df["dti"]) and out["other"].equals(df["other"])     # all non-positive / not listed: untouched
