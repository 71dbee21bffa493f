"""Runs the REFERENCE's own preprocessing (imported from /root/reference with boto3 stubbed out) on a
small synthetic raw-schema frame and stores input + outputs as a golden fixture.  Run here only."""
import io
import os
import sys
import types
from contextlib import redirect_stdout
from datetime import datetime

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
FIXED_TODAY = datetime(2025, 1, 15)


def raw_frame(n=400, seed=4):
    rng = np.random.default_rng(seed)
    pick = lambda vals, p=None: rng.choice(vals, size=n, p=p)  # noqa: E731
    df = pd.DataFrame({
        "Unnamed: 0": np.arange(n), "id": np.arange(n) + 1000,
        "loan_amnt": rng.integers(7, 400, n) * 100.0, "funded_amnt": rng.integers(7, 400, n) * 100.0,
        "term": pick([" 36 months", " 60 months"]), "int_rate": [f"{x:.2f}%" for x in rng.uniform(5, 31, n)],
        "installment": rng.uniform(22, 1700, n).round(2), "grade": pick(list("ABCDEFG")),
        "sub_grade": pick(["A1", "B2", "C3"]), "emp_title": pick(["nurse", "teacher", None]),
        "emp_length": pick(["< 1 year", "1 year", "5 years", "10+ years", None], [0.1, 0.2, 0.3, 0.3, 0.1]),
        "home_ownership": pick(["RENT", "MORTGAGE", "OWN", "ANY"]), "annual_inc": rng.choice([0.0, 30000, 66000, 120000], n),
        "verification_status": pick(["Verified", "Source Verified", "Not Verified"]),
        "issue_d": pick(["Jan-2016", "Feb-2018"]), "loan_status": pick(
            ["Fully Paid", "Current", "Charged Off", "Late (31-120 days)", "In Grace Period",
             "Does not meet the credit policy. Status:Fully Paid"], [0.5, 0.3, 0.12, 0.03, 0.03, 0.02]),
        "pymnt_plan": "n", "url": "x", "purpose": pick(["debt_consolidation", "credit_card", "car", "other", "wedding"]),
        "title": "t", "zip_code": "123xx", "addr_state": pick(["CA", "NY"]),
        "dti": np.where(rng.random(n) < 0.02, np.nan, rng.uniform(-1, 40, n).round(2)),
        "earliest_cr_line": pick(["Aug-2003", "Dec-1999", "Jan-2011", None], [0.4, 0.3, 0.28, 0.02]),
        "fico_range_low": rng.integers(128, 170, n) * 5.0, "fico_range_high": rng.integers(128, 170, n) * 5.0 + 4,
        "open_acc": rng.poisson(11, n).astype(float), "revol_util": [None if u < 0.01 else f"{u * 100:.1f}%" for u in rng.random(n)],
        "total_acc": rng.poisson(24, n).astype(float), "initial_list_status": pick(["w", "f"]),
        "total_pymnt": rng.uniform(0, 40000, n).round(2), "recoveries": 0.0, "collection_recovery_fee": 0.0,
        "last_pymnt_d": pick(["Jan-2019", None]), "next_pymnt_d": None, "last_credit_pull_d": pick(["Mar-2020", None]),
        "mths_since_last_delinq": np.where(rng.random(n) < 0.5, np.nan, rng.integers(0, 80, n)),
        "mths_since_recent_revol_delinq": np.where(rng.random(n) < 0.68, np.nan, rng.integers(0, 80, n)),
        "il_util": np.where(rng.random(n) < 0.4, np.nan, rng.uniform(0, 120, n)),
        "all_util": rng.uniform(0, 100, n), "mths_since_recent_bc_dlq": np.where(rng.random(n) < 0.75, np.nan, 5.0),
        "inq_last_12m": np.where(rng.random(n) < 0.3, np.nan, rng.poisson(2, n)),
        "open_acc_6m": np.where(rng.random(n) < 0.3, np.nan, rng.poisson(1, n)),
        "chargeoff_within_12_mths": np.where(rng.random(n) < 0.01, np.nan, 0.0),
        "acc_now_delinq": 0.0, "pub_rec_bankruptcies": rng.choice([0.0, 1.0], n, p=[0.9, 0.1]),
        "num_rev_accts": rng.poisson(14, n).astype(float), "application_type": pick(["Individual", "Joint App"], [0.9, 0.1]),
        "hardship_flag": "N", "hardship_status": pick([None, "ACTIVE", "BROKEN", "COMPLETED"], [0.9, 0.04, 0.03, 0.03]),
        "debt_settlement_flag": "N", "mostly_missing": np.where(rng.random(n) < 0.9, np.nan, 1.0),
    })
    return pd.concat([df, df.iloc[:5]], ignore_index=True)      # duplicates for step 8


def main():
    sys.modules["boto3"] = types.SimpleNamespace(client=lambda *a, **k: None)
    sys.path.insert(0, "/root/reference/src/data_preprocessing")
    import clean_data as ref_clean
    import feature_engineering as ref_fe

    class _FixedDT(datetime):
        @classmethod
        def today(cls):
            return FIXED_TODAY
    ref_fe.datetime = _FixedDT
    raw = raw_frame()
    with redirect_stdout(io.StringIO()):
        cleaned = ref_clean.clean_data_flow(raw.copy())
        # the notebook semantics of step 7 (the in-place fill is a no-op under pandas copy-on-write)
        for col in ("inq_last_12m", "open_acc_6m", "chargeoff_within_12_mths"):
            if col in cleaned.columns:
                cleaned[col] = cleaned[col].fillna(0)
        cl2 = ref_fe.clean_lending_data(cleaned.copy())
        tree, nn = ref_fe.feature_engineer_lending_data(cl2.copy())
    for name, frame in (("raw", raw), ("cleaned", cleaned), ("cl2", cl2), ("tree", tree), ("nn", nn)):
        frame.to_pickle(os.path.join(HERE, f"prep_{name}.pkl"))
    print({k: v.shape for k, v in (("raw", raw), ("cleaned", cleaned), ("cl2", cl2), ("tree", tree), ("nn", nn))})


if __name__ == "__main__":
    main()
