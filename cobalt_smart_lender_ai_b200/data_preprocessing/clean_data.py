"""Mirror of the reference's cleaning entry points (src/data_preprocessing/clean_data.py) with the same
signatures: ``drop_columns_with_missing_values`` (:31-41) and ``clean_data_flow`` (:87-158).

String parsing and row/column filtering have no arithmetic density, so they stay pandas on the host
(SURVEY.md 8a-14); what differs from the reference is only I/O (no S3 client at import time, quiet by
default) and one repaired step: ``df[col].fillna(0, inplace=True)`` (:143) is a silent no-op under pandas
copy-on-write, so the *intended* fill (what the reference's notebook does) is applied by assignment."""
from __future__ import annotations

import pandas as pd


def drop_columns_with_missing_values(df: pd.DataFrame, threshold_percentage: float = 70.0, verbose: bool = False) -> pd.DataFrame:
    null_percent = (df.isnull().sum() / len(df)) * 100
    columns_to_drop = null_percent[null_percent > threshold_percentage].index.tolist()
    if verbose:
        print(f"[INFO] Dropping columns with >{threshold_percentage}% missing: {columns_to_drop}")
    return df.drop(columns=columns_to_drop)


def clean_data_flow(df: pd.DataFrame, verbose: bool = False) -> pd.DataFrame:
    """Steps 1-9 of the reference flow, in the reference's order."""
    out = df.drop(columns=["Unnamed: 0.1", "Unnamed: 0"], errors="ignore")
    out = out.dropna(subset=out.columns[out.isnull().sum() < 10])
    if "hardship_status" in out.columns:
        out["hardship_status"] = out["hardship_status"].fillna("No Hardship")
    if "term" in out.columns:
        out["term"] = out["term"].str.replace(" months", "").astype(int)
    if "int_rate" in out.columns:
        out["int_rate"] = out["int_rate"].str.replace("%", "").astype(float) / 100
    out = drop_columns_with_missing_values(out, threshold_percentage=70.0, verbose=verbose)
    out = out.drop(columns=[c for c in ("next_pymnt_d", "last_pymnt_d", "last_credit_pull_d",
                                        "mths_since_recent_revol_delinq", "il_util", "all_util",
                                        "mths_since_recent_bc_dlq") if c in out.columns])
    for col in ("inq_last_12m", "open_acc_6m", "chargeoff_within_12_mths"):
        if col in out.columns:
            out[col] = out[col].fillna(0)
    out = out.drop_duplicates()
    return out
