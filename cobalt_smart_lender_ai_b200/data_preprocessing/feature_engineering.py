"""Mirror of the reference's encoding entry points (src/data_preprocessing/feature_engineering.py) with
the same signatures: ``clean_lending_data`` (:44-101) and ``feature_engineer_lending_data`` (:103-184).

These are element-wise / dictionary encodes feeding the binned matrix.  Differences from
the reference: no S3 client at import time; the per-element ``Series.apply(lambda ...log1p...)`` loop
(:134-139) is the vectorised ``np.where(x > 0, log1p(x), x)`` (same values); the run date used for
``earliest_cr_line_days`` (:77) can be injected (``today=``) so results are reproducible; the input frame
is not mutated."""
from __future__ import annotations

from datetime import datetime

import numpy as np
import pandas as pd
from sklearn.preprocessing import LabelEncoder

COLUMNS_LOG = [
    'loan_amnt', 'funded_amnt', 'funded_amnt_inv', 'int_rate', 'installment', 'annual_inc', 'dti',
    'fico_range_low', 'fico_range_high', 'mths_since_last_delinq', 'open_acc', 'total_acc',
    'total_pymnt', 'total_pymnt_inv', 'total_rec_prncp', 'total_rec_int', 'total_rec_late_fee',
    'last_pymnt_amnt', 'acc_now_delinq', 'tot_coll_amt', 'tot_cur_bal', 'total_rev_hi_lim',
    'earliest_cr_line_days', 'acc_open_past_24mths', 'avg_cur_bal', 'bc_open_to_buy',
    'mo_sin_old_rev_tl_op', 'mo_sin_rcnt_rev_tl_op', 'mo_sin_rcnt_tl', 'mort_acc',
    'mths_since_recent_bc', 'mths_since_recent_inq', 'mths_since_recent_revol_delinq',
    'num_accts_ever_120_pd', 'num_actv_bc_tl', 'num_actv_rev_tl', 'num_bc_sats', 'num_bc_tl',
    'num_il_tl', 'num_op_rev_tl', 'num_rev_accts', 'num_rev_tl_bal_gt_0', 'num_sats',
    'num_tl_op_past_12m', 'pub_rec_bankruptcies', 'tot_hi_cred_lim', 'total_bal_ex_mort',
    'total_bc_limit', 'total_il_high_credit_limit', 'revol_util'
]
ONE_HOT_COLUMNS = ['grade', 'home_ownership', 'verification_status', 'purpose', 'application_type', 'hardship_status']
LOAN_STATUS_MAP = {'Fully Paid': 0, 'Current': 0, 'Issued': 0, 'In Grace Period': 0, 'Late (16-30 days)': 0,
                   'Late (31-120 days)': 1, 'Charged Off': 1, 'Default': 1}


def clean_lending_data(df: pd.DataFrame, today: datetime | None = None) -> pd.DataFrame:
    df = df.copy()
    leakage_cols = ['recoveries', 'collection_recovery_fee', 'debt_settlement_flag']
    useless_cols = ['id', 'url', 'title', 'zip_code', 'addr_state', 'emp_title', 'issue_d', 'initial_list_status',
                    'hardship_flag', 'sub_grade', 'next_pymnt_d', 'last_credit_pull_d', 'pymnt_plan']
    df = df.drop(columns=leakage_cols + useless_cols, errors='ignore')
    df = df.dropna(thresh=df.shape[1] - 20)
    df['emp_length'] = df['emp_length'].replace('< 1 year', '0')
    df['emp_length_num'] = pd.to_numeric(df['emp_length'].str.extract(r'(\d+)')[0], errors='coerce')
    df = df.drop(columns=['emp_length'])
    df['revol_util'] = df['revol_util'].str.replace('%', '', regex=False).astype(float) / 100
    today = today or datetime.today()
    if 'earliest_cr_line' in df.columns:
        d = pd.to_datetime(df['earliest_cr_line'], format='%b-%Y', errors='coerce')
        df['earliest_cr_line_days'] = (today - d).dt.days
        df = df.drop(columns=['earliest_cr_line'])
    if 'loan_status' in df.columns:
        df['loan_default'] = df['loan_status'].map(LOAN_STATUS_MAP)
        df = df.drop(columns=['loan_status'])
    return df


def log1p_positive(df: pd.DataFrame, columns=COLUMNS_LOG) -> pd.DataFrame:
    """log1p on the values that are > 0 of every listed column that has at least one such value."""
    out = df.copy()
    for col in columns:
        if col not in out.columns:
            continue
        x = out[col]
        if x.notnull().sum() == 0 or (x.dropna() <= 0).all():
            continue
        v = x.to_numpy(dtype=np.float64, na_value=np.nan)
        with np.errstate(invalid="ignore", divide="ignore"):
            out[col] = np.where(v > 0, np.log1p(np.where(v > 0, v, 0.0)), v)
    return out


def feature_engineer_lending_data(df: pd.DataFrame) -> tuple[pd.DataFrame, pd.DataFrame]:
    df_log = log1p_positive(df)
    df_log_tree = pd.get_dummies(df_log.copy(), columns=ONE_HOT_COLUMNS, drop_first=True)

    df_log_nn = df_log.copy()
    null_counts = df_log_nn.isnull().sum()
    for col in null_counts[null_counts > 0].index.tolist():
        if col == 'dti' or not np.issubdtype(df_log_nn[col].dtype, np.number):
            continue
        df_log_nn[col + '_NA'] = df_log_nn[col].isnull().astype(int)
        df_log_nn[col] = df_log_nn[col].fillna(df_log_nn[col].median())
    df_log_nn['no_income'] = ((df_log_nn['annual_inc'].isnull()) | (df_log_nn['annual_inc'] == 0)).astype(int)
    df_log_nn['dti_NA'] = df_log['dti'].isnull().astype(int)
    df_log_nn['dti'] = df_log_nn['dti'].fillna(df_log_nn['dti'].median())
    for col in df_log_nn.select_dtypes(include=['object', 'category', 'string']).columns.tolist():
        le = LabelEncoder()
        df_log_nn[col] = le.fit_transform(df_log_nn[col].astype(str).fillna("missing"))
    return df_log_tree, df_log_nn
