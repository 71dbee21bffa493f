"""Mirror of the reference's training job (src/model_train_test/model_tree_train_test.py::main, :73-242)
on the B200 path, minus its I/O: the S3 download/upload, the plots and the metrics JSON are the caller's
business (out of scope), everything between them is kept step for step --

  2. drop the known leakage columns (:82-87)            5. RFE down to exactly 20 features (:111-124)
  3. 80/20 split, random_state 22 (:92-98)              6. RandomizedSearchCV, 20 draws x 3 folds, roc_auc (:130-166)
  4. scale_pos_weight = neg / pos on y_train (:103-105) 7. evaluation on the test split (:171-179)

``XGBClassifier`` is this package's drop-in; ``RFE`` and ``RandomizedSearchCV`` are this package's classes with
sklearn's surface and results, which prepare (upload, sketch, bin) a training matrix once and run the many fits of
the step on the resident bins.  ``n_jobs`` is accepted and ignored by them: every fit is GPU bound, and loky workers
would each open their own CUDA context on the same device.
"""
from __future__ import annotations

import logging

import pandas as pd
from sklearn.metrics import classification_report, confusion_matrix, roc_auc_score
from sklearn.model_selection import StratifiedKFold, train_test_split

from ..feature_selection import RFE          # sklearn.feature_selection.RFE's surface; bins once
from ..model_selection import RandomizedSearchCV   # sklearn's surface; every fold is sketched and binned once
from ..sklearn import XGBClassifier

LEAKAGE_COLS = [
    'total_rec_late_fee', 'total_rec_prncp', 'out_prncp', 'last_pymnt_amnt', 'last_pymnt_d',
    'funded_amnt_inv', 'funded_amnt', 'out_prncp_inv', 'total_pymnt', 'total_pymnt_inv',
    'last_pymnt_d_days', 'last_credit_pull_d_days', 'issue_d_days', 'total_rec_int'
]
PARAM_DISTRIBUTIONS = {
    'n_estimators': [100, 200, 300],
    'max_depth': [3, 5, 7, 9],
    'learning_rate': [0.01, 0.05, 0.1],
    'subsample': [0.8, 1.0],
    'colsample_bytree': [0.5, 0.8, 1.0],
    'gamma': [0, 1, 5],
}


def train_pipeline(df_tree: pd.DataFrame, n_features_to_select: int = 20, n_iter: int = 20, cv: int = 3,
                   n_jobs: int = 1, param_distributions: dict | None = None, device: str = "cuda") -> dict:
    """Steps 2-7 of the reference's ``main()``; returns the tuned model, the selected features and the metrics."""
    df_tree = df_tree.drop(columns=LEAKAGE_COLS, errors='ignore')
    X = df_tree.drop('loan_default', axis=1)
    y = df_tree['loan_default']
    X_train, X_test, y_train, y_test = train_test_split(X, y, test_size=0.2, random_state=22)
    logging.info(f"Train shape: {X_train.shape}, Test shape: {X_test.shape}")

    neg_count = (y_train == 0).sum()
    pos_count = (y_train == 1).sum()
    scale_pos_weight = neg_count / pos_count
    logging.info(f"scale_pos_weight={scale_pos_weight:.4f}")

    base_model = XGBClassifier(eval_metric='logloss', use_label_encoder=False, scale_pos_weight=scale_pos_weight,
                               random_state=42, device=device)
    rfe = RFE(estimator=base_model, n_features_to_select=n_features_to_select, step=1)
    rfe.fit(X_train, y_train)
    selected_features = X_train.columns[rfe.support_].tolist()
    logging.info(f"Selected {len(selected_features)} features: {selected_features}")
    X_train_sel = X_train[selected_features]
    X_test_sel = X_test[selected_features]

    xgb_base = XGBClassifier(eval_metric='logloss', use_label_encoder=False, scale_pos_weight=scale_pos_weight,
                             random_state=78, device=device)
    random_search = RandomizedSearchCV(estimator=xgb_base, param_distributions=param_distributions or PARAM_DISTRIBUTIONS,
                                       n_iter=n_iter, scoring='roc_auc', cv=StratifiedKFold(cv), verbose=0,
                                       n_jobs=n_jobs, random_state=22)
    random_search.fit(X_train_sel, y_train)
    logging.info(f"Best score (AUC): {random_search.best_score_}")
    logging.info(f"Best params: {random_search.best_params_}")
    best_model_tree = random_search.best_estimator_

    y_pred_test = best_model_tree.predict(X_test_sel)
    y_proba_test = best_model_tree.predict_proba(X_test_sel)[:, 1]
    metrics = {
        "auc": float(roc_auc_score(y_test, y_proba_test)),
        "classification_report": classification_report(y_test, y_pred_test, output_dict=True),
        "best_params": random_search.best_params_,
        "confusion_matrix": confusion_matrix(y_test, y_pred_test).tolist(),
    }
    logging.info(f"ROC AUC: {metrics['auc']:.4f}")
    return {"model": best_model_tree, "selected_features": selected_features, "metrics": metrics,
            "feature_importances": dict(zip(X_train_sel.columns, best_model_tree.feature_importances_.tolist())),
            "cv_best_score": float(random_search.best_score_)}
