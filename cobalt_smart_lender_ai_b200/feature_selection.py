"""``sklearn.feature_selection.RFE`` for the reference's feature-selection step
(src/model_train_test/model_tree_train_test.py:111-118: ``RFE(estimator=XGBClassifier(...), n_features_to_select=20,
step=1).fit(X_train, y_train)``; only ``support_`` is read, :120).

sklearn's RFE refits the estimator on a column subset at every step: 130 fits from 150 to 20 features, each one
uploading, sketching and binning the matrix again.  Cut points and bin codes of a feature do not depend on the other
features, so this class bins ONCE (``XGBClassifier._prepare``), and every elimination step is just
``n_estimators`` boosting rounds on the resident bins with a feature mask (the colsample mechanism of the split scan).
Trees, importances and therefore ``support_`` / ``ranking_`` are identical to sklearn's RFE around the same estimator
(checked in tests/test_gpu_parity.py).  Anything this shortcut does not cover -- another estimator,
``colsample_bytree < 1`` (its masks depend on the column count), an ``importance_getter`` -- goes to sklearn's RFE.
"""
from __future__ import annotations

import numpy as np
from sklearn.base import BaseEstimator, MetaEstimatorMixin, clone
from sklearn.feature_selection import RFE as _SkRFE
from sklearn.feature_selection import SelectorMixin

from .booster import Booster
from .sklearn import XGBClassifier


class RFE(SelectorMixin, MetaEstimatorMixin, BaseEstimator):
    def __init__(self, estimator, *, n_features_to_select=None, step=1, verbose=0, importance_getter="auto"):
        self.estimator = estimator
        self.n_features_to_select = n_features_to_select
        self.step = step
        self.verbose = verbose
        self.importance_getter = importance_getter

    def _fast_path(self) -> bool:
        e = self.estimator
        return (isinstance(e, XGBClassifier) and float(e.colsample_bytree) == 1.0 and self.importance_getter == "auto")

    def fit(self, X, y, **fit_params):
        if not self._fast_path() or fit_params:
            sk = _SkRFE(self.estimator, n_features_to_select=self.n_features_to_select, step=self.step,
                        verbose=self.verbose, importance_getter=self.importance_getter).fit(X, y, **fit_params)
            self.support_, self.ranking_, self.n_features_, self.estimator_ = sk.support_, sk.ranking_, sk.n_features_, sk.estimator_
            self.n_features_in_ = sk.n_features_in_
            if hasattr(sk, "feature_names_in_"):
                self.feature_names_in_ = sk.feature_names_in_
            self.fits_, self.binned_once_ = None, False
            return self
        est = clone(self.estimator)
        c = est._prepare(X, y)
        f = c["f"]
        # the same bookkeeping as sklearn's RFE._fit (sklearn/feature_selection/_rfe.py), step by step
        if self.n_features_to_select is None:
            n_select = f // 2
        elif isinstance(self.n_features_to_select, (int, np.integer)):
            n_select = int(self.n_features_to_select)
        else:
            n_select = int(f * self.n_features_to_select)
        step = int(max(1, self.step * f)) if 0.0 < self.step < 1.0 else int(self.step)
        if step <= 0:
            raise ValueError("Step must be >0")
        support = np.ones(f, dtype=bool)
        ranking = np.ones(f, dtype=int)
        fits = 0
        while support.sum() > n_select:
            features = np.arange(f)[support]
            forest = est._boost(c, allowed=support.astype(np.uint8))
            fits += 1
            v = _raw_importances(forest, f, est.importance_type or "gain")[features]
            tot = v.sum()                                    # summed over the remaining columns, in their order: the
            imp = v / tot if tot > 0 else v                  # very numbers feature_importances_ of a refit would hold
            importances = imp ** 2                           # sklearn squares what feature_importances_ returns
            ranks = np.ravel(np.argsort(importances, kind="stable"))   # ties: first column first, as sklearn (>= 1.6)
            threshold = min(step, int(support.sum()) - n_select)
            support[features[ranks][:threshold]] = False
            ranking[np.logical_not(support)] += 1
        self.support_, self.ranking_, self.n_features_ = support, ranking, int(support.sum())
        self.n_features_in_ = f
        if c["names"] is not None:
            self.feature_names_in_ = np.array(c["names"], dtype=object)
        self.fits_, self.binned_once_ = fits, True
        del c
        Xa = X.loc[:, support] if hasattr(X, "loc") else np.asarray(X)[:, support]
        self.estimator_ = clone(self.estimator).fit(Xa, y)   # like sklearn: the final estimator sees the reduced matrix
        return self

    def _get_support_mask(self):
        return self.support_

    def __sklearn_tags__(self):
        tags = super().__sklearn_tags__()
        tags.input_tags.allow_nan = True                     # the trees route missing values themselves
        return tags

    def predict(self, X):
        return self.estimator_.predict(self.transform(X))

    def predict_proba(self, X):
        return self.estimator_.predict_proba(self.transform(X))


def _raw_importances(forest: dict, f: int, importance_type: str) -> np.ndarray:
    """Per-column importance of a forest over all f columns, before normalisation (features never split on get 0)."""
    b = Booster(forest, f, None, 0.5)
    score = b.get_score(importance_type=importance_type)
    return np.array([score.get(f"f{i}", 0.0) for i in range(f)], dtype=np.float32)
