"""``sklearn.model_selection.RandomizedSearchCV`` for the reference's tuning step
(src/model_train_test/model_tree_train_test.py:148-159: 20 draws x ``StratifiedKFold(3)``, ``scoring='roc_auc'``,
``random_state=22``; read afterwards: ``best_score_``, ``best_params_``, ``best_estimator_``, :161-166).

sklearn fits every (candidate, fold) pair from scratch: 60 uploads, sketches and binnings of the same three training
folds.  Cut points and bin codes depend on the rows of the fold and on ``max_bin`` only, so this class prepares every
fold ONCE (``XGBClassifier._prepare``: H2D, K1, K2), keeps its held-out rows on the device, and runs each candidate as
boosting rounds on the resident bins followed by a device-resident scoring pass.  Candidates, folds, fits and scores
are the ones sklearn's class would produce (same ``ParameterSampler``, same splitter, same scorer input), so
``cv_results_['mean_test_score']``, ``best_params_`` and ``best_score_`` are identical (checked in
tests/test_gpu_parity.py).  Anything else -- another estimator, another scorer, ``fit_params`` -- goes to sklearn's class.

With one process per GPU (``torch.distributed`` initialised, every process holding the same X and y) the fold x
candidate grid is dealt over the ranks: each rank runs WHOLE fits on its own GPU (``parallel.local_only``: no
collective inside a fit), the score matrix is summed over ranks at the end, and every rank refits the winner -- the
reference's ``n_jobs=-1`` process fan-out (:155) with GPUs in place of CPU workers.  Scores are bit-identical to the
single-GPU run because every fit is the same single-GPU fit.
"""
from __future__ import annotations

import numpy as np
import torch
from scipy.stats import rankdata
from sklearn.base import BaseEstimator, MetaEstimatorMixin, clone, is_classifier
from sklearn.metrics import roc_auc_score
from sklearn.model_selection import ParameterSampler
from sklearn.model_selection import RandomizedSearchCV as _SkRandomizedSearchCV
from sklearn.model_selection import check_cv
from sklearn.utils import _safe_indexing

from . import parallel
from .booster import Booster
from .sklearn import XGBClassifier, _as_matrix, _to_device

_PREPARE_KEYS = ("max_bin", "scale_pos_weight", "base_score", "device")   # what XGBClassifier._prepare reads


class RandomizedSearchCV(MetaEstimatorMixin, BaseEstimator):
    def __init__(self, estimator, param_distributions, *, n_iter=10, scoring=None, n_jobs=None, refit=True, cv=None,
                 verbose=0, random_state=None):
        self.estimator = estimator
        self.param_distributions = param_distributions
        self.n_iter = n_iter
        self.scoring = scoring
        self.n_jobs = n_jobs
        self.refit = refit
        self.cv = cv
        self.verbose = verbose
        self.random_state = random_state

    def _fast_path(self) -> bool:
        return isinstance(self.estimator, XGBClassifier) and self.scoring == "roc_auc"

    def _delegate(self, X, y, **fit_params):
        sk = _SkRandomizedSearchCV(self.estimator, self.param_distributions, n_iter=self.n_iter, scoring=self.scoring,
                                   n_jobs=self.n_jobs, refit=self.refit, cv=self.cv, verbose=self.verbose,
                                   random_state=self.random_state).fit(X, y, **fit_params)
        for k in ("cv_results_", "best_index_", "best_params_", "best_score_", "n_splits_"):
            setattr(self, k, getattr(sk, k))
        if self.refit:
            self.best_estimator_ = sk.best_estimator_
        self.prepared_folds_ = None
        return self

    def fit(self, X, y, **fit_params):
        if not self._fast_path() or fit_params:
            return self._delegate(X, y, **fit_params)
        cv = check_cv(self.cv, y, classifier=is_classifier(self.estimator))
        splits = list(cv.split(X, y))
        candidates = list(ParameterSampler(self.param_distributions, self.n_iter, random_state=self.random_state))
        scores = np.zeros((len(candidates), len(splits)), np.float64)
        prepared = 0
        d = parallel._dist()
        world, rank = (d.get_world_size(), d.get_rank()) if d is not None else (1, 0)
        # jobs in (fold, candidate) order, dealt in contiguous runs: a rank touches as few folds as possible
        jobs = [(k, ci) for k in range(len(splits)) for ci in range(len(candidates))]
        lo, hi = parallel.shard_rows(len(jobs), rank, world)
        mine = jobs[lo:hi]
        with parallel.local_only():
            for k, (train, test) in enumerate(splits):
                todo = [ci for kk, ci in mine if kk == k]
                if not todo:
                    continue
                X_tr, y_tr = _safe_indexing(X, train), _safe_indexing(y, train)
                X_te, y_te = _safe_indexing(X, test), np.asarray(_safe_indexing(y, test))
                ctxs: dict = {}                                  # one prepared fold per distinct (max_bin, spw, ...) setting
                Xte_d = None
                for ci in todo:
                    est = clone(self.estimator).set_params(**candidates[ci])
                    key = tuple(repr(getattr(est, a)) for a in _PREPARE_KEYS)
                    if key not in ctxs:
                        ctxs[key] = est._prepare(X_tr, y_tr)
                        prepared += 1
                    c = ctxs[key]
                    forest = est._boost(c)
                    booster = Booster(forest, c["f"], c["names"], c["base_score"])
                    with torch.cuda.device(c["dev"]):
                        if Xte_d is None:
                            Xte_d = _to_device(_as_matrix(X_te)[0], c["dev"])
                        proba = booster.predict_device(Xte_d).cpu().numpy()
                    scores[ci, k] = roc_auc_score(y_te, proba)   # what the 'roc_auc' scorer computes from predict_proba[:, 1]
                del ctxs, Xte_d
        if d is not None:                                        # every job ran on exactly one rank: the sum assembles the matrix
            dev = self.estimator._cuda_device() if d.get_backend() == "nccl" else "cpu"
            t = torch.from_numpy(scores).to(dev)
            d.all_reduce(t)
            scores = t.cpu().numpy()
        self.jobs_per_rank_ = len(mine)
        mean = scores.mean(axis=1)
        rank = np.asarray(rankdata(-mean, method="min"), dtype=np.int32)
        self.cv_results_ = {"params": candidates, "mean_test_score": mean, "std_test_score": scores.std(axis=1),
                            "rank_test_score": rank}
        for k in range(len(splits)):
            self.cv_results_[f"split{k}_test_score"] = scores[:, k]
        for name in sorted({p for c in candidates for p in c}):
            self.cv_results_[f"param_{name}"] = np.ma.masked_all(len(candidates), dtype=object)
            for ci, c in enumerate(candidates):
                if name in c:
                    self.cv_results_[f"param_{name}"][ci] = c[name]
        self.best_index_ = int(rank.argmin())
        self.best_params_ = candidates[self.best_index_]
        self.best_score_ = float(mean[self.best_index_])
        self.n_splits_ = len(splits)
        self.prepared_folds_ = prepared
        if self.refit:
            with parallel.local_only():                          # every rank keeps its own copy of the winner
                self.best_estimator_ = clone(self.estimator).set_params(**self.best_params_).fit(X, y)
        return self

    def predict(self, X):
        return self.best_estimator_.predict(X)

    def predict_proba(self, X):
        return self.best_estimator_.predict_proba(X)
