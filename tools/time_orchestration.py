#!/usr/bin/env python
"""Wall time of the reference script's two many-fit steps at a realistic size: this package's RFE /
RandomizedSearchCV (matrix prepared once, fits on resident bins) against sklearn's classes around the same estimator.
Usage: python tools/time_orchestration.py [rows] [features]"""
import os
import sys
import time

import numpy as np
import pandas as pd

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sklearn.feature_selection import RFE as SkRFE  # noqa: E402
from sklearn.model_selection import RandomizedSearchCV as SkSearch  # noqa: E402
from sklearn.model_selection import StratifiedKFold  # noqa: E402

from cobalt_smart_lender_ai_b200.feature_selection import RFE  # noqa: E402
from cobalt_smart_lender_ai_b200.model_selection import RandomizedSearchCV  # noqa: E402
from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier  # noqa: E402
from cobalt_smart_lender_ai_b200.synth import make_lending  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
f = int(sys.argv[2]) if len(sys.argv) > 2 else 60
X, y, names = make_lending(n, f, device="cuda")
df = pd.DataFrame(X.cpu().numpy(), columns=names)
yt = pd.Series(y.cpu().numpy().astype(int))
del X
base = XGBClassifier(eval_metric="logloss", scale_pos_weight=6.75, random_state=42, n_estimators=100)
base.fit(df.iloc[:20000], yt.iloc[:20000])                      # warm-up: library load, CUDA context


def timed(fn):
    t0 = time.perf_counter()
    r = fn()
    return r, time.perf_counter() - t0


steps = 10
fast, t_fast = timed(lambda: RFE(estimator=base, n_features_to_select=f - steps, step=1).fit(df, yt))
slow, t_slow = timed(lambda: SkRFE(estimator=base, n_features_to_select=f - steps, step=1).fit(df, yt))
assert np.array_equal(fast.ranking_, slow.ranking_)
print(f"RFE {f} -> {f - steps} (step 1, 100 trees per fit) on {n} rows: prepared once {t_fast:.2f} s, sklearn refits {t_slow:.2f} s "
      f"({t_slow / t_fast:.2f}x), identical ranking")
dist = {"n_estimators": [100, 200], "max_depth": [3, 5, 7], "learning_rate": [0.05, 0.1], "subsample": [0.8, 1.0],
        "colsample_bytree": [0.5, 0.8, 1.0], "gamma": [0, 1, 5]}
kw = dict(param_distributions=dist, n_iter=6, scoring="roc_auc", cv=StratifiedKFold(3), random_state=22, refit=False)
fs, t_fast = timed(lambda: RandomizedSearchCV(estimator=base, **kw).fit(df, yt))
ss, t_slow = timed(lambda: SkSearch(estimator=base, n_jobs=1, **kw).fit(df, yt))
assert np.array_equal(fs.cv_results_["mean_test_score"], ss.cv_results_["mean_test_score"])
print(f"RandomizedSearchCV 6 draws x 3 folds on {n} rows x {f}: prepared folds {t_fast:.2f} s, sklearn {t_slow:.2f} s "
      f"({t_slow / t_fast:.2f}x), identical scores")
