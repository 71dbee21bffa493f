#!/usr/bin/env python
"""Wall time of the reference's real job shape (src/model_train_test/model_tree_train_test.py:73-179: RFE from F to 20
features with a 100-tree model, then RandomizedSearchCV 20 draws x 3 folds + refit) on synthetic data of the engineered
schema, on 1 GPU or -- under torch.distributed.run -- with the search's fold x candidate grid dealt over the GPUs.

  python tools/time_training_job.py [rows] [features]
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/time_training_job.py [rows] [features]

RFE is a chain (every elimination depends on the previous fit's importances), so with N GPUs every rank runs it on its
own GPU; the 60 cross-validation fits are independent and are spread."""
import os
import sys
import time

import numpy as np
import pandas as pd
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sklearn.model_selection import StratifiedKFold, train_test_split  # noqa: E402

from cobalt_smart_lender_ai_b200 import parallel  # noqa: E402
from cobalt_smart_lender_ai_b200.feature_selection import RFE  # noqa: E402
from cobalt_smart_lender_ai_b200.model_selection import RandomizedSearchCV  # noqa: E402
from cobalt_smart_lender_ai_b200.model_train_test.model_tree_train_test import PARAM_DISTRIBUTIONS  # noqa: E402
from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier  # noqa: E402
from cobalt_smart_lender_ai_b200.synth import make_lending  # noqa: E402

world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=__import__("datetime").timedelta(seconds=120))
rank = int(os.environ.get("RANK", "0"))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_200_000
f = int(sys.argv[2]) if len(sys.argv) > 2 else 104
X, y, names = make_lending(n, f, device=f"cuda:{local}")
df = pd.DataFrame(X.cpu().numpy(), columns=names)
yt = pd.Series(y.cpu().numpy().astype(int))
del X
X_train, X_test, y_train, y_test = train_test_split(df, yt, test_size=0.2, random_state=22)
spw = float((y_train == 0).sum() / (y_train == 1).sum())
dev = f"cuda:{local}"
XGBClassifier(n_estimators=4, device=dev).fit(X_train.iloc[:20000], y_train.iloc[:20000])     # warm-up


def timed(fn):
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    t0 = time.perf_counter()
    r = fn()
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    return r, time.perf_counter() - t0


def run_rfe():
    with parallel.local_only():
        return RFE(estimator=XGBClassifier(eval_metric="logloss", scale_pos_weight=spw, random_state=42, device=dev),
                   n_features_to_select=20, step=1).fit(X_train, y_train)


rfe, t_rfe = timed(run_rfe)
sel = X_train.columns[rfe.support_].tolist()
search, t_search = timed(lambda: RandomizedSearchCV(
    estimator=XGBClassifier(eval_metric="logloss", scale_pos_weight=spw, random_state=78, device=dev),
    param_distributions=PARAM_DISTRIBUTIONS, n_iter=20, scoring="roc_auc", cv=StratifiedKFold(3), random_state=22).fit(X_train[sel], y_train))
if rank == 0:
    from sklearn.metrics import roc_auc_score
    auc = roc_auc_score(y_test, search.predict_proba(X_test[sel])[:, 1])
    trees = sum(c["n_estimators"] for c in search.cv_results_["params"]) * 3 + search.best_params_["n_estimators"]
    print(f"training job, {n} rows x {f} features, {world} GPU(s): RFE {f}->20 ({f - 20 + 1} fits x 100 trees) {t_rfe:.2f} s | "
          f"search 20 x 3 + refit ({trees} trees, {search.jobs_per_rank_} of 60 fits on rank 0) {t_search:.2f} s | "
          f"total {t_rfe + t_search:.2f} s | best cv auc {search.best_score_:.5f} test auc {auc:.5f} | best {search.best_params_}",
          flush=True)
if world > 1:
    torch.distributed.destroy_process_group()
