"""Worker of tests/test_multi_gpu.py (one process per GPU): the tuning step's fold x candidate grid dealt over the ranks
(whole fits per GPU, no collective inside a fit) must give exactly the scores of the same search on one GPU."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sklearn.model_selection import StratifiedKFold  # noqa: E402

from cobalt_smart_lender_ai_b200 import parallel  # noqa: E402
from cobalt_smart_lender_ai_b200.model_selection import RandomizedSearchCV  # noqa: E402
from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier  # noqa: E402
from cobalt_smart_lender_ai_b200.synth import make_lending  # noqa: E402


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=__import__("datetime").timedelta(seconds=120))
    rank, world = torch.distributed.get_rank(), torch.distributed.get_world_size()
    X, y, _ = make_lending(30000, 24, seed=21)
    X, y = X.numpy(), y.numpy().astype(int)
    base = XGBClassifier(scale_pos_weight=6.0, random_state=78, device=f"cuda:{local}")
    dist = {"n_estimators": [5, 10], "max_depth": [3, 5], "learning_rate": [0.1, 0.3], "subsample": [0.8, 1.0],
            "colsample_bytree": [0.5, 1.0], "gamma": [0, 1]}
    kw = dict(param_distributions=dist, n_iter=5, scoring="roc_auc", cv=StratifiedKFold(3), random_state=22, refit=True)
    spread = RandomizedSearchCV(estimator=base, **kw).fit(X, y)
    with parallel.local_only():                                   # the same search, all 15 fits on this GPU
        alone = RandomizedSearchCV(estimator=base, **kw).fit(X, y)
    ok = (spread.jobs_per_rank_ < 15 and alone.jobs_per_rank_ == 15
          and np.array_equal(spread.cv_results_["mean_test_score"], alone.cv_results_["mean_test_score"])
          and all(np.array_equal(spread.cv_results_[f"split{k}_test_score"], alone.cv_results_[f"split{k}_test_score"]) for k in range(3))
          and spread.best_params_ == alone.best_params_
          and np.array_equal(spread.predict_proba(X[:500]), alone.predict_proba(X[:500])))
    flag = torch.tensor([0 if ok else 1], device="cuda")
    torch.distributed.all_reduce(flag)
    if rank == 0:
        print(f"[search] world={world} jobs on rank 0: {spread.jobs_per_rank_} of 15 -> {'OK' if int(flag.item()) == 0 else 'MISMATCH'}", flush=True)
    code = int(flag.item())
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(1 if code else 0)


if __name__ == "__main__":
    main()
