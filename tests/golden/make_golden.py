"""Regenerates the committed golden vectors.  Run HERE (the container that has
/root/reference); the GPU box only ever reads the .npz files.

  python tests/golden/make_golden.py

1. fixture_forest.npz -- the reference's shipped model
   (/root/reference/src/api/models/xgb_model_tree.pkl, loaded at
   src/api/cobalt_fast_api.py:45) decoded WITHOUT unpickling (oracle/fixture.py):
   the xgboost-3.0.0 forest in wire format (a-9) plus its training parameters.
   It pins the oracle's gain formula, leaf formula, hessian additivity, the tree
   container and the traversal rule.
2. oracle_small.npz -- the oracle's own outputs on a small seeded synthetic input, so
   that a drift of the oracle itself is caught by the CPU suite.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))


def fixture():
    from oracle.fixture import forest_soa, load_fixture_model
    m = load_fixture_model("/root/reference/src/api/models/xgb_model_tree.pkl")
    f = forest_soa(m["Model"])
    tp = m["Config"]["learner"]["gradient_booster"]["tree_train_param"]
    obj = m["Model"]["learner"]["objective"]
    np.savez_compressed(
        os.path.join(HERE, "fixture_forest.npz"),
        tree_offsets=f["tree_offsets"], left=f["left_children"], right=f["right_children"],
        parent=f["parents"], split_index=f["split_indices"], split_cond=f["split_conditions"],
        default_left=f["default_left"], base_weight=f["base_weights"], loss_chg=f["loss_changes"],
        sum_hess=f["sum_hessian"], feature_names=np.array(f["feature_names"]),
        base_score=np.float32(f["base_score"]), eta=np.float32(tp["eta"]), gamma=np.float32(tp["gamma"]),
        reg_lambda=np.float32(tp["lambda"]), min_child_weight=np.float32(tp["min_child_weight"]),
        max_depth=np.int32(tp["max_depth"]), max_bin=np.int32(tp["max_bin"]),
        scale_pos_weight=np.float32(obj["reg_loss_param"]["scale_pos_weight"]),
        version=np.array(m["Model"]["version"], np.int32))


def oracle_small():
    from cobalt_smart_lender_ai_b200.synth import make_lending
    from oracle import oracle as O
    X, y, _ = make_lending(3000, 24, seed=11)
    X, y = X.numpy(), y.numpy()
    cuts, nc, mv = O.compute_cuts(X, 256)
    bins = O.bin_matrix(X, cuts, nc, ldb=32)
    bp = O.bin_ptr_of(nc)
    spw = float((y == 0).sum() / (y == 1).sum())
    sg, sh = O.quant_shifts(max(1.0, spw))
    tp = O.TreeParams(eta=0.3, max_depth=4, sg=sg, sh=sh)
    margin = np.zeros(len(y), np.float32)
    fo = O.train(bins, 24, bp, cuts, mv, y, None, spw, tp, 5, margin, want_rows=True, want_hist=True)
    np.savez_compressed(
        os.path.join(HERE, "oracle_small.npz"), X=X, y=y, cuts=cuts, n_cuts=nc, min_vals=mv, bins=bins,
        spw=np.float64(spw), sg=np.int32(sg), sh=np.int32(sh), margin=margin,
        **{f"forest_{k}": v for k, v in fo.items() if isinstance(v, np.ndarray)})


if __name__ == "__main__":
    fixture()
    oracle_small()
    for fn in sorted(os.listdir(HERE)):
        print(fn, os.path.getsize(os.path.join(HERE, fn)))
