"""Drop-in for ``xgboost.XGBClassifier`` on the surface the reference uses
(SURVEY.md 8b): construct / clone / set_params / fit / predict / predict_proba /
feature_importances_ / get_booster().get_score / pickling.

Reference call sites: src/model_train_test/model_tree_train_test.py:111-118 (RFE base),
:132-159 (RandomizedSearchCV base), :171-172 (predict / predict_proba), :197
(feature_importances_), :217 (joblib.dump); src/api/cobalt_fast_api.py:45,91,135-136.

Everything numeric runs in libcobalt_gbdt.so on the GPU; there is no CPU path, so ``fit``
and ``predict*`` raise ``CobaltError`` without a CUDA device.  No CUDA state lives in the
estimator object itself (an unfitted clone pickles into worker processes safely).
"""
from __future__ import annotations

import math
import threading

import numpy as np
import torch
from sklearn.base import BaseEstimator, ClassifierMixin

from . import _lib, ops
from .binned import BinnedMatrix
from .booster import Booster, base_margin_of
from .trainer import GpuTrainer

_LOCK = threading.Lock()   # predict_proba may be called from FastAPI's thread pool


def _hash32(seed: int, a: int, b: int) -> int:
    """Host copy of the library's counter hash (colsample feature draws; R-colsample)."""
    M = 0xFFFFFFFF
    rot = lambda x: ((x << 13) | (x >> 19)) & M  # noqa: E731
    h = (seed ^ 0x9E3779B9) & M
    h ^= (a * 0x85EBCA6B) & M; h = rot(h); h = (h * 5 + 0xE6546B64) & M
    h ^= ((b & M) * 0xCC9E2D51) & M; h = rot(h); h = (h * 5 + 0xE6546B64) & M
    h ^= ((b >> 32) * 0x1B873593) & M
    h ^= h >> 16; h = (h * 0x85EBCA6B) & M; h ^= h >> 13; h = (h * 0xC2B2AE35) & M; h ^= h >> 16
    return h


def colsample_masks(n_trees: int, n_features: int, colsample: float, seed: int) -> np.ndarray | None:
    """Per-tree feature subsets: the k = max(1, int(colsample * F)) features with the smallest
    hash(seed, tree, feature).  Own generator: parity with upstream's sampler is statistical only."""
    if colsample >= 1.0:
        return None
    k = max(1, int(colsample * n_features))
    masks = np.zeros((n_trees, n_features), np.uint8)  # row t masks tree t
    for t in range(n_trees):
        keys = [(_hash32(seed ^ 0x5bd1e995, t, j), j) for j in range(n_features)]
        keys.sort()
        masks[t, [j for _, j in keys[:k]]] = 1
    return masks


def estimate_base_score(n_pos_w: float, n_neg_w: float) -> float:
    """R3: one Newton step from margin 0 on the weighted data, then sigmoid, stored as f32."""
    sg = 0.5 * n_neg_w - 0.5 * n_pos_w
    sh = 0.25 * (n_pos_w + n_neg_w)
    m0 = -sg / sh if sh > 0 else 0.0
    return float(np.float32(1.0 / (1.0 + math.exp(-m0))))


_STREAM_BYTES = 256 << 20     # host matrices above this size are scored in double-buffered chunks


def _as_matrix(X):
    """DataFrame (float/int/bool columns) / ndarray / CUDA tensor -> (array-like f32, names or None)."""
    names = None
    if hasattr(X, "columns") and hasattr(X, "to_numpy"):
        names = [str(c) for c in X.columns]
        X = X.to_numpy(dtype=np.float32, na_value=np.nan) if hasattr(X, "isna") else X.to_numpy(dtype=np.float32)
    if isinstance(X, torch.Tensor):
        if X.dim() != 2:
            raise ValueError("X must be 2-dimensional")
        return X.to(torch.float32), names
    X = np.asarray(X)
    if X.ndim != 2:
        raise ValueError("X must be 2-dimensional")
    return np.ascontiguousarray(X, dtype=np.float32), names


def _to_device(a, device) -> torch.Tensor:
    if isinstance(a, torch.Tensor):
        return a.to(device).contiguous()
    t = torch.from_numpy(a)
    try:
        t = t.pin_memory()
    except Exception:
        pass
    return t.to(device, non_blocking=True)


class XGBClassifier(ClassifierMixin, BaseEstimator):
    """Binary logistic histogram GBDT with xgboost's estimator interface and defaults.

    ``max_bin``: bin codes are one byte and code 255 is reserved for missing values, so a feature gets at most 255
    real bins: ``max_bin=256`` (xgboost's default and the shipped model's setting) is accepted and means 255 here (a
    declared deviation); larger values are refused rather than silently clamped."""

    def __init__(self, n_estimators=100, max_depth=6, learning_rate=0.3, subsample=1.0, colsample_bytree=1.0,
                 gamma=0.0, scale_pos_weight=1.0, random_state=0, eval_metric=None, use_label_encoder=None,
                 max_bin=256, min_child_weight=1.0, reg_lambda=1.0, reg_alpha=0.0, base_score=None,
                 missing=np.nan, n_jobs=None, device="cuda", objective="binary:logistic", tree_method="hist",
                 importance_type="gain"):
        self.n_estimators = n_estimators
        self.max_depth = max_depth
        self.learning_rate = learning_rate
        self.subsample = subsample
        self.colsample_bytree = colsample_bytree
        self.gamma = gamma
        self.scale_pos_weight = scale_pos_weight
        self.random_state = random_state
        self.eval_metric = eval_metric
        self.use_label_encoder = use_label_encoder
        self.max_bin = max_bin
        self.min_child_weight = min_child_weight
        self.reg_lambda = reg_lambda
        self.reg_alpha = reg_alpha
        self.base_score = base_score
        self.missing = missing
        self.n_jobs = n_jobs
        self.device = device
        self.objective = objective
        self.tree_method = tree_method
        self.importance_type = importance_type

    # ------------------------------------------------------------------ helpers
    def __sklearn_tags__(self):
        tags = super().__sklearn_tags__()
        tags.input_tags.allow_nan = True     # missing values are routed by the trees (as xgboost's estimator declares)
        return tags

    def _cuda_device(self) -> torch.device:
        if not torch.cuda.is_available():
            raise _lib.CobaltError("XGBClassifier (cobalt B200 path) needs a CUDA device; there is no CPU fallback")
        dev = torch.device(self.device if self.device not in (None, "gpu") else "cuda")
        if dev.type != "cuda":
            raise _lib.CobaltError(f"device={self.device!r}: only CUDA devices are supported")
        if dev.index is None:
            dev = torch.device("cuda", torch.cuda.current_device())
        return dev

    def _validate_params(self):
        if self.objective not in ("binary:logistic", None):
            raise ValueError("only objective='binary:logistic' is implemented")
        if not (isinstance(self.missing, float) and math.isnan(self.missing)):
            raise ValueError("only missing=nan is implemented")
        if not 1 <= int(self.max_depth) <= 9:
            raise ValueError("max_depth must be in [1, 9]")
        if not 2 <= int(self.max_bin) <= 256:
            raise ValueError("max_bin must be in [2, 256] (one-byte bin codes; 256 means 255 real bins + the missing code)")

    # ------------------------------------------------------------------ fit
    def _prepare(self, X, y, sample_weight=None) -> dict:
        """Everything of fit() that does not depend on which features may be split on: H2D, label statistics, K1 cuts,
        K2 bins (+ the position-major copy).  feature_selection.RFE calls it once and then boosts many times."""
        self._validate_params()
        Xa, names = _as_matrix(X)
        dev = self._cuda_device()
        y_np = np.asarray(y.to_numpy() if hasattr(y, "to_numpy") else (y.cpu().numpy() if isinstance(y, torch.Tensor) else y))
        classes = np.unique(y_np)
        if not set(classes.tolist()) <= {0, 1}:
            raise ValueError(f"labels must be in {{0, 1}}, got {classes}")
        if len(y_np) != Xa.shape[0]:
            raise ValueError("X and y have inconsistent lengths")
        y_u8 = np.ascontiguousarray(y_np, dtype=np.uint8)
        w_np = None if sample_weight is None else np.ascontiguousarray(sample_weight, dtype=np.float32)
        n, f = Xa.shape
        spw = float(self.scale_pos_weight)

        from . import parallel
        world = parallel.world_size()
        with torch.cuda.device(dev):
            Xd = _to_device(Xa, dev)
            yd = _to_device(y_u8, dev)
            wd = None if w_np is None else _to_device(w_np, dev)
            # label statistics, weight range and row offset: global when rows are sharded over ranks
            pos_w = float((w_np[y_u8 == 1].astype(np.float64).sum() if w_np is not None else (y_u8 == 1).sum()) * spw)
            neg_w = float(w_np[y_u8 == 0].astype(np.float64).sum() if w_np is not None else (y_u8 == 0).sum())
            wmax_local = float(w_np.max()) if w_np is not None and len(w_np) else 1.0
            pos_w, neg_w, wmax_s, row_offset = parallel.global_stats(pos_w, neg_w, wmax_local, n, dev if world > 1 else None)
            base_score = float(self.base_score) if self.base_score is not None else estimate_base_score(pos_w, neg_w)
            sg, sh = ops.quant_shifts(max(1.0, spw) * wmax_s)
            cuts, n_cuts, min_vals = parallel.sketch_cuts_global(Xd, int(self.max_bin))
            bm = BinnedMatrix.from_dense(Xd, cuts, n_cuts)
            del Xd
            bin_ptr = ops.bin_ptr_of(n_cuts)
        return {"dev": dev, "names": names, "n": n, "f": f, "spw": spw, "yd": yd, "wd": wd, "base_score": base_score,
                "sg": sg, "sh": sh, "row_offset": row_offset, "cuts": cuts, "min_vals": min_vals, "bm": bm, "bin_ptr": bin_ptr}

    def _boost(self, c: dict, allowed: np.ndarray | None = None) -> dict:
        """n_estimators rounds on prepared bins -> forest (host arrays).  allowed: u8 [f], features a split may use
        (None: all); combined with the per-tree colsample masks."""
        f, dev = c["f"], c["dev"]
        T = int(self.n_estimators)
        seed = int(self.random_state or 0)
        masks = colsample_masks(T, f, float(self.colsample_bytree), seed)
        if allowed is not None:
            a = np.ascontiguousarray(allowed, np.uint8).reshape(1, f)
            masks = np.repeat(a, max(T, 1), axis=0) if masks is None else (masks & a)
        with torch.cuda.device(dev):
            masks_d = None if masks is None else torch.from_numpy(np.ascontiguousarray(masks)).to(dev)
            tr = GpuTrainer(c["bm"], c["yd"], c["wd"], f, c["bin_ptr"], c["cuts"], c["min_vals"], max_depth=int(self.max_depth),
                            eta=float(self.learning_rate), reg_lambda=float(self.reg_lambda),
                            reg_alpha=float(self.reg_alpha), gamma=float(self.gamma),
                            min_child_weight=float(self.min_child_weight), scale_pos_weight=c["spw"], sg=c["sg"], sh=c["sh"],
                            subsample=float(self.subsample), seed=seed, max_trees=max(T, 1),
                            base_margin=base_margin_of(c["base_score"]), row_offset=c["row_offset"])
            tr.boost(T, masks_d)
            forest = tr.forest_numpy()
            self._gpu_launches = tr.launch_count()
            tr.close()
        return forest

    def fit(self, X, y, sample_weight=None):
        c = self._prepare(X, y, sample_weight)
        forest = self._boost(c)
        f, names = c["f"], c["names"]
        self._Booster = Booster(forest, f, names, c["base_score"],
                                attributes={"sg": c["sg"], "sh": c["sh"], "scale_pos_weight": c["spw"]})
        self.classes_ = np.array([0, 1])
        self.n_classes_ = 2
        self.n_features_in_ = f
        if names is not None:
            self.feature_names_in_ = np.array(names, dtype=object)
        elif hasattr(self, "feature_names_in_"):
            del self.feature_names_in_
        return self

    # ------------------------------------------------------------------ predict
    def get_booster(self) -> Booster:
        if not hasattr(self, "_Booster"):
            from sklearn.exceptions import NotFittedError
            raise NotFittedError("need to call fit or load_model beforehand")
        return self._Booster

    def _check_names(self, names):
        want = self.get_booster().feature_names
        if want is not None and names is not None and list(names) != list(want):
            missing = [c for c in want if c not in names]
            extra = [c for c in names if c not in want]
            raise ValueError("feature_names mismatch: " + (f"expected {missing} in input data; " if missing else "") +
                             (f"training data did not have the following fields: {extra}" if extra else
                              "columns are in a different order"))

    def _proba_device(self, X) -> torch.Tensor:
        Xa, names = _as_matrix(X)
        self._check_names(names)
        dev = self._cuda_device()
        with _LOCK, torch.cuda.device(dev):
            if isinstance(Xa, np.ndarray) and Xa.nbytes > _STREAM_BYTES:
                return self._proba_streamed(Xa, dev)
            Xd = _to_device(Xa, dev)
            return self.get_booster().predict_device(Xd)

    def _proba_streamed(self, Xa: np.ndarray, dev) -> torch.Tensor:
        """Large host matrices: rows go to the GPU in chunks on a copy stream, double buffered, while the previous
        chunk is scored -- the H2D copy (the slower of the two by far) hides the kernels entirely."""
        n, f = Xa.shape
        rows = max(1, _STREAM_BYTES // (4 * f))
        src = torch.from_numpy(Xa)
        pinned = src.is_pinned()
        cur = torch.cuda.current_stream(dev)
        cs = torch.cuda.Stream(dev)
        bufs = [torch.empty((rows, f), dtype=torch.float32, device=dev) for _ in range(2)]
        stage = None if pinned else [torch.empty((rows, f), dtype=torch.float32).pin_memory() for _ in range(2)]
        copied = [torch.cuda.Event() for _ in range(2)]
        done = [torch.cuda.Event() for _ in range(2)]
        out = torch.empty(n, dtype=torch.float32, device=dev)
        booster = self.get_booster()
        for k, r0 in enumerate(range(0, n, rows)):
            r1, b = min(n, r0 + rows), k & 1
            m = r1 - r0
            if stage is not None:
                if k >= 2:
                    copied[b].synchronize()                 # the staging buffer has left for the device
                stage[b][:m].copy_(src[r0:r1])
            with torch.cuda.stream(cs):
                if k >= 2:
                    cs.wait_event(done[b])                  # the device buffer has been scored
                bufs[b][:m].copy_(src[r0:r1] if stage is None else stage[b][:m], non_blocking=True)
                copied[b].record(cs)
            cur.wait_event(copied[b])
            out[r0:r1] = booster.predict_device(bufs[b][:m])
            done[b].record(cur)
        cur.synchronize()
        return out

    def predict_proba(self, X) -> np.ndarray:
        p = self._proba_device(X).cpu().numpy()
        return np.stack([1.0 - p, p], axis=1).astype(np.float32)

    def predict(self, X) -> np.ndarray:
        p = self._proba_device(X).cpu().numpy()
        return (p > 0.5).astype(np.int64)

    @property
    def feature_importances_(self) -> np.ndarray:
        b = self.get_booster()
        score = b.get_score(importance_type=self.importance_type or "gain")
        names = b._names()
        v = np.array([score.get(nm, 0.0) for nm in names], dtype=np.float32)
        tot = v.sum()
        return v / tot if tot > 0 else v

    def load_fixture_forest(self, forest: dict, feature_names, base_score: float = 0.5):
        """Adopt a forest in wire format (e.g. the reference's shipped model)."""
        self._Booster = Booster(forest, len(feature_names), list(feature_names), base_score)
        self.classes_ = np.array([0, 1]); self.n_classes_ = 2
        self.n_features_in_ = len(feature_names)
        self.feature_names_in_ = np.array(list(feature_names), dtype=object)
        return self
