"""TEST INFRASTRUCTURE -- numpy/ctypes front end of the CPU oracle (oracle/gbdt_oracle.c).

NOT PRODUCT CODE: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs import this module.  See the header of gbdt_oracle.c for what
it restates (xgboost 3.0.0 ``hist`` semantics behind the reference's
``XGBClassifier.fit/predict_proba`` call sites, src/model_train_test/model_tree_train_test.py:111-172,
src/api/cobalt_fast_api.py:90-91) and for the parity-pin status.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
QBITS = 17  # |q_g| <= 2^QBITS per row (R4); must equal the product's CB_QBITS


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "gbdt_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


class _TParams(C.Structure):
    _fields_ = [("eta", C.c_float), ("lambda_", C.c_float), ("alpha", C.c_float), ("gamma", C.c_float),
                ("mcw", C.c_float), ("max_depth", C.c_int), ("sg", C.c_int), ("sh", C.c_int)]


class _Split(C.Structure):
    _fields_ = [("loss_chg", C.c_float), ("feature", C.c_int), ("split_bin", C.c_int),
                ("default_left", C.c_int), ("threshold", C.c_float), ("GL", C.c_int64),
                ("HL", C.c_int64), ("GR", C.c_int64), ("HR", C.c_int64)]


class _Forest(C.Structure):
    _fields_ = [("left", C.c_void_p), ("right", C.c_void_p), ("parent", C.c_void_p),
                ("split_index", C.c_void_p), ("split_bin", C.c_void_p), ("split_cond", C.c_void_p),
                ("default_left", C.c_void_p), ("base_weight", C.c_void_p), ("loss_chg", C.c_void_p),
                ("sum_hess", C.c_void_p), ("n_nodes", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.orc_calc_gain.restype = C.c_double
        _lib.orc_calc_gain.argtypes = [C.c_double, C.c_double, C.c_float, C.c_float, C.c_float]
        _lib.orc_calc_weight.restype = C.c_double
        _lib.orc_calc_weight.argtypes = [C.c_double, C.c_double, C.c_float, C.c_float, C.c_float]
        _lib.orc_partition.restype = C.c_int64
        _lib.orc_hash.restype = C.c_uint32
        _lib.orc_hash.argtypes = [C.c_uint32, C.c_uint32, C.c_uint64]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def num_threads() -> int:
    return int(lib().orc_num_threads())


def set_threads(n: int) -> None:
    lib().orc_set_threads(C.c_int(n))


@dataclass
class TreeParams:
    eta: float = 0.3
    reg_lambda: float = 1.0
    reg_alpha: float = 0.0
    gamma: float = 0.0
    min_child_weight: float = 1.0
    max_depth: int = 6
    sg: int = 0
    sh: int = 0

    def c(self) -> _TParams:
        return _TParams(self.eta, self.reg_lambda, self.reg_alpha, self.gamma, self.min_child_weight,
                        self.max_depth, self.sg, self.sh)


def quant_shifts(wmax: float) -> tuple[int, int]:
    """R4: sg with wmax * 2^sg <= 2^QBITS; hessians (<= wmax/4) get two more bits."""
    e = math.frexp(float(wmax))[1]  # wmax = m * 2^e, 0.5 <= m < 1
    if float(wmax) == 2.0 ** (e - 1):
        e -= 1
    sg = QBITS - e
    return sg, sg + 2


def estimate_base_score(y: np.ndarray, spw: float, w: np.ndarray | None = None) -> float:
    """R3: one Newton step from margin 0 over the weighted data, then sigmoid."""
    y = np.asarray(y)
    wt = np.where(y > 0, np.float64(spw), 1.0)
    if w is not None:
        wt = wt * np.asarray(w, np.float64)
    sg = float(np.sum((0.5 - y.astype(np.float64)) * wt))
    sh = float(np.sum(0.25 * wt))
    m0 = -sg / sh if sh > 0 else 0.0
    # upstream stores base_score as an f32 probability
    return float(np.float32(1.0 / (1.0 + math.exp(-m0))))


def base_margin_of(base_score: float) -> float:
    """logit of the stored probability, rounded to f32 (the initial margin)."""
    p = float(base_score)
    return float(np.float32(-math.log(1.0 / p - 1.0)))


def compute_cuts(X: np.ndarray, max_bin: int = 256):
    X = np.ascontiguousarray(X, np.float32)
    n, f = X.shape
    cuts = np.empty((f, 256), np.float32)
    n_cuts = np.empty(f, np.int32)
    min_vals = np.empty(f, np.float32)
    lib().orc_cuts(_p(X), C.c_int64(n), C.c_int(f), C.c_int64(X.strides[0] // 4), C.c_int(max_bin),
                   _p(cuts), _p(n_cuts), _p(min_vals))
    return cuts, n_cuts, min_vals


def bin_ptr_of(n_cuts: np.ndarray) -> np.ndarray:
    bp = np.zeros(len(n_cuts) + 1, np.int32)
    np.cumsum(n_cuts, out=bp[1:])
    return bp


def bin_matrix(X: np.ndarray, cuts: np.ndarray, n_cuts: np.ndarray, ldb: int | None = None) -> np.ndarray:
    X = np.ascontiguousarray(X, np.float32)
    n, f = X.shape
    ldb = ldb or f
    bins = np.empty((n, ldb), np.uint8)
    lib().orc_bin(_p(X), C.c_int64(n), C.c_int(f), C.c_int64(f), _p(cuts), _p(n_cuts), _p(bins),
                  C.c_int64(ldb))
    return bins


def sigmoid(x: np.ndarray) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(x)
    lib().orc_sigmoid(_p(x), C.c_int64(x.size), _p(out))
    return out


def keep_threshold(subsample: float) -> int:
    return 0 if subsample >= 1.0 else int(math.floor(float(subsample) * 4294967296.0))


def grad(margin, y, w, spw, sg, sh, subsample=1.0, seed=0, tree=0, row_offset=0) -> np.ndarray:
    margin = np.ascontiguousarray(margin, np.float32)
    y = np.ascontiguousarray(y, np.uint8)
    w = None if w is None else np.ascontiguousarray(w, np.float32)
    q = np.empty((margin.size, 2), np.int32)
    lib().orc_grad(_p(margin), _p(y), _p(w), C.c_int64(margin.size), C.c_float(spw), C.c_int(sg), C.c_int(sh),
                   C.c_uint32(keep_threshold(subsample)), C.c_uint32(seed), C.c_uint32(tree),
                   C.c_int64(row_offset), _p(q))
    return q


def hist(bins: np.ndarray, f: int, q: np.ndarray, row_idx: np.ndarray | None, bin_ptr: np.ndarray) -> np.ndarray:
    bins = np.ascontiguousarray(bins)
    S = int(bin_ptr[f])
    out = np.zeros((S, 2), np.int64)
    cnt = bins.shape[0] if row_idx is None else len(row_idx)
    ri = None if row_idx is None else np.ascontiguousarray(row_idx, np.uint32)
    lib().orc_hist(_p(bins), C.c_int64(bins.shape[1]), C.c_int(f), _p(np.ascontiguousarray(q, np.int32)), _p(ri),
                   C.c_int64(cnt), _p(np.ascontiguousarray(bin_ptr, np.int32)), _p(out))
    return out


def scan(hist_node, bin_ptr, cuts, min_vals, Gt, Ht, tp: TreeParams, feature_mask=None) -> dict:
    f = len(bin_ptr) - 1
    sp = _Split()
    cp = tp.c()
    fm = None if feature_mask is None else np.ascontiguousarray(feature_mask, np.uint8)
    lib().orc_scan(_p(np.ascontiguousarray(hist_node, np.int64)), _p(np.ascontiguousarray(bin_ptr, np.int32)),
                   C.c_int(f), _p(np.ascontiguousarray(cuts, np.float32)),
                   _p(np.ascontiguousarray(min_vals, np.float32)), C.c_int64(int(Gt)), C.c_int64(int(Ht)),
                   C.byref(cp), _p(fm), C.byref(sp))
    return {k: getattr(sp, k) for k, _ in _Split._fields_}


def partition(bins, seg, feature, split_bin, default_left):
    bins = np.ascontiguousarray(bins)
    seg = np.ascontiguousarray(seg, np.uint32)
    out = np.empty_like(seg)
    nl = lib().orc_partition(_p(bins), C.c_int64(bins.shape[1]), _p(seg), C.c_int64(len(seg)), C.c_int(feature),
                             C.c_int(split_bin), C.c_int(default_left), _p(out))
    return out, int(nl)


_FOREST_KEYS = [("left", np.int32), ("right", np.int32), ("parent", np.int32), ("split_index", np.int32),
                ("split_bin", np.int32), ("split_cond", np.float32), ("default_left", np.uint8),
                ("base_weight", np.float32), ("loss_chg", np.float32), ("sum_hess", np.float32)]


def train(bins, f, bin_ptr, cuts, min_vals, y, w, spw, tp: TreeParams, n_trees, margin,
          subsample=1.0, seed=0, feature_masks=None, tree_offset=0, want_rows=False, want_hist=False):
    """Runs n_trees boosting rounds; updates ``margin`` in place; returns forest dict."""
    bins = np.ascontiguousarray(bins)
    n = bins.shape[0]
    max_nodes = (1 << (tp.max_depth + 1)) - 1
    arrs = {k: np.zeros((n_trees, max_nodes), dt) for k, dt in _FOREST_KEYS}
    arrs["n_nodes"] = np.zeros(n_trees, np.int32)
    fo = _Forest(*[_p(arrs[k]) for k, _ in _FOREST_KEYS], _p(arrs["n_nodes"]))
    rows = np.empty(n, np.uint32) if want_rows else None
    hd = np.zeros((int(bin_ptr[f]), 2), np.int64) if want_hist else None
    y = np.ascontiguousarray(y, np.uint8)
    w = None if w is None else np.ascontiguousarray(w, np.float32)
    fm = None if feature_masks is None else np.ascontiguousarray(feature_masks, np.uint8)
    assert margin.dtype == np.float32 and margin.flags.c_contiguous
    cp = tp.c()
    rc = lib().orc_train(_p(bins), C.c_int64(bins.shape[1]), C.c_int64(n), C.c_int(f),
                         _p(np.ascontiguousarray(bin_ptr, np.int32)), _p(np.ascontiguousarray(cuts, np.float32)),
                         _p(np.ascontiguousarray(min_vals, np.float32)), _p(y), _p(w), C.c_float(spw),
                         C.byref(cp), C.c_int(n_trees), C.c_int(tree_offset), C.c_uint32(keep_threshold(subsample)),
                         C.c_uint32(seed), _p(fm), _p(margin), C.byref(fo), _p(rows), _p(hd))
    if rc != 0:
        raise MemoryError("orc_train failed")
    arrs["max_nodes"] = max_nodes
    if want_rows:
        arrs["row_order"] = rows
    if want_hist:
        arrs["root_hist"] = hd
    return arrs


def flatten_forest(fo: dict) -> dict:
    """[T][max_nodes] padded arrays -> concatenated arrays + tree offsets (wire format a-9)."""
    nn = fo["n_nodes"]
    offs = np.zeros(len(nn) + 1, np.int64)
    np.cumsum(nn, out=offs[1:])
    out = {"tree_offsets": offs}
    for k, _ in _FOREST_KEYS:
        out[k] = np.concatenate([fo[k][t, :nn[t]] for t in range(len(nn))]) if len(nn) else fo[k].reshape(-1)
    return out


def tree_shap(X, tree_offsets, left, right, split_index, split_cond, default_left, sum_hess, base_margin=0.0):
    """R13: path-dependent TreeSHAP.  Returns f32 [n, f+1]; the last column is the expected value (incl. base_margin)."""
    X = np.ascontiguousarray(X, np.float32)
    n, f = X.shape
    out = np.empty((n, f + 1), np.float32)
    rc = lib().orc_tree_shap(_p(X), C.c_int64(n), C.c_int(f), C.c_int64(f),
                             _p(np.ascontiguousarray(tree_offsets, np.int64)), C.c_int(len(tree_offsets) - 1),
                             _p(np.ascontiguousarray(left, np.int32)), _p(np.ascontiguousarray(right, np.int32)),
                             _p(np.ascontiguousarray(split_index, np.int32)),
                             _p(np.ascontiguousarray(split_cond, np.float32)),
                             _p(np.ascontiguousarray(default_left, np.uint8)),
                             _p(np.ascontiguousarray(sum_hess, np.float32)), C.c_float(base_margin), _p(out))
    if rc != 0:
        raise ValueError("tree_shap: a tree is deeper than 12 levels")
    return out


def shapley_brute_force(x, tree_offsets, left, right, split_index, split_cond, default_left, sum_hess):
    """The DEFINITION TreeSHAP computes, by enumeration (f64; small forests only): Shapley values of
    v(S) = E[f(x) | x_S], the expectation taken down the tree with cover-weighted branches for features not in S.
    Returns (phi [f], expected value)."""
    import itertools
    import math
    x = np.asarray(x, np.float64)
    f = len(x)
    phi = np.zeros(f)
    base = 0.0
    for t in range(len(tree_offsets) - 1):
        o = int(tree_offsets[t])

        def ev(nid, S):
            if left[o + nid] == -1:
                return float(split_cond[o + nid])
            j = int(split_index[o + nid])
            l, r = int(left[o + nid]), int(right[o + nid])
            if j in S:
                v = x[j]
                go_left = bool(default_left[o + nid]) if np.isnan(v) else bool(np.float32(v) < split_cond[o + nid])
                return ev(l if go_left else r, S)
            cl, cr = float(sum_hess[o + l]), float(sum_hess[o + r])
            return (ev(l, S) * cl + ev(r, S) * cr) / (cl + cr)
        cnt = int(tree_offsets[t + 1]) - o
        used = sorted({int(split_index[o + i]) for i in range(cnt) if left[o + i] != -1})
        m = len(used)
        base += ev(0, frozenset())
        for i in used:
            rest = [u for u in used if u != i]
            for k in range(m):
                wgt = math.factorial(k) * math.factorial(m - k - 1) / math.factorial(m)
                for S in itertools.combinations(rest, k):
                    S = frozenset(S)
                    phi[i] += wgt * (ev(0, S | {i}) - ev(0, S))
    return phi, base


def predict_margin(X, tree_offsets, left, right, split_index, split_cond, default_left, base_margin=0.0):
    X = np.ascontiguousarray(X, np.float32)
    n, f = X.shape
    out = np.empty(n, np.float32)
    lib().orc_predict_margin(_p(X), C.c_int64(n), C.c_int(f), C.c_int64(f),
                             _p(np.ascontiguousarray(tree_offsets, np.int64)), C.c_int(len(tree_offsets) - 1),
                             _p(np.ascontiguousarray(left, np.int32)), _p(np.ascontiguousarray(right, np.int32)),
                             _p(np.ascontiguousarray(split_index, np.int32)),
                             _p(np.ascontiguousarray(split_cond, np.float32)),
                             _p(np.ascontiguousarray(default_left, np.uint8)), C.c_float(base_margin), _p(out))
    return out


def calc_gain(G, H, lam=1.0, alpha=0.0, mcw=0.0) -> float:
    return float(lib().orc_calc_gain(float(G), float(H), lam, alpha, mcw))


def calc_weight(G, H, lam=1.0, alpha=0.0, mcw=0.0) -> float:
    return float(lib().orc_calc_weight(float(G), float(H), lam, alpha, mcw))


def fit_reference(X, y, *, n_estimators=100, max_depth=6, learning_rate=0.3, scale_pos_weight=1.0, gamma=0.0,
                  reg_lambda=1.0, reg_alpha=0.0, min_child_weight=1.0, max_bin=256, subsample=1.0, seed=0,
                  feature_masks=None, base_score=None, sample_weight=None, ldb=None, want_rows=False,
                  want_hist=False):
    """The whole reference path on the CPU: cuts -> bins -> boosting rounds (R1..R10)."""
    X = np.ascontiguousarray(X, np.float32)
    y = np.ascontiguousarray(y, np.uint8)
    n, f = X.shape
    cuts, n_cuts, min_vals = compute_cuts(X, max_bin)
    bins = bin_matrix(X, cuts, n_cuts, ldb=ldb or (f + 15) // 16 * 16)
    bp = bin_ptr_of(n_cuts)
    wmax = 1.0 if sample_weight is None else float(np.max(sample_weight))
    sg, sh = quant_shifts(max(1.0, float(scale_pos_weight)) * wmax)
    if base_score is None:
        base_score = estimate_base_score(y, scale_pos_weight, sample_weight)
    tp = TreeParams(eta=float(np.float32(learning_rate)), reg_lambda=reg_lambda, reg_alpha=reg_alpha, gamma=gamma,
                    min_child_weight=min_child_weight, max_depth=max_depth, sg=sg, sh=sh)
    margin = np.full(n, base_margin_of(base_score), np.float32)
    fo = train(bins, f, bp, cuts, min_vals, y, sample_weight, float(scale_pos_weight), tp, n_estimators, margin,
               subsample=subsample, seed=seed, feature_masks=feature_masks, want_rows=want_rows, want_hist=want_hist)
    return {"cuts": cuts, "n_cuts": n_cuts, "min_vals": min_vals, "bins": bins, "bin_ptr": bp, "forest": fo,
            "flat": flatten_forest(fo), "margin": margin, "base_score": base_score, "sg": sg, "sh": sh, "tp": tp}


def predict_proba_reference(fit: dict, X) -> np.ndarray:
    ff = fit["flat"]
    m = predict_margin(X, ff["tree_offsets"], ff["left"], ff["right"], ff["split_index"], ff["split_cond"],
                       ff["default_left"], base_margin_of(fit["base_score"]))
    return sigmoid(m)
