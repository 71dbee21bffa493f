"""Host-side forest container in xgboost's wire format and its scoring
entry point.  Mirrors the parts of ``xgboost.Booster`` the reference touches:
``get_booster().get_score(importance_type='gain')`` (src/api/cobalt_fast_api.py:135-136),
``feature_names`` and prediction behind ``predict_proba`` (src/api/cobalt_fast_api.py:91)."""
from __future__ import annotations

import math
import os

import numpy as np
import torch

from . import _lib, ops

_FIELDS = ("left", "right", "parent", "split_index", "split_cond", "default_left", "base_weight", "loss_chg",
           "sum_hess")


def base_margin_of(base_score: float) -> float:
    p = float(base_score)
    return float(np.float32(-math.log(1.0 / p - 1.0)))


def pack_forest(tree_offsets, left, right, split_index, split_cond, default_left):
    """Wire-format arrays -> packed 8-byte nodes for cb_predict.
    node = cond bits | meta << 32; meta = feature | default_left << 15 | left_child << 16 (0 = leaf)."""
    left = np.asarray(left, np.int64)
    right = np.asarray(right, np.int64)
    internal = left >= 0
    if np.any(right[internal] != left[internal] + 1):
        raise _lib.CobaltError("forest nodes must allocate children consecutively (right = left + 1)")
    sizes = np.diff(np.asarray(tree_offsets, np.int64))
    if sizes.size and sizes.max() > 65535:
        raise _lib.CobaltError("a tree has more than 65535 nodes")
    feat = np.asarray(split_index, np.int64)
    if internal.any() and feat[internal].max() > 32767:
        raise _lib.CobaltError("feature index above 32767")
    meta = np.where(internal, feat | (np.asarray(default_left, np.int64) << 15) | (left << 16), 0).astype(np.uint64)
    cond = np.ascontiguousarray(split_cond, np.float32).view(np.uint32).astype(np.uint64)
    nodes = (cond | (meta << np.uint64(32))).astype(np.uint64)
    return nodes.view(np.int64), np.asarray(tree_offsets, np.int32)


def heap_forest(tree_offsets, left, right, split_index, split_cond, default_left, max_depth_cap: int = 10):
    """Wire-format forest -> complete trees in heap order (cb_predict_heap).  Where a branch ends
    above the common depth the leaf is carried down through pass-through nodes whose two children
    are the same.  Returns (thr f32 [T,NI], meta u16 [T,NI], leaf f32 [T,NL], depth) or None when
    a tree is deeper than max_depth_cap."""
    offs = np.asarray(tree_offsets, np.int64)
    T = len(offs) - 1
    left = np.asarray(left, np.int64)
    right = np.asarray(right, np.int64)
    feat = np.asarray(split_index, np.int64)
    cond = np.asarray(split_cond, np.float32)
    dl = np.asarray(default_left, np.int64)
    if T == 0:
        return None
    cur = offs[:-1].copy()[:, None]                      # [T, nodes at this level] absolute node ids
    levels = []
    depth = 0
    while True:
        internal = left[cur] >= 0
        if not internal.any():
            break
        if depth >= max_depth_cap:
            return None
        levels.append((cur, internal))
        base = np.repeat(offs[:-1][:, None], cur.shape[1], axis=1)
        lch = np.where(internal, base + left[cur], cur)
        rch = np.where(internal, base + right[cur], cur)
        cur = np.stack([lch, rch], axis=2).reshape(T, -1)
        depth += 1
    if depth == 0:                                       # forest of single-leaf trees: one pass-through level
        levels.append((cur, np.zeros_like(cur, bool)))
        cur = np.concatenate([cur, cur], axis=1)
        depth = 1
    NI = (1 << depth) - 1
    thr = np.zeros((T, NI), np.float32)
    meta = np.zeros((T, NI), np.uint16)
    pos = 0
    for ids, internal in levels:
        w = ids.shape[1]
        thr[:, pos:pos + w] = np.where(internal, cond[ids], np.float32(0.0))
        meta[:, pos:pos + w] = np.where(internal, feat[ids] | (dl[ids] << 15), 0).astype(np.uint16)
        pos += w
    return thr, meta, np.ascontiguousarray(cond[cur]), depth


def q8_forest(thr: np.ndarray, meta: np.ndarray, leaf: np.ndarray, depth: int, n_features: int, plane_base: int = 0):
    """heap_forest output -> what cb_predict_q8 needs: the forest's threshold table (per feature: distinct split_cond
    values ascending, then +inf) for cb_bin_pack_u8, and the u32 forest [T, 2 << depth] (1-based heap: nodes, then
    leaves).  "x < t_k" <=> "code <= k" with code = number of thresholds <= x.  None when a feature has more than
    254 distinct thresholds or the feature offsets do not fit 16 bits."""
    T, NI = thr.shape
    NL = 1 << depth
    if n_features > 255:
        return None
    feat = (meta & 0x7FFF).astype(np.int64)
    dl = (meta >> 15).astype(np.int64)
    # internal (real) nodes are those whose two subtrees differ; pass-through nodes carry thr 0 / meta 0 and may go
    # either way, so they simply take part as "feature 0, threshold 0.0" comparisons when they are not filtered here
    cuts = np.full((n_features, 256), np.inf, np.float32)
    n_cuts = np.ones(n_features, np.int32)
    kidx = np.zeros((T, NI), np.int64)
    for j in range(n_features):
        sel = feat == j
        if not sel.any():
            continue
        vals = thr[sel]
        if np.isnan(vals).any():
            return None
        u, inv = np.unique(vals, return_inverse=True)        # -0.0 == 0.0 collapse, like the f32 compare
        if len(u) > 254:
            return None
        cuts[j, :len(u)] = u
        n_cuts[j] = len(u) + 1                               # + the +inf sentinel: codes run 0 .. len(u)
        kidx[sel] = inv.reshape(-1)
    span = kidx + dl                                         # plane 1 (default left) holds code + 1, missing = 0
    off = plane_base + (dl * n_features + feat) * 128        # shared-window address of the (plane, feature) column
    if off.max(initial=0) + 127 > 0xFFFF or span.max(initial=0) > 255:
        return None
    words = np.zeros((T, 2 * NL), np.uint32)
    words[:, 1:NL] = ((span << 16) | off).astype(np.uint32)
    words[:, NL:] = np.ascontiguousarray(leaf, np.float32).view(np.uint32)
    return cuts, n_cuts, words


class Booster:
    def __init__(self, forest: dict, n_features: int, feature_names=None, base_score: float = 0.5,
                 attributes: dict | None = None):
        self.forest = {k: np.ascontiguousarray(forest[k]) for k in _FIELDS if k in forest}
        self.forest["tree_offsets"] = np.asarray(forest["tree_offsets"], np.int64)
        if "split_bin" in forest:
            self.forest["split_bin"] = np.ascontiguousarray(forest["split_bin"])
        self.n_features = int(n_features)
        self.feature_names = list(feature_names) if feature_names is not None else None
        self.base_score = float(base_score)
        self.attributes = dict(attributes or {})
        self._dev_cache = {}

    # ---- pickling: host arrays only (no CUDA state crosses process boundaries)
    def __getstate__(self):
        d = dict(self.__dict__)
        d["_dev_cache"] = {}
        return d

    def num_boosted_rounds(self) -> int:
        return len(self.forest["tree_offsets"]) - 1

    def num_features(self) -> int:
        return self.n_features

    def _names(self):
        return self.feature_names if self.feature_names is not None else [f"f{i}" for i in range(self.n_features)]

    def get_score(self, fmap: str = "", importance_type: str = "weight") -> dict:
        """Per-feature importance; features never split on are omitted (xgboost semantics, R12)."""
        fo = self.forest
        internal = fo["left"] >= 0
        feat = fo["split_index"][internal].astype(np.int64)
        cnt = np.bincount(feat, minlength=self.n_features).astype(np.float64)
        gain = np.bincount(feat, weights=fo["loss_chg"][internal].astype(np.float64), minlength=self.n_features)
        cover = np.bincount(feat, weights=fo["sum_hess"][internal].astype(np.float64), minlength=self.n_features)
        if importance_type == "weight":
            val = cnt
        elif importance_type == "gain":
            val = np.divide(gain, cnt, out=np.zeros_like(gain), where=cnt > 0)
        elif importance_type == "total_gain":
            val = gain
        elif importance_type == "cover":
            val = np.divide(cover, cnt, out=np.zeros_like(cover), where=cnt > 0)
        elif importance_type == "total_cover":
            val = cover
        else:
            raise ValueError(f"Unknown importance type: {importance_type}")
        names = self._names()
        return {names[j]: float(val[j]) for j in range(self.n_features) if cnt[j] > 0}

    def _device_forest(self, device):
        key = str(device)
        if key not in self._dev_cache:
            fo = self.forest
            nodes, toff = pack_forest(fo["tree_offsets"], fo["left"], fo["right"], fo["split_index"], fo["split_cond"],
                                      fo["default_left"])
            self._dev_cache[key] = (torch.from_numpy(nodes).to(device), torch.from_numpy(toff).to(device))
        return self._dev_cache[key]

    def predict_device(self, X: torch.Tensor, output_margin: bool = False) -> torch.Tensor:
        """X: CUDA f32 [n, n_features].  Returns p (or the margin) as a CUDA f32 [n] tensor."""
        if X.shape[1] != self.n_features:
            raise ValueError(f"Feature shape mismatch, expected: {self.n_features}, got {X.shape[1]}")
        q8 = self._device_q8(X.device)
        heap = self._device_heap(X.device) if q8 is None else None
        if q8 is not None:
            cuts, n_cuts, words, depth, plane_base = q8
            m, p = ops.predict_q8(X, cuts, n_cuts, words, depth, plane_base, base_margin_of(self.base_score),
                                  self.num_boosted_rounds(), want_margin=output_margin, want_proba=not output_margin)
        elif heap is not None:
            nodes_h, leaf, depth = heap
            m, p = ops.predict_heap(X, nodes_h, leaf, depth, base_margin_of(self.base_score),
                                    self.num_boosted_rounds(), want_margin=output_margin, want_proba=not output_margin)
        else:                                            # very deep trees: pointer-chasing kernel
            nodes, toff = self._device_forest(X.device)
            m, p = ops.predict(X, nodes, toff, base_margin_of(self.base_score), want_margin=output_margin,
                               want_proba=not output_margin)
        return m if output_margin else p

    def max_depth(self) -> int:
        """Depth of the deepest tree; also checks that children carry larger ids than their parent."""
        fo = self.forest
        offs = fo["tree_offsets"]
        left, right = np.asarray(fo["left"], np.int64), np.asarray(fo["right"], np.int64)
        depth = np.zeros(len(left), np.int32)
        best = 0
        for t in range(len(offs) - 1):
            o, e = int(offs[t]), int(offs[t + 1])
            for i in range(o, e):
                if left[i] >= 0:
                    if left[i] <= i - o or right[i] <= i - o:
                        raise ValueError("forest numbering: a child id does not follow its parent")
                    depth[o + left[i]] = depth[o + right[i]] = depth[i] + 1
                    best = max(best, int(depth[i]) + 1)
        return best

    def predict_contribs_device(self, X: torch.Tensor) -> torch.Tensor:
        """Path-dependent TreeSHAP on the GPU (what xgboost's pred_contribs / shap.TreeExplainer return):
        CUDA f32 [n, n_features + 1], last column = expected value; rows sum to the margin."""
        if X.shape[1] != self.n_features:
            raise ValueError(f"Feature shape mismatch, expected: {self.n_features}, got {X.shape[1]}")
        key = "shap:" + str(X.device)
        if key not in self._dev_cache:
            fo = self.forest
            if "sum_hess" not in fo:
                raise ValueError("TreeSHAP needs the node covers (sum_hess)")
            dt = {"tree_offsets": np.int64, "left": np.int32, "right": np.int32, "split_index": np.int32,
                  "split_cond": np.float32, "default_left": np.uint8, "sum_hess": np.float32}
            with torch.cuda.device(X.device):
                dev = {k: torch.from_numpy(np.ascontiguousarray(fo[k], d)).to(X.device) for k, d in dt.items()}
            self._dev_cache[key] = (dev, self.max_depth())
        dev, depth = self._dev_cache[key]
        return ops.tree_shap(X, dev, base_margin_of(self.base_score), depth)

    def _device_q8(self, device):
        """Quantised-input form of the forest (the fastest scorer), or None when it does not apply."""
        key = "q8:" + str(device)
        if key not in self._dev_cache:
            self._dev_cache[key] = None
            fo = self.forest
            h = None
            if os.environ.get("COBALT_PREDICT", "q8") == "q8" and self.n_features <= 255:
                h = heap_forest(fo["tree_offsets"], fo["left"], fo["right"], fo["split_index"], fo["split_cond"],
                                fo["default_left"])
            if h is not None:
                thr, meta, leaf, depth = h
                with torch.cuda.device(device):
                    chunk, plane_base = ops.predict_q8_plan(self.n_features, depth)
                q = q8_forest(thr, meta, leaf, depth, self.n_features, plane_base) if chunk >= 8 else None
                if q is not None:
                    cuts, n_cuts, words = q
                    T = words.shape[0]
                    Tp = (T + chunk - 1) // chunk * chunk
                    pad = np.zeros((Tp - T, words.shape[1]), np.uint32)
                    pad[:, 1:1 << depth] = plane_base                # padding trees: valid addresses, leaves 0.0
                    words = np.concatenate([words, pad], 0)
                    with torch.cuda.device(device):
                        self._dev_cache[key] = (torch.from_numpy(cuts).to(device), torch.from_numpy(n_cuts).to(device),
                                                torch.from_numpy(words.view(np.int32)).to(device), depth, plane_base)
        return self._dev_cache[key]

    def _device_heap(self, device):
        key = "heap:" + str(device)
        if key not in self._dev_cache:
            fo = self.forest
            h = heap_forest(fo["tree_offsets"], fo["left"], fo["right"], fo["split_index"], fo["split_cond"],
                            fo["default_left"])
            if h is not None and self.n_features > 32767:
                h = None
            chunk, rows = ops.predict_heap_plan(self.n_features, h[3]) if h is not None else (0, 0)
            if h is None or chunk < 8:
                self._dev_cache[key] = None
            else:
                thr, meta, leaf, depth = h
                T = thr.shape[0]
                Tp = (T + chunk - 1) // chunk * chunk
                xoff = (meta.astype(np.uint64) & np.uint64(0x7FFF)) * np.uint64(rows * 4)
                hi = xoff | ((meta.astype(np.uint64) >> np.uint64(15)) << np.uint64(31))
                nodes = thr.view(np.uint32).astype(np.uint64) | (hi << np.uint64(32))
                pad = lambda a: np.concatenate([a, np.zeros((Tp - T,) + a.shape[1:], a.dtype)], 0)  # noqa: E731
                with torch.cuda.device(device):
                    self._dev_cache[key] = (torch.from_numpy(pad(nodes).view(np.int64)).to(device),
                                            torch.from_numpy(pad(leaf)).to(device), depth)
        return self._dev_cache[key]
