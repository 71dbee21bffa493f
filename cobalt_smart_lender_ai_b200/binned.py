"""The binned training matrix as the trainer wants it: K2's u8 codes in the column placement the histogram kernel
is fastest on, plus the position-major copy the row partition reads.

Replaces the implicit ``GHistIndexMatrix`` build inside ``XGBClassifier.fit`` (reference call sites
src/model_train_test/model_tree_train_test.py:118,159).  One step goes beyond a plain byte-per-feature matrix:
the reference's feature engineering turns six categoricals into one-hot dummies with ``pd.get_dummies(...,
drop_first=True)`` (src/data_preprocessing/feature_engineering.py:142-147), so about a third of the engineered
columns are 0/1 indicators of which at most one per categorical is set.  ``detect_onehot_groups`` finds such sets
on the data itself (2-bin features that are never set together and never missing), and a set of k dummies is
stored as ONE code byte (0 = none set, m + 1 = member m set): one pair of shared-memory atomics per row instead
of k pairs, and 150 engineered columns fit one 128-byte line per row.  The trees are unchanged bit for bit: a
member's bin-1 sums are what the kernel accumulated, its bin-0 sums are the node totals minus those (exact int64),
and the partition decodes "member m set" from the code byte.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib, ops
from ._lib import check

MAX_CANDIDATES = 64


def _all_reduce_sum(t: torch.Tensor) -> None:
    from . import parallel
    d = parallel._dist()
    if d is not None:
        d.all_reduce(t)


def detect_onehot_groups(X: torch.Tensor, cuts: torch.Tensor, n_cuts: torch.Tensor, distributed: bool = True):
    """Sets of 2-bin features that are mutually exclusive on the rows of X (over ALL ranks when row-sharded, so every
    rank plans the same layout).  Returns a list of lists of feature ids (each of length >= 2), possibly empty."""
    ops._need_cuda(X, cuts, n_cuts)
    nc = n_cuts.cpu().numpy()
    cand = np.flatnonzero(nc == 2).astype(np.int32)[:MAX_CANDIDATES]
    if len(cand) < 2:
        return []
    k = len(cand)
    cand_d = torch.from_numpy(cand).to(X.device)
    counts = torch.zeros(k * k + k, dtype=torch.int64, device=X.device)
    n, f = X.shape
    check(_lib.load().cb_onehot_pair_counts(ops._ptr(X), n, f, ops._ld(X), ops._ptr(cuts), ops._ptr(cand_d), k,
                                            ops._ptr(counts), C.c_void_p(counts.data_ptr() + 8 * k * k), ops._stream()))
    if distributed:
        _all_reduce_sum(counts)
    c = counts.cpu().numpy()
    pair = c[:k * k].reshape(k, k)
    pair = pair + pair.T                                  # the kernel fills i < j only
    miss = c[k * k:]
    groups: list[list[int]] = []                          # first fit, in feature order: deterministic
    for i in range(k):
        if miss[i]:
            continue
        for g in groups:
            if len(g) < 254 and all(pair[i, j] == 0 for j in g):
                g.append(i)
                break
        else:
            groups.append([i])
    return [[int(cand[i]) for i in g] for g in groups if len(g) >= 2]


class BinnedMatrix:
    """bins (K2 output in the planned placement), bins_cm (position-major copy), and the placement itself."""

    def __init__(self, bins, bins_cm, pos, pos_host, grp_ptr, grp_feat, groups, blocked, cf, n_rows, n_features, ldb):
        self.bins, self.bins_cm, self.pos, self.pos_host = bins, bins_cm, pos, pos_host
        self.grp_ptr, self.grp_feat, self.groups = grp_ptr, grp_feat, groups
        self.blocked, self.cf = bool(blocked), bool(cf)
        self.n_rows, self.n_features, self.ldb = int(n_rows), int(n_features), int(ldb)

    @property
    def n_groups(self) -> int:
        return len(self.groups)

    @classmethod
    def from_dense(cls, X: torch.Tensor, cuts: torch.Tensor, n_cuts: torch.Tensor, *, group_onehot: bool = True,
                   with_cm: bool = True, ldb: int | None = None, groups: list | None = None) -> "BinnedMatrix":
        """Detects one-hot groups (unless given or disabled), plans the placement, runs K2 and the transpose."""
        ops._need_cuda(X, cuts, n_cuts)
        n, f = X.shape
        if groups is None:
            groups = detect_onehot_groups(X, cuts, n_cuts) if group_onehot else []
        want_ldb = ldb
        for attempt in (groups, []):
            n_cols = f - sum(len(g) for g in attempt) + len(attempt)
            ldb = want_ldb or ops.padded_ldb(n_cols)
            try:
                pos_np, cf = ops.plan_layout(n_cuts, f, ldb=ldb, groups=attempt)
            except _lib.CobaltError:
                if not attempt:
                    raise
                continue                                   # grouped placement does not fit: plain columns
            groups = attempt
            break
        if not cf:
            import warnings
            warnings.warn(f"cobalt: {int(n_cuts.sum().item())} bins over {f} columns have no conflict-free shared-memory "
                          "image (about 24 000 bin slots fit); the tiled histogram kernel will run, several times slower per "
                          "row -- consider a smaller max_bin", RuntimeWarning, stacklevel=2)
        dev = X.device
        pos = torch.from_numpy(pos_np).to(dev)
        gp, gf = ops.group_arrays(groups)
        grp_ptr = None if gp is None else torch.from_numpy(gp).to(dev)
        grp_feat = None if gf is None else torch.from_numpy(gf).to(dev)
        bins = ops.bin_pack(X, cuts, n_cuts, ldb=ldb, pos_feature=pos, blocked=cf, grp_ptr=grp_ptr, grp_feat=grp_feat)
        bins_cm = ops.bin_transpose(bins, blocked=cf) if with_cm else None
        return cls(bins, bins_cm, pos, pos_np, grp_ptr, grp_feat, groups, cf and ldb > 128, cf, n, f, ldb)

    def natural_codes(self) -> np.ndarray:
        """u8 [n_rows, n_features] codes in the natural feature order on the host -- what a byte-per-feature K2 would
        have produced (test and bench helper: this is the matrix the CPU oracle reads)."""
        b = self.bins.cpu().numpy()
        n, ldb = self.n_rows, self.ldb
        if self.blocked and ldb > 128:                        # block-major -> row-major
            flat, parts, off = b.reshape(-1), [], 0
            for k0 in range(0, ldb, 128):
                w = min(128, ldb - k0)
                parts.append(flat[off: off + n * w].reshape(n, w))
                off += n * w
            b = np.concatenate(parts, axis=1)
        out = np.full((n, self.n_features), 255, np.uint8)
        for p, col in enumerate(self.pos_host):
            if col >= 0:
                out[:, col] = b[:, p]
            elif col <= -2:
                for m, ft in enumerate(self.groups[-2 - col]):
                    out[:, ft] = (b[:, p] == m + 1).astype(np.uint8)
        return out
