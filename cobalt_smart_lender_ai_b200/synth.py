"""Synthetic loans of the engineered Lending-Club tree schema.

The raw dataset is not available offline (DVC pointers only,
data/1-raw/lending-club-2007-2020Q3/*.dvc), so every measurement runs on data drawn
here.  The column mix follows the engineered tree matrix the reference trains on
(src/data_preprocessing/feature_engineering.py:118-147 and the notebook statistics in
SURVEY.md Appendix B): log1p'd skewed amounts, small counts, FICO-like integers,
``term``, and ``drop_first`` one-hot dummies, with the missing-value pattern of the
cleaned sample.

The generator is counter based: every value is a pure function of
(seed, global row, column), so any row range can be produced on any rank or device and
row-sharded runs see exactly the rows a single-GPU run sees.  Integer hashing is exact
on CPU and GPU; the float transforms may differ in the last ulp between devices, so
parity tests always hand the *same* array to the oracle and to the CUDA path.
"""
from __future__ import annotations

import math

import torch

_M32 = 0xFFFFFFFF
DATA_SEED = 20240611
ONE_HOT_GROUPS = (("grade", 6), ("home_ownership", 5), ("verification_status", 2), ("purpose", 13),
                  ("application_type", 1), ("hardship_status", 4))


def _hash_u32(rows: torch.Tensor, col: int, stream: int, seed: int) -> torch.Tensor:
    """32-bit mix of (row, col, stream, seed); int64 tensors, every product < 2^63."""
    x = (rows * 0x2545F491 + (col * 0x3C6EF35F + stream * 0x1B873593 + seed) % (1 << 32)) & _M32
    for _ in range(3):
        x = (((x >> 16) ^ x) * 0x45D9F3B) & _M32
    return (x >> 16) ^ x


def _uniform(rows, col, stream, seed):
    return (_hash_u32(rows, col, stream, seed).to(torch.float64) + 0.5) / 4294967296.0


def _normal(rows, col, seed):
    u1 = _uniform(rows, col, 1, seed)
    u2 = _uniform(rows, col, 2, seed)
    return torch.sqrt(-2.0 * torch.log(u1)) * torch.cos(2.0 * math.pi * u2)


def _poisson(rows, col, seed, lam: float):
    u = _uniform(rows, col, 3, seed)
    k = torch.zeros_like(u)
    pk = math.exp(-lam)
    cdf = pk
    for i in range(1, 40):
        k = k + (u > cdf).to(u.dtype)
        pk = pk * lam / i
        cdf += pk
    return k


def schema(f: int):
    """Column plan for an f-column matrix: list of (name, kind, params)."""
    n_cont = max(1, int(round(0.45 * f)))
    n_cnt = max(1, int(round(0.20 * f)))
    n_fico = min(3, max(0, f - n_cont - n_cnt))
    n_term = 1 if f - n_cont - n_cnt - n_fico > 0 else 0
    n_hot = f - n_cont - n_cnt - n_fico - n_term
    cols = []
    for i in range(n_cont):
        mu = 6.0 + 5.0 * ((i * 37) % 11) / 10.0          # log-scale location 6..11
        sg = 0.4 + 0.9 * ((i * 13) % 7) / 6.0
        cols.append((f"amt_{i}", "cont", (mu, sg)))
    for i in range(n_cnt):
        lam = 0.3 + 7.7 * ((i * 29) % 10) / 9.0
        cols.append((f"cnt_{i}", "count", (lam, i % 2 == 0)))   # every other one log1p'd
    for i in range(n_fico):
        cols.append((f"fico_{i}", "fico", (700.0 - 20.0 * i, 35.0)))
    if n_term:
        cols.append(("term", "term", (0.24,)))
    g = 0
    while n_hot > 0:
        name, k = ONE_HOT_GROUPS[g % len(ONE_HOT_GROUPS)]
        k = min(k, n_hot)
        for j in range(k):
            cols.append((f"{name}{g // len(ONE_HOT_GROUPS)}_{j + 1}", "onehot", (g, j, k)))
        n_hot -= k
        g += 1
    assert len(cols) == f
    return cols


def _missing_plan(cols):
    """column -> (probability, sharing key): one jointly-missing block of 11 count/amount columns
    at 0.296, one column at 0.68, one at 0.128, a few at 0.02-0.07."""
    plan = {}
    cand = [i for i, c in enumerate(cols) if c[1] in ("cont", "count")]
    block = cand[len(cand) // 2: len(cand) // 2 + 11]
    for i in block:
        plan[i] = (0.296, 1000)              # shared key -> jointly missing
    rest = [i for i in cand if i not in plan]
    singles = [(0.68,), (0.128,), (0.0695,), (0.055,), (0.054,), (0.029,), (0.024,)]
    for k, (p,) in enumerate(singles):
        if k * 5 + 1 < len(rest):
            plan[rest[k * 5 + 1]] = (p, 2000 + k)
    return plan


def _column(rows, ci, col, seed):
    _name, kind, prm = col
    if kind == "cont":
        mu, sg = prm
        return torch.log1p(torch.exp(mu + sg * _normal(rows, ci, seed)))
    if kind == "count":
        lam, logged = prm
        k = _poisson(rows, ci, seed, lam)
        return torch.log1p(k) if logged else k
    if kind == "fico":
        mu, sd = prm
        v = torch.round((mu + sd * _normal(rows, ci, seed)) / 5.0) * 5.0
        return torch.clamp(v, 300.0, 850.0)
    if kind == "term":
        return torch.where(_uniform(rows, ci, 4, seed) < prm[0], 60.0, 36.0).to(torch.float64)
    if kind == "onehot":
        g, j, k = prm
        u = _uniform(rows, 100000 + g, 5, seed)      # one draw per group -> mutually exclusive
        # category c in 0..k with geometric-ish weights; dummy j+1 fires when c == j+1
        wts = [0.62 ** c for c in range(k + 1)]
        tot = sum(wts)
        lo = sum(wts[: j + 1]) / tot
        hi = sum(wts[: j + 2]) / tot
        return ((u >= lo) & (u < hi)).to(torch.float64)
    raise ValueError(kind)


def _score(Xd: torch.Tensor, cols, rows, seed):
    """Latent default score: sparse linear + 3 pairwise interactions + noise (before intercept)."""
    f = len(cols)
    z = torch.zeros(Xd.shape[0], dtype=torch.float64, device=Xd.device)
    act = list(range(0, f, max(1, f // 24)))
    for k, j in enumerate(act):
        x = torch.nan_to_num(Xd[:, j], nan=0.0)
        kind = cols[j][1]
        if kind == "cont":
            s = (x - 8.5) / 2.0
        elif kind == "count":
            s = (x - 1.0) / 1.5
        elif kind == "fico":
            s = (x - 700.0) / 35.0 * -1.5
        elif kind == "term":
            s = (x - 36.0) / 24.0
        else:
            s = x * 1.5
        beta = (0.9 - 0.06 * k) * (1 if k % 3 else -1)
        z = z + beta * s
        if torch.isnan(Xd[:, j]).any():
            z = z + 0.35 * torch.isnan(Xd[:, j]).to(torch.float64)
    for a, b in ((0, 1), (2, f // 2), (3, f - 1)):
        xa = torch.nan_to_num(Xd[:, a % f], nan=0.0)
        xb = torch.nan_to_num(Xd[:, b % f], nan=0.0)
        z = z + 0.15 * torch.tanh(xa - 8.0) * torch.tanh(xb - 1.0)
    return z + 0.8 * _normal(rows, 200001, seed)


_INTERCEPT_CACHE: dict = {}


def _intercept(f: int, seed: int, pos_rate: float) -> float:
    """Calibrated on global rows 0..65535 (always the same rows, whatever the shard) on CPU."""
    key = (f, seed, pos_rate)
    if key not in _INTERCEPT_CACHE:
        rows = torch.arange(0, 65536, dtype=torch.int64)
        Xd, cols = _features(rows, f, seed)
        z = _score(Xd, cols, rows, seed)
        lo, hi = -30.0, 30.0
        for _ in range(60):
            mid = 0.5 * (lo + hi)
            if torch.sigmoid(z + mid).mean().item() > pos_rate:
                hi = mid
            else:
                lo = mid
        _INTERCEPT_CACHE[key] = 0.5 * (lo + hi)
    return _INTERCEPT_CACHE[key]


def _features(rows, f, seed):
    cols = schema(f)
    miss = _missing_plan(cols)
    Xd = torch.empty((rows.shape[0], f), dtype=torch.float64, device=rows.device)
    for ci, col in enumerate(cols):
        v = _column(rows, ci, col, seed)
        if ci in miss:
            p, key = miss[ci]
            v = torch.where(_uniform(rows, key, 6, seed) < p, float("nan"), v)
        Xd[:, ci] = v
    return Xd, cols


def make_lending(n: int, f: int, seed: int = DATA_SEED, row_offset: int = 0, device="cpu",
                 pos_rate: float = 0.129, chunk: int = 1 << 20):
    """Returns (X float32 [n, f], y uint8 [n], column names) for global rows
    [row_offset, row_offset + n)."""
    b = _intercept(f, seed, pos_rate)
    X = torch.empty((n, f), dtype=torch.float32, device=device)
    y = torch.empty(n, dtype=torch.uint8, device=device)
    cols = schema(f)
    for s in range(0, n, chunk):
        e = min(n, s + chunk)
        rows = torch.arange(row_offset + s, row_offset + e, dtype=torch.int64, device=device)
        Xd, _ = _features(rows, f, seed)
        z = _score(Xd, cols, rows, seed) + b
        u = _uniform(rows, 200002, 7, seed)
        y[s:e] = (u < torch.sigmoid(z)).to(torch.uint8)
        X[s:e] = Xd.to(torch.float32)
    return X, y, [c[0] for c in cols]


def random_forest_like(n_trees: int, depth: int, f: int, seed: int = 78, device="cpu", grid: int = 100):
    """A random full-depth forest in the wire format (for scoring benchmarks without training).
    Thresholds are drawn around the schema's typical values so both branches are exercised, on a grid of
    2 * grid + 1 values per feature: a forest grown by the 256-bin hist method can only split at its <= 255 cut
    values per feature, and the scorer relies on that (grid=0: continuous thresholds)."""
    g = torch.Generator().manual_seed(seed)
    n_int = (1 << depth) - 1
    n_nodes = (1 << (depth + 1)) - 1
    cols = schema(f)
    left = torch.full((n_trees, n_nodes), -1, dtype=torch.int32)
    right = torch.full((n_trees, n_nodes), -1, dtype=torch.int32)
    idx = torch.arange(n_int, dtype=torch.int32)
    left[:, :n_int] = 2 * idx + 1
    right[:, :n_int] = 2 * idx + 2
    feat = torch.randint(0, f, (n_trees, n_nodes), generator=g, dtype=torch.int32)
    base = torch.tensor([{"cont": 8.5, "count": 1.0, "fico": 700.0, "term": 48.0, "onehot": 1.0}[c[1]] for c in cols])
    spread = torch.tensor([{"cont": 2.0, "count": 1.0, "fico": 40.0, "term": 0.0, "onehot": 0.0}[c[1]] for c in cols])
    z = torch.randn((n_trees, n_nodes), generator=g)
    if grid > 0:
        z = torch.clamp(torch.round(z * (grid / 3.0)), -grid, grid) / (grid / 3.0)
    thr = base[feat.long()] + spread[feat.long()] * z
    leafv = 0.05 * torch.randn((n_trees, n_nodes), generator=g)
    cond = torch.where(left >= 0, thr, leafv).to(torch.float32)
    dl = torch.randint(0, 2, (n_trees, n_nodes), generator=g, dtype=torch.uint8)
    feat = torch.where(left >= 0, feat, torch.zeros_like(feat))
    offs = torch.arange(n_trees + 1, dtype=torch.int64) * n_nodes
    cover = torch.zeros((n_trees, n_nodes), dtype=torch.float32)          # node covers for TreeSHAP: leaves draw one,
    cover[:, n_int:] = 0.25 * torch.randint(1, 200, (n_trees, n_nodes - n_int), generator=g).to(torch.float32)
    for i in range(n_int - 1, -1, -1):                                    # parents add their children
        cover[:, i] = cover[:, 2 * i + 1] + cover[:, 2 * i + 2]
    out = {"tree_offsets": offs, "left": left.reshape(-1), "right": right.reshape(-1),
           "split_index": feat.reshape(-1), "split_cond": cond.reshape(-1), "default_left": dl.reshape(-1),
           "sum_hess": cover.reshape(-1)}
    return {k: v.to(device) for k, v in out.items()}
