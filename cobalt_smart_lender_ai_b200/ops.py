"""Torch-tensor front end of the C ABI: device memory and streams come from torch, every
computation is a call into libcobalt_gbdt.so.  All tensors must live on a CUDA device."""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _lib
from ._lib import CB_CUT_STRIDE, CB_QBITS, Split, TreeParams, check


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t: torch.Tensor | None) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def _ld(X: torch.Tensor) -> int:
    """Row stride in elements; a single-row tensor may carry any stride for its size-1 dimension."""
    return int(X.stride(0)) if X.shape[0] > 1 else int(X.shape[1])


def _need_cuda(*ts: torch.Tensor) -> None:
    if not torch.cuda.is_available():
        raise _lib.CobaltError("no CUDA device: this package has no CPU path")
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.CobaltError("expected CUDA tensors")


def padded_ldb(n_columns: int) -> int:
    """Row stride of the u8 bin matrix: whole 128-byte column blocks plus a last block padded to a power of two
    (>= 16 bytes), so that no block's row ever straddles a 128-byte line (row gathers move whole aligned pieces)."""
    full, rest = divmod(int(n_columns), 128)
    w = 0
    if rest:
        w = 16
        while w < rest:
            w *= 2
    return 128 * full + w


def quant_shifts(wmax: float) -> tuple[int, int]:
    """R4: wmax * 2^sg <= 2^CB_QBITS; hessians (<= wmax / 4) carry two more fractional bits."""
    e = math.frexp(float(wmax))[1]
    if float(wmax) == 2.0 ** (e - 1):
        e -= 1
    sg = CB_QBITS - e
    return sg, sg + 2


def keep_threshold(subsample: float) -> int:
    return 0 if subsample >= 1.0 else int(math.floor(float(subsample) * 4294967296.0))


def workspace(nbytes: int, device) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 256) + 256, dtype=torch.uint8, device=device)


def _aligned(ws: torch.Tensor) -> tuple[C.c_void_p, int]:
    p = ws.data_ptr()
    a = (p + 255) & ~255
    return C.c_void_p(a), ws.numel() - (a - p)


def sketch_cuts(X: torch.Tensor, max_bin: int = 256):
    """K1.  X: f32 [n, f] (row stride may exceed f).  Returns cuts [f,256], n_cuts [f], min_vals [f]."""
    _need_cuda(X)
    assert X.dtype == torch.float32 and X.dim() == 2 and X.stride(1) == 1
    n, f = X.shape
    lib = _lib.load()
    need = C.c_size_t()
    check(lib.cb_sketch_workspace_bytes(n, f, C.byref(need)))
    ws = workspace(need.value, X.device)
    cuts = torch.empty((f, CB_CUT_STRIDE), dtype=torch.float32, device=X.device)
    n_cuts = torch.empty(f, dtype=torch.int32, device=X.device)
    min_vals = torch.empty(f, dtype=torch.float32, device=X.device)
    wp, wb = _aligned(ws)
    with torch.cuda.nvtx.range("cobalt:sketch"):
        check(lib.cb_sketch_cuts(_ptr(X), n, f, _ld(X), max_bin, _ptr(cuts), _ptr(n_cuts), _ptr(min_vals), wp, wb,
                                 _stream()))
    return cuts, n_cuts, min_vals


def bin_ptr_of(n_cuts: torch.Tensor) -> torch.Tensor:
    bp = torch.zeros(n_cuts.numel() + 1, dtype=torch.int32, device=n_cuts.device)
    bp[1:] = torch.cumsum(n_cuts, 0)
    return bp


def group_arrays(groups):
    """[[feature, ...], ...] -> (grp_ptr int32 [G + 1], grp_feat int32 [members]) host arrays, or (None, None)."""
    import numpy as np
    if not groups:
        return None, None
    ptr = np.zeros(len(groups) + 1, np.int32)
    ptr[1:] = np.cumsum([len(g) for g in groups])
    return ptr, np.asarray([ft for g in groups for ft in g], np.int32)


def plan_layout(n_cuts, n_features: int, ldb: int | None = None, groups=None):
    """Column placement of the binned matrix (host computation inside the library).  groups: one-hot feature groups
    that share one code byte each (binned.detect_onehot_groups) or None.
    Returns (pos_feature int32 [ldb] on the host: >= 0 feature, -1 padding, <= -2 group -2 - v; conflict_free flag)."""
    import numpy as np
    nc = np.ascontiguousarray(n_cuts.cpu().numpy() if isinstance(n_cuts, torch.Tensor) else n_cuts, dtype=np.int32)
    gp, gf = group_arrays(groups)
    n_cols = n_features - (0 if gf is None else len(gf)) + (0 if gp is None else len(gp) - 1)
    ldb = ldb or padded_ldb(n_cols)
    pos = np.empty(ldb, np.int32)
    flag = C.c_int32(0)
    vp = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)  # noqa: E731
    check(_lib.load().cb_plan_layout(vp(nc), n_features, vp(gp), vp(gf), 0 if gp is None else len(gp) - 1, ldb, vp(pos),
                                     C.byref(flag)))
    return pos, bool(flag.value)


def bin_pack(X: torch.Tensor, cuts: torch.Tensor, n_cuts: torch.Tensor, ldb: int | None = None,
             pos_feature: torch.Tensor | None = None, blocked: bool = False, grp_ptr: torch.Tensor | None = None,
             grp_feat: torch.Tensor | None = None) -> torch.Tensor:
    """K2.  Returns n * ldb bytes as a u8 [n, ldb] tensor; byte p of a row holds column pos_feature[p]
    (natural order when None; group columns need grp_ptr / grp_feat).  blocked=True stores column blocks of 128 bytes
    as separate arrays (see cb_bin_pack_u8) -- the tensor shape then only carries n and ldb."""
    _need_cuda(X, cuts, n_cuts, pos_feature, grp_ptr, grp_feat)
    assert X.dtype == torch.float32 and X.stride(1) == 1
    n, f = X.shape
    ldb = ldb or (padded_ldb(f) if pos_feature is None else int(pos_feature.numel()))
    assert pos_feature is None or (pos_feature.dtype == torch.int32 and pos_feature.numel() == ldb)
    bins = torch.empty((n, ldb), dtype=torch.uint8, device=X.device)
    check(_lib.load().cb_bin_pack_u8(_ptr(X), n, f, _ld(X), _ptr(cuts), _ptr(n_cuts), _ptr(pos_feature), _ptr(grp_ptr),
                                     _ptr(grp_feat), _ptr(bins), ldb, int(bool(blocked)), _stream()))
    return bins


def bin_transpose(bins: torch.Tensor, blocked: bool = False) -> torch.Tensor:
    """Position-major copy u8 [ldb, n] of a binned matrix (speeds up the trainer's row partition)."""
    _need_cuda(bins)
    n, ldb = bins.shape
    out = torch.empty((ldb, n), dtype=torch.uint8, device=bins.device)
    check(_lib.load().cb_bin_transpose_u8(_ptr(bins), n, ldb, int(bool(blocked)), _ptr(out), _stream()))
    return out


def sigmoid(margin: torch.Tensor) -> torch.Tensor:
    _need_cuda(margin)
    out = torch.empty_like(margin)
    check(_lib.load().cb_sigmoid(_ptr(margin), margin.numel(), _ptr(out), _stream()))
    return out


def grad_logistic(margin, y, weight, spw, sg, sh, subsample=1.0, seed=0, tree=0, row_offset=0):
    """K3.  Returns (gpair int32 [n,2], totals int64 [2])."""
    _need_cuda(margin, y)
    n = margin.numel()
    gp = torch.empty((n, 2), dtype=torch.int32, device=margin.device)
    tot = torch.zeros(2, dtype=torch.int64, device=margin.device)
    check(_lib.load().cb_grad_logistic_q(_ptr(margin), _ptr(y), _ptr(weight), n, float(spw), sg, sh,
                                         keep_threshold(subsample), seed, tree, row_offset, _ptr(gp), _ptr(tot),
                                         _stream()))
    return gp, tot


def hist_build(bins, n_features, gpair, row_idx, bin_ptr, total_bins, count=None) -> torch.Tensor:
    """K4.  Returns int64 [total_bins, 2]."""
    _need_cuda(bins, gpair, bin_ptr)
    lib = _lib.load()
    cnt = int(count if count is not None else (bins.shape[0] if row_idx is None else row_idx.numel()))
    need = C.c_size_t()
    check(lib.cb_hist_workspace_bytes(cnt, total_bins, C.byref(need)))
    ws = workspace(need.value, bins.device)
    hist = torch.empty((total_bins, 2), dtype=torch.int64, device=bins.device)
    wp, wb = _aligned(ws)
    check(lib.cb_hist_build(_ptr(bins), bins.stride(0), n_features, _ptr(gpair), _ptr(row_idx), cnt, _ptr(bin_ptr),
                            total_bins, _ptr(hist), wp, wb, _stream()))
    return hist


def hist_build_cf(bm, gpair, row_idx, bin_ptr, total_bins, count=None) -> torch.Tensor:
    """K4 on one node through the conflict-free kernel the trainer uses.  bm: binned.BinnedMatrix.  int64 [total_bins, 2]."""
    _need_cuda(bm.bins, gpair, bin_ptr)
    lib = _lib.load()
    cnt = int(count if count is not None else (bm.n_rows if row_idx is None else row_idx.numel()))
    need = C.c_size_t()
    check(lib.cb_hist_build_cf_workspace_bytes(cnt, bm.n_features, C.byref(need)))
    ws = workspace(need.value, bm.bins.device)
    hist = torch.zeros((total_bins, 2), dtype=torch.int64, device=bm.bins.device)
    wp, wb = _aligned(ws)
    check(lib.cb_hist_build_cf_sync(_ptr(bm.bins), bm.n_rows, bm.ldb, int(bm.blocked), _ptr(bm.pos), _ptr(bm.grp_ptr),
                                    _ptr(bm.grp_feat), bm.n_groups, bm.n_features, _ptr(gpair), _ptr(row_idx), cnt,
                                    _ptr(bin_ptr), total_bins, _ptr(hist), wp, wb, _stream()))
    return hist


def hist_subtract(parent, built) -> torch.Tensor:
    out = torch.empty_like(parent)
    check(_lib.load().cb_hist_subtract(_ptr(parent), _ptr(built), _ptr(out), parent.numel(), _stream()))
    return out


def split_scan(hist, bin_ptr, cuts, min_vals, totals, tp: TreeParams, feature_mask=None) -> dict:
    """K5 on one node.  Returns the split record as a dict (synchronises to read it)."""
    _need_cuda(hist, bin_ptr, cuts, min_vals, totals)
    lib = _lib.load()
    f = bin_ptr.numel() - 1
    need = C.c_size_t()
    check(lib.cb_split_scan_workspace_bytes(f, C.byref(need)))
    ws = workspace(need.value, hist.device)
    out = torch.zeros(C.sizeof(Split), dtype=torch.uint8, device=hist.device)
    wp, wb = _aligned(ws)
    check(lib.cb_split_scan(_ptr(hist), _ptr(bin_ptr), f, _ptr(cuts), _ptr(min_vals), _ptr(totals), C.byref(tp),
                            _ptr(feature_mask), _ptr(out), wp, wb, _stream()))
    rec = Split.from_buffer_copy(out.cpu().numpy().tobytes())
    return {k: getattr(rec, k) for k, _ in Split._fields_ if k != "pad_"}


def partition_stable(bins, seg: torch.Tensor, feature: int, split_bin: int, default_left: int):
    """K6 on one segment.  Returns (seg_out u32-as-int32 tensor, n_left)."""
    _need_cuda(bins, seg)
    lib = _lib.load()
    cnt = seg.numel()
    need = C.c_size_t()
    check(lib.cb_partition_workspace_bytes(cnt, C.byref(need)))
    ws = workspace(need.value, bins.device)
    out = torch.empty_like(seg)
    nl = torch.zeros(1, dtype=torch.int64, device=bins.device)
    wp, wb = _aligned(ws)
    check(lib.cb_partition_stable(_ptr(bins), bins.stride(0), _ptr(seg), cnt, feature, split_bin, default_left,
                                  _ptr(out), _ptr(nl), wp, wb, _stream()))
    return out, int(nl.item())


def leaf_update(margin: torch.Tensor, rows: torch.Tensor | None, leaf_value: float, count: int | None = None) -> None:
    """K7 on one leaf's segment: margin[rows[i]] += leaf_value (in place)."""
    _need_cuda(margin, rows)
    cnt = int(count if count is not None else (margin.numel() if rows is None else rows.numel()))
    check(_lib.load().cb_leaf_update(_ptr(margin), _ptr(rows), cnt, float(leaf_value), _stream()))


def predict(X: torch.Tensor, nodes: torch.Tensor, tree_off: torch.Tensor, base_margin: float,
            want_margin: bool = False, want_proba: bool = True):
    """K8.  nodes: packed int64 [n_nodes]; tree_off: int32 [n_trees+1]."""
    _need_cuda(X, nodes, tree_off)
    assert X.dtype == torch.float32 and X.stride(1) == 1
    n, f = X.shape
    m = torch.empty(n, dtype=torch.float32, device=X.device) if want_margin else None
    p = torch.empty(n, dtype=torch.float32, device=X.device) if want_proba else None
    check(_lib.load().cb_predict(_ptr(X), n, f, _ld(X), _ptr(nodes), _ptr(tree_off), tree_off.numel() - 1,
                                 float(base_margin), _ptr(m), _ptr(p), _stream()))
    return m, p


def predict_heap_plan(n_features: int, depth: int) -> tuple[int, int]:
    """(trees per shared-memory chunk, rows per CTA tile) of the complete-tree scorer; (0, 0) = does not fit."""
    c, r = C.c_int(0), C.c_int(0)
    rc = _lib.load().cb_predict_heap_plan(n_features, depth, C.byref(c), C.byref(r))
    return (c.value, r.value) if rc == 0 else (0, 0)


def predict_heap(X: torch.Tensor, nodes: torch.Tensor, leaf: torch.Tensor, depth: int, base_margin: float,
                 n_trees: int, want_margin: bool = False, want_proba: bool = True):
    """K8, complete-tree form.  nodes int64 [Tp, 2^d-1] (packed), leaf f32 [Tp, 2^d], padded to whole chunks."""
    _need_cuda(X, nodes, leaf)
    assert X.dtype == torch.float32 and X.stride(1) == 1
    n, f = X.shape
    m = torch.empty(n, dtype=torch.float32, device=X.device) if want_margin else None
    p = torch.empty(n, dtype=torch.float32, device=X.device) if want_proba else None
    check(_lib.load().cb_predict_heap(_ptr(X), n, f, _ld(X), _ptr(nodes), _ptr(leaf), int(n_trees), depth,
                                      float(base_margin), _ptr(m), _ptr(p), _stream()))
    return m, p


def q8_ldc(n_features: int) -> int:
    """Row stride of the code matrix cb_predict_q8 wants: the smallest 4 * odd >= n_features."""
    w = (n_features + 3) // 4
    return 4 * (w | 1)


def predict_q8_plan(n_features: int, depth: int) -> tuple[int, int]:
    """(trees per shared-memory chunk, shared-window address of the code planes) of the quantised-input scorer;
    (0, 0) = does not fit / not applicable."""
    if n_features > 255 or depth > 10:
        return 0, 0
    c, b = C.c_int(0), C.c_uint32(0)
    rc = _lib.load().cb_predict_q8_plan(n_features, depth, q8_ldc(n_features), C.byref(c), C.byref(b))
    return (c.value, b.value) if rc == 0 else (0, 0)


def predict_q8(X: torch.Tensor, cuts: torch.Tensor, n_cuts: torch.Tensor, forest: torch.Tensor, depth: int,
               plane_base: int, base_margin: float, n_trees: int, want_margin: bool = False, want_proba: bool = True,
               batch_rows: int = 1 << 19, events: list | None = None):
    """K8 on quantised inputs: K2 against the forest's threshold table, then the byte-tile walk.  Rows go through
    in batches so that the code matrix (80 MB per batch) is still in L2 when the walk reads it."""
    _need_cuda(X, cuts, n_cuts, forest)
    assert X.dtype == torch.float32 and X.stride(1) == 1
    n, f = X.shape
    ldc = q8_ldc(f)
    m = torch.empty(n, dtype=torch.float32, device=X.device) if want_margin else None
    p = torch.empty(n, dtype=torch.float32, device=X.device) if want_proba else None
    lib = _lib.load()
    codes = torch.empty((min(n, batch_rows), ldc), dtype=torch.uint8, device=X.device)
    torch.cuda.nvtx.range_push("cobalt:predict")
    for r0 in range(0, n, batch_rows):
        r1 = min(n, r0 + batch_rows)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if events is not None else None
        if ev:
            ev[0].record()
        check(lib.cb_bin_rows_u8(_ptr(X[r0:r1]), r1 - r0, f, _ld(X), _ptr(cuts), _ptr(n_cuts), _ptr(codes), ldc, _stream()))
        if ev:
            ev[1].record()
        check(lib.cb_predict_q8(_ptr(codes), r1 - r0, f, ldc, _ptr(forest), int(n_trees), depth, float(base_margin),
                                int(plane_base), _ptr(m[r0:r1]) if m is not None else None,
                                _ptr(p[r0:r1]) if p is not None else None, _stream()))
        if ev:
            ev[2].record()
            events.append(ev)
    torch.cuda.nvtx.range_pop()
    return m, p


def tree_shap(X: torch.Tensor, forest: dict, base_margin: float, max_depth: int, batch_rows: int | None = None) -> torch.Tensor:
    """K10.  forest: dict of CUDA tensors tree_offsets i64, left/right/split_index i32, split_cond f32,
    default_left u8, sum_hess f32.  Returns f32 [n, f + 1] (last column: expected value incl. base_margin)."""
    _need_cuda(X, *forest.values())
    assert X.dtype == torch.float32 and X.stride(1) == 1
    n, f = X.shape
    T = forest["tree_offsets"].numel() - 1
    n_nodes = forest["left"].numel()
    lib = _lib.load()
    if batch_rows is None:                               # keep the per-(row, tree) slices under ~256 MB
        batch_rows = max(1, (1 << 26) // max(1, T * (f + 1)))
    out = torch.empty((n, f + 1), dtype=torch.float32, device=X.device)
    need = C.c_size_t()
    check(lib.cb_tree_shap_workspace_bytes(min(n, batch_rows), f, T, n_nodes, C.byref(need)))
    ws = workspace(need.value, X.device)
    wp, wb = _aligned(ws)
    for r0 in range(0, n, batch_rows):
        r1 = min(n, r0 + batch_rows)
        check(lib.cb_tree_shap(_ptr(X[r0:r1]), r1 - r0, f, _ld(X), _ptr(forest["tree_offsets"]), T, n_nodes,
                               _ptr(forest["left"]), _ptr(forest["right"]), _ptr(forest["split_index"]),
                               _ptr(forest["split_cond"]), _ptr(forest["default_left"]), _ptr(forest["sum_hess"]),
                               float(base_margin), int(max_depth), _ptr(out[r0:r1]), wp, wb, _stream()))
    return out
