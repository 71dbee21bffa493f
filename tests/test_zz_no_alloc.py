"""Collected last on purpose (file name): a failure here cannot hide the other tests in a -x run.

cb_split_scan_noalloc is the split-gain scan for callers that cannot hand the library a workspace --
for instance a scoring/training service that reserves all of its device memory when it starts and must
never call cudaMalloc afterwards, or a binding that can pass nothing but its own input and output
buffers.  Its scratch is a fixed array inside the library, so the test checks these things through the
C ABI: it returns exactly what cb_split_scan returns AND what the CPU oracle's scan (R6/R7) returns, it
keeps doing so on a second call that uses fewer features than the first (left-over records must not
leak), and it refuses inputs larger than its built-in scratch."""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _case(n, f, seed):
    from cobalt_smart_lender_ai_b200.synth import make_lending
    X, y, _ = make_lending(n, f, seed=seed)
    X, y = X.numpy(), y.numpy()
    cuts, nc, mv = O.compute_cuts(X, 256)
    bins = O.bin_matrix(X, cuts, nc)
    bp = O.bin_ptr_of(nc)
    q = O.grad(np.random.default_rng(seed).normal(size=n).astype(np.float32), y, None, 6.75, 14, 16)
    return O.hist(bins, f, q, None, bp), bp, cuts, mv, q.astype(np.int64).sum(0)


def _noalloc(h, bp, cuts, mv, tot, tp, mask=None):
    from cobalt_smart_lender_ai_b200 import _lib
    from cobalt_smart_lender_ai_b200._lib import Split
    lib = _lib.load()
    d = [_dev(a) for a in (h, bp, cuts, mv, tot)]
    md = None if mask is None else _dev(mask)
    out = torch.zeros(C.sizeof(Split), dtype=torch.uint8, device="cuda")
    rc = lib.cb_split_scan_noalloc(C.c_void_p(d[0].data_ptr()), C.c_void_p(d[1].data_ptr()), len(bp) - 1,
                                   C.c_void_p(d[2].data_ptr()), C.c_void_p(d[3].data_ptr()),
                                   C.c_void_p(d[4].data_ptr()), C.byref(tp),
                                   C.c_void_p(0 if md is None else md.data_ptr()), C.c_void_p(out.data_ptr()),
                                   C.c_void_p(torch.cuda.current_stream().cuda_stream))
    if rc != 0:
        return rc, lib.cb_last_error().decode()
    rec = Split.from_buffer_copy(out.cpu().numpy().tobytes())
    return 0, {k: getattr(rec, k) for k, _ in Split._fields_ if k != "pad_"}


def test_noalloc_split_scan_matches_workspace_version_twice_in_a_row():
    from cobalt_smart_lender_ai_b200 import ops
    from cobalt_smart_lender_ai_b200._lib import TreeParams
    tp = TreeParams(np.float32(0.3), 1.0, 0.0, 0.0, 1.0, 6, 14, 16)
    # first call: 60 features; second call: 20 features (fewer records than the scratch still holds)
    for n, f, seed, mask in ((20000, 60, 1, None), (8000, 20, 2, (np.arange(20) % 3 != 0).astype(np.uint8))):
        h, bp, cuts, mv, tot = _case(n, f, seed)
        want = ops.split_scan(_dev(h), _dev(bp), _dev(cuts), _dev(mv), _dev(tot), tp, None if mask is None else _dev(mask))
        rc, got = _noalloc(h, bp, cuts, mv, tot, tp, mask)
        assert rc == 0, got
        assert want["feature"] >= 0
        for k in ("feature", "split_bin", "default_left", "GL", "HL", "GR", "HR"):
            assert got[k] == want[k], k
        for k in ("loss_chg", "threshold"):            # bit-exact, like every other split test
            assert np.float32(got[k]).view(np.uint32) == np.float32(want[k]).view(np.uint32), k
        # ... and against the oracle itself, not only against the sibling entry point
        otp = O.TreeParams(eta=float(np.float32(0.3)), max_depth=6, sg=14, sh=16)
        ref = O.scan(h, bp, cuts, mv, int(tot[0]), int(tot[1]), otp, feature_mask=mask)
        for k in ("feature", "split_bin", "default_left", "GL", "HL", "GR", "HR"):
            assert got[k] == ref[k], k
        for k in ("loss_chg", "threshold"):
            assert np.float32(got[k]).view(np.uint32) == np.float32(ref[k]).view(np.uint32), k


def test_noalloc_split_scan_rejects_oversized_input():
    from cobalt_smart_lender_ai_b200._lib import TreeParams
    tp = TreeParams(np.float32(0.3), 1.0, 0.0, 0.0, 1.0, 6, 14, 16)
    f = 4097
    bp = np.arange(f + 1, dtype=np.int32)
    rc, msg = _noalloc(np.zeros((f, 2), np.int64), bp, np.zeros((f, 256), np.float32), np.zeros(f, np.float32),
                       np.zeros(2, np.int64), tp)
    assert rc == -4 and "4096" in msg
