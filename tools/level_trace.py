#!/usr/bin/env python
"""Where the level tail spends its time: %globaltimer at the start / end of every CTA of the scan, flags, scatter and
final-apply kernels (debug build with -DCB_LEVEL_TRACE; `--build` compiles tools/_trace/libcobalt_gbdt.so here, no GPU
needed).  Per level and kernel: wall span of the grid, duration of the serial extra CTA (block 0 of flags / scatter /
final), median and maximum duration of the other CTAs, and how late the last CTA started."""
import ctypes as C
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cobalt_smart_lender_ai_b200 import _lib  # noqa: E402

TRACE_LIB = os.path.join(ROOT, "tools", "_trace", "libcobalt_gbdt.so")
if "--build" in sys.argv:
    os.makedirs(os.path.dirname(TRACE_LIB), exist_ok=True)
    src = os.path.join(ROOT, "cobalt_smart_lender_ai_b200", "csrc", "cobalt_gbdt.cu")
    subprocess.check_call(["nvcc"] + _lib.NVCC_FLAGS + ["-DCB_LEVEL_TRACE", "-o", TRACE_LIB, src])
    print("built", TRACE_LIB)
    sys.exit(0)

_lib.LIB_PATH = TRACE_LIB
_lib.build = lambda *a, **k: TRACE_LIB

import numpy as np  # noqa: E402
import torch  # noqa: E402
from cobalt_smart_lender_ai_b200 import ops  # noqa: E402
from cobalt_smart_lender_ai_b200.synth import make_lending  # noqa: E402
from cobalt_smart_lender_ai_b200.binned import BinnedMatrix
from cobalt_smart_lender_ai_b200.trainer import GpuTrainer  # noqa: E402

n, f, depth = int(os.environ.get("ROWS", 2_200_000)), 150, 6
X, y, _ = make_lending(n, f, device="cuda")
cuts, n_cuts, min_vals = ops.sketch_cuts(X, 256)
bm = BinnedMatrix.from_dense(X, cuts, n_cuts)
sg, sh = ops.quant_shifts(6.75)
tr = GpuTrainer(bm, y, None, f, ops.bin_ptr_of(n_cuts), cuts, min_vals, max_depth=depth, eta=0.3, scale_pos_weight=6.75,
                sg=sg, sh=sh, max_trees=16, group=False)
lib = _lib.load()
lib.cb_debug_level_trace.restype = C.c_int
lib.cb_debug_level_trace.argtypes = [C.c_void_p, C.c_int]
tr.boost(6, use_graph=False)
stream = torch.cuda.current_stream().cuda_stream
buf = np.zeros((6, 2048, 2), dtype=np.uint64)
names = ["scan", "flags+finish", "scatter+plan", "final_apply", "hist_reduce", "hist_accum"]
_lib.check(lib.cb_debug_level_trace(buf.ctypes.data, 1))
for tree in range(6, 8):
    _lib.check(lib.cb_trainer_round_begin(tr.handle, tree, None, stream))
    for level in range(depth):
        _lib.check(lib.cb_trainer_level_hist(tr.handle, level, stream))
        _lib.check(lib.cb_trainer_level_split(tr.handle, level, stream))
        _lib.check(lib.cb_debug_level_trace(buf.ctypes.data, 1))
        t = buf.astype(np.int64)
        for k in (5, 4, 0, 1, 2, 3):
            a = t[k][(t[k][:, 0] > 0) & (t[k][:, 1] > 0)]
            if not len(a):
                continue
            t0 = a[:, 0].min()
            dur = (a[:, 1] - a[:, 0]) / 1e3
            special = f"block0 {dur[0]:5.1f} us (ends at {(a[0, 1] - t0) / 1e3:5.1f})" if k in (1, 2, 3) else ""
            rest = dur[1:] if k in (1, 2, 3) and len(dur) > 1 else dur
            print(f"tree {tree} level {level} {names[k]:13s}: ctas {len(a):4d} span {(a[:, 1].max() - t0) / 1e3:6.1f} us | "
                  f"cta med {np.median(rest):5.1f} max {rest.max():5.1f} us | last start +{(a[:, 0].max() - t0) / 1e3:5.1f} us | {special}")
    _lib.check(lib.cb_trainer_round_end(tr.handle, stream))
torch.cuda.synchronize()
