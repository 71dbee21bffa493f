#!/usr/bin/env python
"""Times the pieces of the scoring path separately (CUDA events): K2 against the forest's threshold table,
the quantised walk, and the f32 walk.  Usage: python tools/time_scoring.py [rows]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cobalt_smart_lender_ai_b200 import _lib, ops  # noqa: E402
from cobalt_smart_lender_ai_b200.booster import Booster  # noqa: E402
from cobalt_smart_lender_ai_b200.synth import make_lending, random_forest_like  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
f, depth, trees = 150, 6, 1000
dev = torch.device("cuda")
X = torch.empty((n, f), dtype=torch.float32, device=dev)
for b in range(0, n, 1 << 20):
    e = min(n, b + (1 << 20))
    X[b:e] = make_lending(e - b, f, row_offset=10_000_000 + b, device=dev)[0]
fo = random_forest_like(trees, depth, f)
bst = Booster({k: v.numpy() for k, v in fo.items()}, f, None, 0.5)
cuts, n_cuts, words, depth, plane_base = bst._device_q8(dev)
lib = _lib.load()
ldc = ops.q8_ldc(f)
codes = torch.empty((n, ldc), dtype=torch.uint8, device=dev)
out = torch.empty(n, dtype=torch.float32, device=dev)
st = torch.cuda.current_stream().cuda_stream


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


t_bin0 = timed(lambda: _lib.check(lib.cb_bin_pack_u8(X.data_ptr(), n, f, f, cuts.data_ptr(), n_cuts.data_ptr(), None, None, None,
                                                     codes.data_ptr(), ldc, 0, st)))
ref = codes.clone()
codes.zero_()
t_bin = timed(lambda: _lib.check(lib.cb_bin_rows_u8(X.data_ptr(), n, f, f, cuts.data_ptr(), n_cuts.data_ptr(),
                                                    codes.data_ptr(), ldc, st)))
assert torch.equal(codes[:, :f], ref[:, :f]), "cb_bin_rows_u8 != cb_bin_pack_u8"
print(f"general K2 kernel {t_bin0:.2f} ms, slab kernel {t_bin:.2f} ms (identical codes)")
t_q8 = timed(lambda: _lib.check(lib.cb_predict_q8(codes.data_ptr(), n, f, ldc, words.data_ptr(), trees, depth, 0.0,
                                                  plane_base, None, out.data_ptr(), st)))
os.environ["COBALT_PREDICT"] = "f32"
bst2 = Booster({k: v.numpy() for k, v in fo.items()}, f, None, 0.5)
t_f32 = timed(lambda: bst2.predict_device(X))
os.environ["COBALT_PREDICT"] = "q8"
t_all = timed(lambda: bst.predict_device(X))
print(f"rows {n}: bin_pack {t_bin:.2f} ms | predict_q8 {t_q8:.2f} ms | both via Booster {t_all:.2f} ms | f32 walk {t_f32:.2f} ms")
print(f"per 10M rows: bin {t_bin * 1e7 / n:.1f} q8 {t_q8 * 1e7 / n:.1f} f32 {t_f32 * 1e7 / n:.1f} ms")
