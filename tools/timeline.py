#!/usr/bin/env python
"""Live (warm-cache, back-to-back) per-launch timeline of one boosting round at the BASELINE size."""
import collections
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cobalt_smart_lender_ai_b200 import ops  # noqa: E402
from cobalt_smart_lender_ai_b200.synth import make_lending  # noqa: E402
from cobalt_smart_lender_ai_b200.binned import BinnedMatrix
from cobalt_smart_lender_ai_b200.trainer import GpuTrainer  # noqa: E402

from cobalt_smart_lender_ai_b200 import parallel  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_200_000
f = 150
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
if world > 1:
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
r0, r1 = parallel.shard_rows(n, rank, world)
X, y, _ = make_lending(r1 - r0, f, row_offset=r0, device="cuda")
cuts, n_cuts, min_vals = parallel.sketch_cuts_global(X, 256)
bm = BinnedMatrix.from_dense(X, cuts, n_cuts)
spw = 6.75
sg, sh = ops.quant_shifts(spw)
tr = GpuTrainer(bm, y, None, f, ops.bin_ptr_of(n_cuts), cuts, min_vals, max_depth=6, eta=0.3, scale_pos_weight=spw,
                sg=sg, sh=sh, max_trees=40, group=None if world > 1 else False,
                row_offset=r0)
tr.boost(10, use_graph=False)
agg = collections.OrderedDict()
rounds = 5
for _ in range(rounds):
    tl = tr.profile_round()
    for k, v in tl:
        agg.setdefault(k, []).append(v)
tot = sum(sum(v) for v in agg.values()) / rounds
if rank != 0:
    sys.exit(0)
for k, v in agg.items():
    per = len(v) // rounds
    print(f"{k:14s} x{per}  {1e3 * sum(v) / rounds:8.1f} us/round  {100 * sum(v) / rounds / tot:5.1f}%   per launch: "
          + " ".join(f"{1e3 * sum(v[i::per]) / rounds:.1f}" for i in range(per)))
print(f"round total {1e3 * tot:.1f} us")
