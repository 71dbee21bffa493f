"""Worker of tests/test_multi_gpu.py (one process per GPU under torch.distributed.run): trains on row shards in
both collective modes and checks, on rank 0, that the forest equals the oracle's fit on the whole data."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cobalt_smart_lender_ai_b200 import ops, parallel  # noqa: E402
from cobalt_smart_lender_ai_b200.binned import BinnedMatrix  # noqa: E402
from cobalt_smart_lender_ai_b200.synth import make_lending  # noqa: E402
from cobalt_smart_lender_ai_b200.trainer import GpuTrainer  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    local = int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=__import__("datetime").timedelta(seconds=120))
    rank, world = torch.distributed.get_rank(), torch.distributed.get_world_size()
    ok = True
    for n, f, depth, trees in ((40000, 150, 6, 6), (9001, 37, 4, 5), (5003, 5, 3, 3)):   # last: fewer features than ranks at world 8
        Xf, yf, _ = make_lending(n, f, seed=13)
        b, e = parallel.shard_rows(n, rank, world)
        X, y = Xf[b:e].cuda(), yf[b:e].cuda()
        spw = float((yf == 0).sum() / (yf == 1).sum())
        ref = O.fit_reference(Xf.numpy(), yf.numpy(), n_estimators=trees, max_depth=depth, scale_pos_weight=spw) if rank == 0 else None
        for mode in ("peer", "nccl"):
            os.environ["COBALT_MULTIGPU"] = mode
            cuts, n_cuts, min_vals = parallel.sketch_cuts_global(X, 256)
            bm = BinnedMatrix.from_dense(X, cuts, n_cuts)        # one-hot groups are agreed on over all ranks
            sg, sh = ops.quant_shifts(max(1.0, spw))
            tr = GpuTrainer(bm, y, None, f, ops.bin_ptr_of(n_cuts), cuts, min_vals, max_depth=depth, eta=0.3,
                            scale_pos_weight=spw, sg=sg, sh=sh, max_trees=trees, row_offset=b)
            assert tr.peer_mode == (mode == "peer")
            print(f"[mgpu] rank {rank} trainer ready mode={mode}", flush=True)
            tr.boost(trees, use_graph=os.environ.get("MGPU_GRAPH", "1") == "1")
            torch.cuda.synchronize()
            print(f"[mgpu] rank {rank} boosted mode={mode}", flush=True)
            fo = tr.forest_numpy()
            margin = tr.margin.cpu().numpy()
            if rank == 0:
                good = np.array_equal(cuts.cpu().numpy().view(np.uint32), ref["cuts"].view(np.uint32))
                for k in ("tree_offsets", "left", "right", "split_index", "split_bin", "default_left"):
                    good &= np.array_equal(fo[k], ref["flat"][k])
                for k in ("split_cond", "base_weight", "loss_chg", "sum_hess"):
                    good &= np.array_equal(fo[k].view(np.uint32), ref["flat"][k].view(np.uint32))
                good &= np.array_equal(margin.view(np.uint32), ref["margin"][b:e].view(np.uint32))
                print(f"[mgpu] n={n} f={f} mode={mode} world={world} graph_error={tr.graph_error} -> {'OK' if good else 'MISMATCH'}", flush=True)
                ok &= bool(good)
            tr.close()
            torch.distributed.barrier()
    flag = torch.tensor([0 if ok else 1], device="cuda")
    torch.distributed.broadcast(flag, 0)
    code = int(flag.item())
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(code)        # skip interpreter teardown: process groups that were captured in CUDA graphs can stall it


if __name__ == "__main__":
    main()
