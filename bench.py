#!/usr/bin/env python
"""bench.py -- the reference's headline metric on B200: boosting rounds/s (and scoring
rows/s) of histogram-GBDT on synthetic 2.2M x 150 Lending-Club-schema loans.

  python bench.py --gpus N --steps K --warmup W            # this repo (CUDA, one rank per GPU)
  python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port)

A "step" is one boosting round (K3 + per level K4/K9/K5/K6 + K7) over the whole 2.2M x 150
binned matrix at depth 6 -- BASELINE.json configs[1]; with N > 1 the same 2.2M rows are
sharded over the ranks (configs[2], strong scaling) and each level's histogram buffer is
all-reduced over NCCL.  One JSON line is printed by rank 0.  See DESIGN.md "measurement".
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
_OUT = sys.stdout

N_ROWS, N_FEAT, DEPTH, ROUNDS_CFG, ETA = 2_200_000, 150, 6, 500, 0.3
SCORE_ROWS, SCORE_TREES = 10_000_000, 1000
WORKLOAD = "synthetic Lending-Club schema 2.2M x 150, 256 bins, depth 6 (BASELINE.json configs[1])"


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock + throttle reasons sampled every 2 ms while the timed region runs.

    NVML in-process (the library nvidia-smi itself reads; handle opened BEFORE the timed region): starting an
    `nvidia-smi -lms` process inside a 100-ms timed region enumerates all 8 GPUs of the box under a driver lock and
    was seen to stall the launches of every rank.  Falls back to the nvidia-smi loop when pynvml is missing."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int, period_s: float = 0.002):
        self.samples, self.proc, self.index, self.nv, self.h, self.source = [], None, index, None, None, None
        self.period = period_s
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.replace(",", "").isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nv, self.source = pynvml, "nvml"
        except Exception:
            self.nv = None

    def _poll_nvml(self):
        nv = self.nv
        bits = [nv.nvmlClocksThrottleReasonHwSlowdown, nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                nv.nvmlClocksThrottleReasonSwThermalSlowdown, nv.nvmlClocksThrottleReasonSwPowerCap]
        while not self._stop.is_set():
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.samples.append((mhz, self.max_mhz, [bool(r & b) for b in bits]))
            except Exception:
                pass
            self._stop.wait(self.period)

    def _poll_smi(self):
        for l in self.proc.stdout:
            parts = [x.strip() for x in l.split(",")]
            if len(parts) < 6:
                continue
            try:
                self.samples.append((float(parts[0]), float(parts[1]),
                                     [v.lower().startswith("active") for v in parts[2:6]]))
            except ValueError:
                continue

    def __enter__(self):
        if self.nv is not None:
            self.th = threading.Thread(target=self._poll_nvml, daemon=True)
            self.th.start()
            return self
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi"
            self.th = threading.Thread(target=self._poll_smi, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None
        return self

    def __exit__(self, *a):
        if self.nv is not None:
            self._stop.set()
            self.th.join(timeout=2)
        elif self.proc:
            time.sleep(0.25)
            self.proc.terminate()
            self.th.join(timeout=2)

    def summary(self):
        sm = sorted(x[0] for x in self.samples)
        mx = [x[1] for x in self.samples]
        reasons = {nm for x in self.samples for nm, on in zip(self.NAMES, x[2]) if on}
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "source": self.source}


def _dist_init(n_gpus: int):
    import torch
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        torch.cuda.set_device(local)
        import datetime
        # a short watchdog: a desynchronised run must end in minutes, not after the default 10-minute collective timeout
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local),
                                             timeout=datetime.timedelta(seconds=90))
    else:
        torch.cuda.set_device(0)
    return world, rank, local


def _timed(fn, world, dev):
    """barrier + synchronize on both sides, CUDA events on the launching stream, max over ranks."""
    import torch
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    return float(ms.item())


def _xgboost_probe() -> str:
    """BASELINE.md section 3, baseline 1: the reference's own CPU path is the xgboost package -- probed at run start."""
    try:
        import xgboost  # noqa: F401
        return f"importable ({xgboost.__version__})"
    except Exception as e:
        return f"unavailable on this box ({type(e).__name__})"


def _xgboost_rounds_per_s(X, y, spw, steps, depth):
    """The reference's real CPU path, if the box has it: rounds/s of xgboost `hist` between a 2-round and a
    (2 + steps)-round fit (its sketch and binning are not charged to the rounds).  None when xgboost is absent or fails."""
    try:
        import xgboost
        kw = dict(tree_method="hist", max_bin=256, max_depth=depth, learning_rate=ETA, scale_pos_weight=spw,
                  n_jobs=os.cpu_count(), eval_metric="logloss")
        ts = []
        for k in (2, 2 + steps):
            t0 = time.perf_counter()
            xgboost.XGBClassifier(n_estimators=k, **kw).fit(X, y)
            ts.append(time.perf_counter() - t0)
        return steps / max(ts[1] - ts[0], 1e-9)
    except Exception:
        return None


def _config(n_rows, steps, warmup, world):
    """the `config` object: the workload, identical in both arms (the driver compares them)"""
    return {"workload": WORKLOAD, "rows": n_rows, "features": N_FEAT, "max_bin": 256, "max_depth": DEPTH,
            "rounds_window": f"rounds {warmup + 1}..{warmup + steps} of one fit from margin 0 (early rounds scan the most rows per level)",
            "l2_policy": "inputs larger than L2 (binned matrix vs 126 MB L2)" if world == 1 else
            "binned shard may fit L2 at N>=4 (strong scaling of one matrix)",
            "parallelism": f"rows sharded x{world}, one exchange of the level's histograms per tree level" if world > 1 else "single GPU"}


def _slice_forest(flat: dict, k: int) -> dict:
    """first k trees of a flat wire-format forest"""
    offs = flat["tree_offsets"][: k + 1]
    out = {"tree_offsets": offs}
    for key, v in flat.items():
        if key not in ("tree_offsets", "n_nodes"):
            out[key] = v[: int(offs[-1])]
    return out


def _forest_equal(a: dict, b: dict) -> bool:
    import numpy as np
    ok = True
    for k in ("tree_offsets", "left", "right", "split_index", "split_bin", "default_left"):
        ok &= bool(np.array_equal(a[k], b[k]))
    for k in ("split_cond", "base_weight", "loss_chg", "sum_hess"):
        ok &= bool(np.array_equal(np.asarray(a[k], np.float32).view(np.uint32), np.asarray(b[k], np.float32).view(np.uint32)))
    return ok


def _forest_parity(O, fo_gpu: dict, oracle_parts: list, k: int, n_feat: int, dev, n_rows: int) -> dict:
    """GPU forest (first k trees) vs the oracle's k trees grown from the same bins and start margin: bitwise equality,
    then max |dp| and |dAUC| of the two forests on 65 536 held-out synthetic rows (targets 1e-5 / 1e-4)."""
    import numpy as np
    import torch
    from sklearn.metrics import roc_auc_score

    from cobalt_smart_lender_ai_b200.booster import Booster
    from cobalt_smart_lender_ai_b200.synth import make_lending
    flats = [O.flatten_forest(p) for p in oracle_parts]
    offs = [0]
    for fl in flats:
        offs += [int(o) + offs[-1] for o in fl["tree_offsets"][1:]]
    ref = {"tree_offsets": np.asarray(offs, np.int64)}
    for key in flats[0]:
        if key != "tree_offsets":
            ref[key] = np.concatenate([f[key] for f in flats])
    got = _slice_forest(fo_gpu, k)
    equal = _forest_equal(got, ref)
    Xh, yh, _ = make_lending(65536, n_feat, row_offset=n_rows + 1_000_000, device=dev)
    p_gpu = Booster(got, n_feat, None, 0.5).predict_device(Xh).cpu().numpy()
    m_cpu = O.predict_margin(Xh.cpu().numpy(), ref["tree_offsets"], ref["left"], ref["right"], ref["split_index"],
                             ref["split_cond"], ref["default_left"], 0.0)
    p_cpu = O.sigmoid(m_cpu)
    yy = yh.cpu().numpy()
    return {"config": f"{n_rows} x {n_feat} (the benchmarked matrix), trees 1..{k}", "trees": k, "forest_equal": bool(equal),
            "max_dp": float(np.max(np.abs(p_gpu - p_cpu))), "dauc": float(abs(roc_auc_score(yy, p_gpu) - roc_auc_score(yy, p_cpu))),
            "held_out_rows": 65536, "targets": {"max_dp": 1e-5, "dauc": 1e-4}}


def _hgb_baseline(n_rows: int, n_feat: int, depth: int, dev) -> dict:
    """sklearn HistGradientBoostingClassifier on the host cores, full matrix, a bounded number of iterations.
    An independent hist-GBDT (different binning, no parity claim); rounds/s = extra iterations / extra seconds between a
    2-iteration and an 8-iteration fit, so its one-off binning is not charged to the rounds."""
    import numpy as np
    from sklearn.ensemble import HistGradientBoostingClassifier as HGB

    from cobalt_smart_lender_ai_b200.synth import make_lending
    X, y, _ = make_lending(n_rows, n_feat, device=dev)
    X, y = X.cpu().numpy(), y.cpu().numpy()
    kw = dict(max_depth=depth, max_leaf_nodes=None, learning_rate=ETA, l2_regularization=1.0, min_samples_leaf=1,
              max_bins=255, early_stopping=False)
    ts = []
    for it in (2, 8):
        t0 = time.perf_counter()
        m = HGB(max_iter=it, **kw).fit(X, y)
        ts.append(time.perf_counter() - t0)
    ns = min(n_rows, 1_000_000)
    t0 = time.perf_counter()
    m.predict_proba(X[:ns])
    tpred = time.perf_counter() - t0
    return {"value": round(6 / max(ts[1] - ts[0], 1e-9), 3), "unit": "rounds/s", "cores": os.cpu_count(), "kind": "sklearn-HGB stand-in",
            "sample": f"full {n_rows} x {n_feat}; (8-iteration fit - 2-iteration fit) / 6; fits took {ts[0]:.1f} s and {ts[1]:.1f} s",
            "predict_rows_per_s_8_trees": round(ns / tpred, 0)}


def _config_c0(O, dev) -> dict:
    """BASELINE.json configs[0]: 100k x 100, the training job's default classifier (100 trees, depth 6, eta 0.3),
    train + predict on the CPU (oracle port and sklearn HGB) and through this package's estimator from host arrays;
    the GPU forest must equal the oracle's."""
    import numpy as np
    import torch
    from sklearn.ensemble import HistGradientBoostingClassifier as HGB
    from sklearn.metrics import roc_auc_score

    from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier
    from cobalt_smart_lender_ai_b200.synth import make_lending
    n, f, T = 100_000, 100, 100
    X, y, _ = make_lending(n, f, seed=7, device=dev)
    X, y = X.cpu().numpy(), y.cpu().numpy()
    spw = float((y == 0).sum() / (y == 1).sum())
    out = {"workload": "100k x 100, 100 trees, depth 6, eta 0.3 (configs[0])"}
    t0 = time.perf_counter()
    ref = O.fit_reference(X, y, n_estimators=T, max_depth=6, scale_pos_weight=spw)
    t_fit = time.perf_counter() - t0
    t0 = time.perf_counter()
    p_ref = O.predict_proba_reference(ref, X)
    t_pr = time.perf_counter() - t0
    out["oracle_port"] = {"fit_rounds_per_s": round(T / t_fit, 2), "predict_rows_per_s": round(n / t_pr, 0), "cores": O.num_threads()}
    t0 = time.perf_counter()
    m = HGB(max_iter=T, max_depth=6, max_leaf_nodes=None, learning_rate=0.3, l2_regularization=1.0, min_samples_leaf=1,
            max_bins=255, early_stopping=False, class_weight={0: 1.0, 1: spw}).fit(X, y)
    t_fit = time.perf_counter() - t0
    t0 = time.perf_counter()
    p_h = m.predict_proba(X)[:, 1]
    t_pr = time.perf_counter() - t0
    out["sklearn_hgb"] = {"fit_rounds_per_s": round(T / t_fit, 2), "predict_rows_per_s": round(n / t_pr, 0), "cores": os.cpu_count(),
                          "train_auc": round(float(roc_auc_score(y, p_h)), 5)}
    clf = XGBClassifier(n_estimators=T, max_depth=6, scale_pos_weight=spw, device=str(dev))
    clf.fit(X[:1000], y[:1000])                       # library warm-up (module load, first-launch costs)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    clf.fit(X, y)
    torch.cuda.synchronize()
    t_fit = time.perf_counter() - t0
    t0 = time.perf_counter()
    p = clf.predict_proba(X)[:, 1]
    t_first = time.perf_counter() - t0            # includes packing the forest for the scorer (once per fitted model)
    t0 = time.perf_counter()
    p = clf.predict_proba(X)[:, 1]
    t_pr = time.perf_counter() - t0
    out["b200"] = {"fit_rounds_per_s": round(T / t_fit, 2), "predict_rows_per_s": round(n / t_pr, 0),
                   "first_predict_seconds": round(t_first, 4),
                   "what": "XGBClassifier.fit / predict_proba from host arrays (H2D, sketch, binning, D2H included); the first "
                           "predict of a fitted model also packs its forest for the scorer"}
    out["parity"] = {"forest_equal": _forest_equal(clf.get_booster().forest, ref["flat"]),
                     "max_dp": float(np.max(np.abs(p - p_ref))),
                     "dauc": float(abs(roc_auc_score(y, p) - roc_auc_score(y, p_ref))), "train_auc": round(float(roc_auc_score(y, p_ref)), 5)}
    return out


def run_ours(args):
    import ctypes as C

    import numpy as np
    import torch

    from cobalt_smart_lender_ai_b200 import _lib, ops, parallel
    from cobalt_smart_lender_ai_b200.binned import BinnedMatrix
    from cobalt_smart_lender_ai_b200.booster import Booster
    from cobalt_smart_lender_ai_b200.sklearn import XGBClassifier
    from cobalt_smart_lender_ai_b200.synth import make_lending, random_forest_like
    from cobalt_smart_lender_ai_b200.trainer import GpuTrainer

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this package has no CPU path)")
    world, rank, local = _dist_init(args.gpus)
    dev = torch.device("cuda", local if world > 1 else 0)
    peak, peak_src = _peaks()
    lib = _lib.load()
    n_rows = args.rows
    r0, r1 = parallel.shard_rows(n_rows, rank, world)
    X, y, _names = make_lending(r1 - r0, N_FEAT, row_offset=r0, device=dev)
    if args.schema == "continuous":       # every column continuous: 150 x 255 bins do not fit one shared-memory image,
        g = torch.Generator(device=dev).manual_seed(1234 + rank)        # so this measures the tiled fallback kernel
        X = X + 1e-3 * torch.randn(X.shape, generator=g, device=dev)
        X[X != X] = 0.0
    n_pos = torch.tensor([float((y == 1).sum())], dtype=torch.float64, device=dev)
    if world > 1:
        torch.distributed.all_reduce(n_pos)
    spw = float((n_rows - n_pos.item()) / n_pos.item())
    sg, sh = ops.quant_shifts(max(1.0, spw))

    # ---- one-off preparation (K1, K2), timed for the record, not part of `value`
    t_prep = _timed(lambda: None, world, dev)
    cuts = n_cuts = min_vals = bm = None

    def prep():
        nonlocal cuts, n_cuts, min_vals, bm
        cuts, n_cuts, min_vals = parallel.sketch_cuts_global(X, 256)
        bm = BinnedMatrix.from_dense(X, cuts, n_cuts, group_onehot=not args.no_groups)
    t_prep = _timed(prep, world, dev)
    cf_flag = bm.cf
    bin_ptr = ops.bin_ptr_of(n_cuts)
    S = int(bin_ptr[-1].item())
    total_rounds = args.warmup + args.steps + 4
    tr = GpuTrainer(bm, y, None, N_FEAT, bin_ptr, cuts, min_vals, max_depth=DEPTH, eta=ETA, scale_pos_weight=spw,
                    sg=sg, sh=sh, max_trees=total_rounds, row_offset=r0)
    tr.boost(args.warmup)
    l0 = tr.launch_count()
    # NVML calls take a driver-wide lock: with one process per GPU only rank 0 samples, and less often
    with (ClockSampler(dev.index or 0, 0.002 if world == 1 else 0.004) if rank == 0 else contextlib.nullcontext()) as clk:
        ms = _timed(lambda: tr.boost(args.steps), world, dev)
    launches = tr.launch_count() - l0
    rounds_per_s = args.steps / (ms / 1e3)

    # ---- roofline of the dominant kernel (hist_accum): CUDA events inside the library around every launch
    graph_state = {"captured": bool(tr._graphs and all(g is not None for g in tr._graphs.values())), "error": tr.graph_error}
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    acc_ms = red_ms = 0.0
    per_level = {}
    alg_bytes = 0
    n_launch = 0
    # stepping a round level by level with a host synchronisation in between is only possible when the ranks meet in a
    # host-issued collective (NCCL); with the exchange fused into the kernels (peer mode) a rank that waits on the host
    # would leave its peers spinning on the device, so that mode reports no per-launch timing
    if not tr.peer_mode:
        lib.cb_trainer_enable_timing(tr.handle, 1)
        for _ in range(3):
            _lib.check(lib.cb_trainer_round_begin(tr.handle, tr.n_trees, None, st))
            for level in range(DEPTH):
                _lib.check(lib.cb_trainer_level_hist(tr.handle, level, st))
                a, r_, rows, nb = C.c_float(), C.c_float(), C.c_int64(), C.c_int()
                _lib.check(lib.cb_trainer_last_hist_timing_sync(tr.handle, C.byref(a), C.byref(r_), C.byref(rows), C.byref(nb)))
                if world > 1:
                    torch.distributed.all_reduce(tr._level_view(level), group=tr.group)
                _lib.check(lib.cb_trainer_level_split(tr.handle, level, st))
                acc_ms += a.value; red_ms += r_.value; n_launch += 1
                per_level.setdefault(level, []).append((int(rows.value), int(nb.value), round(a.value * 1e3, 1)))
                # R*(F + 8 + 4) + nodes_built*S*16 (S = total bins of the layout, not F*256);
                # the row-index term is 0 at the root
                alg_bytes += rows.value * (N_FEAT + 8 + (0 if level == 0 else 4)) + nb.value * S * 16
            _lib.check(lib.cb_trainer_round_end(tr.handle, st))
            tr.n_trees += 1
        torch.cuda.synchronize()
        lib.cb_trainer_enable_timing(tr.handle, 0)
    timed_launches = n_launch > 0
    if not timed_launches:
        acc_ms, n_launch, alg_bytes = 1.0, 1, 0
    achieved = alg_bytes / (acc_ms / 1e3) / 1e9
    traffic, traffic_src, traffic_ratio = None, None, None
    try:      # per-launch DRAM bytes of the same kernel from the committed ncu --set full capture
        rf = [f for f in ("r02_roofline.json", "r01_roofline.json") if os.path.exists(os.path.join(ROOT, "profiles", f))][0]
        with open(os.path.join(ROOT, "profiles", rf)) as fh:
            rfj = json.load(fh)
        traffic = int(rfj["dram_bytes_per_launch_mean"]) if world == 1 else None
        traffic_ratio = rfj.get("traffic_over_algorithmic")
        traffic_src = (f"profiles/{rf}: ncu --set full capture of this command's histogram launches (committed); its own "
                       f"launches' DRAM bytes / algorithmic bytes = {traffic_ratio}")
    except Exception:
        pass
    roofline = {"kernel": "hist_accum_cf_kernel" if cf_flag else "hist_accum_kernel", "bound": "hbm", "achieved": round(achieved, 1), "peak": peak,
                "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": traffic, "traffic_source": traffic_src,
                "traffic_over_algorithmic_in_capture": traffic_ratio, "peak_source": peak_src,
                "frac_accum_plus_reduce": round(alg_bytes / ((acc_ms + red_ms) / 1e3) / 1e9 / peak, 4),
                "avg_launch_ms": round(acc_ms / n_launch, 4), "launches_timed": n_launch,
                "alg_bytes_per_launch": int(alg_bytes / n_launch), "hist_reduce_ms_per_launch": round(red_ms / n_launch, 4),
                "hist_share_of_round": round((acc_ms + red_ms) / 3 / (ms / args.steps), 3),
                "per_level_rows_nodes_us": {str(k): v[-1] for k, v in per_level.items()}}
    if not timed_launches:
        roofline = {"kernel": roofline["kernel"], "bound": "hbm", "achieved": None, "peak": peak, "unit": "GB/s", "frac": None,
                    "traffic": None, "note": "no per-launch timing in peer mode (the exchange is inside the kernels; a host "
                    "synchronisation between the halves of a level would stall the peers)"}
    auc = None
    try:
        from sklearn.metrics import roc_auc_score
        if world == 1:
            auc = float(roc_auc_score(y[:200000].cpu().numpy(), tr.margin[:200000].cpu().numpy()))
    except Exception:
        pass
    bins_cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        bins_cpu = bm.natural_codes()                      # the natural column order the oracle reads
    cuts_cpu, nc_cpu, mv_cpu = cuts.cpu().numpy(), n_cuts.cpu().numpy(), min_vals.cpu().numpy()
    y_cpu = y.cpu().numpy()
    peer_mode = bool(tr.peer_mode)
    fo_gpu = tr.forest_numpy() if bins_cpu is not None else None
    hist_info = tr.layout_info()
    tr.close()
    layout = {"bytes_per_row": bm.ldb, "onehot_groups": [len(g) for g in bm.groups], "conflict_free": bool(bm.cf), "histogram_kernel": hist_info}
    del tr, bm

    # ---- e2e: the public API with HOST buffers (H2D of X, sketch, binning, ROUNDS_CFG rounds, D2H of the forest)
    e2e = None
    if not args.no_e2e:
        Xh = torch.empty(X.shape, dtype=torch.float32, pin_memory=True)
        Xh.copy_(X)
        Xn = Xh.numpy()
        clf = XGBClassifier(n_estimators=args.e2e_rounds, max_depth=DEPTH, learning_rate=ETA, scale_pos_weight=spw,
                            device=str(dev))
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        clf.fit(Xn, y_cpu)
        torch.cuda.synchronize()
        t_fit = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            torch.distributed.all_reduce(t_fit, op=torch.distributed.ReduceOp.MAX)
        fo = clf.get_booster().forest
        d2h = sum(int(v.nbytes) for v in fo.values())
        e2e = {"value": round(args.e2e_rounds / float(t_fit.item()), 2), "unit": "rounds/s",
               "h2d_bytes_per_step": int((Xn.nbytes + y_cpu.nbytes) / args.e2e_rounds),
               "d2h_bytes_per_step": int(d2h / args.e2e_rounds), "rounds": args.e2e_rounds,
               "fit_seconds": round(float(t_fit.item()), 3),
               "what": "XGBClassifier.fit(host X, host y): H2D + sketch + binning + rounds + forest D2H"}
        del clf, Xh, Xn
    del X

    # ---- scoring (BASELINE.json configs[4]): 10M x 150 through a 1000-tree depth-6 ensemble, rows sharded
    scoring = None
    if not args.no_scoring:
        s0, s1 = parallel.shard_rows(args.score_rows, rank, world)
        fo = random_forest_like(SCORE_TREES, DEPTH, N_FEAT)
        booster = Booster({k: v.numpy() for k, v in fo.items()}, N_FEAT, None, 0.5)
        Xs = torch.empty((s1 - s0, N_FEAT), dtype=torch.float32, device=dev)
        step = 1 << 20
        for b in range(0, s1 - s0, step):
            e = min(s1 - s0, b + step)
            Xs[b:e] = make_lending(e - b, N_FEAT, row_offset=10_000_000 + s0 + b, device=dev)[0]
        booster.predict_device(Xs[:4096])
        out = None

        def score():
            nonlocal out
            for _ in range(args.score_steps):
                out = booster.predict_device(Xs)
        for _ in range(2):
            booster.predict_device(Xs)
        ms_s = _timed(score, world, dev) / args.score_steps
        rows_s = args.score_rows / (ms_s / 1e3)
        sbytes = (s1 - s0) * (4 * N_FEAT + 4)
        visits = args.score_rows * SCORE_TREES * DEPTH
        q8 = booster._device_q8(dev)
        split = None
        if q8 is not None:                     # per-kernel split of one pass: CUDA events around the two launches of every batch
            evs = []
            cuts_q, nc_q, words_q, depth_q, pbase_q = q8
            ops.predict_q8(Xs, cuts_q, nc_q, words_q, depth_q, pbase_q, 0.0, SCORE_TREES, events=evs)
            torch.cuda.synchronize()
            t_codes = sum(e[0].elapsed_time(e[1]) for e in evs)
            t_walk = sum(e[1].elapsed_time(e[2]) for e in evs)
            nloc = s1 - s0
            split = {"bin_rows_q8_kernel": {"ms": round(t_codes, 3), "alg_bytes": nloc * (4 * N_FEAT + N_FEAT),
                                            "hbm_frac": round(nloc * 5 * N_FEAT / (t_codes / 1e3) / 1e9 / peak, 4)},
                     "predict_q8_kernel": {"ms": round(t_walk, 3), "node_visits_per_s": round(nloc * SCORE_TREES * DEPTH / (t_walk / 1e3), 0),
                                           "note": "reads the codes from L2 (batches of 512k rows); bound by shared-memory wavefronts and issue"}}
        straffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "r02_predict_roofline.json")) as fh:
                straffic = json.load(fh)
        except Exception:
            pass
        sc = {"metric": "scoring rows/s", "value": round(rows_s, 0), "ms_per_pass": round(ms_s, 3),
              "workload": f"{args.score_rows} rows x {N_FEAT}, {SCORE_TREES} trees depth {DEPTH} (configs[4])",
              "roofline": {"kernel": "bin_rows_q8_kernel + predict_q8_kernel" if q8 is not None else "predict_heap_kernel",
                           "bound": "hbm (contract); binding resource is shared-memory lookups + ALU issue", "achieved": round(sbytes / (ms_s / 1e3) / 1e9, 1),
                           "peak": peak, "unit": "GB/s", "frac": round(sbytes / (ms_s / 1e3) / 1e9 / peak, 4),
                           "traffic": straffic, "per_kernel": split},
              "node_visits_per_s": round(visits / (ms_s / 1e3), 0)}
        if rank == 0 and not args.no_cpu:      # parity on a sample of the benchmarked rows and forest
            from oracle import oracle as O
            ns_ = min(65536, s1 - s0)
            fo_np = booster.forest
            m_gpu = booster.predict_device(Xs[:ns_], output_margin=True).cpu().numpy()
            m_cpu = O.predict_margin(Xs[:ns_].cpu().numpy(), fo_np["tree_offsets"], fo_np["left"], fo_np["right"], fo_np["split_index"],
                                     fo_np["split_cond"], fo_np["default_left"], 0.0)
            sc["parity"] = {"rows": ns_, "trees": SCORE_TREES, "margins_bitwise_equal": bool(np.array_equal(m_gpu.view(np.uint32), m_cpu.view(np.uint32))),
                            "max_dp": float(np.max(np.abs(O.sigmoid(m_gpu) - O.sigmoid(m_cpu))))}
        if not args.no_e2e:
            Xsh = torch.empty(Xs.shape, dtype=torch.float32, pin_memory=True)
            Xsh.copy_(Xs)
            clf = XGBClassifier(device=str(dev)).load_fixture_forest(booster.forest, [f"f{i}" for i in range(N_FEAT)], 0.5)
            Xsn = Xsh.numpy()
            clf.predict_proba(Xsn[:4096])
            if world > 1:
                torch.distributed.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            pr = clf.predict_proba(Xsn)
            t_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
            if world > 1:
                torch.distributed.all_reduce(t_s, op=torch.distributed.ReduceOp.MAX)
            sc["e2e"] = {"value": round(args.score_rows / float(t_s.item()), 0), "unit": "rows/s",
                         "h2d_bytes_per_step": int(Xsn.nbytes), "d2h_bytes_per_step": int(pr.nbytes // 2),
                         "what": "XGBClassifier.predict_proba(host X)"}
            del Xsh, Xsn
        # ---- TreeSHAP (the /predict explanation): same forest, a bounded number of rows
        if rank == 0 and args.shap_rows > 0:
            ns = min(args.shap_rows, s1 - s0)
            booster.predict_contribs_device(Xs[:64])
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            phi = booster.predict_contribs_device(Xs[:ns])
            e1.record()
            torch.cuda.synchronize()
            ms_shap = e0.elapsed_time(e1)
            shap = {"rows_per_s": round(ns / (ms_shap / 1e3), 1), "rows": ns, "ms": round(ms_shap, 2), "trees": SCORE_TREES,
                    "additivity_max_err": float((phi.double().sum(1) - booster.predict_device(Xs[:ns], output_margin=True).double()).abs().max())}
            if not args.no_cpu:
                from oracle import oracle as O
                nc_ = min(ns, 256)
                fo_np = booster.forest
                t0 = time.perf_counter()
                O.tree_shap(Xs[:nc_].cpu().numpy(), fo_np["tree_offsets"], fo_np["left"], fo_np["right"], fo_np["split_index"],
                            fo_np["split_cond"], fo_np["default_left"], fo_np["sum_hess"], 0.0)
                shap["cpu_rows_per_s"] = round(nc_ / (time.perf_counter() - t0), 1)
                shap["cpu_sample"] = f"{nc_} rows, oracle R13 with OpenMP on {O.num_threads()} threads"
            sc["tree_shap"] = shap
        scoring = sc
        del Xs

    # ---- CPU baseline: the oracle port on this box's host cores, same bins, bounded number of rounds
    cpu = parity = hgb = c0 = None
    if rank == 0 and bins_cpu is not None:
        from oracle import oracle as O
        O.set_threads(os.cpu_count() or 1)
        tp = O.TreeParams(eta=float(np.float32(ETA)), max_depth=DEPTH, sg=sg, sh=sh)
        margin = np.zeros(len(y_cpu), np.float32)
        bp = O.bin_ptr_of(nc_cpu)
        fo_a = O.train(bins_cpu, N_FEAT, bp, cuts_cpu, mv_cpu, y_cpu, None, spw, tp, 1, margin)
        t0 = time.perf_counter()
        fo_b = O.train(bins_cpu, N_FEAT, bp, cuts_cpu, mv_cpu, y_cpu, None, spw, tp, args.cpu_rounds, margin, tree_offset=1)
        dt = time.perf_counter() - t0
        cpu = {"value": round(args.cpu_rounds / dt, 3), "unit": "rounds/s", "cores": O.num_threads(),
               "kind": "port", "sample": f"{args.cpu_rounds} rounds on the full {n_rows} x {N_FEAT} binned matrix "
               f"(rounds 2..{args.cpu_rounds + 1}), oracle/gbdt_oracle.c with OpenMP; reference xgboost unavailable on this box"}
        # ---- parity ON THE BENCHMARKED CONFIGURATION: the GPU's first trees (same bins, same start margin 0) against
        # the oracle's, bit for bit; then both forests on a held-out slice of fresh synthetic rows
        K = args.cpu_rounds + 1
        parity = _forest_parity(O, fo_gpu, [fo_a, fo_b], K, N_FEAT, dev, n_rows)
        del bins_cpu
        # ---- second CPU baseline (SURVEY 8d "always"): sklearn HistGradientBoostingClassifier, all cores, and config C0
        if not args.no_hgb:
            hgb = _hgb_baseline(n_rows, N_FEAT, DEPTH, dev)
            c0 = _config_c0(O, dev)

    if rank == 0:
        line = {"metric": "boosting rounds/s", "value": round(rounds_per_s, 2), "unit": "rounds/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4),
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64 fixed-point (f32 gradients, f64 gains)",
                "data": "synthetic", "config": _config(n_rows, args.steps, args.warmup, world),
                "detail": {"total_bins": S, "effective_bins_per_feature": 255, "layout": layout, "cuda_graph": graph_state,
                           "exchange": ("NVLink peer memory inside the split scan" if peer_mode else "NCCL allreduce") if world > 1 else None},
                "reference_xgboost": _xgboost_probe(),
                "roofline": roofline, "cpu_baseline": cpu, "cpu_baseline_hgb": hgb, "parity": parity, "c0": c0,
                "e2e": e2e, "gpu_launches": int(launches),
                "clocks": clk.summary(), "scoring": scoring, "prep_ms": round(t_prep, 2), "train_auc_200k": auc}
        print(json.dumps(line), file=_OUT, flush=True)
    if world > 1:
        torch.distributed.barrier()               # rank 0 has the extra legs (TreeSHAP, printing): leave together
        torch.distributed.destroy_process_group()


def run_reference(args):
    """The reference's CPU path for this metric: oracle port (xgboost is not installable here), all host threads
    (torch.distributed.run exports OMP_NUM_THREADS=1: the thread count is set explicitly), one step = one boosting
    round on the SAME rows x features x depth as our arm (no row sampling unless --ref-rows asks for it)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import numpy as np
    import torch

    from cobalt_smart_lender_ai_b200.synth import make_lending
    from oracle import oracle as O
    O.set_threads(os.cpu_count() or 1)
    n = min(args.rows, args.ref_rows) if args.ref_rows > 0 else args.rows
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    X, y, _ = make_lending(n, N_FEAT, device=dev)
    X, y = X.cpu().numpy(), y.cpu().numpy()
    if dev == "cpu":
        # no GPU (this arm's contract test): the data was generated by torch's CPU thread pool, whose OpenMP runtime is
        # not the port's -- two pools fighting over the cores slowed the port 8x.  Park torch's, then size the port's
        # again (where the two share one runtime the first call would otherwise have set the port to one thread too).
        torch.set_num_threads(1)
        O.set_threads(os.cpu_count() or 1)
    spw = float((y == 0).sum() / (y == 1).sum())
    t0 = time.perf_counter()
    cuts, nc, mv = O.compute_cuts(X, 256)
    bins = O.bin_matrix(X, cuts, nc, ldb=160)
    t_prep = time.perf_counter() - t0
    bp = O.bin_ptr_of(nc)
    sg, sh = O.quant_shifts(max(1.0, spw))
    tp = O.TreeParams(eta=float(np.float32(ETA)), max_depth=DEPTH, sg=sg, sh=sh)
    margin = np.zeros(n, np.float32)
    O.train(bins, N_FEAT, bp, cuts, mv, y, None, spw, tp, args.warmup, margin)
    t0 = time.perf_counter()
    O.train(bins, N_FEAT, bp, cuts, mv, y, None, spw, tp, args.steps, margin, tree_offset=args.warmup)
    dt = time.perf_counter() - t0
    scale = n / args.rows
    val = args.steps / dt * scale
    xgb_val = _xgboost_rounds_per_s(X, y, spw, args.steps, DEPTH)
    kind = "port"
    if xgb_val is not None:                   # the box has the reference's own implementation: that is the reference arm
        val, kind, scale = xgb_val * scale, "reference", scale
    if xgb_val is not None:
        sample = (f"xgboost hist, n_jobs={os.cpu_count()}, {n} x {N_FEAT}: ({2 + args.steps}-round fit - 2-round fit) / {args.steps}")
    elif n == args.rows:
        sample = (f"every step is one round on the full {n} x {N_FEAT} binned matrix; oracle/gbdt_oracle.c with OpenMP on "
                  f"{O.num_threads()} threads; reference xgboost unavailable on this box")
    else:
        sample = (f"{n} of {args.rows} rows per step (value = measured rounds/s x {scale:.3f}, linear in rows); "
                  f"oracle/gbdt_oracle.c with OpenMP; reference xgboost unavailable on this box")
    line = {"impl": "reference", "metric": "boosting rounds/s", "value": round(val, 3), "unit": "rounds/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3 / scale, 2),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int64 fixed-point (f32 gradients, f64 gains)",
            "data": "synthetic", "config": _config(args.rows, args.steps, args.warmup, args.gpus),
            "cpu_baseline": {"value": round(val, 3), "unit": "rounds/s", "cores": O.num_threads(), "kind": kind, "sample": sample},
            "e2e": {"value": round(val, 3), "unit": "rounds/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "raw_rounds_per_s_on_sample": round(args.steps / dt, 3), "prep_seconds": round(t_prep, 2),
            "reference_xgboost": _xgboost_probe()}
    print(json.dumps(line), file=_OUT, flush=True)


def _quiet_stdout():
    """Libraries (NCCL prints its version) may write to fd 1; the contract is ONE JSON line on stdout.
    Point fd 1 at stderr for the run and return a writer on the original stdout."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    return os.fdopen(saved, "w")


def main():
    global _OUT, DEPTH, WORKLOAD
    _OUT = _quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=N_ROWS)
    ap.add_argument("--depth", type=int, default=DEPTH, help="tree depth; 8 with --rows 20000000 --gpus 8 is configs[3]")
    ap.add_argument("--e2e-rounds", type=int, default=ROUNDS_CFG)
    ap.add_argument("--score-rows", type=int, default=SCORE_ROWS)
    ap.add_argument("--score-steps", type=int, default=3)
    ap.add_argument("--shap-rows", type=int, default=4096, help="rows explained with TreeSHAP in the scoring leg (0: skip)")
    ap.add_argument("--cpu-rounds", type=int, default=8)
    ap.add_argument("--ref-rows", type=int, default=0, help="reference arm: rows per step (0 = the full --rows, the default)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-scoring", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--schema", default="lending", choices=["lending", "continuous"],
                    help="continuous: jitter every column so that all 150 have 255 bins (measures the tiled fallback histogram kernel)")
    ap.add_argument("--no-groups", action="store_true", help="one byte per feature (no one-hot group columns)")
    ap.add_argument("--no-hgb", action="store_true", help="skip the sklearn HistGradientBoosting baseline and config C0")
    args = ap.parse_args()
    if args.schema != "lending":
        DEPTH = args.depth
        WORKLOAD = f"{args.rows / 1e6:g}M x {N_FEAT} all-continuous columns (150 x 255 bins), depth {args.depth}: NOT a BASELINE config, fallback-kernel measurement"
    elif args.depth != DEPTH or args.rows != N_ROWS:
        DEPTH = args.depth
        WORKLOAD = (f"synthetic Lending-Club schema {args.rows / 1e6:g}M x {N_FEAT}, 256 bins, depth {DEPTH}"
                    + (" (BASELINE.json configs[3])" if (args.rows, DEPTH) == (20_000_000, 8) else ""))
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
