"""Boosting driver: owns the device buffers of one fit and steps the fused C-ABI trainer.

Replaces the body of ``XGBClassifier.fit`` (reference call sites
src/model_train_test/model_tree_train_test.py:118,159).  Python only sequences calls; all
arithmetic, split decisions and row movement happen in libcobalt_gbdt.so on the device.
With ``world_size > 1`` rows are sharded across ranks and the per-level histogram buffer is
summed with ``torch.distributed.all_reduce`` (NCCL over NVLink) between the two halves of a
level -- the path's one collective.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _lib, ops
from ._lib import TrainBuffers, TrainConfig, TreeParams, check

FOREST_FIELDS = (("left", torch.int32), ("right", torch.int32), ("parent", torch.int32),
                 ("split_index", torch.int32), ("split_bin", torch.int32), ("split_cond", torch.float32),
                 ("default_left", torch.uint8), ("base_weight", torch.float32), ("loss_chg", torch.float32),
                 ("sum_hess", torch.float32))


class GpuTrainer:
    def __init__(self, bins, y: torch.Tensor, weight: torch.Tensor | None, n_features: int,
                 bin_ptr: torch.Tensor, cuts: torch.Tensor, min_vals: torch.Tensor, *, max_depth: int = 6,
                 eta: float = 0.3, reg_lambda: float = 1.0, reg_alpha: float = 0.0, gamma: float = 0.0,
                 min_child_weight: float = 1.0, scale_pos_weight: float = 1.0, sg: int, sh: int,
                 subsample: float = 1.0, seed: int = 0, max_trees: int = 100, base_margin: float = 0.0,
                 row_offset: int = 0, group=None, pos_feature: torch.Tensor | None = None, blocked: bool = False,
                 bins_cm: torch.Tensor | None = None):
        grp_ptr = grp_feat = None
        n_groups = 0
        if hasattr(bins, "natural_codes"):               # a binned.BinnedMatrix carries its own placement
            bm = bins
            bins, pos_feature, blocked, bins_cm = bm.bins, bm.pos, bm.blocked, bm.bins_cm
            grp_ptr, grp_feat, n_groups = bm.grp_ptr, bm.grp_feat, bm.n_groups
        if not bins.is_cuda:
            raise _lib.CobaltError("GpuTrainer needs CUDA tensors: this package has no CPU path")
        self.lib = _lib.load()
        self.dev = bins.device
        self.group, self.world = None, 1
        from . import parallel
        if group is not False and parallel._dist() is not None:
            w = torch.distributed.get_world_size(group)
            if w > 1:
                self.world, self.group = w, group
        n, ldb = bins.shape
        total_bins = int(bin_ptr[-1].item())
        self.n, self.f, self.max_depth, self.max_trees = n, n_features, max_depth, max_trees
        self.max_nodes = (1 << (max_depth + 1)) - 1
        self.bins, self.y, self.weight = bins, y.to(torch.uint8).contiguous(), weight
        self.bin_ptr, self.cuts, self.min_vals = bin_ptr, cuts, min_vals
        self.pos_feature = pos_feature
        self.bins_cm = bins_cm
        self._grp = (grp_ptr, grp_feat)                  # keep the device arrays alive
        self.margin = torch.full((n,), float(base_margin), dtype=torch.float32, device=self.dev)
        self.forest = {k: torch.zeros((max_trees, self.max_nodes), dtype=dt, device=self.dev) for k, dt in FOREST_FIELDS}
        self.n_nodes = torch.zeros(max_trees, dtype=torch.int32, device=self.dev)
        tp = TreeParams(float(np.float32(eta)), float(reg_lambda), float(reg_alpha), float(gamma),
                        float(min_child_weight), int(max_depth), int(sg), int(sh))
        self.cfg = TrainConfig(n, ldb, int(row_offset), n_features, total_bins, max_trees, self.world, tp,
                               float(scale_pos_weight), ops.keep_threshold(subsample), int(seed) & 0xFFFFFFFF,
                               int(bool(blocked)), int(n_groups))
        need = C.c_size_t()
        check(self.lib.cb_trainer_workspace_bytes(C.byref(self.cfg), C.byref(need)))
        self.ws = ops.workspace(need.value, self.dev)
        wp, wb = ops._aligned(self.ws)
        self._ws_base = wp.value
        p = lambda t: None if t is None else t.data_ptr()  # noqa: E731
        self.buffers = TrainBuffers(p(self.bins), p(self.y), p(self.weight), p(self.margin), p(self.bin_ptr),
                                    p(self.cuts), p(self.min_vals), p(self.pos_feature), p(self.bins_cm), *[p(self.forest[k]) for k, _ in FOREST_FIELDS],
                                    p(self.n_nodes), p(grp_ptr), p(grp_feat))
        h = C.c_void_p()
        check(self.lib.cb_trainer_create(C.byref(self.cfg), C.byref(self.buffers), wp, wb, C.byref(h)))
        self.handle = h
        self.peer_mode = False
        self._sym_own = None
        self._sym_peers = []
        # measured on 8 x B200: NCCL's small-message all-reduce is as fast as or faster than
        # the peer-memory exchange at 2, 4 and 8 ranks, so NCCL is the default and "peer" is opt-in
        mode = os.environ.get("COBALT_MULTIGPU", "nccl")
        if mode not in ("peer", "nccl"):
            raise ValueError(f"COBALT_MULTIGPU={mode!r}: expected peer or nccl")
        if self.world > 1 and mode == "peer":
            self._setup_peers()
        self.n_trees = 0
        self._level_views = {}
        self._graphs = {}
        self._graph_launches = 0
        self._launches_per_round = 0
        self.graph_error = None

    def _setup_peers(self) -> None:
        """Maps every rank's symmetric region into this process (CUDA IPC) and switches the library to the
        fused form of the collective: no all-reduce calls, peers read each other's level buffers."""
        dist = torch.distributed
        rank = dist.get_rank(self.group)
        need = C.c_size_t()
        check(self.lib.cb_trainer_sym_bytes(C.byref(self.cfg), C.byref(need)))
        own, handle = C.c_void_p(), (C.c_ubyte * 64)()
        check(self.lib.cb_sym_alloc(need.value, C.byref(own), handle))
        self._sym_own = own
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=self.group)
        regions = (C.c_void_p * self.world)()
        for r in range(self.world):
            if r == rank:
                regions[r] = own.value
            else:
                p = C.c_void_p()
                buf = (C.c_ubyte * 64).from_buffer_copy(handles[r])
                check(self.lib.cb_sym_open(buf, C.byref(p)))
                self._sym_peers.append(p)
                regions[r] = p.value
        check(self.lib.cb_trainer_set_peers(self.handle, rank, self.world, regions))
        torch.cuda.synchronize()
        dist.barrier(group=self.group)
        self.peer_mode = True

    def check_peers(self) -> None:
        """Device-side waits are bounded (a peer GPU's flag; the meeting point of a node's histogram CTAs): a wait
        that gave up leaves a flag behind, and the forest is then not to be trusted."""
        flag = C.c_int(0)
        check(self.lib.cb_trainer_peer_status(self.handle, C.byref(flag)))
        if flag.value:
            raise _lib.CobaltError("a device-side wait timed out during training (code %d); results are invalid" % flag.value)

    def close(self):
        if getattr(self, "handle", None):
            torch.cuda.synchronize()
            self._graphs = {}                    # captured rounds reference this trainer's buffers
            if self.peer_mode:
                torch.cuda.synchronize()
                try:
                    torch.distributed.barrier(group=self.group)     # nobody may still be reading our region
                except Exception:
                    pass
                for p in self._sym_peers:
                    self.lib.cb_sym_close(p)
                self.lib.cb_sym_free(self._sym_own)
                self._sym_peers, self._sym_own, self.peer_mode = [], None, False
            self.lib.cb_trainer_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _level_view(self, level: int) -> torch.Tensor:
        v = self._level_views.get(level)
        if v is None:
            ptr, nv = C.c_void_p(), C.c_int64()
            check(self.lib.cb_trainer_level_buffer(self.handle, level, C.byref(ptr), C.byref(nv)))
            off = ptr.value - self.ws.data_ptr()
            v = self.ws[off: off + nv.value * 8].view(torch.int64)
            self._level_views[level] = v
        return v

    def _enqueue_round(self, tree_index: int, feature_masks: torch.Tensor | None) -> None:
        """One boosting round on the current stream.  tree_index < 0 continues the device-side counter."""
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        fm = None if feature_masks is None else C.c_void_p(feature_masks.data_ptr())
        if self.world == 1 or self.peer_mode:
            check(self.lib.cb_trainer_boost_round(self.handle, tree_index, fm, st))
        else:
            check(self.lib.cb_trainer_round_begin(self.handle, tree_index, fm, st))
            for level in range(self.max_depth):
                check(self.lib.cb_trainer_level_hist(self.handle, level, st))
                with torch.cuda.nvtx.range("cobalt:allreduce"):
                    torch.distributed.all_reduce(self._level_view(level), group=self.group)
                check(self.lib.cb_trainer_level_split(self.handle, level, st))
            check(self.lib.cb_trainer_round_end(self.handle, st))

    def boost_one(self, feature_masks: torch.Tensor | None = None) -> None:
        """feature_masks: u8 [max_trees, f] on the device (row t masks tree t) or None."""
        if self.n_trees >= self.max_trees:
            raise _lib.CobaltError("forest capacity exhausted")
        self._enqueue_round(self.n_trees, feature_masks)
        self.n_trees += 1

    def boost(self, n_rounds: int, feature_masks: torch.Tensor | None = None, use_graph: bool | None = None) -> None:
        """n_rounds boosting rounds.  From the second round on, a CUDA graph of one round (kernels and,
        when sharded, the NCCL all-reduces) is replayed: the launch sequence of a round is fixed and
        the forest row comes from a device-side counter, so no host work remains per round."""
        if self.n_trees + n_rounds > self.max_trees:
            raise _lib.CobaltError("forest capacity exhausted")
        if feature_masks is not None:
            assert feature_masks.dtype == torch.uint8 and feature_masks.shape == (self.max_trees, self.f)
        if use_graph is None:
            use_graph = os.environ.get("COBALT_CUDA_GRAPH", "1") != "0"
        done = 0
        if use_graph and n_rounds >= 4:
            self.boost_one(feature_masks)          # eager: warms up NCCL and fixes the counter
            done = 1
            graph = self._capture_round(feature_masks)
            if graph is not None:
                for _ in range(n_rounds - 1):
                    graph.replay()
                self.n_trees += n_rounds - 1
                self._graph_launches += self._launches_per_round * (n_rounds - 1)
                done = n_rounds
        for _ in range(n_rounds - done):
            self.boost_one(feature_masks)

    def _capture_round(self, feature_masks):
        key = None if feature_masks is None else feature_masks.data_ptr()
        if key in self._graphs:
            return self._graphs[key]
        graph = None
        try:
            l0 = self.launch_count_raw()
            g = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream(device=self.dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.graph(g, stream=side, capture_error_mode="thread_local"):
                self._enqueue_round(-1, feature_masks)
            torch.cuda.current_stream().wait_stream(side)
            self._launches_per_round = self.launch_count_raw() - l0
            self._graph_launches -= self._launches_per_round      # the capture pass launched nothing
            graph = g
        except Exception as e:  # capture unsupported (e.g. the collective): stay eager, loudly
            self.graph_error = repr(e)
            torch.cuda.synchronize()
        self._graphs[key] = graph
        return graph

    def profile_round(self, feature_masks: torch.Tensor | None = None):
        """Runs ONE eager round with a CUDA event after every launch; returns [(step name, ms), ...]."""
        check(self.lib.cb_trainer_profile_begin(self.handle))
        self.boost_one(feature_masks)
        ms = (C.c_float * 96)()
        names = (C.c_char_p * 96)()
        n = C.c_int(0)
        check(self.lib.cb_trainer_profile_end_sync(self.handle, ms, names, 96, C.byref(n)))
        return [(names[i].decode(), float(ms[i])) for i in range(n.value)]

    def layout_info(self) -> dict:
        a, b, d = C.c_int(), C.c_int(), C.c_int()
        check(self.lib.cb_trainer_layout_info(self.handle, C.byref(a), C.byref(b), C.byref(d)))
        return {"conflict_free": bool(a.value), "smem_rows": b.value, "smem_bytes": d.value}

    def launch_count_raw(self) -> int:
        return int(self.lib.cb_trainer_launch_count(self.handle))

    def launch_count(self) -> int:
        """Kernels launched so far, counting graph replays."""
        return self.launch_count_raw() + self._graph_launches

    def row_order(self) -> torch.Tensor | None:
        ptr = C.c_void_p()
        check(self.lib.cb_trainer_row_order(self.handle, C.byref(ptr)))
        if not ptr.value:
            return None
        off = ptr.value - self.ws.data_ptr()
        return self.ws[off: off + self.n * 4].view(torch.int32)

    def node_hist(self, nid: int) -> torch.Tensor:
        ptr = C.c_void_p()
        check(self.lib.cb_trainer_node_hist(self.handle, nid, C.byref(ptr)))
        off = ptr.value - self.ws.data_ptr()
        s2 = 2 * self.cfg.total_bins
        return self.ws[off: off + s2 * 8].view(torch.int64).view(-1, 2)

    def forest_numpy(self) -> dict:
        """Padded [T][max_nodes] device arrays -> flat wire-format arrays on the host."""
        self.check_peers()
        t = self.n_trees
        nn = self.n_nodes[:t].cpu().numpy()
        offs = np.zeros(t + 1, np.int64)
        np.cumsum(nn, out=offs[1:])
        out = {"tree_offsets": offs, "n_nodes": nn}
        for k, _ in FOREST_FIELDS:
            a = self.forest[k][:t].cpu().numpy()
            out[k] = np.concatenate([a[i, :nn[i]] for i in range(t)]) if t else a.reshape(-1)
        return out
