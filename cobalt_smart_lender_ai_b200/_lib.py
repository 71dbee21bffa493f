"""ctypes binding of libcobalt_gbdt.so (include/cobalt_gbdt.h).

There is deliberately no fallback: if the shared object cannot be built or loaded, or a
call returns a non-zero status, a ``CobaltError`` is raised.  Nothing on this path ever
routes to a CPU implementation.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
import threading

_PKG = os.path.dirname(os.path.abspath(__file__))
_CSRC = os.path.join(_PKG, "csrc")
LIB_PATH = os.path.join(_PKG, "libcobalt_gbdt.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]

CB_QBITS = 17
CB_MISSING_BIN = 255
CB_CUT_STRIDE = 256


class CobaltError(RuntimeError):
    pass


class TreeParams(C.Structure):
    _fields_ = [("eta", C.c_float), ("reg_lambda", C.c_float), ("reg_alpha", C.c_float), ("gamma", C.c_float),
                ("min_child_weight", C.c_float), ("max_depth", C.c_int32), ("sg", C.c_int32), ("sh", C.c_int32)]


class Split(C.Structure):
    _fields_ = [("loss_chg", C.c_float), ("feature", C.c_int32), ("split_bin", C.c_int32),
                ("default_left", C.c_int32), ("threshold", C.c_float), ("pad_", C.c_int32),
                ("GL", C.c_int64), ("HL", C.c_int64), ("GR", C.c_int64), ("HR", C.c_int64)]


class TrainConfig(C.Structure):
    _fields_ = [("n_rows", C.c_int64), ("ldb", C.c_int64), ("row_offset", C.c_int64), ("n_features", C.c_int32),
                ("total_bins", C.c_int32), ("max_trees", C.c_int32), ("world_size", C.c_int32),
                ("tree", TreeParams), ("scale_pos_weight", C.c_float), ("keep_thresh", C.c_uint32),
                ("seed", C.c_uint32), ("blocked", C.c_int32), ("n_groups", C.c_int32)]


class TrainBuffers(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in (
        "bins", "y", "weight", "margin", "bin_ptr", "cuts", "min_vals", "pos_feature", "bins_cm", "left", "right", "parent", "split_index",
        "split_bin", "split_cond", "default_left", "base_weight", "loss_chg", "sum_hess", "n_nodes", "grp_ptr", "grp_feat")]


def sources() -> list[str]:
    return sorted(os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cu", ".cuh"))) + [
        os.path.join(os.path.dirname(_PKG), "include", "cobalt_gbdt.h")]


def build(force: bool = False, verbose: bool = False) -> str:
    """Compiles csrc/cobalt_gbdt.cu for sm_100a with nvcc (cross-compiles without a GPU).

    Safe with one process per GPU: an flock serialises the builders, the staleness check is repeated under the lock, and
    nvcc writes to a temporary file that is renamed into place, so no rank ever dlopens a half-written library."""
    srcs = sources()

    def fresh() -> bool:
        return os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs)
    if not force and fresh():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise CobaltError("libcobalt_gbdt.so is missing or stale and nvcc was not found to rebuild it")
    import fcntl
    with open(LIB_PATH + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and fresh():          # another process built it while we waited
                return LIB_PATH
            tmp = f"{LIB_PATH}.tmp.{os.getpid()}"
            cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", tmp,
                                                                                  os.path.join(_CSRC, "cobalt_gbdt.cu")]
            res = subprocess.run(cmd, capture_output=True, text=True)
            if res.returncode != 0:
                if os.path.exists(tmp):
                    os.unlink(tmp)
                raise CobaltError("nvcc failed:\n" + res.stderr[-4000:])
            os.replace(tmp, LIB_PATH)
            if verbose:
                print(res.stderr)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB_PATH


_lock = threading.Lock()
_lib = None

_VP, _I, _I64, _F, _U32, _SZ = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_uint32, C.c_size_t
_SIGS = {
    "cb_version": (C.c_int, []),
    "cb_last_error": (C.c_char_p, []),
    "cb_device_info": (C.c_int, [C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "cb_sketch_workspace_bytes": (C.c_int, [_I64, _I, C.POINTER(_SZ)]),
    "cb_sketch_cuts": (C.c_int, [_VP, _I64, _I, _I64, _I, _VP, _VP, _VP, _VP, _SZ, _VP]),
    "cb_onehot_pair_counts": (C.c_int, [_VP, _I64, _I, _I64, _VP, _VP, _I, _VP, _VP, _VP]),
    "cb_plan_layout": (C.c_int, [_VP, _I, _VP, _VP, _I, _I64, _VP, C.POINTER(C.c_int32)]),
    "cb_bin_pack_u8": (C.c_int, [_VP, _I64, _I, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _I64, _I, _VP]),
    "cb_bin_transpose_u8": (C.c_int, [_VP, _I64, _I64, _I, _VP, _VP]),
    "cb_grad_logistic_q": (C.c_int, [_VP, _VP, _VP, _I64, _F, _I, _I, _U32, _U32, _U32, _I64, _VP, _VP, _VP]),
    "cb_sigmoid": (C.c_int, [_VP, _I64, _VP, _VP]),
    "cb_hist_workspace_bytes": (C.c_int, [_I64, _I, C.POINTER(_SZ)]),
    "cb_hist_build": (C.c_int, [_VP, _I64, _I, _VP, _VP, _I64, _VP, _I, _VP, _VP, _SZ, _VP]),
    "cb_hist_build_cf_workspace_bytes": (C.c_int, [_I64, _I, C.POINTER(_SZ)]),
    "cb_hist_build_cf_sync": (C.c_int, [_VP, _I64, _I64, _I, _VP, _VP, _VP, _I, _I, _VP, _VP, _I64, _VP, _I, _VP, _VP, _SZ, _VP]),
    "cb_hist_subtract": (C.c_int, [_VP, _VP, _VP, _I64, _VP]),
    "cb_split_scan_workspace_bytes": (C.c_int, [_I, C.POINTER(_SZ)]),
    "cb_split_scan": (C.c_int, [_VP, _VP, _I, _VP, _VP, _VP, C.POINTER(TreeParams), _VP, _VP, _VP, _SZ, _VP]),
    "cb_split_scan_noalloc": (C.c_int, [_VP, _VP, _I, _VP, _VP, _VP, C.POINTER(TreeParams), _VP, _VP, _VP]),
    "cb_partition_workspace_bytes": (C.c_int, [_I64, C.POINTER(_SZ)]),
    "cb_partition_stable": (C.c_int, [_VP, _I64, _VP, _I64, _I, _I, _I, _VP, _VP, _VP, _SZ, _VP]),
    "cb_leaf_update": (C.c_int, [_VP, _VP, _I64, _F, _VP]),
    "cb_predict": (C.c_int, [_VP, _I64, _I, _I64, _VP, _VP, _I, _F, _VP, _VP, _VP]),
    "cb_predict_heap_plan": (C.c_int, [_I, _I, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "cb_predict_heap": (C.c_int, [_VP, _I64, _I, _I64, _VP, _VP, _I, _I, _F, _VP, _VP, _VP]),
    "cb_tree_shap_workspace_bytes": (C.c_int, [_I64, _I, _I, _I64, C.POINTER(C.c_size_t)]),
    "cb_tree_shap": (C.c_int, [_VP, _I64, _I, _I64, _VP, _I, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _F, _I, _VP, _VP,
                               C.c_size_t, _VP]),
    "cb_predict_q8_plan": (C.c_int, [_I, _I, _I64, C.POINTER(C.c_int), C.POINTER(C.c_uint32)]),
    "cb_bin_rows_u8": (C.c_int, [_VP, _I64, _I, _I64, _VP, _VP, _VP, _I64, _VP]),
    "cb_predict_q8": (C.c_int, [_VP, _I64, _I, _I64, _VP, _I, _I, _F, C.c_uint32, _VP, _VP, _VP]),
    "cb_trainer_workspace_bytes": (C.c_int, [C.POINTER(TrainConfig), C.POINTER(_SZ)]),
    "cb_trainer_create": (C.c_int, [C.POINTER(TrainConfig), C.POINTER(TrainBuffers), _VP, _SZ, C.POINTER(_VP)]),
    "cb_trainer_destroy": (None, [_VP]),
    "cb_trainer_layout_info": (C.c_int, [_VP, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "cb_trainer_boost_round": (C.c_int, [_VP, _I, _VP, _VP]),
    "cb_trainer_round_begin": (C.c_int, [_VP, _I, _VP, _VP]),
    "cb_trainer_level_hist": (C.c_int, [_VP, _I, _VP]),
    "cb_trainer_level_split": (C.c_int, [_VP, _I, _VP]),
    "cb_trainer_round_end": (C.c_int, [_VP, _VP]),
    "cb_trainer_level_buffer": (C.c_int, [_VP, _I, C.POINTER(_VP), C.POINTER(_I64)]),
    "cb_sym_alloc": (C.c_int, [_SZ, C.POINTER(_VP), _VP]),
    "cb_sym_open": (C.c_int, [_VP, C.POINTER(_VP)]),
    "cb_sym_close": (C.c_int, [_VP]),
    "cb_sym_free": (C.c_int, [_VP]),
    "cb_trainer_sym_bytes": (C.c_int, [C.POINTER(TrainConfig), C.POINTER(_SZ)]),
    "cb_trainer_set_peers": (C.c_int, [_VP, _I, _I, C.POINTER(_VP)]),
    "cb_trainer_peer_status": (C.c_int, [_VP, C.POINTER(C.c_int)]),
    "cb_trainer_row_order": (C.c_int, [_VP, C.POINTER(_VP)]),
    "cb_trainer_node_hist": (C.c_int, [_VP, _I, C.POINTER(_VP)]),
    "cb_trainer_launch_count": (C.c_int64, [_VP]),
    "cb_trainer_profile_begin": (C.c_int, [_VP]),
    "cb_trainer_profile_end_sync": (C.c_int, [_VP, C.POINTER(C.c_float), C.POINTER(C.c_char_p), _I, C.POINTER(C.c_int)]),
    "cb_trainer_enable_timing": (C.c_int, [_VP, _I]),
    "cb_trainer_last_hist_timing_sync": (C.c_int, [_VP, C.POINTER(C.c_float), C.POINTER(C.c_float),
                                                    C.POINTER(_I64), C.POINTER(C.c_int)]),
}
EXPORTS = tuple(_SIGS)


def load():
    """Loads (building first if needed) the shared object; raises CobaltError when impossible."""
    global _lib
    with _lock:
        if _lib is None:
            path = build()
            try:
                lib = C.CDLL(path)
            except OSError as e:  # pragma: no cover
                raise CobaltError(f"cannot load {path}: {e}") from e
            for name, (res, args) in _SIGS.items():
                fn = getattr(lib, name, None)
                if fn is None:
                    raise CobaltError(f"{path} does not export {name}")
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().cb_last_error()
        raise CobaltError(f"libcobalt_gbdt status {rc}: {msg.decode() if msg else ''}")
