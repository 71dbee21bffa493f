"""Reads an xgboost model file without xgboost: the reference's serving layer
loads ``models/xgb_model_tree.pkl`` with joblib (src/api/cobalt_fast_api.py:45), which needs the
xgboost package to unpickle.  Here the pickle is never executed: its opcode stream is scanned for
the booster's raw bytes (a UBJSON document), which are decoded into the forest arrays the B200
scorer consumes.  Also accepts a bare ``.ubj`` file saved by ``Booster.save_model``."""
from __future__ import annotations

import pickletools
import struct

import numpy as np

_SCALARS = {"i": ">b", "U": ">B", "I": ">h", "l": ">i", "L": ">q", "d": ">f", "D": ">d"}
_NP = {"i": ">i1", "U": ">u1", "I": ">i2", "l": ">i4", "L": ">i8", "d": ">f4", "D": ">f8"}


def decode(buf: bytes):
    """Minimal UBJSON decoder (typed/counted containers included); returns plain Python + numpy."""
    mv = memoryview(buf)
    pos = 0

    def scalar(tag: str):
        nonlocal pos
        fmt = _SCALARS[tag]
        (v,) = struct.unpack_from(fmt, mv, pos)
        pos += struct.calcsize(fmt)
        return v

    def read_len():
        nonlocal pos
        tag = chr(mv[pos])
        pos += 1
        return scalar(tag)

    def read_str():
        nonlocal pos
        n = read_len()
        s = bytes(mv[pos:pos + n]).decode("utf-8")
        pos += n
        return s

    def value(tag: str | None = None):
        nonlocal pos
        if tag is None:
            tag = chr(mv[pos])
            pos += 1
        if tag in _SCALARS:
            return scalar(tag)
        if tag == "S":
            return read_str()
        if tag == "C":
            pos += 1
            return chr(mv[pos - 1])
        if tag == "T":
            return True
        if tag == "F":
            return False
        if tag in "ZN":
            return None
        if tag == "[":
            typ = cnt = None
            if chr(mv[pos]) == "$":
                typ = chr(mv[pos + 1])
                pos += 2
            if chr(mv[pos]) == "#":
                pos += 1
                cnt = read_len()
            if typ is not None:
                dt = np.dtype(_NP[typ])
                arr = np.frombuffer(mv, dtype=dt, count=cnt, offset=pos).astype(dt.newbyteorder("="))
                pos += cnt * dt.itemsize
                return arr
            if cnt is not None:
                return [value() for _ in range(cnt)]
            out = []
            while chr(mv[pos]) != "]":
                out.append(value())
            pos += 1
            return out
        if tag == "{":
            out = {}
            cnt = None
            if chr(mv[pos]) == "#":
                pos += 1
                cnt = read_len()
            if cnt is not None:
                for _ in range(cnt):
                    k = read_str()
                    out[k] = value()
                return out
            while chr(mv[pos]) != "}":
                k = read_str()
                out[k] = value()
            pos += 1
            return out
        raise ValueError(f"UBJSON: unexpected tag {tag!r} at byte {pos - 1}")

    return value()


class NoBoosterPayload(ValueError):
    """The file holds no xgboost booster buffer (it may be a pickle of this package's own estimator)."""


def booster_bytes(path: str) -> bytes:
    """The xgboost Booster's raw buffer inside a pickle (largest bytes constant that opens a UBJSON object), or the
    file itself.  The pickle is only scanned opcode by opcode, never executed."""
    with open(path, "rb") as fh:
        data = fh.read()
    if data[:1] == b"{":
        return data
    best = b""
    try:
        for op, arg, _ in pickletools.genops(data):
            if (arg is not None and op.name.startswith(("BINBYTES", "SHORT_BINBYTES", "BYTEARRAY")) and len(arg) > len(best)
                    and bytes(arg[:1]) == b"{" and b"learner" in bytes(arg[:4096])):
                best = bytes(arg)
    except Exception:          # not a plain pickle stream (e.g. joblib's interleaved array payloads): keep what was found
        pass
    if not best:
        raise NoBoosterPayload(f"{path}: no booster payload found")
    return best


def forest_from_model(doc: dict) -> tuple[dict, list, float]:
    """(forest in wire format, feature names, base_score) from a decoded xgboost model document."""
    learner = (doc.get("Model") or doc)["learner"]
    trees = learner["gradient_booster"]["model"]["trees"]
    n = [int(t["tree_param"]["num_nodes"]) for t in trees]
    offs = np.zeros(len(trees) + 1, np.int64)
    np.cumsum(n, out=offs[1:])
    cat = lambda key, dt: np.concatenate([np.asarray(t[key], dt) for t in trees])  # noqa: E731
    forest = {"tree_offsets": offs, "left": cat("left_children", np.int32), "right": cat("right_children", np.int32),
              "parent": cat("parents", np.int32), "split_index": cat("split_indices", np.int32),
              "split_cond": cat("split_conditions", np.float32), "default_left": cat("default_left", np.uint8),
              "base_weight": cat("base_weights", np.float32), "loss_chg": cat("loss_changes", np.float32),
              "sum_hess": cat("sum_hessian", np.float32)}
    if any(len(t.get("categories", [])) for t in trees):
        raise ValueError("categorical splits are not supported")
    names = list(learner.get("feature_names") or [])
    nf = int(learner["learner_model_param"]["num_feature"])
    if not names:
        names = [f"f{i}" for i in range(nf)]
    obj = learner.get("objective", {}).get("name", "binary:logistic")
    if obj != "binary:logistic":
        raise ValueError(f"objective {obj!r} is not supported")
    return forest, names, float(learner["learner_model_param"]["base_score"])


def load_xgboost_pickle(path: str, device: str = "cuda"):
    """The reference's shipped model -> an XGBClassifier on the B200 scoring path."""
    from .sklearn import XGBClassifier
    forest, names, base = forest_from_model(decode(booster_bytes(path)))
    return XGBClassifier(device=device).load_fixture_forest(forest, names, base)


# ------------------------------------------------------------------------------------ export
def _w_int(out: bytearray, v: int) -> None:
    if -128 <= v < 128:
        out += b"i" + struct.pack(">b", v)
    elif -32768 <= v < 32768:
        out += b"I" + struct.pack(">h", v)
    elif -2 ** 31 <= v < 2 ** 31:
        out += b"l" + struct.pack(">i", v)
    else:
        out += b"L" + struct.pack(">q", v)


def _w_key(out: bytearray, s: str) -> None:
    b = s.encode("utf-8")
    _w_int(out, len(b))
    out += b


def encode(obj) -> bytes:
    """UBJSON encoder for the subset xgboost writes (objects, lists, strings, ints, floats, typed arrays)."""
    out = bytearray()

    def emit(v):
        nonlocal out
        if isinstance(v, dict):
            out += b"{"
            for k, x in v.items():
                _w_key(out, k)
                emit(x)
            out += b"}"
        elif isinstance(v, np.ndarray):
            tag = {"f4": b"d", "f8": b"D", "i4": b"l", "i8": b"L", "u1": b"U", "i1": b"i", "i2": b"I"}[v.dtype.str[1:]]
            out += b"[$" + tag + b"#"
            _w_int(out, int(v.size))
            out += v.astype(v.dtype.newbyteorder(">")).tobytes()
        elif isinstance(v, (list, tuple)):
            out += b"["
            for x in v:
                emit(x)
            out += b"]"
        elif isinstance(v, bool):
            out += b"T" if v else b"F"
        elif isinstance(v, (int, np.integer)):
            _w_int(out, int(v))
        elif isinstance(v, (float, np.floating)):
            out += b"D" + struct.pack(">d", float(v))
        elif isinstance(v, str):
            out += b"S"
            _w_key(out, v)
        elif v is None:
            out += b"Z"
        else:
            raise TypeError(type(v))

    emit(obj)
    return bytes(out)


def model_document(forest: dict, feature_names, base_score: float) -> dict:
    """A forest in wire format -> the xgboost model document layout."""
    offs = forest["tree_offsets"]
    trees = []
    for t in range(len(offs) - 1):
        a, b = int(offs[t]), int(offs[t + 1])
        n = b - a
        trees.append({
            "base_weights": np.asarray(forest["base_weight"][a:b], np.float32), "categories": np.zeros(0, np.int32),
            "categories_nodes": np.zeros(0, np.int32), "categories_segments": np.zeros(0, np.int64),
            "categories_sizes": np.zeros(0, np.int64), "default_left": np.asarray(forest["default_left"][a:b], np.uint8),
            "id": t, "left_children": np.asarray(forest["left"][a:b], np.int32),
            "loss_changes": np.asarray(forest["loss_chg"][a:b], np.float32),
            "parents": np.asarray(forest["parent"][a:b], np.int32),
            "right_children": np.asarray(forest["right"][a:b], np.int32),
            "split_conditions": np.asarray(forest["split_cond"][a:b], np.float32),
            "split_indices": np.asarray(forest["split_index"][a:b], np.int32),
            "split_type": np.zeros(n, np.uint8), "sum_hessian": np.asarray(forest["sum_hess"][a:b], np.float32),
            "tree_param": {"num_deleted": "0", "num_feature": str(len(feature_names)), "num_nodes": str(n),
                           "size_leaf_vector": "1"}})
    T = len(trees)
    return {"learner": {
        "attributes": {}, "feature_names": list(feature_names), "feature_types": ["float"] * len(feature_names),
        "gradient_booster": {"model": {"gbtree_model_param": {"num_parallel_tree": "1", "num_trees": str(T)},
                                       "iteration_indptr": np.arange(T + 1, dtype=np.int32),
                                       "tree_info": np.zeros(T, np.int32), "trees": trees}, "name": "gbtree"},
        "learner_model_param": {"base_score": repr(float(base_score)), "boost_from_average": "1", "num_class": "0",
                                "num_feature": str(len(feature_names)), "num_target": "1"},
        "objective": {"name": "binary:logistic", "reg_loss_param": {"scale_pos_weight": "1"}}},
        "version": [3, 0, 0]}


def save_model(booster, path: str) -> None:
    """Writes a Booster as an xgboost-layout ``.ubj`` file."""
    with open(path, "wb") as fh:
        fh.write(encode(model_document(booster.forest, booster._names(), booster.base_score)))
