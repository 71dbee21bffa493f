"""TEST INFRASTRUCTURE (oracle side) -- decoder for the reference's shipped model file.

Reads ``src/api/models/xgb_model_tree.pkl`` of the reference WITHOUT unpickling it
(unpickling would import ``xgboost.sklearn``, which is not installed here): the pickle
opcode stream is walked with ``pickletools.genops`` and the single large BINBYTES
payload (the UBJSON booster) is handed to a small UBJSON
reader.  Used only by ``tests/golden/make_golden.py`` to produce the committed golden
vectors; the product has its own, independent UBJSON reader
(``cobalt_smart_lender_ai_b200/ubjson.py``).

Reference call sites that load this file: src/api/cobalt_fast_api.py:45 (joblib.load).
"""
from __future__ import annotations

import pickletools
import struct
from typing import Any

import numpy as np

_INT_FMT = {b"i": (">b", 1), b"U": (">B", 1), b"I": (">h", 2), b"l": (">i", 4), b"L": (">q", 8)}
_ARR_DT = {b"i": ">i1", b"U": ">u1", b"I": ">i2", b"l": ">i4", b"L": ">i8", b"d": ">f4", b"D": ">f8"}


class _UBJ:
    def __init__(self, buf: bytes):
        self.b = buf
        self.p = 0

    def _tag(self) -> bytes:
        t = self.b[self.p:self.p + 1]
        self.p += 1
        return t

    def _int(self, tag: bytes) -> int:
        fmt, n = _INT_FMT[tag]
        v = struct.unpack_from(fmt, self.b, self.p)[0]
        self.p += n
        return v

    def _str(self) -> str:
        n = self._int(self._tag())
        s = self.b[self.p:self.p + n].decode("utf-8")
        self.p += n
        return s

    def value(self, tag: bytes | None = None) -> Any:
        tag = tag or self._tag()
        if tag in _INT_FMT:
            return self._int(tag)
        if tag == b"d":
            v = struct.unpack_from(">f", self.b, self.p)[0]
            self.p += 4
            return v
        if tag == b"D":
            v = struct.unpack_from(">d", self.b, self.p)[0]
            self.p += 8
            return v
        if tag == b"S":
            return self._str()
        if tag == b"T":
            return True
        if tag == b"F":
            return False
        if tag == b"Z":
            return None
        if tag == b"C":
            return chr(self._int(b"U"))
        if tag == b"{":
            return self._obj()
        if tag == b"[":
            return self._arr()
        raise ValueError(f"UBJSON: unknown tag {tag!r} at {self.p - 1}")

    def _obj(self) -> dict:
        out = {}
        while self.b[self.p:self.p + 1] != b"}":
            k = self._str()
            out[k] = self.value()
        self.p += 1
        return out

    def _arr(self):
        if self.b[self.p:self.p + 1] == b"$":
            self.p += 1
            et = self._tag()
            assert self._tag() == b"#"
            cnt = self._int(self._tag())
            dt = np.dtype(_ARR_DT[et])
            a = np.frombuffer(self.b, dtype=dt, count=cnt, offset=self.p)
            self.p += cnt * dt.itemsize
            return a.astype(dt.newbyteorder("="))
        if self.b[self.p:self.p + 1] == b"#":
            self.p += 1
            cnt = self._int(self._tag())
            return [self.value() for _ in range(cnt)]
        out = []
        while self.b[self.p:self.p + 1] != b"]":
            out.append(self.value())
        self.p += 1
        return out


def booster_bytes_from_pickle(path: str) -> bytes:
    """Largest bytes-like constant in the pickle stream = the UBJSON booster."""
    best = b""
    with open(path, "rb") as fh:
        data = fh.read()
    for op, arg, _pos in pickletools.genops(data):
        if op.name in ("BINBYTES", "BINBYTES8", "SHORT_BINBYTES", "BYTEARRAY8") and arg is not None:
            if len(arg) > len(best):
                best = bytes(arg)
    if not best:
        raise ValueError("no bytes payload found in pickle")
    return best


def load_fixture_model(path: str) -> dict:
    return _UBJ(booster_bytes_from_pickle(path)).value()


def forest_soa(model: dict) -> dict:
    """Flatten the decoded booster into concatenated per-node arrays + tree offsets."""
    learner = model["learner"] if "learner" in model else model["Model"]["learner"]
    trees = learner["gradient_booster"]["model"]["trees"]
    keys_i = ["left_children", "right_children", "parents", "split_indices"]
    keys_f = ["split_conditions", "base_weights", "loss_changes", "sum_hessian"]
    out: dict[str, Any] = {}
    offs = np.zeros(len(trees) + 1, np.int64)
    for t, tr in enumerate(trees):
        offs[t + 1] = offs[t] + int(tr["tree_param"]["num_nodes"])
    out["tree_offsets"] = offs
    for k in keys_i:
        out[k] = np.concatenate([np.asarray(tr[k], np.int32) for tr in trees])
    for k in keys_f:
        out[k] = np.concatenate([np.asarray(tr[k], np.float32) for tr in trees])
    out["default_left"] = np.concatenate([np.asarray(tr["default_left"], np.uint8) for tr in trees])
    out["feature_names"] = list(learner["feature_names"])
    out["base_score"] = float(learner["learner_model_param"]["base_score"])
    out["num_feature"] = int(learner["learner_model_param"]["num_feature"])
    return out
