"""ctypes binding of oracle/fpgrowth_ref.cpp — TEST INFRASTRUCTURE ONLY (see that file's
header; parity unpinned upstream). Builds the library with ``make`` on first use when the
prebuilt ``_build/liboracle_fpgrowth.so`` is absent."""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Dict, FrozenSet, Tuple

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_fpgrowth.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "fpgrowth_ref.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-s"], check=True)
    return _SO


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(build())
        lib.oracle_fpgrowth_run.restype = ctypes.c_void_p
        lib.oracle_fpgrowth_run.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                            ctypes.c_int32, ctypes.c_int64, ctypes.c_int64,
                                            ctypes.c_int, ctypes.c_int, ctypes.c_int]
        for name, res in (("num_itemsets", ctypes.c_uint64), ("checksum", ctypes.c_uint64),
                          ("total_items", ctypes.c_int64), ("overflow", ctypes.c_int)):
            f = getattr(lib, "oracle_fpgrowth_" + name)
            f.restype, f.argtypes = res, [ctypes.c_void_p]
        lib.oracle_fpgrowth_level_counts.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        lib.oracle_fpgrowth_copy.argtypes = [ctypes.c_void_p] * 4
        lib.oracle_fpgrowth_free.argtypes = [ctypes.c_void_p]
        lib.oracle_rules.restype = ctypes.c_int64
        lib.oracle_rules.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_double, ctypes.c_int,
                                     ctypes.c_void_p]
        lib.oracle_rules_last_index_seconds.restype = ctypes.c_double
        lib.oracle_rules_last_index_seconds.argtypes = []
        lib.oracle_rules_last_splits.restype = ctypes.c_int64
        lib.oracle_rules_last_splits.argtypes = []
        lib.oracle_rules_strided.restype = ctypes.c_int64
        lib.oracle_rules_strided.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_double,
                                             ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        _lib = lib
    return _lib


class Mined:
    """level_counts[k] = number of frequent k-itemsets; checksum is order-independent
    (same formula as ``itemset_checksum`` below); offsets/items/counts only if stored."""

    def __init__(self, n, checksum, level_counts, offsets=None, items=None, counts=None):
        self.n, self.checksum, self.level_counts = n, checksum, level_counts
        self.offsets, self.items, self.counts = offsets, items, counts

    def as_dict(self) -> Dict[FrozenSet[int], int]:
        return {frozenset(self.items[self.offsets[i]:self.offsets[i + 1]].tolist()): int(self.counts[i])
                for i in range(self.n)}


def fpgrowth(offsets: np.ndarray, items: np.ndarray, n_items: int, t1: int, tk: int,
             max_len: int = 0, store: bool = True, n_threads: int = 0,
             rules_min_confidence: float | None = None) -> Mined:
    """With ``rules_min_confidence`` the association rules are generated too (needs
    store=True); the result then carries ``n_rules`` and ``rule_checksum``."""
    lib = _load()
    offsets = np.ascontiguousarray(offsets, dtype=np.int64)
    items = np.ascontiguousarray(items, dtype=np.int32)
    h = lib.oracle_fpgrowth_run(offsets.ctypes.data, items.ctypes.data, len(offsets) - 1,
                                n_items, t1, tk, max_len, int(store), n_threads)
    try:
        if lib.oracle_fpgrowth_overflow(h):
            raise RuntimeError("oracle: single path longer than 40 items cannot be enumerated")
        n = lib.oracle_fpgrowth_num_itemsets(h)
        lv = np.zeros(64, dtype=np.uint64)
        lib.oracle_fpgrowth_level_counts(h, lv.ctypes.data)
        out = Mined(int(n), int(lib.oracle_fpgrowth_checksum(h)), lv.astype(np.int64))
        out.n_rules = out.rule_checksum = None
        if rules_min_confidence is not None:
            cs = ctypes.c_uint64(0)
            out.n_rules = int(lib.oracle_rules(h, len(offsets) - 1, float(rules_min_confidence),
                                               n_threads, ctypes.byref(cs)))
            if out.n_rules < 0:
                raise RuntimeError("oracle: a subset of a frequent itemset is missing")
            out.rule_checksum = int(cs.value)
        if store:
            tot = lib.oracle_fpgrowth_total_items(h)
            out.offsets = np.zeros(n + 1, dtype=np.int64)
            out.items = np.zeros(max(tot, 1), dtype=np.int32)[:tot]
            out.counts = np.zeros(max(n, 1), dtype=np.int64)[:n]
            lib.oracle_fpgrowth_copy(h, out.offsets.ctypes.data, out.items.ctypes.data,
                                     out.counts.ctypes.data)
        return out
    finally:
        lib.oracle_fpgrowth_free(h)


_M64 = (1 << 64) - 1


def _mix64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def itemset_checksum(offsets: np.ndarray, items: np.ndarray, counts: np.ndarray) -> int:
    """Order-independent checksum of an itemset table (same arithmetic as Sink::emit in
    fpgrowth_ref.cpp): sum over itemsets of mix(sum_i mix(item_i+1) ^ mix(count*P + k))."""
    with np.errstate(over="ignore"):
        offsets = np.asarray(offsets, dtype=np.int64)
        k = np.diff(offsets).astype(np.uint64)
        hi = _mix64(np.asarray(items, dtype=np.int64).astype(np.uint64) + np.uint64(1))
        cs = np.concatenate([[np.uint64(0)], np.cumsum(hi, dtype=np.uint64)])
        h = cs[offsets[1:]] - cs[offsets[:-1]]
        c = np.asarray(counts, dtype=np.int64).astype(np.uint64)
        tot = _mix64(h ^ _mix64(c * np.uint64(0x100000001B3) + k))
        return int(np.sum(tot, dtype=np.uint64)) & _M64


def rule_checksum(rules) -> int:
    """Same checksum as oracle_rules() for an iterable of (antecedent ids, consequent ids,
    support, confidence, lift)."""
    tot = 0
    with np.errstate(over="ignore"):
        def h(ids):
            return int(np.sum(_mix64(np.asarray(sorted(ids), dtype=np.int64).astype(np.uint64) + np.uint64(1)),
                              dtype=np.uint64))

        def bits(x):
            return int(np.float64(x).view(np.uint64))

        def m(x):
            return int(_mix64(np.array([x & _M64], dtype=np.uint64))[0])
        for a, c, s, conf, lift in rules:
            v = h(a) ^ m(h(c) + 0x1234) ^ m(bits(conf)) ^ m(bits(lift) + 1) ^ m(bits(s) + 2)
            tot = (tot + m(v)) & _M64
    return tot


def dense_to_csr(onehot: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    rows, cols = np.nonzero(onehot)
    offsets = np.zeros(onehot.shape[0] + 1, dtype=np.int64)
    np.cumsum(np.bincount(rows, minlength=onehot.shape[0]), out=offsets[1:])
    return offsets, cols.astype(np.int32)
