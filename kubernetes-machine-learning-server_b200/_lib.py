"""ctypes binding of ``libkmls_b200.so`` (declared in ``include/kmls_b200.h``).

There is no CPU fallback: if the library is missing, or a call fails, this raises. The
reference job would then fail and be retried by Kubernetes (``restartPolicy: OnFailure``,
reference kubernetes/job.yaml:41), which is the error convention SURVEY.md §8b asks for.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_double, c_int32, c_int64, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# (KMLS_LIB: another build of the same library, for A/B measurements of kernel variants)
LIB_PATH = os.environ.get("KMLS_LIB") or os.path.join(_HERE, "libkmls_b200.so")

_P, _I32, _I64, _SZ, _F64 = c_void_p, c_int32, c_int64, c_size_t, c_double

# name -> (restype, argtypes); mirrors include/kmls_b200.h one to one.
SIGNATURES = {
    "kmls_last_error": (ctypes.c_char_p, []),
    "kmls_version": (ctypes.c_int, []),
    "kmls_item_histogram": (ctypes.c_int, [_P, _I64, _I32, _P, _P]),
    "kmls_bitmap_row_words": (_I64, [_I64]),
    "kmls_encode_bitmap": (ctypes.c_int, [_P, _P, _I64, _P, _I32, _I32, _P, _I64, _P]),
    "kmls_row_popcount": (ctypes.c_int, [_P, _I32, _I64, _P, _P]),
    "kmls_gather_rows": (ctypes.c_int, [_P, _P, _I32, _I64, _P, _P]),
    "kmls_pair_counts": (ctypes.c_int, [_P, _I32, _I64, _P, _P]),
    "kmls_pair_gemm_scratch_bytes": (_SZ, [_I32, _I64]),
    "kmls_pair_counts_i8": (ctypes.c_int, [_P, _I32, _I64, _P, _P, _SZ, _P]),
    "kmls_pair_select": (ctypes.c_int, [_P, _I32, _I32, _I64, _P, _P, _P, _P]),
    "kmls_pair_map_scratch_bytes": (_SZ, [_I64, _I32]),
    "kmls_pair_map": (ctypes.c_int, [_P, _P, _I64, _P, _I32, _I64, _P, _P, _P, _P, _SZ, _P]),
    "kmls_supports_f64": (ctypes.c_int, [_P, _I64, _I64, _P, _P]),
    "kmls_scan_scratch_bytes": (_SZ, [_I64]),
    "kmls_exclusive_scan_i32": (ctypes.c_int, [_P, _I64, _P, _P, _SZ, _P]),
    "kmls_exclusive_scan_i32_noalloc": (ctypes.c_int, [_P, _I64, _P, _P]),
    "kmls_set_capacity": (_I64, [_I64]),
    "kmls_itemset_set_build": (ctypes.c_int, [_P, _I64, _I32, _P, _I64, _P]),
    "kmls_join_count": (ctypes.c_int, [_P, _I64, _I32, _P, _I64, _P, _I32, _I32, _P, _P, _P, _P, _P, _SZ, _P]),
    "kmls_join_write": (ctypes.c_int, [_P, _I64, _I32, _P, _I64, _P, _I32, _I32, _P, _P, _P, _P, _P, _P,
                                       _P, _P, _P, _SZ, _P]),
    "kmls_count_candidates": (ctypes.c_int, [_P, _I64, _P, _I32, _P, _I64, _P, _P, _I64, _P, _P]),
    "kmls_emit_count": (ctypes.c_int, [_P, _I64, _I32, _P, _P, _SZ, _P]),
    "kmls_emit_level": (ctypes.c_int, [_P, _I32, _P, _P, _P, _I64, _I32, _P, _P, _P, _P]),
    "kmls_count_pairs": (ctypes.c_int, [_P, _P, _P, _P, _P, _I64, _P, _P]),
    "kmls_emit_level_px": (ctypes.c_int, [_P, _I32, _P, _P, _P, _P, _I64, _I32, _P, _P, _P, _P, _P, _P, _P]),
    "kmls_row_alloc": (ctypes.c_int, [_P, _I64, _P, _P, _I32, _P, _P]),
    "kmls_project_rows": (ctypes.c_int, [_P, _P, _P, _P, _P, _I64, _P, _P, _P, _I32, _P]),
    "kmls_row_tidlists": (ctypes.c_int, [_P, _I32, _I64, _P, _P, _P]),
    "kmls_project_rows_gather": (ctypes.c_int, [_P, _I64, _P, _P, _I64, _P, _P, _P, _P, _I64, _P]),
    "kmls_trie_index_build": (ctypes.c_int, [_P, _P, _P, _P, _I32, _P, _P]),
    "kmls_rules_parts_per_itemset": (_I64, [_I32]),
    "kmls_rules_count": (ctypes.c_int, [_P, _P, _P, _I32, _I32, _I64, _I64, _P, _I64, _F64, _P, _P,
                                        _P, _P]),
    "kmls_pair_counts_csr": (ctypes.c_int, [_P, _P, _I64, _P, _I32, _P, _P, _P]),
    "kmls_pair_counts_csr_cold": (ctypes.c_int, [_P, _P, _I64, _P, _I32, _I32, _P, _P, _P]),
    "kmls_pair_counts_i8_block": (ctypes.c_int, [_P, _I32, _I64, _I64, _P, _I64, _P, _SZ, _P]),
    "kmls_words_to_host": (ctypes.c_int, [_P, _I32, _P, _P]),
    "kmls_rules_subsets": (ctypes.c_int, [_P, _P, _P, _I32, _I32, _P, _P, _P, _P, _P]),
    "kmls_rules_count_gather": (ctypes.c_int, [_P, _P, _P, _P, _I32, _I32, _P, _P, _I64, _F64, _P, _P, _P, _P]),
    "kmls_rules_write": (ctypes.c_int, [_P, _P, _I32, _I32, _I64, _I64, _I64, _F64, _P, _P, _P, _P, _P, _P, _P]),
    "kmls_itemset_checksum": (ctypes.c_int, [_P, _P, _I64, _I32, _P, _P, _P]),
    "kmls_rules_write_planes": (ctypes.c_int, [_P, _P, _I32, _I32, _I64, _I64, _I64, _F64, _P, _P, _P, _P, _P, _P, _P, _P]),
    "kmls_rule_supports": (ctypes.c_int, [_P, _I64, _P, _I64, _P, _P]),
    "kmls_rule_checksum": (ctypes.c_int, [_P, _P, _P, _I32, _P, _P, _P, _I64, _P, _I64, _P, _P]),
    "kmls_thresholds": (ctypes.c_int, [_F64, _I64, _P, _P]),
    "kmls_mine": (ctypes.c_int, [_P, _P, _I64, _I32, _I64, _I32, _I32, _P, _P, _SZ, _P, _P]),
    "kmls_rules": (ctypes.c_int, [_P, _P, _F64, _I32, _I32, _I32, _I64, _P, _SZ, _P, _P]),
    "kmls_subtree_roots": (ctypes.c_int, [_P, _I64, _P, _P, _P, _P]),
    "kmls_place_level": (ctypes.c_int, [_P, _P, _P, _I64, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "kmls_redg_microbench": (ctypes.c_int, [_P, _I64, _I64, ctypes.c_uint64, _P]),
    "kmls_stream_mix_microbench": (ctypes.c_int, [_P, _P, _P, _I64, _I32, _I32, _P, _P, _P, _P]),
    "kmls_redg_microbench_span": (ctypes.c_int, [_P, _I64, _I64, _I64, ctypes.c_uint64, _P]),
    "kmls_profile_begin": (ctypes.c_int, []),
    "kmls_profile_begin_events": (ctypes.c_int, []),
    "kmls_profile_report": (_I64, [_P, _SZ]),
    "kmls_place_level_compact": (ctypes.c_int, [_P, _P, _P, _I64, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "kmls_rebuild_items": (ctypes.c_int, [_P, _I32, _P, _P, _I64, _P, _P]),
    "kmls_rule_division_selftest": (ctypes.c_int, [_I64, _I64, ctypes.c_uint64, _P, _P]),
}

KMLS_MAX_LEVELS = 32
_LV = KMLS_MAX_LEVELS + 2


class MineOptions(ctypes.Structure):
    """struct kmls_mine_options (include/kmls_b200.h)."""
    _fields_ = [("max_len", c_int32), ("pair_impl", c_int32), ("keep_words", c_int32),
                ("n_parts", c_int32), ("part", c_int32), ("part_block", c_int32),
                ("squeeze_pct", c_int32), ("hot_rows", c_int32),
                ("max_itemsets", c_int64), ("h_tables", c_void_p), ("h_tables_bytes", c_size_t),
                ("allreduce_i32", c_void_p), ("allreduce_ctx", c_void_p)]


# int (*allreduce_i32)(void* ctx, int32_t* d_buf, int64_t n, kmls_stream_t stream)
ALLREDUCE_FN = ctypes.CFUNCTYPE(ctypes.c_int, c_void_p, c_void_p, c_int64, c_void_p)


class MineResultC(ctypes.Structure):
    """struct kmls_mine_result (include/kmls_b200.h)."""
    _fields_ = [("n_playlists_total", c_int64), ("t1", c_int32), ("tk", c_int32), ("thr_k", c_int32),
                ("n_rows", c_int32), ("max_k", c_int32), ("pair_impl_used", c_int32),
                ("n_itemsets", c_int64), ("kernel_launches", c_int64), ("workspace_peak_bytes", c_size_t),
                ("level_size", c_int64 * _LV), ("level_candidates", c_int64 * _LV),
                ("level_row_words", c_int64 * _LV), ("items_off", c_int64 * _LV),
                ("counts_off", c_int64 * _LV), ("parent_off", c_int64 * _LV),
                ("row_items_off", c_int64), ("occurrences_off", c_int64), ("pair_off", c_int64),
                ("h_items_off", c_int64 * _LV), ("h_counts_off", c_int64 * _LV),
                ("h_row_items_off", c_int64), ("h_bytes", c_int64)]


class RulesResultC(ctypes.Structure):
    """struct kmls_rules_result (include/kmls_b200.h)."""
    _fields_ = [("n_rules", c_int64), ("n_itemsets", c_int64), ("n_splits", c_int64),
                ("rule_off", c_int64), ("confidence_off", c_int64), ("lift_off", c_int64),
                ("arena_items_off", c_int64), ("arena_counts_off", c_int64),
                ("workspace_needed_bytes", c_size_t), ("workspace_peak_bytes", c_size_t),
                ("kernel_launches", c_int64), ("pass1_used", c_int32), ("reserved", c_int32)]


_lib = None


class KmlsError(RuntimeError):
    """A call into libkmls_b200.so failed (CUDA error, bad argument, or a buffer that would
    have overflowed — the library never truncates)."""


def load() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise KmlsError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (or `make -C kubernetes-machine-learning-server_b200/csrc`). "
                "There is no CPU fallback for this path.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


# Kernels enqueued by one call of each entry point (for bench.py's `gpu_launches`; CUB scans
# and selections count as the kernels they launch from inside this library).
KERNELS_PER_CALL = {
    "kmls_item_histogram": 1, "kmls_encode_bitmap": 1, "kmls_row_popcount": 1, "kmls_gather_rows": 1,
    "kmls_pair_counts": 1, "kmls_pair_counts_i8": 2, "kmls_pair_counts_i8_block": 2, "kmls_pair_select": 2, "kmls_supports_f64": 1, "kmls_exclusive_scan_i32": 3,
    "kmls_exclusive_scan_i32_noalloc": 3,
    "kmls_itemset_set_build": 1, "kmls_join_count": 4, "kmls_join_write": 4, "kmls_count_candidates": 1,
    "kmls_emit_count": 3, "kmls_emit_level": 1, "kmls_trie_index_build": 1, "kmls_rules_count": 1,
    "kmls_rules_write": 1, "kmls_rule_supports": 1, "kmls_count_pairs": 1, "kmls_emit_level_px": 1, "kmls_row_alloc": 1,
    "kmls_project_rows": 1, "kmls_row_tidlists": 1, "kmls_project_rows_gather": 1,
}

# Optional tracing for bench.py: when TRACE is a list, every call is bracketed by CUDA events
# on the current stream and appended as (name, start_event, end_event, algorithmic_bytes).
TRACE = None
CALLS = {}


def _arg(a):
    # torch tensors are passed as their device (or host) address
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    if isinstance(a, ctypes.Structure):
        return ctypes.addressof(a)
    return a


EXTRA_LAUNCHES = 0   # kernels launched inside the one-call entry points (they report their own count)


def call_rc(name: str, *args):
    """Call an int-returning entry point and hand back (return code, error text) instead of
    raising: for callers that react to a capacity error by retrying with a larger buffer."""
    lib = load()
    if TRACE is not None:
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = getattr(lib, name)(*[_arg(a) for a in args])
        e1.record()
        TRACE.append((name, e0, e1, None))
    else:
        rc = getattr(lib, name)(*[_arg(a) for a in args])
    return rc, (lib.kmls_last_error().decode("utf-8", "replace") if rc != 0 else "")


def call(name: str, *args, algo_bytes=None):
    """Call an int-returning entry point; raise KmlsError with the library's text on failure."""
    lib = load()
    CALLS[name] = CALLS.get(name, 0) + 1
    if TRACE is not None:
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = getattr(lib, name)(*[_arg(a) for a in args])
        e1.record()
        TRACE.append((name, e0, e1, algo_bytes))
    else:
        rc = getattr(lib, name)(*[_arg(a) for a in args])
    if rc != 0:
        msg = lib.kmls_last_error().decode("utf-8", "replace")
        raise KmlsError(f"{name} failed with code {rc}: {msg}")


def kernel_launches() -> int:
    return sum(KERNELS_PER_CALL.get(k, 1) * v for k, v in CALLS.items()) + EXTRA_LAUNCHES


def value(name: str, *args):
    """Call a value-returning helper (sizes)."""
    return getattr(load(), name)(*[_arg(a) for a in args])


def profile_begin(work: bool = True):
    """Per-kernel events inside kmls_mine / kmls_rules from now on; ``work=False``: events only,
    without the data-dependent work figures (which cost extra small kernels and a host
    synchronisation per level) - for timed steps."""
    if work:
        load().kmls_profile_begin()
    else:
        load().kmls_profile_begin_events()


def profile_report():
    """[(kernel, level, ms, algorithmic work)] of every launch kmls_mine / kmls_rules made on
    this thread since profile_begin()."""
    lib = load()
    need = int(lib.kmls_profile_report(None, 0))
    buf = ctypes.create_string_buffer(need + 16)
    lib.kmls_profile_report(buf, need + 16)     # the first call switched recording off; events are kept
    out = []
    for ln in buf.value.decode().splitlines():
        nm, lv, ms, work = ln.rsplit(",", 3)
        out.append((nm, int(lv), float(ms), float(work)))
    return out
