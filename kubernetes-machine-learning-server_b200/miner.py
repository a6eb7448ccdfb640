"""Level-wise frequent-itemset mining on one or several B200s.

Host control loop around the CUDA kernels of ``libkmls_b200.so``. It replaces what the
reference delegates to mlxtend at machine-learning/main.py:267-272
(``TransactionEncoder().fit().transform()`` + ``fpgrowth(df, min_support, use_colnames=True)``)
and produces the same result set, level by level:

    level 1   K_hist -> [allreduce] -> K0 encode -> K1 popcount -> [allreduce]
    level 2   K2 pair counts -> [allreduce] -> pair select
    level k   K3 join+prune -> K4 count -> [allreduce] -> filter/emit   (until empty)

PyTorch supplies device memory, the stream and ``torch.distributed``; every computation on
playlist data is one of this repository's kernels. Multi-GPU: each rank holds a contiguous
range of playlists (bitmap columns); candidate generation is replicated and deterministic,
so the only exchange is one integer sum-allreduce per level.
"""
from __future__ import annotations

import ctypes
import dataclasses
import math
from typing import List, Optional, Tuple

import numpy as np
import torch

from . import _lib


def thresholds(min_support: float, n_transactions: int) -> Tuple[int, int]:
    """The two integer thresholds mlxtend's two float/int tests amount to.

    T1 = min{c : c / float(N) >= min_support}   level 1, fpcommon.setup_fptree
    Tk = math.ceil(min_support * N)             levels >= 2, fpgrowth.fpgrowth
    Computed on the host exactly as Python computes them; the device only ever compares
    integers."""
    if min_support <= 0.0:
        raise ValueError(
            "`min_support` must be a positive number within the interval `(0, 1]`. "
            "Got %s." % min_support)
    n = int(n_transactions)
    tk = int(math.ceil(min_support * n))
    if n == 0:
        return 0, tk
    c = max(0, min(n + 1, tk))
    while c > 0 and (c - 1) / float(n) >= min_support:
        c -= 1
    while c <= n and not (c / float(n) >= min_support):
        c += 1
    return c, tk


@dataclasses.dataclass
class MineResult:
    """Frequent itemsets as device tables. ``levels[k-1] = (items[n_k, k], counts[n_k])``
    where members are ROW indices of the bitmap (ascending support order); ``row_items``
    maps a row to the caller's track id."""
    n_playlists: int
    t1: int
    tk: int
    row_items: torch.Tensor                       # int32[F1]: row -> track id
    levels: List[Tuple[torch.Tensor, torch.Tensor]]
    pair_counts: Optional[torch.Tensor] = None    # int32[F1, F1]
    stats: Optional[dict] = None
    # parents[k-2][i] = index, in level k-1, of level-k itemset i without its last member
    parents: Optional[List[torch.Tensor]] = None
    workspace: Optional[object] = None      # keeps the native path's device memory alive
    occurrences: Optional[torch.Tensor] = None
    host_layout: Optional[dict] = None
    merge_marks: Optional[list] = None      # (phase, CUDA event) pairs of a multi-GPU merge

    @property
    def n_itemsets(self) -> int:
        return sum(int(c.numel()) for _, c in self.levels)

    def to_host(self):
        """(offsets int64[M+1], items int32[sum k] as TRACK ids, counts int64[M]); itemsets
        ordered by size then lexicographically by row, members in row order."""
        offs, items, counts = [np.zeros(1, np.int64)], [], []
        base = 0
        rows = self.row_items.cpu().numpy()
        for k, (it, ct) in enumerate(self.levels, start=1):
            n = int(ct.numel())
            if n == 0:
                continue
            items.append(rows[it.cpu().numpy().reshape(-1)])
            counts.append(ct.cpu().numpy().astype(np.int64))
            offs.append(base + k * np.arange(1, n + 1, dtype=np.int64))
            base += k * n
        if not items:
            return np.zeros(1, np.int64), np.zeros(0, np.int32), np.zeros(0, np.int64)
        return np.concatenate(offs), np.concatenate(items).astype(np.int32), np.concatenate(counts)

    def as_dict(self):
        off, items, counts = self.to_host()
        return {frozenset(items[off[i]:off[i + 1]].tolist()): int(counts[i])
                for i in range(len(counts))}


class _Dev:
    """Small conveniences over torch allocations on the current CUDA device."""

    def __init__(self, device):
        self.device = device
        self.scratch = torch.empty(1 << 20, dtype=torch.uint8, device=device)

    def empty(self, n, dtype):
        return torch.empty(max(int(n), 1), dtype=dtype, device=self.device)[:int(n)]

    _host_words = {}

    def scalar(self, t: torch.Tensor) -> int:
        """One int32 / int64 element of a device tensor, read back through a kernel that writes
        pinned host memory (kmls_words_to_host) instead of ``.item()``: a copy-engine read-back
        queues behind every bulk device-to-host transfer in flight (the level tables a caller
        streams out while mining), which serialises the whole loop behind PCIe."""
        key = (self.device.type, self.device.index)
        h = _Dev._host_words.get(key)
        if h is None:
            h = _Dev._host_words[key] = torch.zeros(8, dtype=torch.int64).pin_memory()
        _lib.call("kmls_words_to_host", t, t.element_size() // 4, h, _stream())
        torch.cuda.current_stream().synchronize()
        return int(h.view(t.dtype)[0])

    def scan_scratch(self, n):
        need = int(_lib.value("kmls_scan_scratch_bytes", int(n)))
        if self.scratch.numel() < need:
            self.scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self.scratch, self.scratch.numel()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _allreduce(t: torch.Tensor, group):
    if group is not None:
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)


def _world(group):
    if group is None:
        return 1
    import torch.distributed as dist
    return dist.get_world_size(group)


@dataclasses.dataclass
class Prepared:
    """Levels 1 and 2 of a playlist set up to the pair counts: everything that does not
    depend on the threshold except through WHICH rows exist. Rows are in ascending-support
    order, so the rows (and the pair sub-matrix) of any higher min_support are a suffix."""
    n_playlists: int
    t1: int
    row_words: int
    n_rows: int
    bitmap: torch.Tensor                 # int32[n_rows * row_words]
    counts1: torch.Tensor                # int32[n_rows] global supports, ascending
    counts1_local: torch.Tensor          # this rank's part
    row_items: torch.Tensor              # int32[n_rows]: row -> track id
    occurrences: torch.Tensor            # int32[n_items]: rows of the input per track (with repeats)
    pair: Optional[torch.Tensor]         # int32[n_rows * n_rows] global pair counts (upper triangle valid)
    pair_local: Optional[torch.Tensor]


def prepare(offsets: torch.Tensor, items: torch.Tensor, n_items: int, min_support: float, *,
            n_playlists_total: Optional[int] = None, group=None, pair_impl: str = "auto",
            with_pairs: bool = True) -> Prepared:
    """K_hist -> [allreduce] -> K0 encode -> K1 popcount -> [allreduce] -> K2 pair counts ->
    [allreduce] for the tracks frequent at ``min_support``."""
    dev = _Dev(offsets.device)
    s = _stream()
    n_local = int(offsets.numel()) - 1
    nnz = int(items.numel())
    if n_playlists_total is None:
        n_playlists_total = n_local
        if group is not None:
            t = torch.tensor([n_local], dtype=torch.int64, device=dev.device)
            _allreduce(t, group)
            n_playlists_total = int(t.item())
    n_total = int(n_playlists_total)
    t1, _ = thresholds(min_support, n_total)
    e32 = dev.empty(0, torch.int32)
    occ = dev.empty(n_items, torch.int32)
    _lib.call("kmls_item_histogram", items, nnz, n_items, occ, s, algo_bytes=4 * nnz)
    _allreduce(occ, group)
    row_words = int(_lib.value("kmls_bitmap_row_words", n_local))
    cand = torch.nonzero(occ >= max(t1, 1)).reshape(-1)              # ascending track id
    if cand.numel() == 0:
        return Prepared(n_total, t1, row_words, 0, e32, e32, e32, e32, occ, None, None)
    order = torch.argsort(occ[cand], stable=True)                    # by (occ, id)
    cand = cand[order]
    n_rows = int(cand.numel())
    row_of_item = torch.full((n_items,), -1, dtype=torch.int32, device=dev.device)
    row_of_item[cand] = torch.arange(n_rows, dtype=torch.int32, device=dev.device)
    bitmap = dev.empty(n_rows * row_words, torch.int32)
    _lib.call("kmls_encode_bitmap", offsets, items, n_local, row_of_item, n_items, n_rows,
              bitmap, row_words, s, algo_bytes=4 * nnz + 8 * n_local + 4 * n_rows * row_words)
    counts1 = dev.empty(n_rows, torch.int32)
    _lib.call("kmls_row_popcount", bitmap, n_rows, row_words, counts1, s,
              algo_bytes=4 * n_rows * row_words)
    counts1_local = counts1.clone() if group is not None else counts1
    _allreduce(counts1, group)
    if not bool(torch.equal(counts1, occ[cand])):
        # some playlist repeated a track: exact supports differ from the occurrence counts,
        # so drop rows that fell below T1 and restore ascending-support order
        keep = torch.nonzero(counts1 >= max(t1, 1)).reshape(-1)
        order = torch.argsort(counts1[keep], stable=True)
        src = keep[order].to(torch.int32)
        n_rows = int(src.numel())
        if n_rows == 0:
            return Prepared(n_total, t1, row_words, 0, e32, e32, e32, e32, occ, None, None)
        bitmap2 = dev.empty(n_rows * row_words, torch.int32)
        _lib.call("kmls_gather_rows", bitmap, src, n_rows, row_words, bitmap2, s)
        same = counts1_local is counts1
        bitmap, counts1, cand = bitmap2, counts1[src.long()].contiguous(), cand[src.long()]
        counts1_local = counts1 if same else counts1_local[src.long()].contiguous()
        row_of_item.fill_(-1)
        row_of_item[cand] = torch.arange(n_rows, dtype=torch.int32, device=dev.device)
    row_items = cand.to(torch.int32)
    pair = pair_local = None
    if with_pairs and n_rows >= 2:
        pair = dev.empty(n_rows * n_rows, torch.int32)
        if pair_impl == "auto":
            # tensor cores once there are enough 128x256 tiles and playlists to feed them ...
            pair_impl = "i8" if (n_rows >= 256 and row_words >= 512) else "popc"
            if pair_impl == "i8" and n_rows <= 16384 and n_local > 0:
                # ... unless the one-hot matrix is so sparse that counting the pairs of each
                # playlist is cheaper: ~0.65 m^2 counter updates per playlist (m = mean number
                # of kept tracks; the factor covers the spread of playlist lengths) at about
                # 2e11 updates/s, against F(F+1) N int8 operations at about 2.2e15/s
                m_mean = dev.scalar(counts1_local.sum()) / n_local
                if 0.65 * m_mean * m_mean * n_local / 2.0e11 < n_rows * (n_rows + 1) * n_local / 2.2e15:
                    pair_impl = "csr"
        if pair_impl == "csr":
            overflow = torch.zeros(1, dtype=torch.int32, device=dev.device)
            _lib.call("kmls_pair_counts_csr", offsets, items, n_local, row_of_item, n_rows, pair,
                      overflow, s)
            if dev.scalar(overflow) != 0:
                pair_impl = "i8" if (n_rows >= 256 and row_words >= 512) else "popc"   # a very long playlist
        if pair_impl == "csr":
            pass
        elif pair_impl == "popc":
            _lib.call("kmls_pair_counts", bitmap, n_rows, row_words, pair, s)
        elif pair_impl == "i8":
            gbytes = int(_lib.value("kmls_pair_gemm_scratch_bytes", n_rows, row_words))
            gscratch = dev.empty((gbytes + 7) // 8, torch.int64)
            _lib.call("kmls_pair_counts_i8", bitmap, n_rows, row_words, pair, gscratch, gbytes, s,
                      algo_bytes=n_rows * (n_rows + 1) * n_local)   # = int8 ops: 2 * N * F(F+1)/2
            del gscratch
        else:
            raise ValueError(f"unknown pair_impl {pair_impl!r}")
        pair_local = pair.clone() if group is not None else pair
        _allreduce(pair, group)
    return Prepared(n_total, t1, row_words, n_rows, bitmap, counts1, counts1_local, row_items, occ,
                    pair, pair_local)


class Workspace:
    """Device memory handed to the one-call entry points (kmls_mine, kmls_rules): one torch
    uint8 tensor that grows on demand. The tables a call returns are VIEWS of this memory:
    they stay valid until the same Workspace is passed to another call."""

    def __init__(self, device, nbytes: int = 0):
        self.device = device
        self.buf = None
        if nbytes:
            self.ensure(nbytes)

    @property
    def nbytes(self) -> int:
        return 0 if self.buf is None else int(self.buf.numel())

    def ensure(self, nbytes: int):
        if self.nbytes < nbytes:
            self.buf = None                       # release first: the two need not coexist
            self.buf = torch.empty(int(nbytes), dtype=torch.uint8, device=self.device)
        return self.buf

    def grow(self, to: Optional[int] = None):
        """Double, or grow to ``to`` bytes (bounded by what the device can still give); False
        when that is no growth."""
        free = torch.cuda.mem_get_info(self.device)[0] + torch.cuda.memory_reserved(self.device) \
            - torch.cuda.memory_allocated(self.device)
        have = self.nbytes
        want = max(2 * have, 1 << 30) if to is None else int(to)
        want = min(want, int(0.95 * (free + have)))
        if want <= have:
            return False
        self.ensure(want)
        return True

    def view(self, byte_off: int, n: int, dtype) -> torch.Tensor:
        size = torch.empty(0, dtype=dtype).element_size()
        return self.buf[byte_off:byte_off + n * size].view(dtype)


_PAIR_IMPL = {"auto": 0, "csr": 1, "popc": 2, "i8": 3}
_PAIR_IMPL_USED = {1: "csr", 2: "popc", 3: "i8", 4: "csr+i8"}   # 4: csr with the hot rows' block on the tensor cores
_SIZE_HINT = {}   # (device index, "mine" | "rules") -> bytes the last call needed: first guess of the next


def mine_native(offsets: torch.Tensor, items: torch.Tensor, n_items: int, min_support: float, *,
                n_playlists_total: Optional[int] = None, max_len: Optional[int] = None,
                max_itemsets: int = 200_000_000, keep_pair_counts: bool = False, pair_impl: str = "auto",
                keep_words: Optional[int] = None, workspace: Optional[Workspace] = None,
                part: Tuple[int, int] = (0, 1), part_block: int = 64,
                host_tables: Optional[torch.Tensor] = None, allreduce=None, retry_sync=None,
                hot_rows: Optional[int] = None) -> MineResult:
    """All levels through ONE call into the library (``kmls_mine``): the level loop, its
    sizing and its buffers live in C++; this function only marshals. ``workspace``: device
    memory for the call (a fresh one by default; the returned tables are views of it, so a
    workspace that is reused invalidates earlier results). A workspace that turns out too small
    is grown and the call repeated - the library never truncates. ``part = (i, n)``: mine only
    the subtrees of the i-th of n interleaved blocks of ``part_block`` frequent pairs (levels 1
    and 2 are always complete): the class partition of a multi-GPU job. ``host_tables``: a
    pinned uint8 tensor the level tables are also streamed into while mining continues.
    ``hot_rows``: with the per-playlist pair counts, the block of this many most frequent tracks
    is counted on the tensor cores instead (None/0 = the library decides, -1 = never)."""
    if min_support <= 0.0:
        thresholds(min_support, 1)  # raises the reference's ValueError
    if not (offsets.is_cuda and items.is_cuda):
        raise _lib.KmlsError("mine_native needs CUDA tensors: this path has no CPU fallback")
    assert offsets.dtype == torch.int64 and items.dtype == torch.int32
    device = offsets.device
    n_local = int(offsets.numel()) - 1
    n_total = n_local if n_playlists_total is None else int(n_playlists_total)
    t1, tk = thresholds(min_support, n_total)
    opt = _lib.MineOptions()
    opt.max_len = int(max_len or 0)
    opt.pair_impl = _PAIR_IMPL[pair_impl]
    if keep_words is None:
        import os
        keep_words = int(os.environ.get("KMLS_KEEP_WORDS", 0)) or None    # tuning switch
    opt.keep_words = int(keep_words or 0)
    import os as _os
    opt.squeeze_pct = int(_os.environ.get("KMLS_SQUEEZE_PCT", 0))               # tuning switch (0 = library default)
    opt.hot_rows = int(hot_rows if hot_rows is not None else _os.environ.get("KMLS_PAIR_HOT_ROWS", 0))
    opt.n_parts, opt.part, opt.part_block = int(part[1]), int(part[0]), int(part_block)
    opt.max_itemsets = int(max_itemsets)
    if host_tables is not None:
        opt.h_tables, opt.h_tables_bytes = host_tables.data_ptr(), int(host_tables.numel())
    ws = workspace if workspace is not None else Workspace(device)
    cb = None
    if allreduce is not None and part[1] > 1:
        # the library hands back an address inside the workspace: wrap it as a tensor view and
        # let the caller's collective (torch.distributed over NCCL) sum it in place
        def _ar(_ctx, ptr, n, _stream):
            try:
                off = int(ptr) - ws.buf.data_ptr()
                allreduce(ws.buf[off:off + 4 * int(n)].view(torch.int32))
                return 0
            except Exception:      # noqa: BLE001 - reported by the library as a failed callback
                return 1
        cb = _lib.ALLREDUCE_FN(_ar)
        opt.allreduce_i32 = ctypes.cast(cb, ctypes.c_void_p).value
    hint_key = (device.index, "mine")
    if ws.nbytes == 0:
        ws.ensure(max(1 << 28, 48 * int(items.numel()), _SIZE_HINT.get(hint_key, 0)))
    out = _lib.MineResultC()
    while True:
        rc, msg = _lib.call_rc("kmls_mine", offsets, items, n_local, int(n_items), n_total, t1, tk, opt,
                               ws.buf, ws.nbytes, out, _stream())
        _lib.CALLS["kmls_mine"] = _lib.CALLS.get("kmls_mine", 0) + 1
        too_small = rc == -2 and "workspace too small" in msg
        if rc != 0 and not too_small:
            raise _lib.KmlsError(f"kmls_mine failed with code {rc}: {msg}")
        # ranks that share collectives inside the call must repeat it TOGETHER or not at all
        again = retry_sync(too_small) if retry_sync is not None else too_small
        if not again:
            break
        if too_small and not ws.grow():
            raise _lib.KmlsError(f"kmls_mine failed with code {rc}: {msg} (and the device has no more memory to give)")
    _lib.EXTRA_LAUNCHES += int(out.kernel_launches) - 1
    _SIZE_HINT[hint_key] = max(_SIZE_HINT.get(hint_key, 0), int(out.workspace_peak_bytes * 1.1) + (1 << 20))
    stats = {"levels": [], "t1": t1, "tk": tk, "thr_k": int(out.thr_k), "n_playlists": n_total,
             "nnz_local": int(items.numel()), "workspace_peak_bytes": int(out.workspace_peak_bytes),
             "pair_impl": _PAIR_IMPL_USED.get(int(out.pair_impl_used)),
             "h_bytes": int(out.h_bytes)}
    e32 = torch.empty(0, dtype=torch.int32, device=device)
    n_rows, max_k = int(out.n_rows), int(out.max_k)
    levels, parents = [], []
    for k in range(1, max_k + 1):
        n = int(out.level_size[k])
        levels.append((ws.view(int(out.items_off[k]), n * k, torch.int32).view(n, k),
                       ws.view(int(out.counts_off[k]), n, torch.int32)))
        if k >= 2:
            parents.append(ws.view(int(out.parent_off[k]), n, torch.int32))
        stats["levels"].append({"k": k, "candidates": int(out.level_candidates[k]), "frequent": n,
                                "row_words_total": int(out.level_row_words[k])})
    if max_k + 1 < _lib.KMLS_MAX_LEVELS and int(out.level_candidates[max_k + 1]):
        stats["levels"].append({"k": max_k + 1, "candidates": int(out.level_candidates[max_k + 1]), "frequent": 0})
    row_items = ws.view(int(out.row_items_off), n_rows, torch.int32) if n_rows else e32
    pair = None
    if keep_pair_counts and int(out.pair_off) >= 0:
        pair = ws.view(int(out.pair_off), n_rows * n_rows, torch.int32).view(n_rows, n_rows)
    res = MineResult(n_total, t1, tk, row_items, levels, pair, stats, parents)
    res.workspace = ws
    res.occurrences = ws.view(int(out.occurrences_off), int(n_items), torch.int32) \
        if int(out.occurrences_off) >= 0 else None
    res.host_layout = None
    if host_tables is not None and n_rows:
        res.host_layout = {"row_items": (int(out.h_row_items_off), n_rows),
                           "levels": [(int(out.h_items_off[k]), int(out.h_counts_off[k]), int(out.level_size[k]))
                                      for k in range(1, max_k + 1)]}
    return res


def mine_csr(offsets: torch.Tensor, items: torch.Tensor, n_items: int, min_support: float, *,
             n_playlists_total: Optional[int] = None, group=None, max_len: Optional[int] = None,
             max_itemsets: int = 200_000_000, keep_pair_counts: bool = False,
             pair_impl: str = "auto", count_impl: str = "projected",
             prune: Optional[bool] = None, first_projection: str = "gather",
             keep_words: Optional[int] = None, prepared: Optional["Prepared"] = None,
             on_level=None, impl: str = "auto", workspace: Optional["Workspace"] = None) -> MineResult:
    """Mine the playlists ``offsets[int64, n+1] / items[int32, nnz]`` (CUDA tensors, CSR by
    playlist, this rank's shard) for itemsets with the reference's thresholds.

    ``n_playlists_total`` is N of the whole job (sum over ranks); it defaults to the local
    count (allreduced when ``group`` is given). ``count_impl``: "projected" counts candidates
    on conditional (column-compacted) bitmaps, "fullwidth" on the full bitmap with the prefix
    staged in shared memory. ``prune`` switches Apriori pruning of the joins (default: on for
    "fullwidth", where a candidate costs N/8 bytes, off for "projected", where probing the
    fingerprint set costs more than counting the few extra candidates - with members in
    ascending-support order the subsets a probe checks drop a RARER member and are almost
    always frequent). ``first_projection``: "gather" or "pext" for the level 1 -> 2 rows."""
    native_ok = (group is None and prepared is None and on_level is None and count_impl == "projected"
                 and not prune and first_projection == "gather"
                 and (keep_words is None or (0 < keep_words <= 1024 and keep_words % 4 == 0)))
    if impl == "native" and not native_ok:
        raise ValueError("impl='native' does not take group / prepared / on_level / fullwidth / prune")
    if impl in ("auto", "native") and native_ok:
        # the default: the whole level loop inside the library (kmls_mine)
        import os
        kw = keep_words if keep_words is not None else int(os.environ.get("KMLS_KEEP_WORDS", 0)) or None
        return mine_native(offsets, items, n_items, min_support, n_playlists_total=n_playlists_total,
                           max_len=max_len, max_itemsets=max_itemsets, keep_pair_counts=keep_pair_counts,
                           pair_impl=pair_impl, keep_words=kw, workspace=workspace)
    if prune is None:
        prune = count_impl != "projected"
    if keep_words is None:
        # rows at most this wide stay in their parent's column space (AND, no bit extract)
        import os
        keep_words = int(os.environ.get("KMLS_KEEP_WORDS", 256))
    if min_support <= 0.0:
        thresholds(min_support, 1)  # raises the reference's ValueError
    if not (offsets.is_cuda and items.is_cuda):
        raise _lib.KmlsError("mine_csr needs CUDA tensors: this path has no CPU fallback")
    assert offsets.dtype == torch.int64 and items.dtype == torch.int32
    dev = _Dev(offsets.device)
    s = _stream()
    n_local = int(offsets.numel()) - 1
    nnz = int(items.numel())
    if n_playlists_total is None:
        n_playlists_total = n_local
        if group is not None:
            t = torch.tensor([n_local], dtype=torch.int64, device=dev.device)
            _allreduce(t, group)
            n_playlists_total = int(t.item())
    n_total = int(n_playlists_total)
    t1, tk = thresholds(min_support, n_total)
    stats = {"levels": [], "t1": t1, "tk": tk, "n_playlists": n_total, "nnz_local": nnz}
    empty_i32 = dev.empty(0, torch.int32)

    parents: List[torch.Tensor] = []

    def result(row_items, levels, pair=None):
        return MineResult(n_total, t1, tk, row_items, levels, pair, stats, parents)

    if n_total == 0 or n_items == 0:
        return result(empty_i32, [])

    # ---- levels 1 and 2 up to the pair counts (reusable across thresholds, see prepare()) ---
    if prepared is None:
        prepared = prepare(offsets, items, n_items, min_support, n_playlists_total=n_total,
                           group=group, pair_impl=pair_impl, with_pairs=max_len is None or max_len > 1)
    elif prepared.t1 > t1 or prepared.n_playlists != n_total:
        raise ValueError("`prepared` was built for a higher min_support or another playlist set")
    stats["row_words"] = row_words = prepared.row_words
    # rows are in ascending-support order: the rows frequent at this threshold are a suffix
    r0 = int(torch.searchsorted(prepared.counts1, torch.tensor([max(t1, 1)], dtype=torch.int32,
                                                               device=dev.device)).item())
    n_rows = prepared.n_rows - r0
    if n_rows <= 0:
        return result(empty_i32, [])
    bitmap = prepared.bitmap[r0 * row_words:]
    counts1 = prepared.counts1[r0:]
    counts1_local = prepared.counts1_local[r0:]
    row_items = prepared.row_items[r0:]
    lvl1_items = torch.arange(n_rows, dtype=torch.int32, device=dev.device).reshape(-1, 1)
    levels = [(lvl1_items, counts1)]
    if on_level is not None:
        on_level(1, lvl1_items, counts1)
    stats["levels"].append({"k": 1, "candidates": n_items, "frequent": n_rows})
    if n_rows < 2 or (max_len is not None and max_len <= 1):
        return result(row_items, levels)
    if prepared.pair is None:
        raise ValueError("`prepared` holds no pair counts (built with with_pairs=False)")
    if r0 == 0:
        pair, pair_local = prepared.pair, prepared.pair_local
    else:
        full = prepared.n_rows
        pair = prepared.pair.view(full, full)[r0:, r0:].contiguous().view(-1)
        pair_local = pair if prepared.pair_local is prepared.pair else \
            prepared.pair_local.view(full, full)[r0:, r0:].contiguous().view(-1)
    thr = tk
    if t1 < tk:
        # mlxtend's root single-path shortcut applies no Tk test:
        # the root FP-tree is one path iff the supports are nested, i.e. every pair count
        # equals the rarer member's support. Rows are in ascending-support order.
        p2 = pair.view(n_rows, n_rows)
        upper = torch.triu(torch.ones_like(p2, dtype=torch.bool), 1)
        if bool(((p2 == counts1.view(-1, 1)) | ~upper).all()):
            thr = t1
    stats["thr_k"] = thr
    row_off = dev.empty(n_rows + 1, torch.int64)
    _lib.call("kmls_pair_select", pair, n_rows, thr, 0, row_off, 0, 0, s)
    n2 = dev.scalar(row_off[n_rows])
    stats["levels"].append({"k": 2, "candidates": n_rows * (n_rows - 1) // 2, "frequent": n2})
    if n2 == 0:
        return result(row_items, levels, pair.view(n_rows, n_rows) if keep_pair_counts else None)
    items_k = dev.empty(n2 * 2, torch.int32)
    counts_k = dev.empty(n2, torch.int32)
    _lib.call("kmls_pair_select", pair, n_rows, thr, n2, row_off, items_k, counts_k, s)
    levels.append((items_k.view(n2, 2), counts_k))
    if on_level is not None:
        on_level(2, items_k, counts_k)
    parents.append(items_k.view(n2, 2)[:, 0].contiguous())
    total = n_rows + n2
    projected = count_impl == "projected"
    more = (max_len is None or 2 < max_len) and n2 >= 2

    def project(arena_old, rowptr_old, roww_old, par_p, par_q, parent_cnt_local, n_new,
                first=False):
        """Rows of a new level from the rows of its parents' level (``first``: from the full
        bitmap, by gathering at the parents' column lists)."""
        roww = dev.empty(n_new, torch.int32)
        _lib.call("kmls_row_alloc", par_p, n_new, parent_cnt_local, 0 if first else roww_old,
                  keep_words, roww, s)
        scratch, sbytes = dev.scan_scratch(n_new)
        rowptr = dev.empty(n_new + 1, torch.int64)
        _lib.call("kmls_exclusive_scan_i32", roww, n_new, rowptr, scratch, sbytes, s)
        words = dev.scalar(rowptr[n_new])
        arena = dev.empty(words, torch.int32)
        if first and first_projection == "gather":
            scratch, sbytes = dev.scan_scratch(n_rows)
            tid_off = dev.empty(n_rows + 1, torch.int64)
            _lib.call("kmls_exclusive_scan_i32", parent_cnt_local, n_rows, tid_off, scratch, sbytes, s)
            tids = dev.empty(dev.scalar(tid_off[n_rows]), torch.int32)
            _lib.call("kmls_row_tidlists", bitmap, n_rows, row_words, tid_off, tids, s,
                      algo_bytes=4 * n_rows * row_words + 4 * int(tids.numel()))
            _lib.call("kmls_project_rows_gather", bitmap, row_words, par_p, par_q, n_new, tids,
                      tid_off, arena, rowptr, words, s, algo_bytes=4 * words)
        else:
            _lib.call("kmls_project_rows", arena_old, rowptr_old, roww_old, par_p, par_q, n_new,
                      arena, rowptr, roww, 0 if first else keep_words, s, algo_bytes=8 * words)
        return arena, rowptr, roww, words

    if projected and more:
        rowptr1 = torch.arange(n_rows, dtype=torch.int64, device=dev.device) * row_words
        roww1 = torch.full((n_rows,), row_words, dtype=torch.int32, device=dev.device)
        it2 = items_k.view(n2, 2)
        par_p, par_q = it2[:, 0].contiguous(), it2[:, 1].contiguous()
        if group is not None:
            cnt_loc_k = pair_local.view(n_rows, n_rows)[par_p.long(), par_q.long()].contiguous()
        else:
            cnt_loc_k = counts_k
        arena_k, rowptr_k, roww_k, words = project(bitmap, rowptr1, roww1, par_p, par_q,
                                                   counts1_local, n2, first=True)
        stats["levels"][-1]["row_words_total"] = words

    # ---- levels k >= 3 -------------------------------------------------------------------
    k = 2
    while (max_len is None or k < max_len) and int(counts_k.numel()) >= 2:
        n = int(counts_k.numel())
        scratch, sbytes = dev.scan_scratch(n)
        if k >= 3 and prune:
            cap = int(_lib.value("kmls_set_capacity", n))
            slots = dev.empty(cap, torch.int64)
            _lib.call("kmls_itemset_set_build", items_k, n, k, slots, cap, s)
        else:
            cap, slots = 0, 0
        kept_cnt = dev.empty(n, torch.int32)
        kept_mask = dev.empty(n, torch.int32)
        kept_off = dev.empty(n + 1, torch.int64)
        pair_arg = pair if (k == 2 and prune) else 0
        par_k = parents[-1]
        _lib.call("kmls_join_count", items_k, n, k, slots, cap, pair_arg, n_rows, thr, par_k, kept_cnt,
                  kept_mask, kept_off, scratch, sbytes, s)
        n_cand = dev.scalar(kept_off[n])
        if n_cand == 0:
            break
        if total + n_cand > max_itemsets * 4:
            raise _lib.KmlsError(f"level {k + 1}: {n_cand} candidates exceed the max_itemsets guard")
        cand_p = dev.empty(n_cand, torch.int32)
        cand_q = dev.empty(n_cand, torch.int32)
        cand_cnt = dev.empty(n_cand, torch.int32)
        if projected:
            _lib.call("kmls_join_write", items_k, n, k, slots, cap, pair_arg, n_rows, thr, par_k,
                      kept_cnt, kept_mask, kept_off, cand_p, cand_q, 0, 0, scratch, sbytes, s)
            n_groups = None
            _lib.call("kmls_count_pairs", arena_k, rowptr_k, roww_k, cand_p, cand_q, n_cand,
                      cand_cnt, s)
            cand_cnt_local = cand_cnt.clone() if group is not None else cand_cnt
        else:
            grp_p = dev.empty(n, torch.int32)
            n_groups_d = dev.empty(1, torch.int64)
            _lib.call("kmls_join_write", items_k, n, k, slots, cap, pair_arg, n_rows, thr, par_k,
                      kept_cnt, kept_mask, kept_off, cand_p, cand_q, grp_p, n_groups_d, scratch, sbytes, s)
            n_groups = dev.scalar(n_groups_d)
            _lib.call("kmls_count_candidates", bitmap, row_words, items_k, k, grp_p, n_groups,
                      kept_off, cand_q, n_cand, cand_cnt, s,
                      algo_bytes=(n_cand + n_groups * k) * row_words * 4)
        _allreduce(cand_cnt, group)
        scratch, sbytes = dev.scan_scratch(n_cand)
        out_off = dev.empty(n_cand + 1, torch.int64)
        _lib.call("kmls_emit_count", cand_cnt, n_cand, thr, out_off, scratch, sbytes, s)
        n_next = dev.scalar(out_off[n_cand])
        stats["levels"].append({"k": k + 1, "candidates": n_cand, "groups": n_groups,
                                "frequent": n_next})
        if n_next == 0:
            break
        total += n_next
        if total > max_itemsets:
            raise _lib.KmlsError(
                f"more than max_itemsets={max_itemsets} frequent itemsets at level {k + 1}; "
                "raise the guard or min_support (results are never truncated)")
        items_n = dev.empty(n_next * (k + 1), torch.int32)
        counts_n = dev.empty(n_next, torch.int32)
        cnt_loc_n = dev.empty(n_next, torch.int32)
        par_p = dev.empty(n_next, torch.int32)
        par_q = dev.empty(n_next, torch.int32)
        _lib.call("kmls_emit_level_px", items_k, k, cand_p, cand_q, cand_cnt,
                  cand_cnt_local if projected else cand_cnt, n_cand, thr, out_off, items_n,
                  counts_n, cnt_loc_n, par_p, par_q, s)
        parents.append(par_p)
        if projected:
            if (max_len is None or k + 1 < max_len) and n_next >= 2:
                arena_k, rowptr_k, roww_k, words = project(arena_k, rowptr_k, roww_k, par_p, par_q,
                                                           cnt_loc_k, n_next)
                stats["levels"][-1]["row_words_total"] = words
            cnt_loc_k = cnt_loc_n
        k += 1
        items_k, counts_k = items_n, counts_n
        levels.append((items_k.view(n_next, k), counts_k))
        if on_level is not None:
            on_level(k, items_k, counts_k)
    return result(row_items, levels, pair.view(n_rows, n_rows) if keep_pair_counts else None)


@dataclasses.dataclass
class RuleTable:
    """Rules as device columns: rule r is about itemset ``itemset[r]`` (global id: level
    tables concatenated in level order); bit t of ``antecedent[r]`` selects its t-th member
    (row order) into the antecedent, the rest is the consequent. ``support`` is not stored by
    the generating pass (it is the itemset's own support): it is gathered on first use."""
    rule: torch.Tensor         # int32[R, 2]: (itemset global id, antecedent bit mask)
    confidence: torch.Tensor   # float64[R]
    lift: torch.Tensor         # float64[R]
    arena_items: torch.Tensor  # int32: all level tables' members (rows), level after level
    level_sizes: List[int]     # index k -> number of k-itemsets (index 0 unused)
    arena_counts: Optional[torch.Tensor] = None
    n_playlists: int = 0
    _support: Optional[torch.Tensor] = None
    workspace: Optional[object] = None     # keeps the native path's device memory alive
    stats: Optional[dict] = None

    @property
    def n_rules(self) -> int:
        return int(self.confidence.numel())

    @property
    def itemset(self) -> torch.Tensor:
        return self.rule[:, 0]

    @property
    def antecedent(self) -> torch.Tensor:
        return self.rule[:, 1]

    @property
    def support(self) -> torch.Tensor:
        if self._support is None:
            out = torch.empty(max(self.n_rules, 1), dtype=torch.float64, device=self.rule.device)[:self.n_rules]
            if self.n_rules:
                _lib.call("kmls_rule_supports", self.rule, self.n_rules, self.arena_counts,
                          self.n_playlists, out, _stream())
            self._support = out
        return self._support


def _tables_as_c(res: MineResult) -> "_lib.MineResultC":
    """A kmls_mine_result naming the tables of ``res`` by ABSOLUTE device address (to be used
    with a NULL base): works for tables from any source (kmls_mine, the Python loop, a merged
    multi-GPU result)."""
    c = _lib.MineResultC()
    c.n_playlists_total = int(res.n_playlists)
    c.t1, c.tk = int(res.t1), int(res.tk)
    c.max_k = len(res.levels)
    c.n_rows = int(res.row_items.numel())
    for k, (it, ct) in enumerate(res.levels, start=1):
        assert it.is_contiguous() and ct.is_contiguous() and it.dtype == torch.int32 and ct.dtype == torch.int32
        c.level_size[k] = int(ct.numel())
        c.items_off[k] = it.data_ptr() if ct.numel() else -1
        c.counts_off[k] = ct.data_ptr() if ct.numel() else -1
        if k >= 2:
            par = res.parents[k - 2]
            assert par.is_contiguous() and par.dtype == torch.int32
            c.parent_off[k] = par.data_ptr() if ct.numel() else -1
    return c


_PASS1 = {"auto": 0, "gather": 1, "search": 2}


def generate_rules(res: MineResult, min_confidence: float, max_rules: int = 4_000_000_000,
                   part: Tuple[int, int] = (0, 1), pass1: str = "auto",
                   workspace: Optional[Workspace] = None, impl: str = "native") -> RuleTable:
    """fp64 confidence/lift for every (itemset, antecedent) split, kept iff
    confidence >= min_confidence (mlxtend ``association_rules`` semantics), through ONE call
    into the library (``kmls_rules``; ``impl="python"`` = the kernel-by-kernel loop below it).
    ``part = (i, n)`` generates only the rules of the i-th of n contiguous slices of the
    itemset table (how ranks of a multi-GPU job divide the rule table among themselves).
    ``pass1``: how the subset counts of a split are found - "gather" (from the previous level's
    pass-1 output; needs the whole table, i.e. ``part == (0, 1)``), "search" (in the subset
    index; works on any slice), "auto" = gather when possible. ``workspace``: device memory for
    the call (the returned columns are views of it); grown and the call repeated when too small."""
    if impl != "native":
        return generate_rules_python(res, min_confidence, max_rules, part, pass1)
    if pass1 not in _PASS1:
        raise ValueError(f"unknown pass1 {pass1!r}")
    device = res.row_items.device
    sizes = [0] + [int(ct.numel()) for _, ct in res.levels]
    e_i32 = torch.empty(0, dtype=torch.int32, device=device)
    e_f64 = torch.empty(0, dtype=torch.float64, device=device)
    if len(res.levels) < 2 or sum(sizes[2:]) == 0:
        arena = torch.cat([it.reshape(-1) for it, _ in res.levels]) if res.levels else e_i32
        return RuleTable(e_i32.view(0, 2), e_f64, e_f64, arena, sizes + [0], None, res.n_playlists)
    if len(res.levels) > 15:
        raise _lib.KmlsError(f"rule generation supports itemsets of at most 15 members, got {len(res.levels)}")
    mined = _tables_as_c(res)
    ws = workspace if workspace is not None else Workspace(device)
    hint_key = (device.index, "rules")
    if ws.nbytes == 0:
        n_all = sum(sizes)
        splits = sum(sizes[k] << (k - 1) for k in range(2, len(sizes))) // (max(part[1], 1) if part[1] > 2 else 1)
        guess = 4 * sum(sizes[k] * k for k in range(1, len(sizes))) + 28 * n_all + 32 * splits + (64 << 20)
        free = torch.cuda.mem_get_info(device)[0] + torch.cuda.memory_reserved(device) - torch.cuda.memory_allocated(device)
        ws.ensure(min(max(guess, _SIZE_HINT.get(hint_key, 0)), int(0.9 * free)))
    out = _lib.RulesResultC()
    while True:
        rc, msg = _lib.call_rc("kmls_rules", mined, 0, float(min_confidence), int(part[0]), int(part[1]),
                               _PASS1[pass1], int(max_rules), ws.buf, ws.nbytes, out, _stream())
        _lib.CALLS["kmls_rules"] = _lib.CALLS.get("kmls_rules", 0) + 1
        if rc == 0:
            break
        if rc == -2 and "workspace too small" in msg:
            want = int(out.workspace_needed_bytes * 1.02) + (16 << 20)
            # the library's figure is a lower bound while it is still sizing: never ask for less than 1.25x
            if ws.grow(max(want, int(ws.nbytes * 1.25))):
                continue
        raise _lib.KmlsError(f"kmls_rules failed with code {rc}: {msg}")
    _lib.EXTRA_LAUNCHES += int(out.kernel_launches) - 1
    _SIZE_HINT[hint_key] = max(_SIZE_HINT.get(hint_key, 0), int(out.workspace_peak_bytes * 1.02) + (16 << 20))
    n_rules = int(out.n_rules)
    if n_rules == 0 and int(out.rule_off) < 0:
        arena = torch.cat([it.reshape(-1) for it, _ in res.levels])
        return RuleTable(e_i32.view(0, 2), e_f64, e_f64, arena, sizes + [0], torch.cat([ct for _, ct in res.levels]),
                         res.n_playlists)
    n_members = sum(sizes[k] * k for k in range(1, len(sizes)))
    table = RuleTable(ws.view(int(out.rule_off), 2 * n_rules, torch.int32).view(n_rules, 2),
                      ws.view(int(out.confidence_off), n_rules, torch.float64),
                      ws.view(int(out.lift_off), n_rules, torch.float64),
                      ws.view(int(out.arena_items_off), n_members, torch.int32), sizes + [0],
                      ws.view(int(out.arena_counts_off), sum(sizes), torch.int32), res.n_playlists)
    table.workspace = ws
    table.stats = {"n_splits": int(out.n_splits), "pass1": {1: "gather", 2: "search"}.get(int(out.pass1_used)),
                   "workspace_peak_bytes": int(out.workspace_peak_bytes)}
    return table


def generate_rules_python(res: MineResult, min_confidence: float, max_rules: int = 4_000_000_000,
                          part: Tuple[int, int] = (0, 1), pass1: str = "auto") -> RuleTable:
    """fp64 confidence/lift for every (itemset, antecedent) split, kept iff
    confidence >= min_confidence (mlxtend ``association_rules`` semantics).
    ``part = (i, n)`` generates only the rules of the i-th of n contiguous slices of the
    itemset table (how ranks of a multi-GPU job divide the rule table among themselves).
    ``pass1``: how the subset counts of a split are found - "gather" (from the previous level's
    pass-1 output; needs the whole table, i.e. ``part == (0, 1)``), "search" (in the subset
    index; works on any slice), "auto" = gather when possible."""
    device = res.row_items.device
    dev = _Dev(device)
    s = _stream()
    max_k = len(res.levels)
    sizes = np.zeros(max_k + 2, dtype=np.int64)
    for k, (_, ct) in enumerate(res.levels, start=1):
        sizes[k] = int(ct.numel())
    n_all = int(sizes.sum())
    e_i32 = dev.empty(0, torch.int32)
    e_f64 = dev.empty(0, torch.float64)
    if max_k < 2 or n_all == 0 or int(sizes[2:].sum()) == 0:
        arena = torch.cat([it.reshape(-1) for it, _ in res.levels]) if res.levels else e_i32
        return RuleTable(e_i32.view(0, 2), e_f64, e_f64, arena, sizes.tolist(), None, res.n_playlists)
    arena_items = torch.cat([it.reshape(-1) for it, _ in res.levels])
    arena_counts = torch.cat([ct for _, ct in res.levels])
    arena_par = torch.cat(res.parents)
    n_x_all = n_all - int(sizes[1])
    nodes = dev.empty(n_all * 2, torch.int64)          # 16 bytes per itemset
    sizes_ptr = sizes.ctypes.data
    _lib.call("kmls_trie_index_build", arena_items, arena_counts, arena_par, sizes_ptr, max_k,
              nodes, s)
    # slice boundaries by work (2^(k-1) splits per k-itemset), not by itemset count, so
    # that the ranks of a multi-GPU job get rule tables of similar size
    def boundary(frac_num, frac_den):
        weights = [int(sizes[k]) << (k - 1) for k in range(2, max_k + 1)]
        target = (sum(weights) * frac_num) // frac_den
        x = 0
        for k, w in zip(range(2, max_k + 1), weights):
            if target >= w:
                target -= w
                x += int(sizes[k])
            else:
                return x + (target >> (k - 1))
        return x
    x_begin = boundary(part[0], part[1])
    x_end = boundary(part[0] + 1, part[1]) if part[0] + 1 < part[1] else n_x_all
    n_x = x_end - x_begin
    rule_cnt = dev.empty(n_x, torch.int32)
    missing = torch.zeros(1, dtype=torch.int32, device=device)
    # never truncate, never die half-way: refuse up front what cannot fit
    splits = 0
    lvl_start = 0
    for k in range(2, max_k + 1):
        lo, hi = max(x_begin, lvl_start), min(x_end, lvl_start + int(sizes[k]))
        if hi > lo:
            splits += (hi - lo) * (1 << (k - 1))       # slots of pass 1's output, 8 bytes each
        lvl_start += int(sizes[k])
    free_bytes = torch.cuda.mem_get_info(device)[0] + torch.cuda.memory_reserved(device) - torch.cuda.memory_allocated(device)
    if splits * 8 > free_bytes:
        raise _lib.KmlsError(
            f"rule generation would examine {splits} (itemset, antecedent) splits and needs {splits * 8 / 1e9:.1f} GB "
            f"of scratch for them, more than the {free_bytes / 1e9:.1f} GB free on this GPU: raise min_support, "
            "shard over more GPUs, or skip the rule table (results are never truncated)")
    # per level: the slice of this part that falls into the level
    plan, lvl_start = [], 0
    for k in range(2, max_k + 1):
        lo, hi = max(x_begin, lvl_start), min(x_end, lvl_start + int(sizes[k]))
        if hi > lo:
            ppi = int(_lib.value("kmls_rules_parts_per_itemset", k))
            if ppi < 0:
                raise _lib.KmlsError(f"rule generation supports itemsets of at most 15 members, got {k}")
            plan.append((k, lo - lvl_start, hi - lo, lo - x_begin, dev.empty((hi - lo) * ppi, torch.int64)))
        lvl_start += int(sizes[k])
    if pass1 == "auto":
        pass1 = "gather" if part == (0, 1) else "search"
    if pass1 == "gather":
        # whole levels in ascending order: every subset count is read from the previous
        # level's output; only the k-1 subsets that keep the last member are searched
        if part != (0, 1):
            raise ValueError("pass1='gather' needs the whole table on this GPU (part=(0, 1))")
        sub_prev, part_prev = 0, 0
        for k, lvl_begin, cnt, out0, part_buf in plan:
            sub = dev.empty(cnt * (k - 1), torch.int32)
            _lib.call("kmls_rules_subsets", arena_items, arena_par, sizes_ptr, max_k, k, nodes,
                      sub_prev, sub, missing, s)
            _lib.call("kmls_rules_count_gather", arena_items, arena_counts, arena_par, sizes_ptr, max_k,
                      k, sub, part_prev, res.n_playlists, float(min_confidence), part_buf,
                      rule_cnt[out0:], missing, s)
            sub_prev, part_prev = sub, part_buf
    elif pass1 == "search":
        for k, lvl_begin, cnt, out0, part_buf in plan:
            _lib.call("kmls_rules_count", arena_items, arena_counts, sizes_ptr, max_k, k, lvl_begin, cnt,
                      nodes, res.n_playlists, float(min_confidence), part_buf, rule_cnt[out0:],
                      missing, s)
    else:
        raise ValueError(f"unknown pass1 {pass1!r}")
    scratch, sbytes = dev.scan_scratch(n_x)
    rule_off = dev.empty(n_x + 1, torch.int64)
    _lib.call("kmls_exclusive_scan_i32", rule_cnt, n_x, rule_off, scratch, sbytes, s)
    n_rules = dev.scalar(rule_off[n_x])
    if dev.scalar(missing) != 0:
        raise _lib.KmlsError("rule generation: a subset of a frequent itemset is missing from the "
                             "itemset table (table not downward closed)")
    if n_rules > max_rules:
        raise _lib.KmlsError(f"{n_rules} rules exceed max_rules={max_rules} (never truncated)")
    o_rule = dev.empty(n_rules * 2, torch.int32)
    o_conf = dev.empty(n_rules, torch.float64)
    o_lift = dev.empty(n_rules, torch.float64)
    if n_rules:
        for k, lvl_begin, cnt, out0, part_buf in plan:
            _lib.call("kmls_rules_write", arena_counts, sizes_ptr, max_k, k, lvl_begin, cnt,
                      res.n_playlists, float(min_confidence), part_buf, rule_off[out0:], o_rule,
                      o_conf, o_lift, missing, s)
        if dev.scalar(missing) != 0:
            raise _lib.KmlsError("rule generation: pass 2 kept a different number of rules than pass 1 "
                                 "sized for some itemset (the rule table would be corrupt)")
    return RuleTable(o_rule.view(n_rules, 2), o_conf, o_lift, arena_items, sizes.tolist(), arena_counts,
                     res.n_playlists)


def itemset_checksum(res: MineResult) -> int:
    """Order-independent checksum of the whole itemset table, computed on the device (same
    arithmetic as ``oracle.fpgrowth_c.itemset_checksum``)."""
    out = torch.zeros(1, dtype=torch.int64, device=res.row_items.device)
    for k, (it, ct) in enumerate(res.levels, start=1):
        if int(ct.numel()):
            _lib.call("kmls_itemset_checksum", it, ct, int(ct.numel()), k, res.row_items, out, _stream())
    return int(out.item()) & ((1 << 64) - 1)


def rule_checksum(res: MineResult, rules: RuleTable) -> int:
    """Order-independent checksum of the whole rule table, computed on the device (same
    arithmetic as the CPU oracle's): compares full-size tables without copying them."""
    out = torch.zeros(1, dtype=torch.int64, device=res.row_items.device)
    sizes = np.asarray(rules.level_sizes, dtype=np.int64)
    if rules.n_rules:
        _lib.call("kmls_rule_checksum", rules.arena_items, rules.arena_counts, sizes.ctypes.data,
                  len(res.levels), rules.rule, rules.confidence, rules.lift, rules.n_rules,
                  res.row_items, res.n_playlists, out, _stream())
    return int(out.item()) & ((1 << 64) - 1)


def rules_to_host(res: MineResult, rules: RuleTable):
    """[(antecedent frozenset, consequent frozenset, support, confidence, lift)] with TRACK
    ids as members — plain Python objects, for artifacts and tests (small tables only)."""
    rows = res.row_items.cpu().numpy()
    arena = rules.arena_items.cpu().numpy()
    sizes = rules.level_sizes
    id_off = np.zeros(len(sizes) + 1, dtype=np.int64)
    item_off = np.zeros(len(sizes) + 1, dtype=np.int64)
    for k in range(1, len(sizes)):
        id_off[k + 1] = id_off[k] + sizes[k]
        item_off[k + 1] = item_off[k] + sizes[k] * k
    out = []
    ids = rules.itemset.cpu().numpy()
    ante = rules.antecedent.cpu().numpy().astype(np.uint32)
    sup, conf, lift = (t.cpu().numpy() for t in (rules.support, rules.confidence, rules.lift))
    for r in range(len(ids)):
        gid = int(ids[r])
        k = int(np.searchsorted(id_off, gid, side="right")) - 1
        b = item_off[k] + (gid - id_off[k]) * k
        members = rows[arena[b:b + k]]
        m = int(ante[r])
        a = frozenset(int(members[t]) for t in range(k) if m >> t & 1)
        c = frozenset(int(members[t]) for t in range(k) if not m >> t & 1)
        out.append((a, c, float(sup[r]), float(conf[r]), float(lift[r])))
    return out
