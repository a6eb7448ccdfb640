"""The reference's plugin boundary for the rule-generation hot path.

``calculate_and_save_fp_growth_fast`` has the name, arguments, return tuple, printed lines
and error behaviour of reference machine-learning/main.py:262-313, so the job's ``__main__``
(main.py:476, :461) and ``rest_api`` (rest_api/app/main.py:74-76, :229-254) keep working
unchanged; behind it the work runs on the GPU through ``libkmls_b200.so``.

Additions that the reference does not have: the frequent-itemset table
and the rule table (antecedents, consequents, support, confidence, lift) are kept on the
returned ``last_result`` / written by ``save_rules``; ``MIN_CONFIDENCE`` (default 0.04, the
value hard-coded in the reference's dormant path, main.py:226-227) selects the rules.
"""
from __future__ import annotations

import os
import pickle
import warnings
from typing import Dict, Iterable, List, Optional, Sequence, Tuple

import numpy as np
import pandas as pd
import torch

from . import _lib, miner

MIN_CONFIDENCE = float(os.getenv("MIN_CONFIDENCE", 0.04))
# multi-GPU: from this many playlists on, the playlists are sharded instead of the candidate space
PLAYLIST_SHARDING_FROM = int(os.getenv("KMLS_PLAYLIST_SHARDING_FROM", 4_000_000))
RULES_FILE = os.getenv("RULES_FILE", "rules.pickle")

# The reference reads a module-global `total_songs` (main.py:49, set at :436, read at :304).
# The patched job passes it as a keyword; this global is the fallback with the same name.
total_songs = None

# Result of the most recent call (itemset tables, rule table, timings) for callers that want
# more than the reference's return value.
last_result: Optional[dict] = None


def _device() -> torch.device:
    if not torch.cuda.is_available():
        raise _lib.KmlsError("no CUDA device: the GPU rule-generation path has no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def encode_transactions(transactions: Sequence[Sequence[str]]):
    """Host half of TransactionEncoder.fit: ``columns_ = sorted(unique names)`` and the
    transactions as ids in CSR. (Bit-packing — ``transform`` — happens on the device.)"""
    lens = np.fromiter((len(t) for t in transactions), dtype=np.int64, count=len(transactions))
    offsets = np.zeros(len(transactions) + 1, dtype=np.int64)
    np.cumsum(lens, out=offsets[1:])
    flat: List = []
    for t in transactions:
        flat.extend(t)
    columns = sorted(set(flat))
    mapping = {name: i for i, name in enumerate(columns)}
    ids = np.fromiter((mapping[x] for x in flat), dtype=np.int32, count=len(flat))
    return columns, offsets, ids


def _shard(offsets: np.ndarray, ids: np.ndarray, rank: int, world: int):
    n = len(offsets) - 1
    lo, hi = (n * rank) // world, (n * (rank + 1)) // world
    o = offsets[lo:hi + 1]
    return (o - o[0]).astype(np.int64), ids[o[0]:o[-1]]


def _dist_group():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist.group.WORLD, dist.get_rank(), dist.get_world_size()
    return None, 0, 1


class HostTables:
    """Level tables streamed to pinned host memory while the GPU keeps mining: every level is
    copied on a side stream as soon as it is final. ``levels[k-1] = (items[n_k, k] as bitmap
    ROW indices, counts[n_k])`` as numpy views of the pinned buffers (valid until the next
    use of the same HostTables), ``row_items`` maps rows to track ids."""

    def __init__(self, device):
        self.device = device
        self.stream = torch.cuda.Stream(device=device)
        self._pinned = {}
        self._pending = []
        self.row_items = None

    def _buf(self, key, t):
        b = self._pinned.get(key)
        if b is None or b.numel() < t.numel():
            b = torch.empty(max(int(t.numel()), 1), dtype=t.dtype).pin_memory()
            self._pinned[key] = b
        return b[:t.numel()]

    # Bulk copies go out in pieces: the mining loop reads a few scalars back per level, and a
    # scalar queued behind one 200 MB transfer on the device-to-host copy engine stalls the
    # loop for the whole transfer (measured: no overlap at all with one copy per table).
    CHUNK = int(os.environ.get("KMLS_D2H_CHUNK_ELEMS", 1 << 20))

    def _copy(self, dst, src):
        n = src.numel()
        for a in range(0, n, self.CHUNK):
            dst[a:a + self.CHUNK].copy_(src[a:a + self.CHUNK], non_blocking=True)

    def begin(self):
        self._pending = []
        self.row_items = None

    def on_level(self, k, items, counts):
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(ev)
            hi = self._buf(("i", k), items)
            hc = self._buf(("c", k), counts)
            self._copy(hi, items.reshape(-1))
            self._copy(hc, counts)
        items.record_stream(self.stream)
        counts.record_stream(self.stream)
        self._pending.append((k, hi, hc))

    def copy_result(self, res):
        """All level tables of a finished result, queued on the side stream."""
        for k, (items, counts) in enumerate(res.levels, start=1):
            if k not in [p[0] for p in self._pending]:
                self.on_level(k, items, counts)

    def finish(self, res):
        with torch.cuda.stream(self.stream):
            self.stream.wait_stream(torch.cuda.current_stream(self.device))
            hr = self._buf(("r", 0), res.row_items)
            hr.copy_(res.row_items, non_blocking=True)
        self.stream.synchronize()
        self.row_items = hr.numpy()
        self.levels = [(hi.numpy().reshape(-1, k), hc.numpy()) for k, hi, hc in self._pending]
        self.nbytes = hr.numel() * 4 + sum(hi.numel() * 4 + hc.numel() * 4 for _, hi, hc in self._pending)
        return self


def mine_ids(offsets: np.ndarray, ids: np.ndarray, n_items: int, min_support: float,
             min_confidence: Optional[float] = None, max_len: Optional[int] = None,
             host_tables: Optional[HostTables] = None, split_rules: bool = False,
             rules_required: bool = True, partition: str = "auto",
             workspaces: Optional[dict] = None, sharded: bool = False,
             n_playlists_total: Optional[int] = None, **mine_kwargs):
    """Host CSR of track ids in, device tables out: (MineResult, RuleTable or None).

    One process: the whole path runs through ``kmls_mine`` / ``kmls_rules``. Under torchrun
    (an initialised process group) every rank passes the same playlists and
    ``partition="classes"`` (default) splits the candidate space over the ranks
    (``multi.mine_partitioned``: every rank uploads the whole set, mines its share of the
    subtrees, and ends up with the complete table); ``partition="playlists"`` gives each rank a
    contiguous range of the playlists and sums every level's counts with an allreduce
    (``miner.mine_csr(group=...)``). With ``host_tables`` the level tables are also copied to
    pinned host memory on a side stream while the rules are generated (see HostTables);
    ``offsets``/``ids`` may be pinned torch tensors to make the upload async. ``split_rules``:
    each rank generates only its share of the rule table. ``rules_required=False`` turns a rule
    table that cannot be built into a warning and ``rules = None`` (the reference-facing entry
    points use this). ``workspaces``: a dict the call keeps its device workspaces in
    (``"mine"``, ``"rules"``) so that repeated calls reuse them; the returned tables are views of
    those workspaces and are overwritten by the next call that is given the same dict.
    ``partition="auto"`` takes "classes" unless the job has 4 million playlists or more: there
    the work that grows with the number of playlists (encoding, pair counts, wide rows)
    outweighs the work per candidate, and sharding the playlists wins (config 5: 25 ms against
    44 ms on 8 GPUs). ``sharded=True``: ``offsets``/``ids`` already are THIS rank's share of the
    playlists (``n_playlists_total`` = the job's count); implies ``partition="playlists"``."""
    device = _device()
    group, rank, world = _dist_group()
    n_total = int(n_playlists_total) if n_playlists_total is not None else len(offsets) - 1
    if sharded:
        partition = "playlists"
    elif partition == "auto":
        partition = "playlists" if n_total >= PLAYLIST_SHARDING_FROM else "classes"
    python_loop = (mine_kwargs.get("impl") == "python" or mine_kwargs.get("count_impl", "projected") != "projected"
                   or bool(mine_kwargs.get("prune")) or mine_kwargs.get("first_projection", "gather") != "gather"
                   or mine_kwargs.get("prepared") is not None or mine_kwargs.get("on_level") is not None)
    by_playlists = world > 1 and (partition == "playlists" or python_loop)
    if by_playlists and not sharded:
        if torch.is_tensor(offsets):
            offsets, ids = offsets.numpy(), ids.numpy()
        offsets, ids = _shard(offsets, ids, rank, world)
    if world > 1 and not by_playlists and torch.is_tensor(ids) and int(ids.numel()) >= (1 << 20):
        # every rank needs ALL playlists on its device, but not all of them over ITS host link:
        # each rank uploads 1/R of the id array and the ranks exchange the pieces over NVLink
        import torch.distributed as dist
        d_off = offsets.to(device, non_blocking=True)
        nnz = int(ids.numel())
        piece = (nnz + world - 1) // world
        piece += (-piece) % 4                                   # keeps every rank's piece 16-byte aligned
        d_ids_all = torch.empty(piece * world, dtype=ids.dtype, device=device)
        lo, hi = min(rank * piece, nnz), min((rank + 1) * piece, nnz)
        send = torch.empty(piece, dtype=ids.dtype, device=device)      # (the tail past nnz is never read)
        send[:hi - lo].copy_(ids[lo:hi], non_blocking=True)
        dist.all_gather_into_tensor(d_ids_all, send, group=group)
        d_ids = d_ids_all[:nnz]
    elif torch.is_tensor(offsets):
        d_off, d_ids = offsets.to(device, non_blocking=True), ids.to(device, non_blocking=True)
    else:
        d_off = torch.from_numpy(np.ascontiguousarray(offsets)).to(device, non_blocking=True)
        d_ids = torch.from_numpy(np.ascontiguousarray(ids)).to(device, non_blocking=True)
    ws = workspaces if workspaces is not None else {}
    if host_tables is not None:
        host_tables.begin()
    if world > 1 and not by_playlists:
        from . import multi
        for k_ in ("count_impl", "prune", "first_projection", "prepared", "on_level", "impl"):
            mine_kwargs.pop(k_, None)
        if "mine" not in ws and workspaces is not None:
            ws["mine"] = miner.Workspace(device)
        res = multi.mine_partitioned(d_off, d_ids, n_items, min_support, group, max_len=max_len,
                                     workspace=ws.get("mine"), **mine_kwargs)
    elif by_playlists:
        if host_tables is not None:
            mine_kwargs["on_level"] = host_tables.on_level
        mine_kwargs.pop("impl", None)
        res = miner.mine_csr(d_off, d_ids, n_items, min_support, n_playlists_total=n_total,
                             group=group, max_len=max_len, impl="python", **mine_kwargs)
    else:
        if "mine" not in ws and workspaces is not None:
            ws["mine"] = miner.Workspace(device)
        res = miner.mine_csr(d_off, d_ids, n_items, min_support, n_playlists_total=n_total,
                             max_len=max_len, workspace=ws.get("mine"), **mine_kwargs)
    if host_tables is not None and not by_playlists:
        host_tables.copy_result(res)        # async, on the side stream: overlaps the rule generation
    rules = None
    if min_confidence is not None:
        try:
            if "rules" not in ws and workspaces is not None:
                ws["rules"] = miner.Workspace(device)
            rules = miner.generate_rules(res, min_confidence,
                                         part=(rank, world) if split_rules else (0, 1),
                                         workspace=ws.get("rules"))
        except _lib.KmlsError as e:
            if rules_required:
                raise
            # The rule table is an ADDITION to the reference's result: when it
            # cannot be built (an itemset of more than 15 members, a table larger than the free
            # device memory or max_rules) the itemsets - all the reference itself computes -
            # must still come back.
            warnings.warn(f"rule table skipped: {e}", RuntimeWarning, stacklevel=2)
            rules = None
    if host_tables is not None:
        host_tables.finish(res)
    return res, rules


RULES_HOST_LIMIT = int(float(os.getenv("KMLS_RULES_HOST_GB", 96)) * (1 << 30))


def rules_to_pinned(rules: miner.RuleTable, cache: Optional[dict] = None, piece: int = 64 << 20):
    """The rule table's columns copied to pinned host memory: ``(rule int32[R, 2], confidence
    float64[R], lift float64[R])`` as numpy views of the pinned buffers (kept in ``cache`` for
    reuse by the next call). 24 bytes per rule, moved in pieces so that other transfers can
    interleave. Refuses, with the size it would need, a table beyond ``KMLS_RULES_HOST_GB``
    (default 96 GB of pinned host memory) instead of exhausting the host."""
    n = rules.n_rules
    need = 24 * n
    if need > RULES_HOST_LIMIT:
        raise _lib.KmlsError(f"the rule table ({n} rules, {need / 1e9:.1f} GB) exceeds the host-memory bound "
                             f"KMLS_RULES_HOST_GB={RULES_HOST_LIMIT / (1 << 30):.0f}: raise the bound, raise "
                             "min_confidence, or keep the table on the device (nothing is truncated)")
    cache = cache if cache is not None else {}
    out = []
    for key, t in (("rule", rules.rule.reshape(-1)), ("confidence", rules.confidence), ("lift", rules.lift)):
        b = cache.get(key)
        if b is None or b.numel() < t.numel() or b.dtype != t.dtype:
            b = cache[key] = torch.empty(max(int(t.numel()), 1), dtype=t.dtype).pin_memory()
        dst = b[:t.numel()]
        step = max(1, piece // t.element_size())
        for a in range(0, int(t.numel()), step):
            dst[a:a + step].copy_(t[a:a + step], non_blocking=True)
        out.append(dst)
    torch.cuda.current_stream().synchronize()
    return out[0].numpy().reshape(-1, 2), out[1].numpy(), out[2].numpy()


def pair_map_csr(res: miner.MineResult):
    """The fold's content as host arrays, built on the device (``kmls_pair_map``): ``(rows,
    indptr, other, support)``: ``rows[r]`` = track id of bitmap row r; the entries of row r are
    ``indptr[r]:indptr[r+1]``; ``other`` = the co-occurring TRACK id, ``support`` = count / N;
    inside a row ordered by support descending, ties by track id."""
    device = res.row_items.device
    n_rows = int(res.row_items.numel())
    rows = res.row_items.cpu().numpy()
    n2 = int(res.levels[1][1].numel()) if len(res.levels) >= 2 else 0
    if n_rows == 0 or n2 == 0:
        return rows, np.zeros(n_rows + 1, np.int64), np.zeros(0, np.int32), np.zeros(0, np.float64)
    it2, ct2 = res.levels[1]
    indptr = torch.empty(n_rows + 1, dtype=torch.int64, device=device)
    other = torch.empty(2 * n2, dtype=torch.int32, device=device)
    support = torch.empty(2 * n2, dtype=torch.float64, device=device)
    nbytes = int(_lib.value("kmls_pair_map_scratch_bytes", n2, n_rows))
    scratch = torch.empty(nbytes, dtype=torch.uint8, device=device)
    _lib.call("kmls_pair_map", it2, ct2, n2, res.row_items, n_rows, res.n_playlists, indptr, other, support,
              scratch, nbytes, torch.cuda.current_stream().cuda_stream)
    return rows, indptr.cpu().numpy(), other.cpu().numpy(), support.cpu().numpy()


def recommendations_from_result(res: miner.MineResult, columns: Sequence[str]
                                ) -> Dict[str, Dict[str, float]]:
    """The fold of main.py:282-296 from the level-1 and level-2 tables: every frequent single
    name is a key (possibly with an empty dict) and d[a][b] = support({a, b}),
    which is what the reference's max over all itemsets containing both evaluates to. The
    pairs arrive as a CSR sorted on the device (``pair_map_csr``), so the dict is two array
    walks: keys in the caller's column order, each inner dict in support-descending order
    (ties by name) - plain ``dict/str/float`` only, as rest_api's image has no numpy."""
    rows, indptr, other, support = pair_map_csr(res)
    names = np.asarray(list(columns), dtype=object)
    other_names = names[other].tolist() if len(other) else []
    support = support.tolist()
    out: Dict[str, Dict[str, float]] = {}
    for r in np.argsort(rows, kind="stable").tolist():
        lo, hi = int(indptr[r]), int(indptr[r + 1])
        out[names[rows[r]]] = dict(zip(other_names[lo:hi], support[lo:hi]))
    return out


def itemsets_frame(res: miner.MineResult, columns: Sequence[str]) -> pd.DataFrame:
    """What ``fpgrowth(df, min_support, use_colnames=True)`` returns: columns ``support``
    (count / N) and ``itemsets`` (frozenset of names); row order is not part of parity."""
    off, items, counts = res.to_host()
    n = res.n_playlists
    return pd.DataFrame({
        "support": [int(c) / n for c in counts],
        "itemsets": [frozenset(columns[i] for i in items[off[j]:off[j + 1]]) for j in range(len(counts))],
    })


def rule_columns(res: miner.MineResult, rules: miner.RuleTable, cache: Optional[dict] = None) -> dict:
    """The rule table as host columns without a Python loop over rules: ``ante_items`` /
    ``ante_off`` and ``cons_items`` / ``cons_off`` (CSR of TRACK ids per rule, members in row
    order), ``support``, ``confidence``, ``lift`` (float64). The three device columns come over
    in pieces through pinned memory (``rules_to_pinned``), the member lists are expanded with
    numpy from the itemset arena."""
    n = rules.n_rules
    rule, conf, lift = rules_to_pinned(rules, cache)
    sup = rules.support.cpu().numpy() if n else np.zeros(0, np.float64)
    rows = res.row_items.cpu().numpy()
    arena = rules.arena_items.cpu().numpy()
    sizes = np.asarray(rules.level_sizes, dtype=np.int64)
    id_off = np.concatenate([[0], np.cumsum(sizes)])                 # id_off[k] = first global id of level k
    item_off = np.concatenate([[0], np.cumsum(sizes * np.arange(len(sizes)))])
    gid = rule[:, 0].astype(np.int64)
    mask = rule[:, 1].astype(np.int64)
    k = np.searchsorted(id_off, gid, side="right") - 1               # itemset size of every rule
    base = item_off[k] + (gid - id_off[k]) * k                       # first member in the arena
    # expand to one entry per (rule, member)
    rep = np.repeat(np.arange(n, dtype=np.int64), k)
    first = np.concatenate([[0], np.cumsum(k)])[:-1] if n else np.zeros(0, np.int64)
    t = np.arange(int(k.sum()), dtype=np.int64) - np.repeat(first, k)   # member position inside its itemset
    member = rows[arena[np.repeat(base, k) + t]]
    in_ante = ((np.repeat(mask, k) >> t) & 1).astype(bool)
    n_ante = np.bincount(rep[in_ante], minlength=n).astype(np.int64) if n else np.zeros(0, np.int64)
    return {"ante_items": member[in_ante].astype(np.int32), "ante_off": np.concatenate([[0], np.cumsum(n_ante)]),
            "cons_items": member[~in_ante].astype(np.int32), "cons_off": np.concatenate([[0], np.cumsum(k - n_ante)]),
            "support": sup, "confidence": np.array(conf, copy=True), "lift": np.array(lift, copy=True)}


def rules_frame(res: miner.MineResult, rules: miner.RuleTable, columns: Sequence[str]) -> pd.DataFrame:
    c = rule_columns(res, rules)
    names = np.asarray(list(columns), dtype=object)
    an, cn = names[c["ante_items"]], names[c["cons_items"]]
    ao, co = c["ante_off"], c["cons_off"]
    return pd.DataFrame({
        "antecedents": [frozenset(an[ao[r]:ao[r + 1]]) for r in range(len(c["support"]))],
        "consequents": [frozenset(cn[co[r]:co[r + 1]]) for r in range(len(c["support"]))],
        "support": c["support"], "confidence": c["confidence"], "lift": c["lift"],
    })


def calculate_and_save_fp_growth_fast(playlist_dict: dict, min_support=0.07, total_songs=None,
                                      min_confidence: Optional[float] = None,
                                      max_len: Optional[int] = None):
    """Drop-in for reference machine-learning/main.py:262-313.

    Returns ``(songs_to_song_sets, info, (songs_without_recommendations, duration))`` with
    ``songs_to_song_sets: dict[str, dict[str, float]]`` made of plain Python objects
    (rest_api's image has neither numpy nor torch). ``min_support <= 0`` raises ValueError as
    mlxtend does; an empty result is ``{}`` (the reference's ``data is None`` branch,
    main.py:274-277, is unreachable). CUDA failures raise (no CPU fallback)."""
    global last_result
    start_time = pd.Timestamp.now()
    if min_support <= 0.0:
        miner.thresholds(min_support, 1)
    transactions = list(playlist_dict.values())
    columns, offsets, ids = encode_transactions(transactions)
    conf = MIN_CONFIDENCE if min_confidence is None else min_confidence
    res, rules = mine_ids(offsets, ids, len(columns), min_support, conf, max_len, rules_required=False)
    songs_to_song_sets = recommendations_from_result(res, columns)

    ts = total_songs if total_songs is not None else globals()["total_songs"]
    if ts is None:
        ts = len(columns)
    songs_without_recommendations = ts - len(songs_to_song_sets.keys())
    print("Songs without recommendations:", songs_without_recommendations)
    end_time = pd.Timestamp.now()
    duration = end_time - start_time
    print(f"Time elapsed in rule generation: {duration}")
    info = f"min_support: {min_support} \tmissing songs: {songs_without_recommendations} \ttime: {duration}"
    runtime_info = (songs_without_recommendations, duration)
    last_result = {"columns": columns, "itemsets": res, "rules": rules, "min_confidence": conf}
    return songs_to_song_sets, info, runtime_info


# --------------------------------------------------------------------------------------
# The rows SURVEY.md section 8(f) marks "next": the callers and data formats either side of the path.
# --------------------------------------------------------------------------------------
def encode_frame(pid, track_name):
    """(f1) Ingest without the Python dict: what ``group_tracks_by_playlist...`` (main.py:195-207)
    + ``list(playlist_dict.values())`` (:265) + ``TransactionEncoder.fit`` produce, straight from the
    CSV's two columns: ``columns`` = sorted unique names, CSR of name ids grouped by ``pid``
    (playlists in ascending pid order, rows of a playlist in file order, duplicates kept)."""
    pid = np.asarray(pid)
    names = pd.Series(np.asarray(track_name, dtype=object))
    cat = pd.Categorical(names, categories=sorted(set(names.tolist())))
    ids = np.asarray(cat.codes, dtype=np.int32)
    order = np.argsort(pid, kind="stable")
    pid_s = pid[order]
    starts = np.flatnonzero(np.concatenate([[True], pid_s[1:] != pid_s[:-1]])) if len(pid_s) else np.zeros(0, int)
    offsets = np.concatenate([starts, [len(pid_s)]]).astype(np.int64)
    return list(cat.categories), offsets, ids[order]


def recommendations_from_frame(pid, track_name, min_support=0.07, total_songs=None,
                               min_confidence: Optional[float] = None):
    """``calculate_and_save_fp_growth_fast`` for the two CSV columns themselves (no dict of
    lists of strings in between): same return tuple, same prints."""
    global last_result
    start_time = pd.Timestamp.now()
    if min_support <= 0.0:
        miner.thresholds(min_support, 1)
    columns, offsets, ids = encode_frame(pid, track_name)
    conf = MIN_CONFIDENCE if min_confidence is None else min_confidence
    res, rules = mine_ids(offsets, ids, len(columns), min_support, conf, rules_required=False)
    out = recommendations_from_result(res, columns)
    ts = total_songs if total_songs is not None else (globals()["total_songs"] or len(columns))
    missing = ts - len(out)
    print("Songs without recommendations:", missing)
    duration = pd.Timestamp.now() - start_time
    print(f"Time elapsed in rule generation: {duration}")
    last_result = {"columns": columns, "itemsets": res, "rules": rules, "min_confidence": conf}
    return out, f"min_support: {min_support} \tmissing songs: {missing} \ttime: {duration}", (missing, duration)


def most_frequent_tracks(columns: Sequence[str], occurrences) -> List[dict]:
    """(f3) ``get_most_frequent_tracks`` (main.py:168-175) from the level-1 kernel's output:
    the reference counts ROWS per track_name (``group_by("track_name").agg(count)``, repeats
    inside a playlist included), which is exactly the occurrence histogram K_hist produces
    (``Prepared.occurrences``). Sorted by count descending (ties by name, the reference's tie
    order is unspecified). ``filter_best_tracks`` (main.py:177-184) then keeps a prefix."""
    occ = np.asarray(occurrences.cpu() if hasattr(occurrences, "cpu") else occurrences)
    order = sorted(range(len(columns)), key=lambda i: (-int(occ[i]), columns[i]))
    return [{"track_name": columns[i], "count": int(occ[i])} for i in order if occ[i] > 0]


def sweep_supports(playlist_dict: dict, supports: Iterable[float], total_songs=None) -> pd.DataFrame:
    """(f4) The reference's support sweep (main.py:452-473: one full
    ``calculate_and_save_fp_growth_fast`` per threshold, logging min_support,
    songs_without_recommendations, duration) with the threshold-independent work done once: the
    bitmap and the pair counts are built at the LOWEST support of the sweep and every threshold
    mines on a suffix of their rows. Returns the reference's results frame."""
    supports = [round(float(x), 3) for x in supports]
    device = _device()
    columns, offsets, ids = encode_transactions(list(playlist_dict.values()))
    d_off = torch.from_numpy(offsets).to(device)
    d_ids = torch.from_numpy(ids).to(device)
    t_prep = pd.Timestamp.now()
    prepared = miner.prepare(d_off, d_ids, len(columns), min(supports))
    torch.cuda.synchronize()
    prep_s = (pd.Timestamp.now() - t_prep).total_seconds()
    ts = total_songs if total_songs is not None else (globals()["total_songs"] or len(columns))
    rows = []
    for s_ in supports:
        t0 = pd.Timestamp.now()
        res = miner.mine_csr(d_off, d_ids, len(columns), s_, prepared=prepared)
        recs = recommendations_from_result(res, columns)
        rows.append({"min_support": s_, "songs_without_recommendations": ts - len(recs),
                     "duration": (pd.Timestamp.now() - t0).total_seconds(),
                     "itemsets": res.n_itemsets})
    out = pd.DataFrame(rows)
    out.attrs["prepare_seconds"] = prep_s
    return out


def filter_best_tracks(track_infos: List[dict], percentile: Optional[float] = None) -> List[dict]:
    """``filter_best_tracks`` (main.py:177-184): the first ``len * TOP_TRACKS_SAVE_PERCENTILE``
    entries of the frequency-sorted list (env knob of main.py:47, default 0.03)."""
    p = float(os.getenv("TOP_TRACKS_SAVE_PERCENTILE", 0.03)) if percentile is None else percentile
    return track_infos[:int(len(track_infos) * p)]


def save_rules(path: str, result: Optional[dict] = None) -> None:
    """Write the extra rule artifact: a dict of plain-Python columns ``antecedents``,
    ``consequents`` (tuples of names, sorted), ``support``, ``confidence``, ``lift`` (floats),
    loadable without numpy/torch. rest_api ignores it. Built from ``rule_columns`` (numpy);
    the only per-rule Python work is making the tuples the pickle consists of. Tables beyond
    the host-memory bound of ``rules_to_pinned`` are refused, never truncated."""
    result = result or last_result
    if result is None or result.get("rules") is None:
        raise ValueError("no rule table to save: run calculate_and_save_fp_growth_fast first")
    c = rule_columns(result["itemsets"], result["rules"])
    names = np.asarray(list(result["columns"]), dtype=object)
    an, cn = names[c["ante_items"]].tolist(), names[c["cons_items"]].tolist()
    ao, co = c["ante_off"].tolist(), c["cons_off"].tolist()
    n = len(c["support"])
    data = {
        "antecedents": [tuple(sorted(an[ao[r]:ao[r + 1]])) for r in range(n)],
        "consequents": [tuple(sorted(cn[co[r]:co[r + 1]])) for r in range(n)],
        "support": c["support"].tolist(), "confidence": c["confidence"].tolist(), "lift": c["lift"].tolist(),
    }
    with open(path, "wb") as f:
        pickle.dump(data, f)
