"""Multi-GPU mining by partitioning the CANDIDATE space (one process per GPU, NCCL).

Every rank holds the whole playlist set (the bitmap of BASELINE config 3 is 0.75 GB, config 5's
5.6 GB: both fit a B200 many times over) and mines levels 1 and 2 completely. From level 3 on
the work is split by frequent pair: the subtree under a pair (all itemsets whose two rarest
members are that pair) needs nothing from any other subtree, so rank r mines the subtrees of
blocks r, r + R, r + 2R, ... of ``block`` consecutive pairs with NO communication at all.
Afterwards the ranks' tables are merged into the job-wide tables - in the order of the root
pair, which is the sorted order a single GPU produces, so the result (and everything derived
from it: rule ids, checksums) is bit-identical to the 1-GPU result. Cost of the merge: one
small allreduce of per-pair itemset counts and one sum-allreduce of (last member, count, parent)
per itemset - 12 bytes each; the member lists are rebuilt locally from the merged parent links.

The alternative - sharding the PLAYLISTS and summing every level's candidate counts with an
allreduce (``miner.mine_csr(group=...)``, BASELINE.json's sketch) - replicates all per-candidate
work on every rank and synchronises the ranks a dozen times per level; it remains available
for inputs whose bitmap does not fit one GPU.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib, miner


def mine_partitioned(offsets: torch.Tensor, items: torch.Tensor, n_items: int, min_support: float,
                     group, *, block: int = 64, workspace: Optional[miner.Workspace] = None,
                     **kwargs) -> miner.MineResult:
    """``miner.mine_native`` over the ranks of ``group`` (every rank passes the WHOLE playlist
    set). Returns, on every rank, the complete job-wide result."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)

    def allreduce(t):
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)

    def any_rank(flag: bool) -> bool:
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=offsets.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
        return bool(t.item())

    res = miner.mine_native(offsets, items, n_items, min_support, part=(rank, world), part_block=block,
                            workspace=workspace, allreduce=allreduce if world > 1 else None,
                            retry_sync=any_rank if world > 1 else None, **kwargs)
    if world == 1:
        return res
    device = offsets.device
    s = torch.cuda.current_stream().cuda_stream
    marks = []

    def mark(name):      # CUDA events between the merge's phases (read by bench.py after the step)
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        marks.append((name, e))

    mark("mined")
    n2 = int(res.levels[1][1].numel()) if len(res.levels) >= 2 else 0
    # deepest level over all ranks
    kk = torch.tensor([len(res.levels)], dtype=torch.int64, device=device)
    dist.all_reduce(kk, op=dist.ReduceOp.MAX, group=group)
    max_k = int(kk.item())
    if max_k < 3 or n2 == 0:
        return res
    # roots and per-root counts of this rank's itemsets, levels 3 .. max_k
    local_cnt = torch.zeros((max_k - 2, n2), dtype=torch.int32, device=device)
    roots = []
    for k in range(3, max_k + 1):
        if k <= len(res.levels):
            par = res.parents[k - 2]
            n = int(par.numel())
            root = torch.empty(max(n, 1), dtype=torch.int32, device=device)[:n]
            _lib.call("kmls_subtree_roots", par, n, roots[-1] if roots else 0, root, local_cnt[k - 3], s)
        else:
            root = torch.empty(0, dtype=torch.int32, device=device)
        roots.append(root)
    global_cnt = local_cnt.clone()
    dist.all_reduce(global_cnt, op=dist.ReduceOp.SUM, group=group)
    sizes = global_cnt.sum(dim=1, dtype=torch.int64).tolist()          # job-wide level sizes (one read-back)
    mark("roots+counts")
    while sizes and sizes[-1] == 0:
        sizes.pop()
    max_k = 2 + len(sizes)
    # one flat buffer for what is exchanged: level k takes 3n int32 (last member, count, parent) -
    # an itemset is its parent plus one member, so the member lists are rebuilt locally afterwards
    total = sum(3 * n for n in sizes)
    flat = torch.zeros(max(total, 1), dtype=torch.int32, device=device)
    scratch = torch.empty(int(_lib.value("kmls_scan_scratch_bytes", n2)), dtype=torch.uint8, device=device)
    g_start = torch.empty(n2 + 1, dtype=torch.int64, device=device)
    l_start = torch.empty(n2 + 1, dtype=torch.int64, device=device)
    levels, parents = list(res.levels[:2]), list(res.parents[:1])
    at, g_prev, merged = 0, None, []
    for k, n_glob in zip(range(3, max_k + 1), sizes):
        last_g = flat[at:at + n_glob]
        ct_g = flat[at + n_glob:at + 2 * n_glob]
        par_g = flat[at + 2 * n_glob:at + 3 * n_glob]
        at += 3 * n_glob
        n_loc = int(roots[k - 3].numel())
        g_here = torch.empty(max(n_loc, 1), dtype=torch.int32, device=device)[:n_loc]
        if n_loc:
            _lib.call("kmls_exclusive_scan_i32", global_cnt[k - 3], n2, g_start, scratch, scratch.numel(), s)
            _lib.call("kmls_exclusive_scan_i32", local_cnt[k - 3], n2, l_start, scratch, scratch.numel(), s)
            it_l, ct_l = res.levels[k - 1]
            _lib.call("kmls_place_level_compact", it_l, ct_l, res.parents[k - 2], n_loc, k, roots[k - 3], g_start,
                      l_start, g_prev if k > 3 else 0, last_g, ct_g, par_g, g_here, s)
        g_prev = g_here
        merged.append((k, n_glob, last_g, ct_g, par_g))
    mark("placed")
    if total:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)      # every entry is written by exactly one rank
    mark("allreduced")
    members = torch.empty(max(sum(n * k for k, n in zip(range(3, max_k + 1), sizes)), 1), dtype=torch.int32,
                          device=device)
    at, prev_items = 0, res.levels[1][0]
    for k, n_glob, last_g, ct_g, par_g in merged:
        it_g = members[at:at + n_glob * k].view(n_glob, k)
        at += n_glob * k
        _lib.call("kmls_rebuild_items", prev_items, k, par_g, last_g, n_glob, it_g, s)
        levels.append((it_g, ct_g))
        parents.append(par_g)
        prev_items = it_g
    mark("rebuilt")
    stats = dict(res.stats or {})
    stats["partition"] = {"kind": "classes", "world": world, "block": block,
                          "local_itemsets": int(sum(int(c.numel()) for _, c in res.levels[2:])),
                          "merged_bytes": int(total) * 4}
    out = miner.MineResult(res.n_playlists, res.t1, res.tk, res.row_items, levels, res.pair_counts, stats, parents)
    out.workspace = (res.workspace, flat, members)   # the first two levels still live in the mining workspace
    out.occurrences = res.occurrences
    out.merge_marks = marks
    return out
