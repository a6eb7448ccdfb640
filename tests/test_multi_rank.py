"""N > 1: playlists are sharded by rank and every level's counts are summed with one
allreduce. The CPU tests run the host-side logic (sharding, thresholds from the global N,
the allreduce helper) on the gloo backend with world size 2, the per-shard counts being
computed with numpy; the GPU test runs the real path on 2 GPUs when the box has them."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _dense(off, ids, n_items):
    d = np.zeros((len(off) - 1, n_items), bool)
    for r in range(len(off) - 1):
        d[r, ids[off[r]:off[r + 1]]] = True
    return d


def _cpu_worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import kmls_b200
    from kmls_b200 import api, miner, synth
    n, t, s = 1501, 40, 0.07
    off, ids = synth.zipf_playlists(n, t, mean_draws=8, alpha=0.6, seed=9)
    o, i = api._shard(off, ids, rank, world)
    local = _dense(o, i, t)
    # global N by allreduce, thresholds from it
    n_tot = torch.tensor([local.shape[0]], dtype=torch.int64)
    miner._allreduce(n_tot, dist.group.WORLD)
    t1, tk = kmls_b200.thresholds(s, int(n_tot.item()))
    c1 = torch.from_numpy(local.sum(0).astype(np.int32))
    miner._allreduce(c1, dist.group.WORLD)
    pair = torch.from_numpy((local.astype(np.int32).T @ local.astype(np.int32)).astype(np.int32))
    miner._allreduce(pair, dist.group.WORLD)
    freq = np.nonzero(c1.numpy() >= t1)[0]
    trip = [(a, b, c) for a in freq for b in freq for c in freq if a < b < c]
    c3 = torch.tensor([int((local[:, a] & local[:, b] & local[:, c]).sum()) for a, b, c in trip],
                      dtype=torch.int32)
    miner._allreduce(c3, dist.group.WORLD)
    if rank == 0:
        full = _dense(off, ids, t)
        ok = (int(n_tot.item()) == n and (c1.numpy() == full.sum(0)).all()
              and (pair.numpy() == full.astype(np.int32).T @ full.astype(np.int32)).all()
              and all(int(c3[j]) == int((full[:, a] & full[:, b] & full[:, c]).sum())
                      for j, (a, b, c) in enumerate(trip))
              and (t1, tk) == kmls_b200.thresholds(s, n))
        out.put(bool(ok))
    dist.destroy_process_group()


def test_sharded_counts_sum_to_global_counts_gloo():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_cpu_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert out.get(timeout=5) is True


def _merge_worker(rank, world, port, out):
    """The class partition's host logic in numpy over gloo: ownership of the frequent pairs by
    interleaved blocks, per-root counts summed over the ranks, their scans, the placement of a
    rank's itemsets at job-wide positions and the sum that completes the tables
    (kmls_b200/multi.py does the same with kmls_subtree_roots / kmls_place_level on the GPU)."""
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from kmls_b200 import synth
    from oracle import fpgrowth_c, mlxtend_restated as M
    n, t, s, block = 4000, 120, 0.02, 3
    off, ids = synth.zipf_playlists(n, t, mean_draws=12, alpha=0.8, seed=4)
    t1, tk = M.thresholds(s, n)
    full = fpgrowth_c.fpgrowth(off, ids, t, t1, tk).as_dict()
    max_k = max(len(k) for k in full)
    tables = [sorted(tuple(sorted(k)) for k in full if len(k) == kk) for kk in range(1, max_k + 1)]
    index = [{x: i for i, x in enumerate(tb)} for tb in tables]
    n2 = len(tables[1])
    # what this rank mined: levels >= 3 restricted to the subtrees of its own pairs, parents local
    own = [[x for x in tables[kk - 1] if (index[1][x[:2]] // block) % world == rank] for kk in range(3, max_k + 1)]
    lidx = [{x: i for i, x in enumerate(tb)} for tb in own]
    roots, local_cnt = [], np.zeros((max_k - 2, n2), np.int64)
    for li, tb in enumerate(own):
        par = np.array([index[1][x[:2]] if li == 0 else lidx[li - 1][x[:-1]] for x in tb], dtype=np.int64)
        root = par if li == 0 else roots[li - 1][par] if len(par) else par
        roots.append(root)
        local_cnt[li] = np.bincount(root, minlength=n2) if len(root) else 0
    g_cnt = torch.from_numpy(local_cnt.copy())
    dist.all_reduce(g_cnt)
    g_cnt = g_cnt.numpy()
    ok = True
    g_prev = None
    for li, tb in enumerate(own):
        kk = li + 3
        n_glob = int(g_cnt[li].sum())
        ok &= n_glob == len(tables[kk - 1])
        g_start = np.concatenate([[0], np.cumsum(g_cnt[li])])[:-1]
        l_start = np.concatenate([[0], np.cumsum(local_cnt[li])])[:-1]
        g = g_start[roots[li]] + (np.arange(len(tb)) - l_start[roots[li]]) if len(tb) else np.zeros(0, np.int64)
        items = np.zeros((n_glob, kk), np.int64)
        par_g = np.zeros(n_glob, np.int64)
        if len(tb):
            items[g] = np.array(tb, dtype=np.int64)
            lpar = np.array([index[1][x[:2]] if li == 0 else lidx[li - 1][x[:-1]] for x in tb], dtype=np.int64)
            par_g[g] = lpar if li == 0 else g_prev[lpar]
        ti, tp = torch.from_numpy(items), torch.from_numpy(par_g)
        dist.all_reduce(ti)
        dist.all_reduce(tp)
        ok &= ti.numpy().tolist() == [list(x) for x in tables[kk - 1]]                    # the sorted job-wide table
        ok &= tp.numpy().tolist() == [index[kk - 2][x[:-1]] for x in tables[kk - 1]]      # job-wide parent links
        g_prev = g
    if rank == 0:
        out.put(bool(ok) and max_k >= 4)
    dist.destroy_process_group()


def test_class_partition_merge_gloo():
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_merge_worker, args=(r, 2, port, out)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert out.get(timeout=5) is True


def _gpu_worker(rank, world, port, out, partition):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import kmls_b200
    from kmls_b200 import miner, synth
    n, t, s, conf = 30011, 900, 0.01, 0.5
    off, ids = synth.zipf_playlists(n, t, mean_draws=30, alpha=0.9, seed=17)
    # every rank passes the whole set: "classes" splits the candidate space, "playlists" the rows
    res, rules = kmls_b200.mine_ids(off, ids, t, s, conf, split_rules=True, partition=partition)
    got = res.as_dict()
    tables = [(it.cpu().numpy(), ct.cpu().numpy()) for it, ct in res.levels]
    parents = [p.cpu().numpy() for p in res.parents]
    cs = miner.rule_checksum(res, rules)
    t_ = torch.tensor([rules.n_rules, cs - (1 << 64) if cs >= (1 << 63) else cs], dtype=torch.int64, device="cuda")
    dist.all_reduce(t_)
    if rank == 0:
        out.put((got, int(t_[0].item()), int(t_[1].item()) & ((1 << 64) - 1), tables, parents,
                 (res.stats or {}).get("partition")))
    dist.destroy_process_group()


@pytest.mark.gpu
@pytest.mark.parametrize("partition", ["classes", "playlists"])
def test_two_gpus_give_the_single_gpu_result(partition):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    sys.path.insert(0, ROOT)
    import kmls_b200
    from kmls_b200 import miner, synth
    from oracle import fpgrowth_c, mlxtend_restated as M
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gpu_worker, args=(r, 2, port, out, partition)) for r in range(2)]
    for p in procs:
        p.start()
    got, n_rules, rule_cs, tables, parents, part_info = out.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    n, t, s, conf = 30011, 900, 0.01, 0.5
    off, ids = synth.zipf_playlists(n, t, mean_draws=30, alpha=0.9, seed=17)
    t1, tk = M.thresholds(s, n)
    ref = fpgrowth_c.fpgrowth(off, ids, t, t1, tk, rules_min_confidence=conf)
    assert got == ref.as_dict()
    assert n_rules == ref.n_rules
    assert rule_cs == ref.rule_checksum
    # the merged tables are the single-GPU tables, entry for entry (order and parent links included)
    one, _ = kmls_b200.mine_ids(off, ids, t, s)
    assert len(tables) == len(one.levels)
    for (it, ct), (it1, ct1) in zip(tables, one.levels):
        assert (it == it1.cpu().numpy()).all() and (ct == ct1.cpu().numpy()).all()
    for p2, p1 in zip(parents, one.parents):
        assert (p2 == p1.cpu().numpy()).all()
    if partition == "classes":
        assert part_info and part_info["world"] == 2
