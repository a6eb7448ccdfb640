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
