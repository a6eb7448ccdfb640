"""The one-call boundary (kmls_mine): the level loop inside the library against the CPU oracle,
against the kernel-by-kernel Python loop, and through ctypes alone."""
import ctypes
import json
import os

import numpy as np
import pytest
import torch

import kmls_b200
from kmls_b200 import _lib, miner, synth
from oracle import fpgrowth_c, mlxtend_restated as M
from conftest import ROOT

pytestmark = pytest.mark.gpu
GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "generated_cases.json")))["cases"]


def dev():
    return torch.device("cuda", 0)


def to_dev(off, items):
    return torch.from_numpy(np.ascontiguousarray(off)).to(dev()), torch.from_numpy(np.ascontiguousarray(items)).to(dev())


def oracle_dict(off, items, n_items, s, max_len=0):
    t1, tk = M.thresholds(s, len(off) - 1)
    return fpgrowth_c.fpgrowth(off, items, n_items, t1, tk, max_len=max_len).as_dict()


@pytest.mark.parametrize("keep_words", [4, 8, 64, 256, 1024])
@pytest.mark.parametrize("pair_impl", ["auto", "csr", "popc"])
def test_native_loop_equals_oracle_and_python_loop(keep_words, pair_impl):
    # small keep_words push every level through the bit-extract route, 256 through the AND route
    off, items = synth.zipf_playlists(9000, 700, mean_draws=30, alpha=0.9, seed=31)
    want = oracle_dict(off, items, 700, 0.012)
    d_off, d_items = to_dev(off, items)
    res = miner.mine_native(d_off, d_items, 700, 0.012, keep_words=keep_words, pair_impl=pair_impl)
    assert res.as_dict() == want
    assert len(res.levels) >= 5
    py = miner.mine_csr(d_off, d_items, 700, 0.012, impl="python")
    for (a, b), (c, d) in zip(res.levels, py.levels):
        assert torch.equal(a, c) and torch.equal(b, d)          # same tables in the same order
    for a, b in zip(res.parents, py.parents):
        assert torch.equal(a, b)


def test_native_loop_wide_rows():
    # rows wider than 256 words at several levels (60k playlists, popular tracks): both routes in one run
    off, items = synth.zipf_playlists(60000, 3000, mean_draws=40, alpha=1.0, seed=5)
    d_off, d_items = to_dev(off, items)
    res = miner.mine_native(d_off, d_items, 3000, 0.004)
    assert res.as_dict() == oracle_dict(off, items, 3000, 0.004)
    assert any(l.get("row_words_total", 0) for l in res.stats["levels"])


def test_native_loop_rows_wider_than_one_segment():
    # 600k playlists: the popular tracks' conditional rows are > 8192 words, so the bit-extract
    # route splits them into several segment units whose outputs must tile the new row exactly
    n, t, s_ = 600_000, 400, 0.02
    off, items = synth.zipf_playlists(n, t, mean_draws=10, alpha=1.0, seed=9)
    d_off, d_items = to_dev(off, items)
    res = miner.mine_native(d_off, d_items, t, s_)
    t1, tk = M.thresholds(s_, n)
    ref = fpgrowth_c.fpgrowth(off, items, t, t1, tk, store=False)
    assert [int(c.numel()) for _, c in res.levels] == ref.level_counts[1:len(res.levels) + 1].tolist()
    assert miner.itemset_checksum(res) == ref.checksum
    assert len(res.levels) >= 4 and max(l.get("row_words_total", 0) for l in res.stats["levels"]) > 0
    # the same result when nothing is squeezed at all and when everything is
    for kw, sq in ((1024, 0), (4, 100)):
        import os
        os.environ["KMLS_SQUEEZE_PCT"] = str(sq)
        try:
            r2 = miner.mine_native(d_off, d_items, t, s_, keep_words=kw)
        finally:
            os.environ.pop("KMLS_SQUEEZE_PCT", None)
        assert miner.itemset_checksum(r2) == ref.checksum


@pytest.mark.parametrize("seed", range(8))
def test_native_loop_random_small(seed):
    rng = np.random.default_rng(100 + seed)
    n, t = int(rng.integers(1, 600)), int(rng.integers(1, 30))
    onehot = rng.random((n, t)) < rng.uniform(0.05, 0.7)
    s = float(rng.choice([0.05, 0.07, 0.1, 0.2, 0.035, 0.14]))
    off, items = fpgrowth_c.dense_to_csr(onehot)
    d_off, d_items = to_dev(off, items)
    for kw in (4, 256):
        assert miner.mine_native(d_off, d_items, t, s, keep_words=kw).as_dict() == M.fpgrowth_counts(onehot, s)


def test_native_loop_duplicates_and_max_len():
    rng = np.random.default_rng(5)
    n, t = 3000, 120
    onehot = rng.random((n, t)) < 0.08
    off, items = fpgrowth_c.dense_to_csr(onehot)
    rows = [rng.permutation(np.tile(items[off[i]:off[i + 1]], 3)) for i in range(n)]
    off3 = np.zeros(n + 1, np.int64)
    np.cumsum([len(r) for r in rows], out=off3[1:])
    items3 = np.concatenate(rows).astype(np.int32)
    d_off, d_items = to_dev(off3, items3)
    assert miner.mine_native(d_off, d_items, t, 0.08).as_dict() == M.fpgrowth_counts(onehot, 0.08)
    d_off, d_items = to_dev(off, items)
    for ml in (1, 2, 3):
        assert miner.mine_native(d_off, d_items, t, 0.08, max_len=ml).as_dict() == \
            oracle_dict(off, items, t, 0.08, max_len=ml)


def test_workspace_grows_and_results_survive():
    off, items = synth.zipf_playlists(9000, 700, mean_draws=30, alpha=0.9, seed=31)
    d_off, d_items = to_dev(off, items)
    ws = miner.Workspace(dev(), 1 << 16)               # far too small: must be grown, never truncated
    res = miner.mine_native(d_off, d_items, 700, 0.012, workspace=ws)
    assert ws.nbytes > (1 << 16)
    assert res.as_dict() == oracle_dict(off, items, 700, 0.012)
    # the bare C call reports the shortage as a capacity error that names what was missing
    small = torch.empty(1 << 16, dtype=torch.uint8, device=dev())
    out, opt = _lib.MineResultC(), _lib.MineOptions()
    t1, tk = miner.thresholds(0.012, 9000)
    rc, msg = _lib.call_rc("kmls_mine", d_off, d_items, 9000, 700, 9000, t1, tk, opt, small, small.numel(), out,
                           torch.cuda.current_stream().cuda_stream)
    assert rc == -2 and "workspace too small" in msg and "more bytes needed" in msg


def test_class_partition_covers_everything_once():
    off, items = synth.zipf_playlists(9000, 700, mean_draws=30, alpha=0.9, seed=31)
    d_off, d_items = to_dev(off, items)
    whole = miner.mine_native(d_off, d_items, 700, 0.012).as_dict()
    for n_parts, block in ((2, 64), (3, 7), (8, 1)):
        seen = {}
        for part in range(n_parts):
            r = miner.mine_native(d_off, d_items, 700, 0.012, part=(part, n_parts), part_block=block)
            d = r.as_dict()
            for key, cnt in d.items():
                if len(key) <= 2:
                    assert whole[key] == cnt                     # levels 1 and 2 are complete in every part
                else:
                    assert key not in seen, "an itemset of size >= 3 mined by two parts"
                    seen[key] = cnt
        assert seen == {k_: v for k_, v in whole.items() if len(k_) >= 3}


def test_host_tables_are_streamed_out():
    off, items = synth.zipf_playlists(9000, 700, mean_draws=30, alpha=0.9, seed=31)
    d_off, d_items = to_dev(off, items)
    host = torch.empty(64 << 20, dtype=torch.uint8).pin_memory()
    res = miner.mine_native(d_off, d_items, 700, 0.012, host_tables=host)
    lay = res.host_layout
    h = host.numpy()
    o, n = lay["row_items"]
    assert (h[o:o + 4 * n].view(np.int32) == res.row_items.cpu().numpy()).all()
    for k, ((io, co, n), (it, ct)) in enumerate(zip(lay["levels"], res.levels), start=1):
        assert (h[io:io + 4 * n * k].view(np.int32).reshape(n, k) == it.cpu().numpy()).all()
        assert (h[co:co + 4 * n].view(np.int32) == ct.cpu().numpy()).all()
    tiny = torch.empty(1024, dtype=torch.uint8).pin_memory()
    with pytest.raises(_lib.KmlsError, match="h_tables too small"):
        miner.mine_native(d_off, d_items, 700, 0.012, host_tables=tiny)


@pytest.mark.parametrize("g", GOLDEN, ids=[g["name"] for g in GOLDEN])
def test_golden_through_ctypes_alone(g):
    """What a C host does: thresholds, one kmls_mine call on a workspace it owns, tables read
    at the offsets the result names. Only device memory and the stream come from torch."""
    if g["generator"] == "zipf":
        off, items = synth.zipf_playlists(**g["params"])
        n_items = g["params"]["n_tracks"]
    else:
        d = synth.mpd_standin(**g["params"])
        playlists = synth.group_by_playlist(d["pid"], d["track_name"])
        columns, off, items = kmls_b200.encode_transactions(list(playlists.values()))
        n_items = len(columns)
    lib = _lib.load()
    t1, tk = ctypes.c_int64(), ctypes.c_int64()
    assert lib.kmls_thresholds(g["min_support"], len(off) - 1, ctypes.addressof(t1), ctypes.addressof(tk)) == 0
    assert (t1.value, tk.value) == (g["t1"], g["tk"])
    d_off, d_items = to_dev(off, items)
    ws = torch.empty(512 << 20, dtype=torch.uint8, device=dev())
    out, opt = _lib.MineResultC(), _lib.MineOptions()
    rc = lib.kmls_mine(d_off.data_ptr(), d_items.data_ptr(), len(off) - 1, n_items, len(off) - 1, t1.value, tk.value,
                       ctypes.addressof(opt), ws.data_ptr(), ws.numel(), ctypes.addressof(out),
                       torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.kmls_last_error()
    sizes = [int(out.level_size[k]) for k in range(1, out.max_k + 1)]
    assert sizes == g["level_counts"][1:]
    assert out.n_itemsets == g["n_itemsets"]
    cs = torch.zeros(1, dtype=torch.int64, device=dev())
    base = ws.data_ptr()
    for k, n in enumerate(sizes, start=1):
        assert lib.kmls_itemset_checksum(base + out.items_off[k], base + out.counts_off[k], n, k,
                                         base + out.row_items_off, cs.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream) == 0
    assert str(int(cs.item()) & ((1 << 64) - 1)) == g["itemset_checksum"]


# ------------------------------------------------------------------ kmls_rules (one call)
@pytest.mark.parametrize("pass1", ["gather", "search"])
def test_rules_one_call_equals_kernel_by_kernel_loop(pass1):
    off, items = synth.zipf_playlists(20000, 2000, mean_draws=35, alpha=1.0, seed=21)
    d_off, d_items = to_dev(off, items)
    res = miner.mine_native(d_off, d_items, 2000, 0.008)
    a = miner.generate_rules(res, 0.3, pass1=pass1)
    b = miner.generate_rules_python(res, 0.3, pass1=pass1)
    assert a.n_rules == b.n_rules > 1000
    assert torch.equal(a.rule, b.rule) and torch.equal(a.confidence, b.confidence) and torch.equal(a.lift, b.lift)
    assert miner.rule_checksum(res, a) == miner.rule_checksum(res, b)
    t1, tk = M.thresholds(0.008, 20000)
    ref = fpgrowth_c.fpgrowth(off, items, 2000, t1, tk, rules_min_confidence=0.3)
    assert a.n_rules == ref.n_rules and miner.rule_checksum(res, a) == ref.rule_checksum
    assert miner.itemset_checksum(res) == ref.checksum


def test_rules_one_call_slices_add_up_and_workspace_grows():
    off, items = synth.zipf_playlists(20000, 2000, mean_draws=35, alpha=1.0, seed=21)
    d_off, d_items = to_dev(off, items)
    res = miner.mine_native(d_off, d_items, 2000, 0.008)
    whole = miner.generate_rules(res, 0.3)
    for pass1, n_parts in (("auto", 3), ("gather", 3), ("auto", 2), ("search", 2)):
        total, cs = 0, 0
        for part in range(n_parts):
            ws = miner.Workspace(dev(), 1 << 16)              # too small: grown to what the call says it needs
            r = miner.generate_rules(res, 0.3, part=(part, n_parts), workspace=ws, pass1=pass1)
            assert ws.nbytes > (1 << 16)
            assert r.stats["pass1"] == {"auto": "gather" if n_parts <= 3 else "search"}.get(pass1, pass1)
            total += r.n_rules
            cs = (cs + miner.rule_checksum(res, r)) & ((1 << 64) - 1)
        assert total == whole.n_rules and cs == miner.rule_checksum(res, whole)
    with pytest.raises(_lib.KmlsError, match="max_rules"):
        miner.generate_rules(res, 0.3, max_rules=10)
    # the bare C call: a capacity error that names the bytes needed
    small = torch.empty(1 << 16, dtype=torch.uint8, device=dev())
    out = _lib.RulesResultC()
    rc, msg = _lib.call_rc("kmls_rules", miner._tables_as_c(res), 0, 0.3, 0, 1, 0, 0, small, small.numel(), out,
                           torch.cuda.current_stream().cuda_stream)
    assert rc == -2 and "workspace too small" in msg


def test_profile_report_names_the_kernels():
    off, items = synth.zipf_playlists(9000, 700, mean_draws=30, alpha=0.9, seed=31)
    d_off, d_items = to_dev(off, items)
    _lib.profile_begin()
    res = miner.mine_native(d_off, d_items, 700, 0.012)
    miner.generate_rules(res, 0.3)
    rec = _lib.profile_report()
    names = {r[0] for r in rec}
    assert {"k_encode_bitmap", "k_expand_units", "k_emit_rows", "k_rules_gather", "k_rules_write"} <= names
    assert names & {"k_count_and", "k_count_project"}
    assert all(ms >= 0 for _, _, ms, _ in rec)
    assert any(nm in ("k_count_and", "k_count_project") and work > 0 for nm, _, _, work in rec)
    assert miner.mine_native(d_off, d_items, 700, 0.012).as_dict() == res.as_dict()   # profiling off again: same result
    # events only (what bench.py's timed steps use): the same launches in the same order, the same
    # results, no data-dependent work figures
    _lib.profile_begin(work=False)
    res2 = miner.mine_native(d_off, d_items, 700, 0.012)
    miner.generate_rules(res2, 0.3)
    rec2 = _lib.profile_report()
    assert [(nm, lv) for nm, lv, _, _ in rec2] == [(nm, lv) for nm, lv, _, _ in rec]
    assert res2.as_dict() == res.as_dict()
    assert all(w2 <= w for (_, _, _, w2), (_, _, _, w) in zip(rec2, rec))


# ---- level 2 hybrid: hot rows' block on the tensor cores, the rest per playlist ------------

def _row_order(off, items, n_items, thr):
    """row_of_item for the tracks with at least thr occurrences, rows in ascending-count order."""
    occ = np.bincount(items, minlength=n_items)
    keep = np.flatnonzero(occ >= thr)
    keep = keep[np.argsort(occ[keep], kind="stable")]
    row_of = np.full(n_items, -1, np.int32)
    row_of[keep] = np.arange(len(keep), dtype=np.int32)
    return row_of, len(keep)


@pytest.mark.parametrize("n_hot", [128, 200, 512])
@pytest.mark.parametrize("stride_extra", [0, 8])
def test_pair_counts_cold_plus_block_equal_csr(n_hot, stride_extra):
    # kmls_pair_counts_csr_cold leaves the hot block zero; kmls_pair_counts_i8_block fills it
    n, t = 40_000, 1500
    off, items = synth.zipf_playlists(n, t, mean_draws=25, alpha=1.0, seed=17)
    row_of, n_rows = _row_order(off, items, t, 40)
    assert n_rows > 600
    d_off, d_items = to_dev(off, items)
    d_row_of = torch.from_numpy(row_of).to(dev())
    s = torch.cuda.current_stream().cuda_stream
    want = torch.empty((n_rows, n_rows), dtype=torch.int32, device=dev())
    ovf = torch.zeros(1, dtype=torch.int32, device=dev())
    _lib.call("kmls_pair_counts_csr", d_off, d_items, n, d_row_of, n_rows, want, ovf, s)
    assert int(ovf.item()) == 0
    got = torch.full((n_rows, n_rows), -1, dtype=torch.int32, device=dev())
    _lib.call("kmls_pair_counts_csr_cold", d_off, d_items, n, d_row_of, n_rows, n_hot, got, ovf, s)
    h0 = n_rows - n_hot
    assert int(got[h0:, h0:].abs().sum().item()) == 0              # the block is left out, and zero
    iu = torch.triu(torch.ones((n_rows, n_rows), dtype=torch.bool, device=dev()))
    outside = iu.clone()
    outside[h0:, h0:] = False
    assert torch.equal(got[outside], want[outside])
    # bitmap of the hot rows with a row stride larger than the words counted
    row_words = int(_lib.value("kmls_bitmap_row_words", n))
    stride = row_words + stride_extra
    bitmap = torch.zeros((n_rows, stride), dtype=torch.int32, device=dev())
    enc = torch.zeros((n_rows, row_words), dtype=torch.int32, device=dev())
    _lib.call("kmls_encode_bitmap", d_off, d_items, n, d_row_of, t, n_rows, enc, row_words, s)
    bitmap[:, :row_words] = enc
    need = int(_lib.value("kmls_pair_gemm_scratch_bytes", n_hot, row_words))
    scratch = torch.empty(need, dtype=torch.uint8, device=dev())
    _lib.call("kmls_pair_counts_i8_block", bitmap[h0:], n_hot, row_words, stride, got[h0:, h0:], n_rows,
              scratch, need, s)
    assert torch.equal(got[iu], want[iu])


@pytest.mark.parametrize("hot_rows", [-1, 128, 300, 4096])
def test_native_loop_hybrid_pair_counts(hot_rows):
    off, items = synth.zipf_playlists(60000, 3000, mean_draws=40, alpha=1.0, seed=5)
    d_off, d_items = to_dev(off, items)
    res = miner.mine_native(d_off, d_items, 3000, 0.004, pair_impl="csr", hot_rows=hot_rows, keep_pair_counts=True)
    assert res.as_dict() == oracle_dict(off, items, 3000, 0.004)
    assert res.stats["pair_impl"] == ("csr" if hot_rows < 0 else "csr+i8")
    base = miner.mine_native(d_off, d_items, 3000, 0.004, pair_impl="popc", keep_pair_counts=True)
    n_rows = base.pair_counts.shape[0]
    iu = torch.triu(torch.ones((n_rows, n_rows), dtype=torch.bool, device=dev()))
    assert torch.equal(res.pair_counts[iu], base.pair_counts[iu])
