"""Parity of the CUDA path (through the C-ABI) against the CPU oracle, the hand-computed
vectors and the committed golden fixtures. Integer work is compared bit-exact; the fp64
columns (support, confidence, lift) are compared for bit equality as well, which is tighter
than the 1e-12 relative tolerance BASELINE.json allows."""
import json
import os

import numpy as np
import pytest
import torch

import kmls_b200
from kmls_b200 import _lib, synth
from oracle import fpgrowth_c, mlxtend_restated as M
from conftest import ROOT, load_hand_vectors

pytestmark = pytest.mark.gpu
CASES = load_hand_vectors()
GOLDEN = json.load(open(os.path.join(ROOT, "tests", "golden", "generated_cases.json")))["cases"]


def dev():
    return torch.device("cuda", 0)


def stream():
    return torch.cuda.current_stream().cuda_stream


def _key(fs):
    return "|".join(sorted(fs))


def oracle_dict(off, items, n_items, s, max_len=0):
    t1, tk = M.thresholds(s, len(off) - 1)
    return fpgrowth_c.fpgrowth(off, items, n_items, t1, tk, max_len=max_len).as_dict()


def rule_checksum(rules):
    import hashlib
    tot = 0
    for a, c, s, conf, lift in rules:
        key = (tuple(sorted(a)), tuple(sorted(c)), float(s).hex(), float(conf).hex(), float(lift).hex())
        tot = (tot + int.from_bytes(hashlib.sha256(repr(key).encode()).digest()[:8], "little")) % (1 << 64)
    return tot


# ------------------------------------------------------------------------- kernels, one by one
def test_item_histogram():
    rng = np.random.default_rng(0)
    for nnz, t in [(0, 5), (1, 1), (7, 3), (100003, 977), (1 << 20, 50000)]:
        items = rng.integers(0, t, nnz).astype(np.int32)
        d = torch.from_numpy(items).to(dev())
        occ = torch.full((t,), -7, dtype=torch.int32, device=dev())
        _lib.call("kmls_item_histogram", d, nnz, t, occ, stream())
        assert (occ.cpu().numpy() == np.bincount(items, minlength=t)).all()


@pytest.mark.parametrize("n,t,dup", [(1, 1, False), (255, 40, True), (256, 40, False), (257, 9000, True),
                                     (5000, 300, True), (70001, 64, False),
                                     (600, 20000, False)])   # 10,000 rows: two row tiles in shared memory
def test_encode_popcount_roundtrip(n, t, dup):
    rng = np.random.default_rng(n)
    lens = rng.integers(0, min(t, 60) + 1, n)
    rows = [rng.choice(t, l, replace=False) for l in lens]
    if dup:
        rows = [np.concatenate([r, r[:2]]) for r in rows]
    off = np.zeros(n + 1, np.int64)
    np.cumsum([len(r) for r in rows], out=off[1:])
    items = (np.concatenate(rows) if off[-1] else np.zeros(0)).astype(np.int32)
    # keep every second track, in a scrambled row order
    kept = rng.permutation(np.arange(0, t, 2))
    row_of = np.full(t, -1, np.int32)
    row_of[kept] = np.arange(len(kept), dtype=np.int32)
    W = int(_lib.value("kmls_bitmap_row_words", n))
    bm = torch.full((len(kept) * W,), 0x5A5A5A5A, dtype=torch.int32, device=dev())
    _lib.call("kmls_encode_bitmap", torch.from_numpy(off).to(dev()), torch.from_numpy(items).to(dev()), n,
              torch.from_numpy(row_of).to(dev()), t, len(kept), bm, W, stream())
    dense = np.zeros((len(kept), W * 32), bool)
    for p, r in enumerate(rows):
        for it in r:
            if row_of[it] >= 0:
                dense[row_of[it], p] = True
    want = np.packbits(dense, axis=1, bitorder="little").view(np.uint32).reshape(len(kept), W)
    got = bm.cpu().numpy().view(np.uint32).reshape(len(kept), W)
    assert (got == want).all()
    cnt = torch.empty(len(kept), dtype=torch.int32, device=dev())
    _lib.call("kmls_row_popcount", bm, len(kept), W, cnt, stream())
    assert (cnt.cpu().numpy() == dense.sum(1)).all()
    # gather rows back into ascending-track order
    src = torch.from_numpy(np.argsort(kept).astype(np.int32)).to(dev())
    out = torch.empty_like(bm)
    _lib.call("kmls_gather_rows", bm, src, len(kept), W, out, stream())
    assert (out.cpu().numpy().view(np.uint32).reshape(len(kept), W) == want[np.argsort(kept)]).all()


def test_row_popcount_long_rows():
    rng = np.random.default_rng(1)
    W = 8 * 700  # > 4096 words: one CTA per row
    bm = rng.integers(0, 2**32, (5, W), dtype=np.uint64).astype(np.uint32)
    d = torch.from_numpy(bm.view(np.int32)).to(dev())
    cnt = torch.empty(5, dtype=torch.int32, device=dev())
    _lib.call("kmls_row_popcount", d, 5, W, cnt, stream())
    want = np.unpackbits(bm.view(np.uint8), axis=1).sum(1)
    assert (cnt.cpu().numpy() == want).all()


@pytest.mark.parametrize("rows,n", [(1, 100), (63, 300), (64, 256), (65, 5000), (200, 70000), (3, 400000)])
def test_pair_counts_and_select(rows, n):
    rng = np.random.default_rng(rows * 7 + n)
    W = int(_lib.value("kmls_bitmap_row_words", n))
    dense = rng.random((rows, n)) < rng.uniform(0.02, 0.6, (rows, 1))
    full = np.zeros((rows, W * 32), bool)
    full[:, :n] = dense
    bm = np.packbits(full, axis=1, bitorder="little").view(np.int32).reshape(rows, W)
    d = torch.from_numpy(bm).to(dev())
    pair = torch.full((rows * rows,), -1, dtype=torch.int32, device=dev())
    _lib.call("kmls_pair_counts", d, rows, W, pair, stream())
    want = dense.astype(np.int64) @ dense.astype(np.int64).T
    got = pair.cpu().numpy().reshape(rows, rows)
    assert (got == want).all()
    thr = int(np.median(want[np.triu_indices(rows, 1)])) if rows > 1 else 1
    row_off = torch.empty(rows + 1, dtype=torch.int64, device=dev())
    _lib.call("kmls_pair_select", pair, rows, thr, 0, row_off, 0, 0, stream())
    ii, jj = np.nonzero(np.triu(want >= thr, 1))
    assert int(row_off[rows].item()) == len(ii)
    if len(ii):
        it2 = torch.empty(len(ii) * 2, dtype=torch.int32, device=dev())
        ct2 = torch.empty(len(ii), dtype=torch.int32, device=dev())
        _lib.call("kmls_pair_select", pair, rows, thr, len(ii), row_off, it2, ct2, stream())
        assert (it2.cpu().numpy().reshape(-1, 2) == np.stack([ii, jj], 1)).all()
        assert (ct2.cpu().numpy() == want[ii, jj]).all()
        sup = torch.empty(len(ii), dtype=torch.float64, device=dev())
        _lib.call("kmls_supports_f64", ct2, len(ii), n, sup, stream())
        assert sup.cpu().numpy().tolist() == [int(c) / n for c in want[ii, jj]]


@pytest.mark.parametrize("variant", ["ss", "ts"])
@pytest.mark.parametrize("rows,n", [(64, 256), (129, 3000), (300, 70000), (1000, 200000), (257, 1024)])
def test_pair_counts_tensor_core_matches_cuda_core(rows, n, variant, monkeypatch):
    """tcgen05 int8 GEMM (A operand from shared memory / from tensor memory) vs the AND+popc
    kernel: bit-equal on the upper triangle."""
    monkeypatch.setenv("KMLS_PAIR_GEMM", variant)
    rng = np.random.default_rng(rows + n)
    W = int(_lib.value("kmls_bitmap_row_words", n))
    dense = rng.random((rows, n)) < rng.uniform(0.01, 0.7, (rows, 1))
    full = np.zeros((rows, W * 32), bool)
    full[:, :n] = dense
    bm = torch.from_numpy(np.packbits(full, axis=1, bitorder="little").view(np.int32).reshape(rows, W)).to(dev())
    ref = torch.empty(rows * rows, dtype=torch.int32, device=dev())
    _lib.call("kmls_pair_counts", bm, rows, W, ref, stream())
    got = torch.full((rows * rows,), -5, dtype=torch.int32, device=dev())
    sb = int(_lib.value("kmls_pair_gemm_scratch_bytes", rows, W))
    scratch = torch.empty(sb // 8 + 1, dtype=torch.int64, device=dev())
    _lib.call("kmls_pair_counts_i8", bm, rows, W, got, scratch, sb, stream())
    torch.cuda.synchronize()
    iu = torch.triu(torch.ones(rows, rows, dtype=torch.bool, device=dev()))
    assert torch.equal(got.view(rows, rows)[iu], ref.view(rows, rows)[iu])


@pytest.mark.parametrize("n,t,kept,max_len,dup", [(1, 5, 5, 4, False), (300, 50, 37, 30, True), (5000, 900, 700, 80, True),
                                                  (20000, 3000, 2500, 120, False)])
def test_pair_counts_from_playlists_match_bitmap_counts(n, t, kept, max_len, dup):
    """Sparse route (one counter update per pair of a playlist) vs the AND+popc kernel on the
    bitmap of the same playlists: equal on the upper triangle and the diagonal; repeated
    tracks count once; tracks without a row are ignored."""
    rng = np.random.default_rng(n + t)
    lens = rng.integers(0, max_len + 1, n)
    rows_ = [rng.choice(t, min(l, t), replace=False) for l in lens]
    if dup:
        rows_ = [np.concatenate([r, r[:3], r[:1]]) for r in rows_]
    off = np.zeros(n + 1, np.int64)
    np.cumsum([len(r) for r in rows_], out=off[1:])
    items = (np.concatenate(rows_) if off[-1] else np.zeros(0)).astype(np.int32)
    keep = rng.permutation(t)[:kept]
    row_of = np.full(t, -1, np.int32)
    row_of[keep] = np.arange(kept, dtype=np.int32)
    d_off, d_items = torch.from_numpy(off).to(dev()), torch.from_numpy(items).to(dev())
    d_row_of = torch.from_numpy(row_of).to(dev())
    W = int(_lib.value("kmls_bitmap_row_words", n))
    bm = torch.empty(kept * W, dtype=torch.int32, device=dev())
    _lib.call("kmls_encode_bitmap", d_off, d_items, n, d_row_of, t, kept, bm, W, stream())
    ref = torch.empty(kept * kept, dtype=torch.int32, device=dev())
    _lib.call("kmls_pair_counts", bm, kept, W, ref, stream())
    got = torch.full((kept * kept,), -3, dtype=torch.int32, device=dev())
    ov = torch.full((1,), 7, dtype=torch.int32, device=dev())
    _lib.call("kmls_pair_counts_csr", d_off, d_items, n, d_row_of, kept, got, ov, stream())
    assert int(ov.item()) == 0
    iu = torch.triu(torch.ones(kept, kept, dtype=torch.bool, device=dev()))
    assert torch.equal(got.view(kept, kept)[iu], ref.view(kept, kept)[iu])


def test_pair_counts_from_playlists_flag_overlong_playlists():
    # one playlist with 1500 distinct kept tracks (> 1024): flagged and skipped, the others counted
    t = 2000
    lists = [np.arange(1500), np.array([1, 2, 3, 2]), np.array([3, 1])]
    off = np.zeros(4, np.int64)
    np.cumsum([len(r) for r in lists], out=off[1:])
    items = np.concatenate(lists).astype(np.int32)
    d_row_of = torch.arange(t, dtype=torch.int32, device=dev())
    got = torch.empty(t * t, dtype=torch.int32, device=dev())
    ov = torch.zeros(1, dtype=torch.int32, device=dev())
    _lib.call("kmls_pair_counts_csr", torch.from_numpy(off).to(dev()), torch.from_numpy(items).to(dev()), 3,
              d_row_of, t, got, ov, stream())
    assert int(ov.item()) == 1
    g = got.view(t, t)
    assert int(g[1, 3]) == 2 and int(g[1, 2]) == 1 and int(g[2, 3]) == 1 and int(g[3, 3]) == 2 and int(g.sum()) == 9
    # and the miner falls back to a dense kernel for such an input
    off2, items2 = synth.zipf_playlists(3000, 1400, mean_draws=20, alpha=0.6, seed=9)
    off2 = np.concatenate([off2, [off2[-1] + 1300]])
    items2 = np.concatenate([items2, np.arange(1300, dtype=np.int32)])
    want = oracle_dict(off2, items2, 1400, 0.004)
    res, _ = kmls_b200.mine_ids(off2, items2, 1400, 0.004, pair_impl="csr")
    assert res.as_dict() == want


def test_exclusive_scan():
    rng = np.random.default_rng(2)
    for n in [0, 1, 31, 1025, 3_000_001]:
        x = rng.integers(0, 1000, n).astype(np.int32)
        d = torch.from_numpy(x).to(dev())
        out = torch.empty(n + 1, dtype=torch.int64, device=dev())
        sb = int(_lib.value("kmls_scan_scratch_bytes", n))
        scratch = torch.empty(sb, dtype=torch.uint8, device=dev())
        _lib.call("kmls_exclusive_scan_i32", d, n, out, scratch, sb, stream())
        want = np.concatenate([[0], np.cumsum(x.astype(np.int64))])
        assert (out.cpu().numpy() == want).all()
    # totals beyond 2^31 must not wrap
    x = torch.full((3_000_000,), 1000, dtype=torch.int32, device=dev())
    out = torch.empty(3_000_001, dtype=torch.int64, device=dev())
    _lib.call("kmls_exclusive_scan_i32", x, 3_000_000, out, scratch, sb, stream())
    assert int(out[-1].item()) == 3_000_000_000


def test_project_rows_and_count_pairs():
    """Projection = numpy boolean compress; pair count = popcount of the AND."""
    rng = np.random.default_rng(3)
    n_cols, n_rows = 5000, 40
    W = int(_lib.value("kmls_bitmap_row_words", n_cols))
    dense = rng.random((n_rows, n_cols)) < rng.uniform(0.01, 0.9, (n_rows, 1))
    dense[7] = False                      # an empty mask row
    dense[8] = True                       # a full mask row
    full = np.zeros((n_rows, W * 32), bool)
    full[:, :n_cols] = dense
    bm = torch.from_numpy(np.packbits(full, axis=1, bitorder="little").view(np.int32).reshape(-1)).to(dev())
    rowptr1 = torch.arange(n_rows, dtype=torch.int64, device=dev()) * W
    roww1 = torch.full((n_rows,), W, dtype=torch.int32, device=dev())
    cnt1 = torch.from_numpy(dense.sum(1).astype(np.int32)).to(dev())
    # new "itemsets": (p, q) pairs sorted by p; p = 20 appears once only (gets no row)
    # (runs of 3 with the same p, cut by the kernel's chunks of 8; p = 13: a run of 19, longer
    # than a chunk)
    pq = [(p, q) for p in range(0, 12) for q in range(p + 1, min(p + 4, n_rows))] + \
        [(13, q) for q in range(14, 33)] + [(20, 30)]
    par_p = torch.tensor([a for a, _ in pq], dtype=torch.int32, device=dev())
    par_q = torch.tensor([b for _, b in pq], dtype=torch.int32, device=dev())
    n = len(pq)
    roww = torch.empty(n, dtype=torch.int32, device=dev())
    _lib.call("kmls_row_alloc", par_p, n, cnt1, 0, 0, roww, stream())
    want_w = [(((int(dense[a].sum()) + 31) // 32 + 3) // 4 * 4) if a != 20 else 0 for a, _ in pq]
    assert roww.cpu().tolist() == want_w
    sb = int(_lib.value("kmls_scan_scratch_bytes", n))
    scratch = torch.empty(sb, dtype=torch.uint8, device=dev())
    rowptr = torch.empty(n + 1, dtype=torch.int64, device=dev())
    _lib.call("kmls_exclusive_scan_i32", roww, n, rowptr, scratch, sb, stream())
    words = int(rowptr[n].item())
    arena = torch.full((max(words, 1),), 0x7BADBEEF, dtype=torch.int32, device=dev())
    _lib.call("kmls_project_rows", bm, rowptr1, roww1, par_p, par_q, n, arena, rowptr, roww, 0, stream())
    got = arena.cpu().numpy().view(np.uint32)
    rp = rowptr.cpu().numpy()
    for i, (a, b) in enumerate(pq):
        if want_w[i] == 0:
            continue
        bits = np.zeros(want_w[i] * 32, bool)
        sel = dense[b][dense[a]]
        bits[:len(sel)] = sel
        want = np.packbits(bits, bitorder="little").view(np.uint32)
        assert (got[rp[i]:rp[i] + want_w[i]] == want).all(), (i, a, b)
    # candidates: consecutive members of the same class
    cand = [(i, j) for i in range(n) for j in range(i + 1, n) if pq[i][0] == pq[j][0] and want_w[i]]
    cp = torch.tensor([i for i, _ in cand], dtype=torch.int32, device=dev())
    cq = torch.tensor([j for _, j in cand], dtype=torch.int32, device=dev())
    cc = torch.empty(len(cand), dtype=torch.int32, device=dev())
    _lib.call("kmls_count_pairs", arena, rowptr, roww, cp, cq, len(cand), cc, stream())
    want_c = [int((dense[pq[i][0]] & dense[pq[i][1]] & dense[pq[j][1]]).sum()) for i, j in cand]
    assert cc.cpu().tolist() == want_c


@pytest.mark.parametrize("count_impl,prune,first,keep", [
    ("projected", True, "gather", None), ("projected", False, "gather", 0), ("projected", False, "pext", 8),
    ("projected-csr", False, "gather", None), ("projected-i8", False, "gather", None),
    ("projected", False, "gather", 100000), ("fullwidth", True, "gather", None),
    ("fullwidth", False, "gather", None)])
def test_counting_variants_agree(count_impl, prune, first, keep):
    off, items = synth.zipf_playlists(9000, 700, mean_draws=30, alpha=0.9, seed=31)
    want = oracle_dict(off, items, 700, 0.012)
    count_impl, _, pair_impl = count_impl.partition("-")
    res, _ = kmls_b200.mine_ids(off, items, 700, 0.012, count_impl=count_impl, prune=prune,
                                first_projection=first, keep_words=keep, pair_impl=pair_impl or "auto")
    assert res.as_dict() == want
    assert len(res.levels) >= 5


# ------------------------------------------------------------------- the whole path, small cases
@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_hand_vectors_through_the_plugin_boundary(case, capsys):
    tx = case["transactions"]
    playlist_dict = {100 + i: t for i, t in enumerate(tx)}
    recs, info, (missing, duration) = kmls_b200.calculate_and_save_fp_growth_fast(
        playlist_dict, case["min_support"], total_songs=10)
    assert recs == case["recommendations"]
    assert all(type(v) is float for d in recs.values() for v in d.values())
    assert missing == 10 - len(recs)
    out = capsys.readouterr().out
    assert "Songs without recommendations:" in out and "Time elapsed in rule generation:" in out
    assert info.startswith(f"min_support: {case['min_support']} \tmissing songs: {missing} \ttime: ")
    last = kmls_b200.api.last_result
    cols = last["columns"]
    got = {_key(cols[i] for i in k): v for k, v in last["itemsets"].as_dict().items()}
    assert got == case["itemsets"]
    frame = kmls_b200.itemsets_frame(last["itemsets"], cols)
    assert {_key(r.itemsets): r.support for r in frame.itertuples()} == \
        {k: v / len(tx) for k, v in case["itemsets"].items()}
    if "rules" in case:
        res, rules = kmls_b200.mine_ids(*kmls_b200.encode_transactions(tx)[1:], len(cols),
                                        case["min_support"], case["rules_min_confidence"])
        rf = kmls_b200.rules_frame(res, rules, cols)
        n = len(tx)
        want = {}
        for r in case["rules"]:
            sx, sa, sc = r["cX"] / n, r["cA"] / n, r["cC"] / n
            want[(r["antecedents"], r["consequents"])] = (sx, sx / sa, (sx / sa) / sc)
        got = {(_key(r.antecedents), _key(r.consequents)): (r.support, r.confidence, r.lift)
               for r in rf.itertuples()}
        assert got == want


def test_empty_and_degenerate_inputs():
    assert kmls_b200.calculate_and_save_fp_growth_fast({}, 0.5)[0] == {}
    assert kmls_b200.calculate_and_save_fp_growth_fast({1: [], 2: []}, 0.5)[0] == {}
    assert kmls_b200.calculate_and_save_fp_growth_fast({1: ["only"]}, 0.5)[0] == {"only": {}}
    with pytest.raises(ValueError):
        kmls_b200.calculate_and_save_fp_growth_fast({1: ["a"]}, 0.0)
    with pytest.raises(ValueError):
        kmls_b200.calculate_and_save_fp_growth_fast({1: ["a"]}, -1.0)


@pytest.mark.parametrize("seed", range(12))
def test_random_small_against_python_restatement(seed):
    rng = np.random.default_rng(seed)
    n, t = int(rng.integers(1, 400)), int(rng.integers(1, 24))
    onehot = rng.random((n, t)) < rng.uniform(0.05, 0.7)
    s = float(rng.choice([0.05, 0.07, 0.1, 0.2, 0.3, 0.035, 0.14]))
    want = M.fpgrowth_counts(onehot, s)
    off, items = fpgrowth_c.dense_to_csr(onehot)
    res, _ = kmls_b200.mine_ids(off, items, t, s)
    assert res.as_dict() == want


def test_duplicates_inside_playlists_change_supports_not_results():
    rng = np.random.default_rng(5)
    n, t = 3000, 120
    onehot = rng.random((n, t)) < 0.08
    off, items = fpgrowth_c.dense_to_csr(onehot)
    # repeat every playlist's items three times, scrambled
    rows = [np.tile(items[off[i]:off[i + 1]], 3) for i in range(n)]
    rows = [rng.permutation(r) for r in rows]
    off3 = np.zeros(n + 1, np.int64)
    np.cumsum([len(r) for r in rows], out=off3[1:])
    items3 = np.concatenate(rows).astype(np.int32)
    s = 0.08
    want = M.fpgrowth_counts(onehot, s)
    res, _ = kmls_b200.mine_ids(off3, items3, t, s)
    assert res.as_dict() == want


def test_max_len():
    off, items = synth.zipf_playlists(2000, 200, mean_draws=20, alpha=0.8, seed=4)
    for ml in (1, 2, 3):
        res, _ = kmls_b200.mine_ids(off, items, 200, 0.03, max_len=ml)
        assert res.as_dict() == oracle_dict(off, items, 200, 0.03, max_len=ml)


# ----------------------------------------------------------------- golden fixtures (oracle-made)
def _golden_input(g):
    if g["generator"] == "zipf":
        off, items = synth.zipf_playlists(**g["params"])
        return off, items, g["params"]["n_tracks"]
    d = synth.mpd_standin(**g["params"])
    playlists = synth.group_by_playlist(d["pid"], d["track_name"])
    columns, off, ids = kmls_b200.encode_transactions(list(playlists.values()))
    return off, ids, len(columns)


@pytest.mark.parametrize("g", GOLDEN, ids=[g["name"] for g in GOLDEN])
def test_golden_fixture(g):
    off, items, n_items = _golden_input(g)
    assert len(off) - 1 == g["n_playlists"] and len(items) == g["nnz"]
    res, rules = kmls_b200.mine_ids(off, items, n_items, g["min_support"], g["min_confidence"])
    assert (res.t1, res.tk) == (g["t1"], g["tk"])
    assert [0] + [int(c.numel()) for _, c in res.levels] == g["level_counts"]
    o, i, c = res.to_host()
    assert str(fpgrowth_c.itemset_checksum(o, i, c)) == g["itemset_checksum"]
    assert rules.n_rules == g["n_rules"]
    if g["n_rules"] <= 250_000:
        assert str(rule_checksum(kmls_b200.rules_to_host(res, rules))) == g["rule_checksum"]
    # and the live oracle on the same input, itemset by itemset
    assert res.as_dict() == oracle_dict(off, items, n_items, g["min_support"])


def test_mpd_standin_recommendations_match_reference_fold():
    d = synth.mpd_standin(seed=1)
    playlists = synth.group_by_playlist(d["pid"], d["track_name"])
    total = len(set(d["track_uri"]))
    recs, _, (missing, _) = kmls_b200.calculate_and_save_fp_growth_fast(playlists, 0.05, total_songs=total)
    want, _ = M.calculate_and_save_fp_growth_fast(playlists, 0.05)
    assert recs == want
    assert missing == total - len(want)
    # plain picklable objects only (rest_api has neither numpy nor torch)
    import pickle
    assert pickle.loads(pickle.dumps(recs)) == want


def test_rules_against_oracle_medium():
    off, items = synth.zipf_playlists(6000, 800, mean_draws=30, alpha=0.9, seed=21)
    s, conf = 0.015, 0.35
    res, rules = kmls_b200.mine_ids(off, items, 800, s, conf)
    n = len(off) - 1
    want_sets = oracle_dict(off, items, 800, s)
    assert res.as_dict() == want_sets
    want = M.association_rules([(v / n, k) for k, v in want_sets.items()], conf)
    got = kmls_b200.rules_to_host(res, rules)
    assert len(got) == len(want)
    assert {(a, c): (x, y, z) for a, c, x, y, z in got} == {(a, c): (x, y, z) for a, c, x, y, z in want}
    # the device-side checksum of the whole table = the C++ oracle's = the formula applied on the host
    t1, tk = M.thresholds(s, n)
    ref = fpgrowth_c.fpgrowth(off, items, 800, t1, tk, rules_min_confidence=conf)
    assert kmls_b200.rule_checksum(res, rules) == ref.rule_checksum == fpgrowth_c.rule_checksum(want)


@pytest.mark.parametrize("n", [97, 127, 255, 1021, 4099, 16381])
def test_rule_doubles_bit_exact_for_awkward_playlist_counts(n):
    """Pass 2 replaces the six IEEE divisions per split by reciprocal + FMA corrections
    (rules.cu, div_by_rcp); every confidence and lift must still be the correctly rounded
    quotient the oracle gets from the CPU's divider. Odd / prime / 2^k-1 playlist counts give
    1/N the largest rounding errors."""
    off, items = synth.zipf_playlists(n, 24, mean_draws=12, alpha=0.3, seed=n)
    s, conf = 0.05, 0.05
    res, rules = kmls_b200.mine_ids(off, items, 24, s, conf)
    t1, tk = M.thresholds(s, n)
    ref = fpgrowth_c.fpgrowth(off, items, 24, t1, tk, rules_min_confidence=conf)
    assert res.n_itemsets == ref.n and rules.n_rules == ref.n_rules > 5000
    assert kmls_b200.rule_checksum(res, rules) == ref.rule_checksum
    if n <= 255:
        want = M.association_rules([(v / n, k) for k, v in ref.as_dict().items()], conf)
        got = kmls_b200.rules_to_host(res, rules)
        assert {(a, c): (x, y, z) for a, c, x, y, z in got} == {(a, c): (x, y, z) for a, c, x, y, z in want}


@pytest.mark.parametrize("n", [3, 97, 255, 65535, 999983, 1000000, 2262292, 10000000, 2**31 - 1])
def test_rule_division_selftest(n):
    """2^28 pseudo-random splits per playlist count: the reciprocal + FMA quotients of pass 2
    are bit-identical to those of __ddiv_rn (s_a and s_c are covered through the confidences
    and lifts computed from them)."""
    out = torch.ones(1, dtype=torch.int64, device=dev())
    _lib.call("kmls_rule_division_selftest", n, 1 << 28, 12345 + n, out, stream())
    assert int(out.item()) == 0


def _rule_rows(*tables):
    """All rules of the given tables as a lexicographically sorted int64[R, 4] array
    (itemset, antecedent mask, bits of confidence, bits of lift)."""
    rows = np.concatenate([np.stack([r.rule[:, 0].cpu().numpy().astype(np.int64),
                                     r.rule[:, 1].cpu().numpy().astype(np.int64),
                                     r.confidence.cpu().numpy().view(np.int64),
                                     r.lift.cpu().numpy().view(np.int64)], 1) for r in tables])
    return rows[np.lexsort(rows.T[::-1])]


def test_pass1_gather_route_equals_index_search():
    """Pass 1 has two routes: subset counts gathered from the previous level's output (one GPU,
    whole levels) or searched in the subset index (any slice of the table, what the ranks of a
    multi-GPU job run). Both, and slices that cut levels at arbitrary places, give the same
    table, which equals the oracle's."""
    n, t, s, conf = 2500, 22, 0.02, 0.3
    off, items = synth.zipf_playlists(n, t, mean_draws=19, alpha=0.2, seed=5)
    res = kmls_b200.mine_csr(torch.from_numpy(off).to(dev()), torch.from_numpy(items).to(dev()), t, s)
    assert len(res.levels) >= 8                      # k = 6, 7, 8: the long-itemset path, chained
    full = kmls_b200.generate_rules(res, conf)
    search = kmls_b200.generate_rules(res, conf, pass1="search")
    want = _rule_rows(full)
    assert len(want) == full.n_rules > 1000000
    assert (_rule_rows(search) == want).all()
    t1, tk = M.thresholds(s, n)
    ref = fpgrowth_c.fpgrowth(off, items, t, t1, tk, rules_min_confidence=conf)
    assert full.n_rules == ref.n_rules
    assert kmls_b200.rule_checksum(res, full) == ref.rule_checksum == kmls_b200.rule_checksum(res, search)
    for n_parts in (2, 3, 7):
        parts = [kmls_b200.generate_rules(res, conf, part=(i, n_parts)) for i in range(n_parts)]
        assert sum(p.n_rules for p in parts) == full.n_rules
        assert (_rule_rows(*parts) == want).all()


def test_two_threshold_split_at_scale():
    # s = 0.07, N = 100k: T1 and Tk differ for many N; counts near both must
    # land on the right side at level 1 (T1) and at levels >= 2 (Tk)
    n = 100
    tx = [["p", "q"]] * 7 + [["r", "q"]] * 8 + [["r"]] + [["q"]] * 35 + [[]] * 49
    assert kmls_b200.thresholds(0.07, n) == (7, 8)
    recs, _, _ = kmls_b200.calculate_and_save_fp_growth_fast(dict(enumerate(tx)), 0.07, total_songs=3)
    assert recs == {"p": {}, "q": {"r": 0.08}, "r": {"q": 0.08}}


# --------------------------------------------------------------- full-size, size-independent checks
def test_properties_at_scale():
    """200k playlists x 20k tracks: (a) level-1 supports equal a numpy bincount of the
    deduplicated input, (b) every itemset's count is <= the count of each of its parents'
    (anti-monotone), (c) level tables are sorted and duplicate-free, (d) the checksum of the
    whole table equals the C++ oracle's."""
    n, t, s = 200_000, 20_000, 0.004
    off, items = synth.zipf_playlists(n, t, mean_draws=50, alpha=1.0, seed=7)
    res, rules = kmls_b200.mine_ids(off, items, t, s, 0.8)
    rows = res.row_items.cpu().numpy()
    c1 = np.bincount(items, minlength=t)  # generator output has no in-playlist duplicates
    assert (res.levels[0][1].cpu().numpy() == c1[rows]).all()
    assert set(rows.tolist()) == set(np.nonzero(c1 >= res.t1)[0].tolist())
    prev = None
    for k, (it, ct) in enumerate(res.levels, start=1):
        it, ct = it.cpu().numpy().astype(np.int64), ct.cpu().numpy()
        assert (ct >= (res.t1 if k == 1 else res.tk)).all()
        if k > 1:
            assert (np.diff(it, axis=1) > 0).all()
            key = np.zeros(len(it), dtype=object)
            order = np.lexsort(it.T[::-1])
            assert (order == np.arange(len(it))).all(), "level table not sorted"
            assert len(np.unique(it, axis=0)) == len(it)
            # parent = itemset minus its last member, must exist with count >= child's
            parent = {tuple(r): c for r, c in zip(prev[0].tolist(), prev[1].tolist())}
            for r, c in list(zip(it.tolist(), ct.tolist()))[:: max(1, len(it) // 5000)]:
                assert parent[tuple(r[:-1])] >= c
        prev = (it, ct)
    t1, tk = M.thresholds(s, n)
    ref = fpgrowth_c.fpgrowth(off, items, t, t1, tk, store=False)
    o, i, c = res.to_host()
    assert res.n_itemsets == ref.n
    assert [int(x.numel()) for _, x in res.levels] == ref.level_counts[1:len(res.levels) + 1].tolist()
    assert fpgrowth_c.itemset_checksum(o, i, c) == ref.checksum
    # rules: confidence >= threshold, support/lift consistent with the integer counts
    conf = rules.confidence.cpu().numpy()
    assert (conf >= 0.8).all() and (conf <= 1.0).all()


def test_guards_fail_loudly_instead_of_truncating():
    off, items = synth.zipf_playlists(4000, 300, mean_draws=25, alpha=0.8, seed=2)
    with pytest.raises(_lib.KmlsError, match="max_itemsets"):
        kmls_b200.mine_ids(off, items, 300, 0.01, max_itemsets=500)
    res, _ = kmls_b200.mine_ids(off, items, 300, 0.01)
    assert res.n_itemsets > 500
    with pytest.raises(_lib.KmlsError, match="max_rules"):
        kmls_b200.generate_rules(res, 0.1, max_rules=10)
