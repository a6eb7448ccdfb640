"""The oracle against the hand-computed vectors, an independent brute force, and itself
(Python restatement vs the C++ restatement). CPU only."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from conftest import load_hand_vectors
from oracle import bruteforce, fpgrowth_c, mlxtend_restated as M

CASES = load_hand_vectors()


def _key(fs):
    return "|".join(sorted(fs))


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_restatement_matches_hand_vectors(case):
    tx = case["transactions"]
    assert M.thresholds(case["min_support"], len(tx)) == (case["t1"], case["tk"])
    te = M.TransactionEncoder().fit(tx)
    onehot = te.transform(tx)
    got = {_key(te.columns_[i] for i in iset): c
           for iset, c in M.fpgrowth_counts(onehot, case["min_support"]).items()}
    assert got == case["itemsets"]
    recs, data = M.calculate_and_save_fp_growth_fast(dict(enumerate(tx)), case["min_support"])
    assert recs == case["recommendations"]
    for s, members in data:
        assert s == case["itemsets"][_key(members)] / len(tx)


@pytest.mark.parametrize("case", CASES, ids=[c["name"] for c in CASES])
def test_cpp_restatement_matches_hand_vectors(case):
    tx = case["transactions"]
    te = M.TransactionEncoder().fit(tx)
    off, items = fpgrowth_c.dense_to_csr(te.transform(tx))
    r = fpgrowth_c.fpgrowth(off, items, len(te.columns_), case["t1"], case["tk"])
    got = {_key(te.columns_[i] for i in iset): c for iset, c in r.as_dict().items()}
    assert got == case["itemsets"]
    assert r.checksum == fpgrowth_c.itemset_checksum(r.offsets, r.items, r.counts)
    assert int(r.level_counts.sum()) == len(case["itemsets"])


def test_rules_hand_vector():
    case = next(c for c in CASES if c["name"] == "triple_with_rules")
    tx = case["transactions"]
    n = len(tx)
    _, data = M.calculate_and_save_fp_growth_fast(dict(enumerate(tx)), case["min_support"])
    rules = M.association_rules(data, case["rules_min_confidence"])
    got = {(_key(a), _key(c)): (s, conf, lift) for a, c, s, conf, lift in rules}
    want = {}
    for r in case["rules"]:
        s_x, s_a, s_c = r["cX"] / n, r["cA"] / n, r["cC"] / n
        conf = s_x / s_a
        want[(r["antecedents"], r["consequents"])] = (s_x, conf, conf / s_c)
    assert got == want


def test_min_support_must_be_positive():
    with pytest.raises(ValueError):
        M.fpgrowth_counts(np.ones((2, 2), bool), 0.0)


def test_threshold_split_values():
    # SURVEY.md F9
    assert M.thresholds(0.07, 100) == (7, 8)
    assert M.thresholds(0.035, 200) == (7, 8)
    assert M.thresholds(0.07, 10_000_000) == (700_000, 700_001)
    assert M.thresholds(0.05, 2246) == (113, 113)
    assert M.thresholds(0.001, 1_000_000) == (1000, 1000)
    for s in np.round(np.arange(0.03, 0.2, 0.0025), 3):
        t1, tk = M.thresholds(float(s), 2246)
        assert t1 == tk


@settings(max_examples=150, deadline=None)
@given(st.integers(1, 48), st.integers(1, 9), st.floats(0.05, 0.95),
       st.sampled_from([0.05, 0.07, 0.1, 0.2, 0.25, 0.3, 0.5, 0.75]), st.integers(0, 2**31 - 1))
def test_restatements_agree_with_brute_force(n, t, density, min_support, seed):
    rng = np.random.default_rng(seed)
    onehot = rng.random((n, t)) < density
    a = M.fpgrowth_counts(onehot, min_support)
    t1, tk = M.thresholds(min_support, n)
    off, items = fpgrowth_c.dense_to_csr(onehot)
    c = fpgrowth_c.fpgrowth(off, items, t, t1, tk).as_dict()
    assert a == c
    b = bruteforce.frequent_itemsets(onehot, min_support)
    if t1 >= tk:
        assert a == b
    else:
        # the only place the restatement may exceed the order-free statement is mlxtend's
        # root single-path shortcut (see hand vector two_thresholds_split_root_single_path_edge)
        assert set(b) <= set(a) and all(a[k] == v for k, v in b.items())
        assert all(a[k] >= t1 for k in set(a) - set(b))


def test_max_len():
    rng = np.random.default_rng(5)
    onehot = rng.random((60, 8)) < 0.5
    full = M.fpgrowth_counts(onehot, 0.1)
    for ml in (1, 2, 3):
        want = {k: v for k, v in full.items() if len(k) <= ml}
        assert M.fpgrowth_counts(onehot, 0.1, max_len=ml) == want
        off, items = fpgrowth_c.dense_to_csr(onehot)
        t1, tk = M.thresholds(0.1, 60)
        assert fpgrowth_c.fpgrowth(off, items, 8, t1, tk, max_len=ml).as_dict() == want
