"""The identities behind the gather route of rule generation (rules.cu, k_rules_gather;
DESIGN.md section 3.1), checked on the CPU with plain dictionaries: every subset count a
level-k itemset needs sits at a computable slot of a level-(k-1) row. This is the algebra
the kernel relies on, restated independently of it (the kernel itself is compared with the
oracle in test_gpu_parity.py)."""
import numpy as np
import pytest

from kmls_b200 import synth
from oracle import fpgrowth_c, mlxtend_restated as M


def bitrev(s, bits):
    r = 0
    for b in range(bits):
        if s >> b & 1:
            r |= 1 << (bits - 1 - b)
    return r


def build_planes(table, counts):
    """planes[k] = (rows, planeA, planeC): rows = sorted k-itemsets (tuples), planes int[n][2^(k-1)]
    in the layout of include/kmls_b200.h: slot s >= 1 = antecedent mask bitrev(s) over the
    first k-1 members, slot 0 = the itemset's own count."""
    planes = {}
    for k, rows in table.items():
        S = 1 << (k - 1)
        pa = np.zeros((len(rows), S), np.int64)
        pc = np.zeros((len(rows), S), np.int64)
        for j, x in enumerate(rows):
            pa[j, 0] = pc[j, 0] = counts[frozenset(x)]
            for s in range(1, S):
                am = bitrev(s, k - 1)
                a = frozenset(x[i] for i in range(k - 1) if am >> i & 1)
                pa[j, s] = counts[a]
                pc[j, s] = counts[frozenset(x) - a]
        planes[k] = (rows, pa, pc)
    return planes


@pytest.mark.parametrize("seed,n,t,s,draws", [(1, 300, 10, 0.05, 8), (2, 500, 12, 0.04, 9), (3, 200, 9, 0.02, 7)])
def test_every_split_count_is_a_slot_of_the_previous_level(seed, n, t, s, draws):
    off, items = synth.zipf_playlists(n, t, mean_draws=draws, alpha=0.3, seed=seed)
    t1, tk = M.thresholds(s, n)
    counts = fpgrowth_c.fpgrowth(off, items, t, t1, tk).as_dict()
    table = {}
    for fs in counts:
        table.setdefault(len(fs), []).append(tuple(sorted(fs)))
    for k in table:
        table[k].sort()
    assert max(table) >= 5
    planes = build_planes(table, counts)
    index = {k: {x: j for j, x in enumerate(rows)} for k, rows in table.items()}
    checked = 0
    for k in sorted(table):
        if k < 2:
            continue
        rows, pa, pc = planes[k]
        _, pa_prev, pc_prev = planes[k - 1]
        S, top = 1 << (k - 1), 1 << (k - 2)          # top = slots per row of level k-1
        for j, x in enumerate(rows):
            row_p = index[k - 1][x[:-1]]             # the parent P = X without its last member
            for sl in range(1, S):
                # count(A): an antecedent of P (even slot) or a consequent of P (odd slot)
                if sl & 1:
                    want_a = pc_prev[row_p, top - 1 - (sl >> 1)]
                else:
                    want_a = pa_prev[row_p, sl >> 1]
                assert pa[j, sl] == want_a, (k, x, sl)
                # count(C): a consequent of X_i = X without p_i, i = k-2-h, h = top bit of the slot
                h = sl.bit_length() - 1
                i = k - 2 - h
                x_i = x[:i] + x[i + 1:]
                assert pc[j, sl] == pc_prev[index[k - 1][x_i], sl - (1 << h)], (k, x, sl)
                checked += 1
    assert checked > 1000


def test_slot_order_is_an_involution_and_keeps_blocks_together():
    for bits in range(1, 11):
        seen = sorted(bitrev(s, bits) for s in range(1 << bits))
        assert seen == list(range(1 << bits))
        assert all(bitrev(bitrev(s, bits), bits) == s for s in range(1 << bits))
