"""Independent brute-force frequent-itemset counter — TEST INFRASTRUCTURE ONLY.

Second opinion for ``oracle/mlxtend_restated.py`` (parity is unpinned upstream, see that
file's header): enumerates itemsets level by level straight from the order-free statement
of mlxtend's result set:

    L1 = {i : c(i) >= T1};  L_k = {X subset of L1, |X| = k : c(X) >= Tk}  for k >= 2

with c(X) counted by scanning every transaction. No tree, no shared code with the
restatement except ``thresholds``. Exponential — small cases only.
"""
from __future__ import annotations

import itertools
from typing import Dict, FrozenSet

import numpy as np

from .mlxtend_restated import thresholds


def frequent_itemsets(onehot: np.ndarray, min_support: float, max_len=None
                      ) -> Dict[FrozenSet[int], int]:
    if min_support <= 0.0:
        raise ValueError("min_support must be positive")
    n, t = onehot.shape
    t1, tk = thresholds(min_support, n)
    rows = [frozenset(np.nonzero(onehot[r])[0].tolist()) for r in range(n)]
    out: Dict[FrozenSet[int], int] = {}
    level = []
    for i in range(t):
        c = sum(1 for r in rows if i in r)
        if c >= t1 and c > 0:
            out[frozenset([i])] = c
            level.append((i,))
    k = 1
    l1 = [x[0] for x in level]
    while level and (max_len is None or k < max_len):
        k += 1
        prev = set(level)
        nxt = []
        for combo in itertools.combinations(l1, k):
            # Apriori prune only for k >= 3 (levels >= 2 share the threshold Tk)
            if k >= 3 and any(tuple(s) not in prev for s in itertools.combinations(combo, k - 1)):
                continue
            fs = frozenset(combo)
            c = sum(1 for r in rows if fs <= r)
            if c >= tk:
                out[fs] = c
                nxt.append(combo)
        level = nxt
    return out
