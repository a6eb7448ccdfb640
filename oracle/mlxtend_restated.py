"""CPU oracle for the rule-generation hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may import
this module. The product path (the package next to this directory) never does: it fails
loudly when its CUDA library is missing.

PARITY UNPINNED UPSTREAM. The arithmetic of the reference's hot path lives in two
third-party packages that are neither vendored in /root/reference nor importable in this
image (no network): ``mlxtend==0.23.3`` (machine-learning/uv.lock:302-303) and, for the
dormant path, ``fpgrowth-py==1.0.0`` (uv.lock:124-125). The reference ships no tests and
no golden vectors. This file therefore restates the *published*
mlxtend 0.23.3 algorithm, anchored on the reference's call sites
(machine-learning/main.py:267-272) and on the glue the reference owns
(main.py:262-313). Its own pins are the hand-computed vectors in ``tests/golden`` and an
independent brute-force counter (``oracle/bruteforce.py``).

Each function cites the reference line (or the mlxtend symbol) it follows.
"""
from __future__ import annotations

import collections
import itertools
import math
from typing import Dict, FrozenSet, Iterable, List, Sequence, Tuple

import numpy as np


# --------------------------------------------------------------------------------------
# mlxtend.preprocessing.TransactionEncoder  (call site: main.py:267-268)
# --------------------------------------------------------------------------------------
class TransactionEncoder:
    """``fit``: columns_ = sorted(set of all items). ``transform``: dense bool[N, T];
    an item repeated inside a transaction sets its cell once."""

    def fit(self, X: Sequence[Sequence]):
        unique = set()
        for tx in X:
            for item in tx:
                unique.add(item)
        self.columns_ = sorted(unique)
        self.columns_mapping_ = {item: i for i, item in enumerate(self.columns_)}
        return self

    def transform(self, X: Sequence[Sequence]) -> np.ndarray:
        array = np.zeros((len(X), len(self.columns_)), dtype=bool)
        for row, tx in enumerate(X):
            for item in tx:
                if item in self.columns_mapping_:
                    array[row, self.columns_mapping_[item]] = True
        return array


# --------------------------------------------------------------------------------------
# mlxtend.frequent_patterns.fpcommon: FPNode / FPTree / setup_fptree
# --------------------------------------------------------------------------------------
class _FPNode:
    __slots__ = ("item", "count", "parent", "children")

    def __init__(self, item, count=0, parent=None):
        self.item = item
        self.count = count
        self.parent = parent
        self.children = {}
        if parent is not None:
            parent.children[item] = self

    def itempath_from_root(self):
        path = []
        if self.item is None:
            return path
        node = self.parent
        while node.item is not None:
            path.append(node.item)
            node = node.parent
        path.reverse()
        return path


class _FPTree:
    def __init__(self, rank=None):
        self.root = _FPNode(None)
        self.nodes = collections.defaultdict(list)
        self.cond_items = []
        self.rank = rank

    def insert_itemset(self, itemset, count=1):
        self.root.count += count
        if len(itemset) == 0:
            return
        index = 0
        node = self.root
        for item in itemset:
            child = node.children.get(item)
            if child is None:
                break
            child.count += count
            node = child
            index += 1
        for item in itemset[index:]:
            child = _FPNode(item, count, node)
            self.nodes[item].append(child)
            node = child

    def is_path(self):
        if len(self.root.children) > 1:
            return False
        for i in self.nodes:
            if len(self.nodes[i]) > 1 or len(self.nodes[i][0].children) > 1:
                return False
        return True

    def conditional_tree(self, cond_item, minsup):
        """Integer test ``count >= minsup`` (the Tk test)."""
        branches = []
        count = collections.defaultdict(int)
        for node in self.nodes[cond_item]:
            branch = node.itempath_from_root()
            branches.append(branch)
            for item in branch:
                count[item] += node.count
        items = [item for item in count if count[item] >= minsup]
        items.sort(key=count.get)
        rank = {item: i for i, item in enumerate(items)}
        cond_tree = _FPTree(rank)
        for idx, branch in enumerate(branches):
            branch = sorted([i for i in branch if i in rank], key=rank.get, reverse=True)
            cond_tree.insert_itemset(branch, self.nodes[cond_item][idx].count)
        cond_tree.cond_items = self.cond_items + [cond_item]
        return cond_tree


def _setup_fptree(itemsets: np.ndarray, min_support: float):
    """fpcommon.setup_fptree: level-1 keep test is the fp64 compare
    ``count / float(N) >= min_support`` (the T1 test)."""
    num_itemsets = itemsets.shape[0]
    with np.errstate(divide="ignore", invalid="ignore"):
        item_support = np.array(np.sum(itemsets, axis=0) / float(num_itemsets)).reshape(-1)
    items = np.nonzero(item_support >= min_support)[0]
    indices = item_support[items].argsort()
    rank = {int(item): i for i, item in enumerate(items[indices])}
    tree = _FPTree(rank)
    for i in range(num_itemsets):
        nonnull = np.where(itemsets[i, :])[0]
        itemset = [int(item) for item in nonnull if int(item) in rank]
        itemset.sort(key=rank.get, reverse=True)
        tree.insert_itemset(itemset)
    return tree, rank


def _fpg_step(tree: _FPTree, minsup: int, max_len):
    """fpgrowth.fpg_step: recursive generator of (count, item-index list)."""
    items = list(tree.nodes.keys())
    if tree.is_path():
        size_remain = len(items) + 1
        if max_len:
            size_remain = max_len - len(tree.cond_items) + 1
        for i in range(1, size_remain):
            for itemset in itertools.combinations(items, i):
                support = min(tree.nodes[it][0].count for it in itemset)
                yield support, tree.cond_items + list(itemset)
    elif not max_len or max_len > len(tree.cond_items):
        for item in items:
            support = sum(node.count for node in tree.nodes[item])
            yield support, tree.cond_items + [item]
    if not tree.is_path() and (not max_len or max_len > len(tree.cond_items)):
        for item in items:
            cond_tree = tree.conditional_tree(item, minsup)
            for sup, iset in _fpg_step(cond_tree, minsup, max_len):
                yield sup, iset


def thresholds(min_support: float, n_transactions: int) -> Tuple[int, int]:
    """(T1, Tk): the two integer thresholds mlxtend's two tests are equivalent to.

    T1 = min{c : c / float(N) >= min_support}   (setup_fptree, fp64 compare)
    Tk = math.ceil(min_support * N)             (fpgrowth, integer compare)
    They can differ (s=0.07, N=100 -> T1=7, Tk=8).
    """
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


def fpgrowth_counts(onehot: np.ndarray, min_support: float, max_len=None
                    ) -> Dict[FrozenSet[int], int]:
    """mlxtend.frequent_patterns.fpgrowth on a bool matrix, returning the integer counts
    keyed by frozenset of column indices (before the ``count / N`` division)."""
    if min_support <= 0.0:
        raise ValueError(
            "`min_support` must be a positive number within the interval `(0, 1]`. "
            "Got %s." % min_support)
    n = onehot.shape[0]
    tree, _ = _setup_fptree(onehot, min_support)
    minsup = math.ceil(min_support * n)
    out: Dict[FrozenSet[int], int] = {}
    for sup, iset in _fpg_step(tree, minsup, max_len):
        out[frozenset(int(i) for i in iset)] = int(sup)
    return out


def fpgrowth(onehot: np.ndarray, columns: Sequence, min_support: float, max_len=None
             ) -> List[Tuple[float, FrozenSet]]:
    """``fpgrowth(df, min_support, use_colnames=True)`` (main.py:272) followed by
    fpcommon.generate_itemsets: support = count / N (Python true division)."""
    n = onehot.shape[0]
    counts = fpgrowth_counts(onehot, min_support, max_len)
    return [(c / n, frozenset(columns[i] for i in iset)) for iset, c in counts.items()]


# --------------------------------------------------------------------------------------
# The reference's own glue: main.py:282-296 (fold) and :304 (bookkeeping)
# --------------------------------------------------------------------------------------
def fold_recommendations(itemsets: Iterable[Tuple[float, FrozenSet]]
                         ) -> Dict[str, Dict[str, float]]:
    """main.py:282-296 verbatim in behaviour: every itemset (singletons included) creates a
    key per member; value = max support over itemsets containing both songs."""
    songs_to_song_sets: Dict = {}
    for support, members in itemsets:
        itemset = set(members)
        for song in itemset:
            other_songs = itemset - {song}
            if song not in songs_to_song_sets:
                songs_to_song_sets[song] = dict()
            current = songs_to_song_sets[song]
            for other in other_songs:
                if other not in current:
                    current[other] = support
                else:
                    current[other] = max(current[other], support)
    return songs_to_song_sets


def calculate_and_save_fp_growth_fast(playlist_dict: dict, min_support=0.07, total_songs=0):
    """main.py:262-313 with the two mlxtend symbols replaced by the restatement above.
    Returns (songs_to_song_sets, itemsets) — timing/printing is the caller's business."""
    transactions = list(playlist_dict.values())
    te = TransactionEncoder()
    te_ary = te.fit(transactions).transform(transactions)
    data = fpgrowth(te_ary, te.columns_, min_support)
    return fold_recommendations(data), data


# --------------------------------------------------------------------------------------
# mlxtend.frequent_patterns.association_rules(df, metric="confidence", min_threshold=c)
# (not executed by the reference; oracle for the extra rule table)
# --------------------------------------------------------------------------------------
def association_rules(itemsets: Iterable[Tuple[float, FrozenSet]], min_confidence: float
                      ) -> List[Tuple[FrozenSet, FrozenSet, float, float, float]]:
    """Rows (antecedents, consequents, support, confidence, lift).

    Operation order follows mlxtend: supports are the already-rounded ``count / N``
    doubles; confidence = sAC / sA; lift = confidence / sC; keep iff
    confidence >= min_confidence.
    """
    support_of = {members: s for s, members in itemsets}
    rules = []
    for members, s_ac in support_of.items():
        k = len(members)
        if k < 2:
            continue
        for r in range(k - 1, 0, -1):
            for combo in itertools.combinations(sorted(members, key=repr), r):
                antecedent = frozenset(combo)
                consequent = members.difference(antecedent)
                s_a = support_of[antecedent]
                s_c = support_of[consequent]
                confidence = s_ac / s_a
                if confidence >= min_confidence:
                    rules.append((antecedent, consequent, s_ac, confidence, confidence / s_c))
    return rules
