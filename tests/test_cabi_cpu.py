"""CPU-side checks of the boundary: the C-ABI library loads and exports every symbol that
include/kmls_b200.h declares, argument validation works without a GPU, and the host logic
(thresholds, encoding, sharding) matches the oracle. No compute calls here."""
import ctypes
import os
import re

import numpy as np
import pytest

import kmls_b200
from kmls_b200 import _lib, api
from oracle import mlxtend_restated as M

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "kmls_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(kmls_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/kmls_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes table and header disagree"


def test_library_is_not_older_than_its_sources():
    """A stale libkmls_b200.so behind a changed header means ctypes calls with the wrong
    argument lists (a crash on the GPU box, not an error): rebuild before anything else."""
    import glob
    srcs = glob.glob(os.path.join(ROOT, "kubernetes-machine-learning-server_b200", "csrc", "*.cu*")) + \
        [os.path.join(ROOT, "include", "kmls_b200.h")]
    newest = max(os.path.getmtime(p) for p in srcs)
    assert os.path.getmtime(_lib.LIB_PATH) >= newest, "run __graft_entry__.build(): library older than sources"


def test_version_and_error_text():
    lib = _lib.load()
    assert lib.kmls_version() == 100
    # an argument error is reported before anything touches a device
    with pytest.raises(_lib.KmlsError, match="row_words"):
        _lib.call("kmls_row_popcount", 0, 4, 3, 0, 0)
    with pytest.raises(_lib.KmlsError, match="power of two"):
        _lib.call("kmls_itemset_set_build", 0, 10, 2, 0, 7, 0)


def test_row_words_padding():
    assert _lib.value("kmls_bitmap_row_words", 1) == 8
    assert _lib.value("kmls_bitmap_row_words", 256) == 8
    assert _lib.value("kmls_bitmap_row_words", 257) == 16
    assert _lib.value("kmls_bitmap_row_words", 1_000_000) == 3907 * 8


@pytest.mark.parametrize("s,n", [(0.07, 100), (0.035, 200), (0.14, 50), (0.07, 10_000_000), (0.05, 2246),
                                 (0.001, 1_000_000), (0.25, 20), (0.5, 4), (1.0, 7), (0.3, 6), (0.0005, 1_000_000)])
def test_thresholds_match_oracle(s, n):
    assert kmls_b200.thresholds(s, n) == M.thresholds(s, n)


def test_thresholds_random_against_definition():
    rng = np.random.default_rng(0)
    for _ in range(2000):
        n = int(rng.integers(1, 5000))
        s = float(rng.choice([rng.random(), round(rng.random(), 2), round(rng.random(), 3)])) or 0.5
        t1, tk = kmls_b200.thresholds(s, n)
        assert tk == int(np.ceil(s * n))
        assert t1 > n or t1 / float(n) >= s
        assert t1 == 0 or not ((t1 - 1) / float(n) >= s)


def test_min_support_must_be_positive():
    with pytest.raises(ValueError):
        kmls_b200.thresholds(0.0, 10)
    with pytest.raises(ValueError):
        kmls_b200.calculate_and_save_fp_growth_fast({1: ["a"]}, 0.0)


def test_encode_transactions_matches_transaction_encoder():
    tx = [["b", "a", "b"], [], ["c", 'q"uo,te', "a"], ["a"]]
    columns, offsets, ids = api.encode_transactions(tx)
    te = M.TransactionEncoder().fit(tx)
    assert columns == te.columns_
    assert offsets.tolist() == [0, 3, 3, 6, 7]
    dense = np.zeros((len(tx), len(columns)), bool)
    for r in range(len(tx)):
        dense[r, ids[offsets[r]:offsets[r + 1]]] = True
    assert (dense == te.transform(tx)).all()


def test_shard_partitions_playlists():
    off = np.array([0, 2, 2, 5, 9, 10], dtype=np.int64)
    ids = np.arange(10, dtype=np.int32)
    seen, total = [], 0
    for r in range(3):
        o, i = api._shard(off, ids, r, 3)
        assert o[0] == 0 and o[-1] == len(i)
        total += len(o) - 1
        seen.extend(i.tolist())
    assert total == 5 and seen == list(range(10))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "kubernetes-machine-learning-server_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in re.sub(r'""".*?"""', "", src, flags=re.S), fn


def test_no_gpu_means_loud_failure():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_lib.KmlsError):
        kmls_b200.calculate_and_save_fp_growth_fast({1: ["a", "b"], 2: ["a"]}, 0.5)
