"""Generates tests/golden/generated_cases.json with the CPU oracle (run from the repo root:
``python tests/golden/make_golden.py``). The reference's own implementation (mlxtend 0.23.3)
is not importable in this image, so these vectors pin the ORACLE's output on seeded inputs;
the hand-computed vectors in hand_vectors.json are what pin the oracle itself."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import kmls_b200  # noqa: E402
from kmls_b200 import synth  # noqa: E402
from oracle import fpgrowth_c, mlxtend_restated as M  # noqa: E402

CASES = [
    # name, generator, params, min_support, min_confidence
    ("zipf_small", "zipf", dict(n_playlists=3000, n_tracks=500, mean_draws=25, alpha=0.9, seed=11), 0.02, 0.5),
    ("zipf_ragged_width", "zipf", dict(n_playlists=1237, n_tracks=300, mean_draws=20, alpha=0.7, seed=12), 0.03, 0.2),
    ("zipf_medium", "zipf", dict(n_playlists=20000, n_tracks=5000, mean_draws=40, alpha=1.0, seed=13), 0.01, 0.6),
    ("mpd_standin_seed1", "mpd", dict(seed=1), 0.05, 0.04),
    ("mpd_standin_seed2", "mpd", dict(seed=2), 0.05, 0.04),
]


def make_input(gen, params):
    if gen == "zipf":
        off, items = synth.zipf_playlists(**params)
        return off, items, params["n_tracks"], None
    d = synth.mpd_standin(**params)
    playlists = synth.group_by_playlist(d["pid"], d["track_name"])
    columns, off, ids = kmls_b200.encode_transactions(list(playlists.values()))
    return off, ids, len(columns), columns


def rule_checksum(rules):
    # order-independent: sum of hashes of (sorted antecedent, sorted consequent, bits of the doubles)
    tot = 0
    for a, c, s, conf, lift in rules:
        key = (tuple(sorted(a)), tuple(sorted(c)), float(s).hex(), float(conf).hex(), float(lift).hex())
        tot = (tot + int.from_bytes(__import__("hashlib").sha256(repr(key).encode()).digest()[:8], "little")) % (1 << 64)
    return tot


def main():
    out = []
    for name, gen, params, s, conf in CASES:
        off, items, n_items, _ = make_input(gen, params)
        n = len(off) - 1
        t1, tk = M.thresholds(s, n)
        r = fpgrowth_c.fpgrowth(off, items, n_items, t1, tk)
        sup = [(int(c) / n, frozenset(r.items[r.offsets[i]:r.offsets[i + 1]].tolist()))
               for i, c in enumerate(r.counts)]
        rules = M.association_rules(sup, conf)
        lv = r.level_counts
        out.append({
            "name": name, "generator": gen, "params": params, "min_support": s, "min_confidence": conf,
            "n_playlists": n, "nnz": int(len(items)), "t1": t1, "tk": tk,
            "level_counts": lv[:int(np.max(np.nonzero(lv)[0], initial=0)) + 1].tolist(),
            "n_itemsets": r.n, "itemset_checksum": str(r.checksum),
            "n_rules": len(rules), "rule_checksum": str(rule_checksum(rules)),
        })
        print(out[-1])
    with open(os.path.join(ROOT, "tests", "golden", "generated_cases.json"), "w") as f:
        json.dump({"_comment": "made by tests/golden/make_golden.py with the CPU oracle", "cases": out}, f, indent=1)


if __name__ == "__main__":
    main()
