"""Generates tests/golden/full_size.json: what the CPU oracle (oracle/fpgrowth_ref.cpp, the C++
restatement of mlxtend's FP-Growth + association_rules) finds on bench.py's FULL-SIZE workloads:
level sizes, the order-independent itemset checksum, the number of rules and the
order-independent rule checksum. bench.py compares every run with these values on the device
(kmls_itemset_checksum / kmls_rule_checksum), at every number of GPUs.

Run from the repo root (minutes to an hour of CPU, tens of GB of RAM for the largest):

    python tests/golden/make_full_size_golden.py [case-name ...]

The reference's own implementation (mlxtend 0.23.3) is not importable in this image, so these
values pin the ORACLE's output ("parity unpinned upstream").
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
from oracle import fpgrowth_c, mlxtend_restated as M  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "full_size.json")

# name -> (workload, min_support override or None, with rules?)
CASES = {
    "zipf_1Mx100k_s0.001": ("zipf_1Mx100k_s0.001", None, True),
    "zipf_200kx20k_s0.004": ("zipf_200kx20k_s0.004", None, True),
    "dense_200kx4k_s0.02": ("dense_200kx4k_s0.02", None, True),
    # BASELINE.json configs 1 and 2: stand-ins for the absent ds1 / ds2 CSVs
    "mpd_standin_seed1_s0.05": ("mpd_standin_seed1_s0.05", None, True),
    "mpd_standin_seed2_s0.05": ("mpd_standin_seed2_s0.05", None, True),
    # BASELINE.json config 4: the low-support sweep on the 1M set (mining only below 0.001: the
    # rule table of the lowest support does not fit anywhere)
    "zipf_1Mx100k_s0.05": ("zipf_1Mx100k_s0.001", 0.05, True),
    "zipf_1Mx100k_s0.02": ("zipf_1Mx100k_s0.001", 0.02, True),
    "zipf_1Mx100k_s0.01": ("zipf_1Mx100k_s0.001", 0.01, True),
    "zipf_1Mx100k_s0.005": ("zipf_1Mx100k_s0.001", 0.005, True),
    "zipf_1Mx100k_s0.002": ("zipf_1Mx100k_s0.001", 0.002, True),
    "zipf_1Mx100k_s0.0005": ("zipf_1Mx100k_s0.001", 0.0005, False),
    # BASELINE.json config 5 (all shards concatenated: the job's whole playlist set)
    "zipf_10Mx2M_s0.001": ("zipf_10Mx2M_s0.001", None, True),
}


def key_of(workload, min_support, n_playlists):
    return f"{workload}|s={min_support:g}|n={n_playlists}"


def run_case(name):
    workload, s_over, with_rules = CASES[name]
    bench.MIN_SUPPORT_OVERRIDE = s_over
    if workload in bench.TOTAL_IS_FIXED:
        off, items, t, s, conf = bench.whole_job_data(workload)
    else:
        off, items, t, s, conf = bench.make_data(workload, shard=0)
    n = len(off) - 1
    t1, tk = M.thresholds(s, n)
    t0 = time.time()
    r = fpgrowth_c.fpgrowth(off, items, t, t1, tk, store=with_rules,
                            rules_min_confidence=conf if with_rules else None)
    sec = time.time() - t0
    lv = r.level_counts
    out = {
        "workload": workload, "min_support": s, "min_confidence": conf if with_rules else None,
        "n_playlists": n, "n_tracks": t, "nnz": int(len(items)), "t1": t1, "tk": tk,
        "level_sizes": lv[1:int(np.max(np.nonzero(lv)[0], initial=0)) + 1].tolist(),
        "n_itemsets": int(r.n), "itemset_checksum": str(r.checksum),
        "n_rules": None if not with_rules else int(r.n_rules),
        "rule_checksum": None if not with_rules else str(r.rule_checksum),
        "oracle": "oracle/fpgrowth_ref.cpp", "oracle_seconds": round(sec, 1), "oracle_threads": os.cpu_count(),
    }
    return key_of(workload, s, n), out


def main():
    names = sys.argv[1:] or list(CASES)
    data = {"_comment": "made by tests/golden/make_full_size_golden.py with the CPU oracle; keys are "
                        "workload|s=min_support|n=playlists of the whole job", "cases": {}}
    if os.path.exists(OUT):
        data = json.load(open(OUT))
    for nm in names:
        k, v = run_case(nm)
        print(k, json.dumps(v), flush=True)
        data["cases"][k] = v
        with open(OUT, "w") as f:
            json.dump(data, f, indent=1)


if __name__ == "__main__":
    main()
