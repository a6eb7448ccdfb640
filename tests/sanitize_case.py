"""Small end-to-end case for compute-sanitizer (memcheck / racecheck / synccheck):
every kernel of the default path runs once, including the tcgen05 pair GEMM."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # repo root
import kmls_b200
from kmls_b200 import synth
from oracle import fpgrowth_c, mlxtend_restated as M

n, t, s, conf = 20000, 600, 0.01, 0.4          # row_words = 632 >= 512 and 300+ rows: GEMM path
off, items = synth.zipf_playlists(n, t, mean_draws=30, alpha=0.8, seed=5)
res, rules = kmls_b200.mine_ids(off, items, t, s, conf)
t1, tk = M.thresholds(s, n)
ref = fpgrowth_c.fpgrowth(off, items, t, t1, tk, rules_min_confidence=conf)
assert res.as_dict() == ref.as_dict() and rules.n_rules == ref.n_rules
print("sanitize case ok:", res.n_itemsets, "itemsets,", rules.n_rules, "rules, rows", int(res.row_items.numel()),
      "pair_impl i8:", int(res.row_items.numel()) >= 256)
