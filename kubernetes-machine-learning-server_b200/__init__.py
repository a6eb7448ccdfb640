"""B200-native replacement for the rule-generation hot path of
diogoneiss/kubernetes-machine-learning-server (machine-learning/main.py:262-313).

``calculate_and_save_fp_growth_fast`` is the reference's own function boundary; the rest are
the lower-level entry points it is built from. All compute runs in ``libkmls_b200.so``
(``include/kmls_b200.h``); there is no CPU fallback.
"""
from . import synth  # noqa: F401  (numpy only)
from ._lib import KmlsError, LIB_PATH  # noqa: F401
from .miner import (MineResult, Prepared, RuleTable, generate_rules, mine_csr, prepare,  # noqa: F401
                    rule_checksum, rules_to_host, thresholds)
from .api import (calculate_and_save_fp_growth_fast, encode_frame, encode_transactions,  # noqa: F401
                  itemsets_frame, mine_ids, most_frequent_tracks, recommendations_from_frame,
                  recommendations_from_result, rules_frame, save_rules, sweep_supports)

__all__ = ["calculate_and_save_fp_growth_fast", "mine_csr", "mine_ids", "generate_rules",
           "thresholds", "encode_transactions", "recommendations_from_result", "itemsets_frame",
           "rules_frame", "save_rules", "rules_to_host", "MineResult", "RuleTable", "KmlsError",
           "synth", "prepare", "Prepared", "encode_frame", "recommendations_from_frame",
           "most_frequent_tracks", "sweep_supports"]
