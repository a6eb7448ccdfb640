"""Counter-update (RED.ADD) rate of this GPU, the peak k_pair_counts_csr's roofline is quoted
against. Run on the GPU box: python tools/redg_peak.py > profiles/r2_redg_peak.json"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmls_b200  # noqa: E402,F401
from kmls_b200 import _lib  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    out = {"gpu": torch.cuda.get_device_name(0), "cases": []}
    s = torch.cuda.current_stream().cuda_stream
    for n_rows in (1024, 5961, 16384):
        words = n_rows * n_rows
        buf = torch.zeros(words, dtype=torch.int32, device=dev)
        n_upd = 1 << 30
        for _ in range(2):
            _lib.call("kmls_redg_microbench", buf, words, n_upd, 1, s)
        torch.cuda.synchronize()
        best = 0.0
        for rep in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.call("kmls_redg_microbench", buf, words, n_upd, 2 + rep, s)
            e1.record()
            torch.cuda.synchronize()
            best = max(best, n_upd / (e0.elapsed_time(e1) / 1e3))
        out["cases"].append({"counter_matrix_rows": n_rows, "bytes": words * 4, "updates": n_upd,
                             "updates_per_s": best})
    # the kernel's pattern: one warp instruction's updates inside one stretch of `span` counters
    words = 5961 * 5961
    buf = torch.zeros(words, dtype=torch.int32, device=dev)
    for span in (32, 256, 1024, 5961):
        n_upd = 1 << 30
        for _ in range(2):
            _lib.call("kmls_redg_microbench_span", buf, words, n_upd, span, 1, s)
        torch.cuda.synchronize()
        best = 0.0
        for rep in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.call("kmls_redg_microbench_span", buf, words, n_upd, span, 2 + rep, s)
            e1.record()
            torch.cuda.synchronize()
            best = max(best, n_upd / (e0.elapsed_time(e1) / 1e3))
        out["cases"].append({"counter_matrix_rows": 5961, "bytes": words * 4, "updates": n_upd,
                             "warp_span_counters": span, "updates_per_s": best})
    out["peak_updates_per_s"] = max(c["updates_per_s"] for c in out["cases"])
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
