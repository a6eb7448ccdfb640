"""Achievable HBM rate of a streaming kernel for several read:write mixes (8-byte elements over
separate arrays) - the roofline of k_rules_write, whose traffic is 1 part reads to 3 parts
writes. Run on the GPU box: python tools/stream_mix_peak.py > profiles/r2_stream_mix_peak.json"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kmls_b200  # noqa: E402,F401
from kmls_b200 import _lib  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    n = 600_000_000          # elements per array: 4.8 GB each, far beyond L2
    arrs = [torch.zeros(n, dtype=torch.int64, device=dev) for _ in range(6)]
    s = torch.cuda.current_stream().cuda_stream
    out = {"gpu": torch.cuda.get_device_name(0), "elements": n, "cases": []}
    for reads, writes in ((1, 0), (0, 1), (1, 1), (1, 3), (2, 3), (0, 3), (3, 1)):
        for _ in range(2):
            _lib.call("kmls_stream_mix_microbench", *arrs[:3], n, reads, writes, *arrs[3:], s)
        torch.cuda.synchronize()
        best = 0.0
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.call("kmls_stream_mix_microbench", *arrs[:3], n, reads, writes, *arrs[3:], s)
            e1.record()
            torch.cuda.synchronize()
            best = max(best, 8.0 * n * (reads + writes) / (e0.elapsed_time(e1) / 1e3) / 1e9)
        out["cases"].append({"reads_per_element": reads, "writes_per_element": writes, "gb_per_s": best})
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
