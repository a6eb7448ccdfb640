"""Per-level shape of one mining pass: candidates, survivors, row widths, class sizes.
Run on the GPU box; the numbers size the fused count+project kernel."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import kmls_b200  # noqa: E402,F401
from kmls_b200 import miner  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "zipf_1Mx100k_s0.001"
    off, items, n_tracks, s, conf = bench.make_data(name, 0)
    dev = torch.device("cuda", 0)
    d_off, d_items = torch.from_numpy(off).to(dev), torch.from_numpy(items).to(dev)
    res = miner.mine_csr(d_off, d_items, n_tracks, s)
    print(json.dumps(res.stats))
    # class structure per level: members per parent
    for k, par in enumerate(res.parents, start=2):
        _, cls = torch.unique_consecutive(par, return_counts=True)
        cls = cls.double()
        n = int(par.numel())
        lone = int((cls == 1).sum())
        cand = int((cls * (cls - 1) / 2).sum())
        print(f"k={k} n={n} classes={int(cls.numel())} lone={lone} mean={cls.mean():.2f} "
              f"max={int(cls.max())} next_candidates={cand}")
        cnt = res.levels[k - 1][1].double()
        print(f"     count quantiles: {[int(torch.quantile(cnt[:: max(1, n // 1000000)], q)) for q in (0.1, 0.5, 0.9, 0.99)]}")


if __name__ == "__main__":
    main()
