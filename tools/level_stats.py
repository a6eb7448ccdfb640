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
    for k, par in enumerate(res.parents, start=2):
        _, cls = torch.unique_consecutive(par, return_counts=True)
        cls = cls.double()
        n = int(par.numel())
        lone = int((cls == 1).sum())
        cand = int((cls * (cls - 1) / 2).sum())
        print(f"k={k} n={n} classes={int(cls.numel())} lone={lone} mean={cls.mean():.2f} "
              f"max={int(cls.max())} next_candidates={cand}")
        cnt = res.levels[k - 1][1].double()
        step = max(1, n // 1000000)
        print(f"     count quantiles: {[int(torch.quantile(cnt[::step], q)) for q in (0.1, 0.5, 0.9, 0.99)]}")
    # subtree sizes under each frequent pair (for the multi-GPU class partition): walk parents up
    if len(res.parents) >= 2:
        n2 = int(res.levels[1][1].numel())
        sub = torch.zeros(n2, dtype=torch.int64, device=dev)
        anc = None
        for k in range(3, len(res.levels) + 1):
            par = res.parents[k - 2].long()
            anc = par if anc is None else anc[par]
            sub += torch.bincount(anc, minlength=n2)
        s_sorted, _ = torch.sort(sub, descending=True)
        tot = int(sub.sum())
        print("pair subtrees: total", tot, "top", s_sorted[:10].tolist(),
              "top-100 share", float(s_sorted[:100].sum()) / max(tot, 1))
        # later-sibling count per pair as a weight estimate: correlation with the true size
        it2 = res.levels[1][0]
        a = it2[:, 0].long()
        _, cls = torch.unique_consecutive(a, return_counts=True)
        ends = torch.cumsum(cls, 0)
        later = torch.repeat_interleave(ends, cls) - torch.arange(n2, device=dev) - 1
        for name_, w in (("later", later.double()), ("later^2", later.double() ** 2)):
            # balance if 8 contiguous blocks are cut by this weight
            cw = torch.cumsum(w, 0)
            cuts = torch.searchsorted(cw, torch.arange(1, 8, device=dev).double() * cw[-1] / 8)
            cs = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), torch.cumsum(sub, 0)])
            b = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), cuts + 1,
                           torch.tensor([n2], device=dev)])
            sizes = (cs[b[1:]] - cs[b[:-1]]).tolist()
            print(f"8 blocks by {name_}: sizes {sizes} max/mean {max(sizes) * 8 / max(tot, 1):.2f}")


if __name__ == "__main__":
    main()
