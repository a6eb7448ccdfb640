"""Per-call device times of one mining + rules pass (CUDA events around every library call),
in call order: which level each kernel spends its time on. Run on the GPU box."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import kmls_b200  # noqa: E402
from kmls_b200 import _lib, miner  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "zipf_1Mx100k_s0.001"
    only = sys.argv[2].split(",") if len(sys.argv) > 2 else None
    off, items, n_tracks, s, conf = bench.make_data(name, 0)
    dev = torch.device("cuda", 0)
    d_off, d_items = torch.from_numpy(off).to(dev), torch.from_numpy(items).to(dev)
    for _ in range(2):
        res = miner.mine_csr(d_off, d_items, n_tracks, s)
        miner.generate_rules(res, conf)
    torch.cuda.synchronize()
    _lib.TRACE = []
    res = miner.mine_csr(d_off, d_items, n_tracks, s)
    miner.generate_rules(res, conf)
    torch.cuda.synchronize()
    tr, _lib.TRACE = _lib.TRACE, None
    seen = {}
    for nm, e0, e1, ab in tr:
        seen[nm] = seen.get(nm, 0) + 1
        if only and not any(o in nm for o in only):
            continue
        ms = e0.elapsed_time(e1)
        extra = f"  {ab / ms / 1e6:8.1f} GB/s" if ab else ""
        print(f"{nm:28s} #{seen[nm]:<3d} {ms:8.3f} ms{extra}")
    print("levels:", [int(c.numel()) for _, c in res.levels])


if __name__ == "__main__":
    main()
