"""Where the end-to-end step (host CSR in, level tables back on the host) spends its time
beyond the device-resident step: wall-clock of the variants, one by one. Run on the GPU box."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import kmls_b200  # noqa: E402
from kmls_b200 import miner  # noqa: E402


def wall(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "zipf_1Mx100k_s0.001"
    off, items, n_tracks, s, conf = bench.make_data(name, 0)
    dev = torch.device("cuda", 0)
    h_off = torch.from_numpy(off).pin_memory()
    h_items = torch.from_numpy(items).pin_memory()
    d_off, d_items = h_off.to(dev), h_items.to(dev)
    ht = kmls_b200.api.HostTables(dev)
    out = {}
    out["h2d_only"] = wall(lambda: (h_off.to(dev, non_blocking=True), h_items.to(dev, non_blocking=True)))
    out["mine_resident"] = wall(lambda: miner.mine_csr(d_off, d_items, n_tracks, s))

    def mine_rules():
        res = miner.mine_csr(d_off, d_items, n_tracks, s)
        return miner.generate_rules(res, conf).n_rules
    out["mine_rules_resident"] = wall(mine_rules)

    def mine_host_tables():
        ht.begin()
        res = miner.mine_csr(d_off, d_items, n_tracks, s, on_level=ht.on_level)
        ht.finish(res)
    out["mine_with_host_tables"] = wall(mine_host_tables)

    def full():
        o = h_off.to(dev, non_blocking=True)
        i = h_items.to(dev, non_blocking=True)
        ht.begin()
        res = miner.mine_csr(o, i, n_tracks, s, on_level=ht.on_level)
        n = miner.generate_rules(res, conf).n_rules
        ht.finish(res)
        return n
    out["full_e2e"] = wall(full)

    def full_no_tables():
        o = h_off.to(dev, non_blocking=True)
        i = h_items.to(dev, non_blocking=True)
        res = miner.mine_csr(o, i, n_tracks, s)
        return miner.generate_rules(res, conf).n_rules
    out["h2d_mine_rules_no_tables"] = wall(full_no_tables)
    for k, v in out.items():
        print(f"{k:32s} {v:8.2f} ms")


if __name__ == "__main__":
    main()
