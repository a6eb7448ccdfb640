#!/usr/bin/env python
"""Benchmark of the rule-generation hot path (mine + rules) on B200.

    python bench.py --gpus N --steps K --warmup W            # this repository's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the CPU algorithm, host cores

A "step" is one pass of the hot path over one batch of synthetic playlists: level-wise
frequent-itemset mining (all levels) followed by rule generation. The workload is
BASELINE.json's config 3: Zipfian playlists, 1M playlists x 100k tracks per GPU, about 60
distinct tracks per playlist, min_support 0.001, min_confidence 0.04 (weak scaling: every
rank holds its own 1M playlists, thresholds follow the global N).

Prints ONE JSON line (rank 0). `value` is the mine+rules wall time in seconds with the
playlists already resident in HBM; `e2e` is the same pass starting from pinned HOST buffers
with the result tables copied back; `roofline` describes the dominant kernel;
`cpu_baseline` is the CPU oracle timed on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (n_playlists per GPU, n_tracks, mean_draws, alpha, min_support, min_confidence)
    "zipf_1Mx100k_s0.001": (1_000_000, 100_000, 72.0, 1.0, 0.001, 0.04),
    "zipf_200kx20k_s0.004": (200_000, 20_000, 50.0, 1.0, 0.004, 0.04),      # quick check
    # a DENSE one-hot matrix (about 5 % ones, the regime of the reference's own ds2 scaled up):
    # here the level-2 contraction is cheaper on the tensor cores than per playlist, so
    # pair_impl="auto" picks the tcgen05 int8 GEMM
    "dense_200kx4k_s0.02": (200_000, 4_000, 215.0, 0.2, 0.02, 0.3),
    # stand-ins for the reference's own (absent) CSVs, shaped after its report's ds2 log:
    # (n_playlists, n_names, -, -, MIN_SUPPORT, MIN_CONFIDENCE); generator synth.mpd_standin(seed)
    # BASELINE.json config 5: 10M playlists x 2M tracks in total, row-sharded over the ranks
    # (strong: the total is fixed, each of R ranks generates 10M/R playlists)
    "zipf_10Mx2M_s0.001": (10_000_000, 2_000_000, 68.0, 1.0, 0.001, 0.04),
    "mpd_standin_seed1_s0.05": (2246, 2123, 0.0, 0.0, 0.05, 0.04),
    "mpd_standin_seed2_s0.05": (2246, 2123, 0.0, 0.0, 0.05, 0.04),
}
TOTAL_IS_FIXED = {"zipf_10Mx2M_s0.001"}   # workloads generated shard by shard (a rank never holds the whole set)
GOLDEN = os.path.join(ROOT, "tests", "golden", "full_size.json")


def golden_for(name, min_support, n_total):
    """The CPU oracle's result for this whole job (tests/golden/full_size.json, made by
    tests/golden/make_full_size_golden.py), or None."""
    try:
        cases = json.load(open(GOLDEN))["cases"]
    except (OSError, ValueError, KeyError):
        return None
    return cases.get(f"{name}|s={min_support:g}|n={n_total}")


def whole_job_data(workload, shards=range(8)):
    """Playlists of a workload that is generated shard by shard (config 5): the concatenation
    of the given generator shards (all 8 = the whole job)."""
    n_total = WORKLOADS[workload][0]
    offs, its = [np.zeros(1, np.int64)], []
    for r in shards:
        o, i, t, s, conf = make_data(workload, shard=r, n_override=n_total // 8)
        offs.append(o[1:] + offs[-1][-1])
        its.append(i)
    return np.concatenate(offs), np.concatenate(its), t, s, conf


def rank_slice(off, items, rank, world):
    """Playlists [rank*N/world, (rank+1)*N/world) of a CSR set (strong scaling: the job's
    playlists are the same at every number of GPUs)."""
    n = len(off) - 1
    lo, hi = (n * rank) // world, (n * (rank + 1)) // world
    o = off[lo:hi + 1]
    return np.ascontiguousarray(o - o[0]), np.ascontiguousarray(items[o[0]:o[-1]])
MIN_SUPPORT_OVERRIDE = None
CPU_SAMPLE_PLAYLISTS = 100_000   # playlists of the CPU baseline's bounded sample
CPU_RULE_STRIDE = 8              # rules are generated for 1/8 of the itemset table and scaled


_REAL_STDOUT = None


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def make_data(name, shard, n_override=None):
    from kmls_b200 import synth
    n, t, draws, alpha, s, conf = WORKLOADS[name]
    if MIN_SUPPORT_OVERRIDE is not None:
        s = MIN_SUPPORT_OVERRIDE
    if name.startswith("mpd_standin"):
        import kmls_b200
        d = synth.mpd_standin(seed=int(name.split("seed")[1].split("_")[0]))
        playlists = synth.group_by_playlist(d["pid"], d["track_name"])
        columns, off, ids = kmls_b200.encode_transactions(list(playlists.values()))
        if n_override:
            off = off[:n_override + 1]
            ids = ids[:off[-1]]
        return off, ids, len(columns), s, conf
    if n_override:
        n = n_override
    cache = os.path.join("/tmp", "kmls_bench_cache")
    os.makedirs(cache, exist_ok=True)
    fn = os.path.join(cache, f"{name}_n{n}_shard{shard}.npz")
    if os.path.exists(fn):
        z = np.load(fn)
        return z["offsets"], z["items"], t, s, conf
    t0 = time.time()
    off, items = synth.zipf_playlists(n, t, draws, alpha, seed=0, shard=shard,
                                      workers=min(16, os.cpu_count() or 1))
    log(f"[bench] generated {n} playlists ({len(items)} rows) in {time.time() - t0:.1f}s")
    try:   # atomic: several ranks of one box may generate the same file at the same time
        tmp = f"{fn}.{os.getpid()}.tmp.npz"
        np.savez(tmp, offsets=off, items=items)
        os.replace(tmp, fn)
    except OSError:
        pass
    return off, items, t, s, conf


def python_reference_run(name):
    """The reference-faithful arm for the small stand-ins: the literal Python restatement of
    mlxtend's TransactionEncoder + fpgrowth driven by the reference's own glue
    (oracle.mlxtend_restated.calculate_and_save_fp_growth_fast = main.py:262-313), one thread,
    exactly how the reference job runs it. Returns seconds."""
    from kmls_b200 import synth
    from oracle import mlxtend_restated as M
    d = synth.mpd_standin(seed=int(name.split("seed")[1].split("_")[0]))
    playlists = synth.group_by_playlist(d["pid"], d["track_name"])
    s = MIN_SUPPORT_OVERRIDE if MIN_SUPPORT_OVERRIDE is not None else WORKLOADS[name][4]
    t0 = time.time()
    recs, data = M.calculate_and_save_fp_growth_fast(playlists, s)
    return time.time() - t0, len(recs), len(data)


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons, sampled during the timed region. Through NVML in this
    process (one handle, opened before the timed region): spawning `nvidia-smi` every 200 ms
    attaches to every GPU of the box each time and stalled the OTHER ranks' launches - a 2-rank
    run measured 109 ms per step with it and 85 ms without. `nvidia-smi` (the recipe's clocks
    line) remains the fallback when NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], threading.Event()
        self.nvml = self.handle = self.max_mhz = None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            try:
                uuid = "GPU-" + str(torch.cuda.get_device_properties(index).uuid)
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.nvml = pynvml
            self.source = "nvml"
            self._sample_nvml()      # first queries (lazy initialisation inside NVML) outside the timed region
        except Exception:
            self.nvml = None
            self.source = "nvidia-smi"

    def _sample_nvml(self):
        p, h = self.nvml, self.handle
        sm = p.nvmlDeviceGetClockInfo(h, p.NVML_CLOCK_SM)
        if self.max_mhz is None:
            self.max_mhz = p.nvmlDeviceGetMaxClockInfo(h, p.NVML_CLOCK_SM)   # constant: asked once
        mx = self.max_mhz
        try:
            r = p.nvmlDeviceGetCurrentClocksEventReasons(h)
        except Exception:
            r = p.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        bits = [p.nvmlClocksThrottleReasonHwSlowdown, p.nvmlClocksThrottleReasonHwThermalSlowdown,
                p.nvmlClocksThrottleReasonSwThermalSlowdown, p.nvmlClocksThrottleReasonSwPowerCap]
        return [str(sm), str(mx), ""] + ["Active" if r & b else "Not Active" for b in bits]

    def run(self):
        while not self.stop_flag.is_set():
            try:
                if self.nvml is not None:
                    self.rows.append(self._sample_nvml())
                else:
                    out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                    self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.05)

    def summary(self):
        self.stop_flag.set()
        if self.ident is not None:       # started (rank 0 only)
            self.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(self.NAMES, r[3:7]) if v == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": reasons, "source": self.source}


# ------------------------------------------------------------------------------ CPU arms
def cpu_sample_run(name, threads):
    """The oracle (C++ restatement of the reference's FP-Growth + association_rules) on a
    bounded sample of the workload, extrapolated to the full size:
      * mining is timed on two samples (50k and 100k playlists of the same generator, same
        min_support, so the itemset structure is the workload's) and extrapolated linearly in
        the number of playlists (FP-tree construction grows with N, enumeration does not);
      * rules are generated for every CPU_RULE_STRIDE-th run of 65,536 itemsets of the 100k
        sample's table (their cost does not depend on N) and scaled by the share of
        (itemset, antecedent) splits that the sample evaluated.
    Checked once against a full-size run (profiles/): see BASELINE.md."""
    from oracle import fpgrowth_c, mlxtend_restated as M
    import ctypes
    n_full = WORKLOADS[name][0]
    lib = fpgrowth_c._load()
    pts = []
    h = None
    for n_s in sorted({min(CPU_SAMPLE_PLAYLISTS // 2, n_full), min(CPU_SAMPLE_PLAYLISTS, n_full)}):
        off, items, t, s, conf = make_data(name, shard=0, n_override=n_s)
        t1, tk = M.thresholds(s, n_s)
        if h is not None:
            lib.oracle_fpgrowth_free(h)
        t0 = time.time()
        h = lib.oracle_fpgrowth_run(off.ctypes.data, items.ctypes.data, n_s, t, t1, tk, 0, 1, threads)
        pts.append((n_s, time.time() - t0))
    n_sets = int(lib.oracle_fpgrowth_num_itemsets(h))
    lv = np.zeros(64, dtype=np.uint64)
    lib.oracle_fpgrowth_level_counts(h, lv.ctypes.data)
    total_splits = sum(int(lv[k]) * ((1 << k) - 2) for k in range(2, 40))
    cs = ctypes.c_uint64(0)
    t0 = time.time()
    n_rules = int(lib.oracle_rules_strided(h, pts[-1][0], conf, threads, CPU_RULE_STRIDE, ctypes.byref(cs)))
    t_rules = time.time() - t0
    rule_scale = total_splits / max(int(lib.oracle_rules_last_splits()), 1)
    t_index = float(lib.oracle_rules_last_index_seconds())   # serial index build: not scaled
    lib.oracle_fpgrowth_free(h)
    if len(pts) == 2 and pts[1][0] > pts[0][0]:
        slope = max((pts[1][1] - pts[0][1]) / (pts[1][0] - pts[0][0]), 0.0)
        t_mine_full = pts[1][1] + slope * (n_full - pts[1][0])
    else:
        t_mine_full = pts[-1][1] * n_full / pts[-1][0]
    est = t_mine_full + t_index + (t_rules - t_index) * rule_scale
    measured = sum(p[1] for p in pts) + t_rules
    return {"measured_s": measured, "mine_s": pts[-1][1], "rules_s": t_rules, "estimate_full_s": est,
            "itemsets": n_sets, "rules_in_slice": n_rules,
            "sample": (f"mining timed on {' and '.join(str(p[0]) for p in pts)} of {n_full} playlists (same generator "
                       f"and min_support) and extrapolated linearly in N to {t_mine_full:.1f} s; rules for every "
                       f"{CPU_RULE_STRIDE}th run of 65,536 itemsets ({t_rules:.1f} s), scaled by the share of "
                       f"(itemset, antecedent) splits evaluated (x{rule_scale:.2f})")}


def try_mlxtend():
    """The reference's own third-party miner (mlxtend 0.23.3, machine-learning/uv.lock:302), if
    this box has it: run it on the hand vectors and diff with the oracle. It is absent from
    every image seen so far; the outcome is reported either way."""
    try:
        import mlxtend  # noqa: F401
        from mlxtend.frequent_patterns import fpgrowth
        from mlxtend.preprocessing import TransactionEncoder
    except Exception as e:      # noqa: BLE001
        return {"available": False, "note": f"mlxtend: unavailable on this box ({type(e).__name__})"}
    import pandas as pd
    vec = json.load(open(os.path.join(ROOT, "tests", "golden", "hand_vectors.json")))
    bad = checked = 0
    for case in vec["cases"]:
        tx, s_ = case.get("transactions"), case.get("min_support")
        if not tx or s_ is None or "itemsets" not in case:
            continue
        te = TransactionEncoder()
        df = pd.DataFrame(te.fit(tx).transform(tx), columns=te.columns_)
        got = {frozenset(r.itemsets): int(round(r.support * len(tx)))
               for r in fpgrowth(df, min_support=s_, use_colnames=True).itertuples()}
        want = {frozenset(k.split("|")): int(v) for k, v in case["itemsets"].items()}
        checked += 1
        bad += got != want
    return {"available": True, "version": getattr(mlxtend, "__version__", "?"),
            "hand_vectors_checked": checked, "hand_vector_mismatches": int(bad)}


def run_reference(args):
    """The reference's algorithm on the host cores: ONE measured pass over the whole job
    (mining + rules), whatever --steps/--warmup say - a pass takes about a minute, and an
    extrapolated figure is not a measurement. The line reports the steps it really ran."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    name = args.workload
    n, t, draws, alpha, s, conf = WORKLOADS[name]
    if MIN_SUPPORT_OVERRIDE is not None:
        s = MIN_SUPPORT_OVERRIDE
    extra = {}
    if name.startswith("mpd_standin"):
        secs = []
        for i in range(args.warmup + args.steps):
            sec, n_keys, n_sets = python_reference_run(name)
            log(f"[bench/reference] step {i}: {sec:.2f}s")
            if i >= args.warmup:
                secs.append(sec)
        v, threads, steps, warmup = float(np.mean(secs)), 1, args.steps, args.warmup
        n_total = n
        sample = ("the whole workload, every step: Python restatement of mlxtend inside the reference's own glue "
                  "(main.py:262-313), 1 thread as the reference runs it; no rule table (the reference computes none)")
    else:
        from oracle import fpgrowth_c, mlxtend_restated as M
        if name in TOTAL_IS_FIXED:
            off, items, t, s, conf = whole_job_data(name)
        else:
            off, items, t, s, conf = make_data(name, shard=0)
            if args.scaling == "weak" and args.gpus > 1:
                extra["not_comparable"] = ("weak scaling: the GPU arm mines --gpus x this many playlists; "
                                           "this arm mines one rank's set")
        n_total = len(off) - 1
        t1, tk = M.thresholds(s, n_total)
        t0 = time.time()
        r = fpgrowth_c.fpgrowth(off, items, t, t1, tk, store=not args.no_rules, n_threads=threads,
                                rules_min_confidence=None if args.no_rules else conf)
        v = time.time() - t0
        steps, warmup = 1, 0
        lv = r.level_counts
        got = {"level_sizes": lv[1:int(np.max(np.nonzero(lv)[0], initial=0)) + 1].tolist(),
               "itemset_checksum": str(r.checksum), "n_rules": r.n_rules,
               "rule_checksum": None if r.rule_checksum is None else str(r.rule_checksum)}
        want = golden_for(name, s, n_total)
        extra["result"] = got
        if want is not None:
            keys = ["level_sizes", "itemset_checksum"] + ([] if args.no_rules else ["n_rules", "rule_checksum"])
            extra["parity"] = {"ok": all(got[k] == want[k] for k in keys), "compared": keys,
                               "source": "tests/golden/full_size.json"}
        log(f"[bench/reference] one full pass: {v:.1f}s on {threads} threads: {got}")
        sample = (f"the WHOLE job measured once ({n_total} playlists, every level, the full rule table): C++ "
                  f"restatement of the reference's algorithm (oracle/fpgrowth_ref.cpp), {threads} threads; "
                  f"--steps {args.steps} --warmup {args.warmup} were requested, 1 pass was run")
    line = {
        "impl": "reference", "metric": "mine+rules wall time", "value": v, "unit": "s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "steps_requested": args.steps,
        "warmup_requested": args.warmup, "ms_per_step": v * 1e3,
        "higher_is_better": False, "scaling": "strong" if (args.scaling == "strong" or name in TOTAL_IS_FIXED) else "weak",
        "vs_baseline": None, "dtype": "int64+f64", "data": "synthetic",
        "config": {"workload": name, "n_playlists": n_total, "n_tracks": t, "mean_draws": draws, "zipf_alpha": alpha,
                   "min_support": s, "min_confidence": None if args.no_rules else conf,
                   "note": "CPU oracle port of the reference algorithm (mlxtend FP-Growth + association_rules "
                           "semantics, C++, all host threads); the reference itself (pure-Python mlxtend) is not "
                           "installable here and cannot hold this workload in memory"},
        "cpu_baseline": {"value": v, "unit": "s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "mlxtend": try_mlxtend(),
    }
    line.update(extra)
    emit(line)


# ------------------------------------------------------------------------------ GPU arm
def partition_of(args):
    """--partition auto: the candidate space is split (classes) unless the job is config 5, whose
    10M playlists make the per-playlist work dominate: there the playlists are sharded."""
    if args.partition != "auto":
        return args.partition
    return "playlists" if WORKLOADS[args.workload][0] >= 4_000_000 else "classes"


def load_job_data(args, rank, world, barrier):
    """(offsets, items, n_tracks, min_support, min_confidence, n_total, strong) of this rank.
    partition "classes": every rank holds the job's WHOLE playlist set (the candidate space is
    what is split); "playlists": a contiguous slice of it."""
    name = args.workload
    fixed = name in TOTAL_IS_FIXED
    strong = fixed or args.scaling == "strong"
    whole = partition_of(args) == "classes" or world == 1
    if fixed:
        # generated in 8 shards; every rank first makes "its" shards (they land in the cache)
        if 8 % world:
            raise SystemExit(f"{name} is generated in 8 shards: --gpus must divide 8")
        mine = range(rank * 8 // world, (rank + 1) * 8 // world)
        off, items, n_tracks, s, conf = whole_job_data(name, mine)
        if whole and world > 1:
            barrier()
            off, items, n_tracks, s, conf = whole_job_data(name)
        return off, items, n_tracks, s, conf, WORKLOADS[name][0], True
    if strong:
        off, items, n_tracks, s, conf = make_data(name, shard=0)
        n_total = len(off) - 1
        if not whole:
            off, items = rank_slice(off, items, rank, world)
        return off, items, n_tracks, s, conf, n_total, True
    # weak: the job is `world` generator shards
    off, items, n_tracks, s, conf = make_data(name, shard=rank)
    n_total = (len(off) - 1) * world
    if whole and world > 1:
        barrier()
        offs, its = [np.zeros(1, np.int64)], []
        for r in range(world):
            o, i, _, _, _ = make_data(name, shard=r)
            offs.append(o[1:] + offs[-1][-1])
            its.append(i)
        off, items = np.concatenate(offs), np.concatenate(its)
    return off, items, n_tracks, s, conf, n_total, False


def run_gpu(args):
    import torch
    import torch.distributed as dist
    import kmls_b200
    from kmls_b200 import _lib, miner, multi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
        group = dist.group.WORLD

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    name = args.workload
    off, items, n_tracks, s, conf, n_total, strong = load_job_data(args, rank, world, barrier)
    n_local = len(off) - 1
    h_off = torch.from_numpy(off).pin_memory()
    h_items = torch.from_numpy(items).pin_memory()
    d_off, d_items = h_off.to(device), h_items.to(device)
    partition = partition_of(args)
    by_playlists = world > 1 and partition == "playlists"
    # device memory of the one-call entry points, kept across steps (the tables a step returns
    # are views of it and are dropped before the next step)
    spaces = {"mine": miner.Workspace(device, int(args.mine_ws_gb * (1 << 30))),
              "rules": miner.Workspace(device, int(args.rules_ws_gb * (1 << 30)))}

    def mine(o, i):
        if by_playlists or args.count_impl != "projected":
            return miner.mine_csr(o, i, n_tracks, s, n_playlists_total=n_total, group=group if by_playlists else None,
                                  pair_impl=args.pair_impl, count_impl=args.count_impl, impl="python")
        if world > 1:
            return multi.mine_partitioned(o, i, n_tracks, s, group, block=args.part_block, workspace=spaces["mine"],
                                          pair_impl=args.pair_impl)
        return miner.mine_native(o, i, n_tracks, s, n_playlists_total=n_total, pair_impl=args.pair_impl,
                                 workspace=spaces["mine"])

    def step(o, i):
        res = mine(o, i)
        if args.no_rules:
            e = torch.empty(0, device=device)
            return res, miner.RuleTable(e.view(0, 2), e, e, e, [])
        rules = miner.generate_rules(res, conf, part=(rank, world), workspace=spaces["rules"])
        return res, rules

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    info = {}
    for i in range(args.warmup):
        t0 = time.time()
        res, rules = step(d_off, d_items)
        torch.cuda.synchronize()
        log(f"[bench] rank {rank} warmup {i}: {time.time() - t0:.3f}s itemsets={res.n_itemsets} "
            f"rules={rules.n_rules} levels={[int(c.numel()) for _, c in res.levels]}")
        merge_ms = None
        if getattr(res, "merge_marks", None):
            mk = res.merge_marks
            merge_ms = {mk[i + 1][0]: mk[i][1].elapsed_time(mk[i + 1][1]) for i in range(len(mk) - 1)}
        info = {"merge_ms": merge_ms, "itemsets": res.n_itemsets, "rules_this_rank": rules.n_rules,
                "levels": [int(c.numel()) for _, c in res.levels], "stats": res.stats,
                "rule_stats": getattr(rules, "stats", None)}
        del res, rules

    # ---- timed region: K steps, inputs resident in HBM ------------------------------------
    # Rank 0 samples its GPU (the line below is rank 0's). Every NVML / nvidia-smi query takes
    # driver-wide locks; with one sampler per rank their stalls add up across the ranks.
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    _lib.CALLS.clear()
    _lib.EXTRA_LAUNCHES = 0
    # the Python-driven loop (playlist sharding) makes hundreds of library calls per step: their
    # per-call events would weigh on the timed region, so that path is traced in extra steps
    trace_in_loop = not by_playlists
    _lib.TRACE = [] if trace_in_loop else None
    # per-kernel CUDA events inside kmls_mine / kmls_rules - events only: the data-dependent work
    # figures cost extra kernels and host synchronisations, they come from one untimed step below
    _lib.profile_begin(work=False)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        res, rules = step(d_off, d_items)
        del res, rules
    e1.record()
    barrier()
    clocks = sampler.summary()
    launches_timed = _lib.kernel_launches()
    if not trace_in_loop:
        _lib.TRACE = []
        for _ in range(args.steps):
            res, rules = step(d_off, d_items)
            del res, rules
        torch.cuda.synchronize()
    trace, _lib.TRACE = _lib.TRACE, None
    prof = _lib.profile_report()
    if prof:
        # algorithmic work per launch from one more step with the full accounting switched on; the
        # launch sequence of a step is deterministic, so record i of a step gets record i's figure
        _lib.profile_begin(work=True)
        res, rules = step(d_off, d_items)
        del res, rules
        torch.cuda.synchronize()
        prof_w = _lib.profile_report()
        per_step = len(prof) // max(args.steps, 1)
        if len(prof_w) == per_step and all(a[0] == b[0] and a[1] == b[1] for a, b in zip(prof, prof_w)):
            prof = [(nm, lv, ms, prof_w[i % per_step][3]) for i, (nm, lv, ms, _) in enumerate(prof)]
        else:
            log(f"[bench] work accounting step has {len(prof_w)} records, timed steps {per_step}: work figures dropped")
    launches = launches_timed
    step_s = max_over_ranks(e0.elapsed_time(e1) / 1e3 / args.steps)

    # ---- per-kernel times and roofline objects ---------------------------------------------
    # `prof`: one record per kernel launch inside the one-call entry points, with the launch's
    # algorithmic work as DESIGN.md section 3 defines it; `trace`: library calls made from Python
    # (the whole kmls_mine / kmls_rules calls, and every kernel of the Python-driven paths).
    per = {}
    for nm, lv, ms, work in prof:
        d = per.setdefault(nm, {"ms": 0.0, "calls": 0, "work": 0.0, "levels": {}})
        d["ms"] += ms
        d["calls"] += 1
        d["work"] += work
        lvd = d["levels"].setdefault(lv, [0.0, 0.0])
        lvd[0] += ms
        lvd[1] += work
    calls = {}
    for nm, a, b, nbytes in trace:
        d = calls.setdefault(nm, {"ms": 0.0, "calls": 0, "work": 0.0})
        d["ms"] += a.elapsed_time(b)
        d["calls"] += 1
        d["work"] += nbytes or 0
        if nm not in ("kmls_mine", "kmls_rules") and not prof:
            per.setdefault(nm, {"ms": 0.0, "calls": 0, "work": 0.0, "levels": {}})
            per[nm]["ms"] += a.elapsed_time(b)
            per[nm]["calls"] += 1
            per[nm]["work"] += nbytes or 0
    if args.trace and rank == 0:
        for nm, lv, ms, work in prof[-(len(prof) // max(args.steps, 1)):]:
            log(f"[trace] {nm:22s} level {lv:2d} {ms:9.3f} ms  {work / max(ms, 1e-9) / 1e6:9.1f} GB/s")
        log("[trace] level stats: " + json.dumps(info.get("stats", {}).get("levels", [])))
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak, peak_src = (peaks["hbm_gbs"], "MEASURED_PEAKS.json hbm_gbs") if "hbm_gbs" in peaks else \
        (6650.0, "fallback of B200_PROFILING.md")
    int8_peak = 2.0 * peaks.get("bf16_tflops", 1590.0)   # TOP/s: int8 dense = 2x the measured bf16 GEMM peak
    try:   # DRAM traffic of one ncu-captured launch per kernel (workload zipf_1Mx100k_s0.001, 1 GPU)
        ncu_traffic = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    except (OSError, ValueError):
        ncu_traffic = {}

    def roof_of(nm):
        d = per[nm]
        if d["ms"] <= 0 or d["work"] <= 0:
            return None
        sec = d["ms"] / 1e3 / args.steps
        work = d["work"] / args.steps
        if nm == "k_pair_gemm_i8":
            ach, pk, unit, src, bound = work / sec / 1e12, int8_peak, "TOP/s", "2 x bf16_tflops of MEASURED_PEAKS.json", "tensor"
        else:
            ach, pk, unit, src, bound = work / sec / 1e9, peak, "GB/s", peak_src, "hbm"
        r = {"kernel": nm, "bound": bound, "achieved": ach, "peak": pk, "unit": unit, "frac": ach / pk,
             "traffic": None, "peak_source": src, "launches_per_step": d["calls"] / args.steps,
             "ms_per_step": sec * 1e3, "algorithmic_work_per_step": work,
             "algorithmic_work_per_launch": work / max(d["calls"] / args.steps, 1e-9)}
        if d["levels"]:
            # the launch that takes longest, on its own: comparable with a one-launch ncu capture
            lv, (ms_l, w_l) = max(d["levels"].items(), key=lambda kv: kv[1][0])
            r["largest_launch"] = {"level": lv, "ms": ms_l / args.steps, "algorithmic_work": w_l / args.steps,
                                   "achieved": w_l / max(ms_l, 1e-9) / (1e9 if bound == "tensor" else 1e6)}
        if nm == "k_pair_counts_csr":
            # this kernel is bound by the rate at which L2 retires counter updates (RED.ADD), not by
            # bytes: quote it against the micro-benchmark's rate as well (tools/redg_peak.py)
            r["atomic_updates_per_s"] = work / 4.0 / sec
            try:
                pk_a = json.load(open(os.path.join(ROOT, "profiles", "r2_redg_peak.json")))["peak_updates_per_s"]
                r["atomic_peak_updates_per_s"] = pk_a
                r["atomic_frac"] = r["atomic_updates_per_s"] / pk_a
                r["atomic_peak_source"] = ("profiles/r2_redg_peak.json (tools/redg_peak.py: the best rate of a kernel that "
                                           "only scatters RED.ADDs - uniformly over a 4 MB matrix; this kernel's addresses are "
                                           "skewed towards popular rows and combine in L2, so the ratio can pass 1; ncu puts "
                                           "its busiest unit, the L1/LSU path, at 68 %)")
            except (OSError, ValueError, KeyError):
                pass
        if nm == "k_rules_write":
            # 1 part plane words read to 3 parts rule columns written: what a trivial streaming kernel
            # with that mix reaches on this GPU (tools/stream_mix_peak.py), next to the copy peak
            try:
                mix = json.load(open(os.path.join(ROOT, "profiles", "r2_stream_mix_peak.json")))["cases"]
                pk_m = next(c_["gb_per_s"] for c_ in mix if (c_["reads_per_element"], c_["writes_per_element"]) == (1, 3))
                r["mix_peak"] = {"read_write_mix": "1:3", "peak": pk_m, "unit": "GB/s", "frac": ach / pk_m,
                                 "source": "profiles/r2_stream_mix_peak.json (tools/stream_mix_peak.py)"}
            except (OSError, ValueError, KeyError, StopIteration):
                pass
        t = ncu_traffic.get(nm)
        if t and name == t.get("workload", "zipf_1Mx100k_s0.001") and world == 1:
            r["traffic"] = t["traffic_bytes"]
            r["traffic_launch"] = t["launch"]
        return r

    rooflines = {nm: roof_of(nm) for nm in per}
    rooflines = {k_: v for k_, v in rooflines.items() if v}
    top = max(per, key=lambda k_: per[k_]["ms"]) if per else None
    roof = rooflines.get(top) or next(iter(rooflines.values()), None)

    # ---- e2e: the public call, pinned host CSR in, tables back on the host ------------------
    host_tables = kmls_b200.api.HostTables(device)

    def e2e_step():
        res, rules = kmls_b200.mine_ids(h_off, h_items, n_tracks, s, None if args.no_rules else conf,
                                        host_tables=host_tables if rank == 0 else None, split_rules=True,
                                        partition=partition, workspaces=spaces, pair_impl=args.pair_impl,
                                        sharded=by_playlists, n_playlists_total=n_total)
        n_rules = rules.n_rules if rules is not None else 0
        torch.cuda.synchronize()
        del res, rules
        return (host_tables.nbytes if rank == 0 else 0), n_rules

    e2e_note = ("kmls_b200.mine_ids(offsets, ids, ...) - the call behind calculate_and_save_fp_growth_fast - with the "
                "playlists in pinned HOST memory: upload, mine + rules, every level table copied back to pinned host "
                "memory on a side stream while the rules are generated (by rank 0 when there are several ranks, the "
                "tables being identical on all of them); the rule table stays on the device (sharded by rank; "
                "e2e_with_rules below moves it too)")
    if name.startswith("mpd_standin") and world == 1:
        # the reference-facing plugin call itself: dict of lists of names in, dict of dicts out
        from kmls_b200 import synth
        dd = synth.mpd_standin(seed=int(name.split("seed")[1].split("_")[0]))
        playlists = synth.group_by_playlist(dd["pid"], dd["track_name"])
        kmls_b200.calculate_and_save_fp_growth_fast(playlists, s, total_songs=len(set(dd["track_uri"])))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            recs, _, _ = kmls_b200.calculate_and_save_fp_growth_fast(
                playlists, s, total_songs=len(set(dd["track_uri"])), min_confidence=conf)
        torch.cuda.synchronize()
        e2e_s = (time.perf_counter() - t0) / args.steps
        h2d = h_off.numel() * 8 + h_items.numel() * 4
        d2h = sum(12 * len(v) for v in recs.values()) + 4 * len(recs)
        e2e_note = ("kmls_b200.calculate_and_save_fp_growth_fast(playlist_dict, min_support): Python dict of lists "
                    "of track names in (name->id encoding on the host included), recommendations dict out")
    else:
        e2e_step()   # untimed: sizes the pinned result buffers
        barrier()
        t0 = time.perf_counter()
        d2h = 0
        for _ in range(args.steps):
            d2h, _ = e2e_step()
        barrier()
        e2e_s = max_over_ranks((time.perf_counter() - t0) / args.steps)
        # per rank: the offsets, and - when every rank needs all playlists - 1/R of the ids (the
        # ranks exchange their pieces over NVLink, api.mine_ids)
        ids_up = h_items.numel() if (world == 1 or by_playlists) else (h_items.numel() + world - 1) // world
        h2d = h_off.numel() * 8 + ids_up * 4

    # ---- parity: one more (untimed) pass, compared on the device with the CPU oracle's
    # committed result for this whole job (tests/golden/full_size.json) - at every N
    parity = None
    if not args.no_parity:
        res, rules = step(d_off, d_items)
        cs_items = miner.itemset_checksum(res)
        n_rules_all, cs_rules = rules.n_rules, 0
        if not args.no_rules:
            cs_rules = miner.rule_checksum(res, rules)
        if world > 1:
            # every rank holds the whole itemset table; the rule table is split by rank
            t = torch.tensor([n_rules_all, cs_rules - (1 << 64) if cs_rules >= (1 << 63) else cs_rules],
                             dtype=torch.int64, device=device)
            dist.all_reduce(t, op=dist.ReduceOp.SUM)      # int64 wraps: sums modulo 2^64
            n_rules_all, cs_rules = int(t[0].item()), int(t[1].item()) & ((1 << 64) - 1)
            t = torch.tensor([cs_items - (1 << 64) if cs_items >= (1 << 63) else cs_items, res.n_itemsets],
                             dtype=torch.int64, device=device)
            tmax, tmin = t.clone(), t.clone()
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
            ranks_agree = bool(torch.equal(tmax, tmin))
        else:
            ranks_agree = True
        got = {"level_sizes": [int(c.numel()) for _, c in res.levels], "itemset_checksum": str(cs_items),
               "n_rules": None if args.no_rules else n_rules_all,
               "rule_checksum": None if args.no_rules else str(cs_rules)}
        del res, rules
        want = golden_for(name, s, n_total)
        if want is None:
            parity = {"ok": None, "note": "no committed oracle result for this job (tests/golden/full_size.json)",
                      "ranks_agree": ranks_agree, "gpu": got}
        else:
            keys = ["level_sizes", "itemset_checksum"]
            if not args.no_rules and want.get("n_rules") is not None and want.get("min_confidence") == conf:
                keys += ["n_rules", "rule_checksum"]
            ok = ranks_agree and all(got[k_] == want[k_] for k_ in keys)
            parity = {"ok": ok, "compared": keys, "ranks_agree": ranks_agree, "gpu": got,
                      "oracle": {k_: want[k_] for k_ in keys},
                      "source": "tests/golden/full_size.json (CPU oracle oracle/fpgrowth_ref.cpp, "
                                f"{want.get('oracle_seconds')} s on {want.get('oracle_threads')} threads)"}
        log(f"[bench] parity: {parity}")

    # ---- e2e with the RULE TABLE brought to the host as well (on request: at config 3 it is
    # 52 GB over PCIe, about a second)
    e2e_rules = None
    if args.rules_to_host and not args.no_rules:
        def rules_step():
            res, rules = kmls_b200.mine_ids(h_off, h_items, n_tracks, s, conf, host_tables=host_tables if rank == 0 else None,
                                            split_rules=True, partition=partition, workspaces=spaces,
                                            pair_impl=args.pair_impl, sharded=by_playlists, n_playlists_total=n_total)
            hr = kmls_b200.api.rules_to_pinned(rules, rules_pinned)
            torch.cuda.synchronize()
            n = rules.n_rules
            del res, rules, hr
            return 24 * n
        rules_pinned = {}
        rules_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(max(1, min(args.steps, 2))):
            rb = rules_step()
        barrier()
        e2e_rules = {"value": max_over_ranks((time.perf_counter() - t0) / max(1, min(args.steps, 2))), "unit": "s",
                     "d2h_rule_bytes_per_step_this_rank": rb,
                     "note": "as e2e, plus this rank's rule table (rule ids, confidence, lift: 24 B per rule) copied "
                             "to pinned host memory in 64 MB pieces"}
        del rules_pinned

    # ---- the dense level-2 path (tcgen05 int8 GEMM) on a workload that takes it: one extra,
    # small, separately reported measurement so that the tensor-core kernel has a number and a
    # parity check in every default run (config 3's one-hot matrix is 1 % dense and takes the
    # per-playlist route)
    dense = None
    if world == 1 and rank == 0 and not args.no_dense and name == "zipf_1Mx100k_s0.001":
        try:
            dname = "dense_200kx4k_s0.02"
            o2, i2, t2, s2, c2 = make_data(dname, shard=0)
            d_o2, d_i2 = torch.from_numpy(o2).to(device), torch.from_numpy(i2).to(device)
            dspaces = {"mine": miner.Workspace(device), "rules": miner.Workspace(device)}
            for _ in range(2):
                r2 = miner.mine_native(d_o2, d_i2, t2, s2, workspace=dspaces["mine"])
                ru2 = miner.generate_rules(r2, c2, workspace=dspaces["rules"])
                del r2, ru2
            torch.cuda.synchronize()
            _lib.profile_begin()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            for _ in range(args.steps):
                r2 = miner.mine_native(d_o2, d_i2, t2, s2, workspace=dspaces["mine"])
                ru2 = miner.generate_rules(r2, c2, workspace=dspaces["rules"])
                if _ < args.steps - 1:
                    del r2, ru2
            ev1.record()
            torch.cuda.synchronize()
            prof2 = _lib.profile_report()
            gemm = [(ms, w) for nm, lv, ms, w in prof2 if nm == "k_pair_gemm_i8"]
            got2 = {"level_sizes": [int(c.numel()) for _, c in r2.levels], "itemset_checksum": str(miner.itemset_checksum(r2)),
                    "n_rules": ru2.n_rules, "rule_checksum": str(miner.rule_checksum(r2, ru2))}
            want2 = golden_for(dname, s2, len(o2) - 1)
            dense = {"workload": dname, "pair_impl": r2.stats.get("pair_impl"), "ms_per_step": ev0.elapsed_time(ev1) / args.steps,
                     "parity": None if want2 is None else all(got2[k_] == want2[k_] for k_ in got2)}
            if gemm:
                ms_g, w_g = sum(g[0] for g in gemm) / len(gemm), sum(g[1] for g in gemm) / len(gemm)
                dense["k_pair_gemm_i8"] = {"ms": ms_g, "int8_ops": w_g, "achieved_tops": w_g / ms_g / 1e9,
                                           "peak_tops": int8_peak, "frac": w_g / ms_g / 1e9 / int8_peak,
                                           "peak_source": "2 x bf16_tflops of MEASURED_PEAKS.json",
                                           "ops_counted": "2 * N * F(F+1)/2: upper triangle incl. diagonal"}
            del r2, ru2, dspaces, d_o2, d_i2
            log(f"[bench] dense extra: {dense}")
        except Exception as e:      # noqa: BLE001 - the extra must never take the headline down
            dense = {"error": repr(e)}

    verify = None
    if args.verify and world == 1:
        from oracle import fpgrowth_c, mlxtend_restated as M
        res, rules = step(d_off, d_items)
        o_, i_, c_ = res.to_host()
        got = {"levels": [int(c.numel()) for _, c in res.levels], "checksum": fpgrowth_c.itemset_checksum(o_, i_, c_),
               "n_rules": rules.n_rules,
               "rule_checksum": 0 if args.no_rules else miner.rule_checksum(res, rules)}
        del res, rules
        t1_, tk_ = M.thresholds(s, n_local)
        t0 = time.time()
        ref = fpgrowth_c.fpgrowth(off, items, n_tracks, t1_, tk_, store=True,
                                  rules_min_confidence=None if args.no_rules else conf)
        want = {"levels": ref.level_counts[1:len(got["levels"]) + 1].tolist(), "checksum": ref.checksum,
                "n_rules": 0 if args.no_rules else ref.n_rules,
                "rule_checksum": 0 if args.no_rules else ref.rule_checksum}
        verify = {"ok": got == want, "gpu": {k: str(v) for k, v in got.items()},
                  "oracle": {k: str(v) for k, v in want.items()}, "oracle_seconds": time.time() - t0,
                  "oracle_threads": os.cpu_count()}
        log(f"[bench] verify: {verify}")

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        threads = os.cpu_count() or 1
        if name.startswith("mpd_standin"):
            sec, n_keys, n_sets = python_reference_run(name)
            cpu = {"value": sec, "unit": "s", "cores": 1, "kind": "port",
                   "sample": "the whole workload: Python restatement of mlxtend inside the reference's own "
                             f"glue (main.py:262-313), single thread as the reference runs; {n_sets} itemsets, "
                             f"{n_keys} keys; no rule table (the reference computes none)"}
        else:
            r = cpu_sample_run(name, threads)
            cpu = {"value": r["estimate_full_s"], "unit": "s", "cores": threads, "kind": "port",
                   "sample": r["sample"], "measured_sample_s": r["measured_s"]}

    if rank == 0:
        n, t, draws, alpha, _, _ = WORKLOADS[name]
        lv = info.get("stats", {}).get("levels", [])
        n_counts = sum(l.get("candidates", 0) for l in lv)
        line = {
            "metric": "mine+rules wall time", "value": step_s, "unit": "s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3,
            "higher_is_better": False, "scaling": "strong" if strong else "weak", "vs_baseline": None,
            "dtype": "u32+f64",
            "data": "synthetic",
            "config": {"workload": name, "n_playlists": n_total, "n_playlists_per_gpu": n_local, "n_tracks": n_tracks, "mean_draws": draws,
                       "zipf_alpha": alpha, "min_support": s, "min_confidence": None if args.no_rules else conf,
                       "parallelism": ("1 GPU" if world == 1 else
                                       (f"candidate space split by frequent pair over {world} ranks (every rank holds all "
                                        f"playlists; blocks of {args.part_block} pairs dealt round-robin), tables merged by "
                                        "one sum-allreduce, rule table split by itemset range"
                                        if not by_playlists else
                                        f"playlists sharded over {world} ranks, one count allreduce per level, rule table "
                                        "split by itemset range")),
                       "l2": ("inputs larger than L2 (CSR 0.5 GB, bitmap 0.75 GB); every step also writes several GB of "
                              "tables and rules, far more than the 126 MB L2, so each step starts cold"
                              if n_local >= 500_000 else
                              "small workload: inputs fit in L2, launch/latency-bound")},
            "clocks": clocks,
            "e2e": {"value": e2e_s, "unit": "s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "note": e2e_note},
            "e2e_with_rules": e2e_rules,
            "phases_ms_per_step": {k_: v_["ms"] / args.steps for k_, v_ in calls.items() if k_ in ("kmls_mine", "kmls_rules")},
            "gpu_launches": launches,
            "roofline": roof,
            "rooflines": {k: v for k, v in rooflines.items() if v},
            "cpu_baseline": cpu,
            "support_counts_per_s": n_counts / step_s if step_s else None,
            "itemsets": info.get("itemsets"), "rules_rank0": info.get("rules_this_rank"),
            "peak_hbm_gb": torch.cuda.max_memory_allocated() / 1e9,
            "level_sizes": info.get("levels"),
            "kernels": {k: {"ms_per_step": v["ms"] / args.steps, "calls_per_step": v["calls"] / args.steps}
                        for k, v in sorted(per.items(), key=lambda kv: -kv[1]["ms"])},
            "partition": info.get("stats", {}).get("partition") or ({"kind": "playlists", "world": world} if by_playlists else None),
            "merge_ms": info.get("merge_ms"),
            "top_kernel": top,
            "parity": parity,
            "dense_extra": dense,
            "verify_full_size": verify,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    if parity is not None and parity["ok"] is False:
        log("[bench] PARITY MISMATCH against the committed oracle result")
        sys.exit(3)


def main():
    import faulthandler
    faulthandler.enable()
    # The contract is ONE JSON line on stdout. Libraries write there too (NCCL prints its
    # version banner to stdout), so everything else is sent to stderr at the descriptor
    # level and only the JSON line goes to the real stdout.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="kmls_b200", choices=["kmls_b200", "reference"])
    ap.add_argument("--workload", default="zipf_1Mx100k_s0.001", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--trace", action="store_true", help="print every library call of the last timed step")
    ap.add_argument("--min-support", type=float, default=None, help="override the workload's min_support (sweeps)")
    ap.add_argument("--no-rules", action="store_true", help="mine only (sweeps at supports whose rule table explodes)")
    ap.add_argument("--verify", action="store_true",
                    help="after the run, mine the FULL workload with the CPU oracle (minutes) and compare level "
                         "sizes, the itemset checksum and the number of rules")
    ap.add_argument("--no-dense", action="store_true", help="skip the extra measurement of the dense (tensor-core) level-2 path")
    ap.add_argument("--no-parity", action="store_true", help="skip the comparison with the committed oracle result")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong (default): the job's playlists are the same at every --gpus N, sliced over the "
                         "ranks; weak: every rank generates its own set of the workload's size")
    ap.add_argument("--partition", default="auto", choices=["auto", "classes", "playlists"],
                    help="multi-GPU: classes (default) = every rank holds all playlists and mines its share of the "
                         "candidate space, no per-level exchange; playlists = row-sharded playlists with one count "
                         "allreduce per level")
    ap.add_argument("--part-block", type=int, default=64, help="frequent pairs per block of the class partition")
    ap.add_argument("--rules-to-host", action="store_true", help="also measure e2e with the rule table copied to the host")
    ap.add_argument("--mine-ws-gb", type=float, default=0.0, help="initial size of the mining workspace (0 = grow on demand)")
    ap.add_argument("--rules-ws-gb", type=float, default=0.0, help="initial size of the rules workspace (0 = grow on demand)")
    ap.add_argument("--no-prune", action="store_true", help="switch Apriori pruning of the joins off")
    ap.add_argument("--pair-impl", default="auto", choices=["auto", "i8", "popc", "csr"])
    ap.add_argument("--count-impl", default="projected", choices=["projected", "fullwidth"])
    args = ap.parse_args()
    global MIN_SUPPORT_OVERRIDE
    MIN_SUPPORT_OVERRIDE = args.min_support
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
