#!/usr/bin/env python
"""bench.py — transactions/sec mined (itemsets + rules + recommendations) on N B200s.

A step is one full pass of the hot path over this rank's shard: level-1 counts, vertical
bitmaps, pair pass, the Apriori level loop, rule generation / prune / sort and per-user
matching, with the frequent itemsets (+ supports) and the recommendations back on the host.

  value : inputs (CSR of dense item ids) already resident in HBM when the timed region starts
  e2e   : the same call with HOST (pinned) CSR buffers, H2D + D2H inside the timed region

Headline workload at every N: BASELINE.json configs[1] per GPU — T40I10D1M-style Quest rows, 1000 items,
min-support 0.5 % — so N GPUs mine N x 1M rows (weak scaling, one int32 all-reduce per level).

The other BASELINE.json configurations are measured in the same run and reported under "extras" (each with its
own timing, digest and device-time split; `--extras none` skips them):
  c3  dense pair stress, 10 M rows x 16384 items, pair pass only (tcgen05 int8 GEMM), rows sharded over the N GPUs
  c4  T20I6D100M, 10 k items, 100 M rows IN TOTAL sharded over the N GPUs (strong scaling: the >= 6x target)
  c5  50 M rows + 1 M users at 0.05 % support, sharded over the N GPUs

Every workload prints a digest of its result (itemsets + supports, rules, recommendations); tests/golden/digests.json
holds the digests the CPU oracle produced for the same seeds (tests/golden/make_digests.py), and the line says whether
they match, and whether all ranks agree.

`--impl reference` times the CPU restatement of the reference (oracle/, "port": the reference is Scala on Spark and
no JVM exists on this image) on a bounded sample of the headline workload, with all host threads.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def _host_threads() -> int:
    """Threads this process may use: the cores it can run on, shared between the ranks of this node."""
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
    return max(1, _cores() // max(local_world, 1))


# torchrun exports OMP_NUM_THREADS=1; the workload generator and the CPU baseline are OpenMP code that must use the
# host's cores, so the count is set explicitly (and printed) before any OpenMP runtime is loaded.  The reference arm
# runs on rank 0 alone and takes every core.
_IS_REF = "reference" in sys.argv
os.environ["OMP_NUM_THREADS"] = str(_cores() if _IS_REF else _host_threads())

import numpy as np  # noqa: E402

from fastapriori_b200.workloads import WORKLOADS, gen_rows, gen_users  # noqa: E402

METRIC = "transactions/sec mined (itemsets+rules)"
UNIT = "transactions/s"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=5)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--workload", default="c2", choices=sorted(WORKLOADS), help="headline workload (default: configs[1])")
    p.add_argument("--rows-per-gpu", type=int, default=0, help="0 = the workload's own size")
    p.add_argument("--users-per-gpu", type=int, default=-1, help="-1 = the workload's own size")
    p.add_argument("--extras", default="c3,c4,c5", help="comma list of c3,c4,c5 or 'none'")
    p.add_argument("--extra-scale", type=float, default=1.0, help="scale the extras' row/user counts (profiling and smoke runs)")
    p.add_argument("--cpu-sample-rows", type=int, default=0, help="rows of the CPU baseline sample (0 = auto)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--skip-e2e", action="store_true", help="profiling runs only: skip the host-buffer arm")
    p.add_argument("--no-digest", action="store_true", help="profiling runs only: skip the digest pass")
    a = p.parse_args()
    a.wl = dict(WORKLOADS[a.workload])
    if a.rows_per_gpu <= 0:
        a.rows_per_gpu = a.wl["rows"]
    if a.users_per_gpu < 0:
        a.users_per_gpu = a.wl["users"]
    a.extras = [] if a.extras in ("none", "") else [e for e in a.extras.split(",") if e]
    return a


def config_dict(wl, rows_per_gpu, users_per_gpu, world, scaling):
    return {
        "workload": f"{wl['name']} synthetic Quest (BASELINE.json {wl['cfg']})" + (" per GPU" if scaling == "weak" else " sharded over the GPUs"),
        "rows_per_gpu": rows_per_gpu, "rows_total": rows_per_gpu * world, "items": wl["n_items"],
        "avg_len": wl["avg_len"], "avg_pattern": wl["avg_pat"], "patterns": wl["n_pat"],
        "min_support": wl["min_support"], "users_per_gpu": users_per_gpu, "users_total": users_per_gpu * world, "seed": wl["seed"],
        "parallelism": f"row-sharded x{world}" if world > 1 else "single GPU",
        "l2": "inputs exceed L2 (CSR %.0f MB per GPU vs 126 MB); no explicit flush" % (rows_per_gpu * (wl["avg_len"] * 4 + 8) / 1e6),
    }


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.proc, self.lines = device, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(self.device)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---- CPU baseline / reference arm ------------------------------------------------------------------------------------
def cpu_run(wl, total_rows, sample_rows, sample_users):
    """ONE measured run of the oracle (CPU restatement of the reference) on the first `sample_rows` rows and
    `sample_users` users of the workload, same relative min-support, all host threads.  Returns
    (sample_rows / seconds, seconds, info): a plain measurement of the sample, nothing is scaled or extrapolated."""
    from oracle import oracle_api
    off, items = gen_rows(wl, total_rows, 0, sample_rows)
    uoff, uitems = gen_users(wl, max(sample_users, 1), 0, sample_users)
    # fast_rules / fast_cands: hashed subset lookups in genRules and a sibling join in genCandidates.  The reference's own
    # linear scans (AssociationRules.scala:136-143) and per-y probes (FastApriori.scala:176-187) do not finish on the
    # ~3 M itemsets this support yields; both flags only make the CPU arm FASTER than the reference's algorithm.  Counting
    # keeps the reference's data structures (weighted de-duplicated rows, one byte per boolean, a full scan per pair).
    r = oracle_api.run_csr(off, items, uoff, uitems, wl["min_support"], fast_rules=True, fast_cands=True, packed=False, light=True)
    secs = r.times["mine"] + r.times["rules"]
    info = {"itemsets": sum(len(c) for _, c in r.levels.values()), "rules": r.n_kept, "max_k": max(r.levels),
            "seconds": round(secs, 2), "omp_threads": oracle_api.threads(),
            "row_dependent_s": round(r.times["level1"] + r.times["bitmaps"] + r.times["pairs"] + r.times["count"], 2),
            "lattice_s": round(r.times["cand"] + r.times["rulegen"] + r.times["sort"], 2), "match_s": round(r.times["match"], 3)}
    return sample_rows / secs, secs, info


def _sample_text(wl, total_rows, total_users, rows, users, info):
    return (f"oracle (reference algorithm; flags fast_rules+fast_cands) on the first {rows} of {total_rows} rows + {users} of {total_users} users "
            f"of the same workload, min-support {wl['min_support']}; value = {rows} sample rows / measured mine+rules seconds of that run "
            f"(text parse excluded), NOT extrapolated to the full row count - the itemset lattice, hence most of the CPU time, does not shrink "
            f"with the sample; {info}")


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle_api
    wl = args.wl
    rows = args.cpu_sample_rows or 10000
    users = max(20, rows // 40)
    # every step is one oracle run on the bounded sample (~15-30 s: the lattice does not shrink with the sample), so the
    # number of runs is capped to keep the whole call within a few minutes; steps_run says how many were timed
    t0 = time.time()
    v, s, info = cpu_run(wl, args.rows_per_gpu, rows, users)
    per_run = time.time() - t0
    budget_runs = max(1, int(150.0 / max(per_run, 1.0)))
    steps_run = max(1, min(args.steps, budget_runs - 1)) if args.warmup > 0 else max(1, min(args.steps, budget_runs))
    vals, secs = ([], []) if args.warmup > 0 else ([v], [s])   # the first run is the warm-up when one was asked for
    while len(vals) < steps_run:
        v, s, info = cpu_run(wl, args.rows_per_gpu, rows, users)
        vals.append(v)
        secs.append(s)
    ms = 1e3 * float(np.mean(secs))
    value = rows / (ms / 1e3)
    cfg = config_dict(wl, args.rows_per_gpu, args.users_per_gpu, max(args.gpus, 1), "weak")
    cfg.update(sample_rows=rows, sample_users=users, sample_fraction=rows / args.rows_per_gpu, extrapolated=False)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "steps_run": steps_run, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32 counts / f64 confidence", "data": "synthetic", "config": cfg,
        "same_config": False,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": oracle_api.threads(), "kind": "port",
                         "sample": _sample_text(wl, args.rows_per_gpu, args.users_per_gpu, rows, users, info)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "CPU restatement of the reference (oracle/), not Spark: no JVM on this image. One step = one measured oracle run on the "
                "sample named in cpu_baseline.sample; value = sample rows / ms_per_step. The reference arm mines on the host cores "
                "whatever --gpus says.",
    }
    print(json.dumps(line), flush=True)


# ---- our arm -----------------------------------------------------------------------------------------------------------
def golden_digests():
    try:
        return json.load(open(os.path.join(ROOT, "tests", "golden", "digests.json")))
    except Exception:
        return {}


class Runner:
    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.args = torch, dist, args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device(f"cuda:{self.local}"))
        from fastapriori_b200 import MiningContext
        self.sc = MiningContext(device=self.local)
        self.eng = self.sc.engine
        self.golden = golden_digests()

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        t = self.torch.tensor([ms], device="cuda", dtype=self.torch.float64)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, fn, steps):
        torch = self.torch
        self.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(self.sc.stream)   # the stream every kernel of the context is launched on
        for _ in range(steps):
            fn()
        b.record(self.sc.stream)
        self.barrier()
        return self.max_over_ranks(a.elapsed_time(b))

    def gather_arrays(self, arr):
        if self.world == 1:
            return [arr]
        out = [None] * self.world
        self.dist.all_gather_object(out, arr)
        return out

    # one workload: generate this rank's shard, time `value` (resident) and `e2e` (pinned host), digest the result
    def run_mining(self, key, wl, rows_per_gpu, users_per_gpu, scaling, steps, warmup, with_clocks=True):
        torch = self.torch
        from fastapriori_b200 import AssociationRules, FastApriori, TokenizedLines
        from fastapriori_b200 import utils as Utils
        world, rank, sc, eng = self.world, self.rank, self.sc, self.eng
        rows_total, users_total = rows_per_gpu * world, users_per_gpu * world
        t_gen = time.time()
        n_items = wl["n_items"]
        # host copies live in pinned memory (e2e arm: async H2D straight from these buffers); generated in slabs so that the
        # generator's own buffers stay small next to the 100 M-row workloads
        lo, hi = rank * rows_per_gpu, (rank + 1) * rows_per_gpu
        slab = 8_000_000
        offs, idss, base = [np.zeros(1, dtype=np.int64)], [], 0
        for s0 in range(lo, hi, slab):
            o, it = gen_rows(wl, rows_total, s0, min(hi, s0 + slab))
            offs.append(o[1:] + base)
            it -= 1
            idss.append(it)
            base += int(o[-1])
        h_off = torch.empty(sum(len(o) for o in offs), dtype=torch.int64, pin_memory=True)
        h_ids = torch.empty(sum(len(i) for i in idss), dtype=torch.int32, pin_memory=True)
        p = 0
        while offs:
            o = offs.pop(0)
            h_off.numpy()[p:p + len(o)] = o
            p += len(o)
        p = 0
        while idss:
            i = idss.pop(0)
            h_ids.numpy()[p:p + len(i)] = i
            p += len(i)
        vocab = [str(i) for i in range(1, n_items + 1)]
        data = TokenizedLines(h_off.numpy(), h_ids.numpy(), vocab)
        # the host-buffer arm ships the ids as the host would keep them for a vocabulary this small: 16 bits each
        # (fa_load_transactions_u16 widens them on the device); offsets stay int64
        data_host, h_ids16 = data, None
        if n_items <= 65536:
            h_ids16 = torch.empty(h_ids.numel(), dtype=torch.int16, pin_memory=True)
            data_host = data.compact(out=h_ids16.numpy().view(np.uint16))
        uoff, uitems = gen_users(wl, users_total, rank * users_per_gpu, (rank + 1) * users_per_gpu)
        users = TokenizedLines.from_int_csr(uoff, uitems, n_items)
        d_off, d_ids = h_off.cuda(), h_ids.cuda()
        resident = (d_off.data_ptr(), d_ids.data_ptr())
        t_gen = time.time() - t_gen
        info, last = {}, {}

        def step(res, fetch):
            fa = FastApriori(wl["min_support"], 0)
            fis, i2r, fi = fa.run(sc, data if res is not None else data_host, fetch=fetch, resident=res)
            ar = AssociationRules(fis, fi, i2r)
            rec = ar.run(sc, users, as_arrays=True)
            n_sets = len(fi) + sum(v["frequent"] for k, v in fa.counters.items() if isinstance(k, int))
            info.update(itemsets=n_sets, max_k=fa.counters["max_k"], rules=ar.counters.get("rules", 0), raw_rules=ar.counters.get("raw", 0),
                        frequent_items=len(fi), recommended=int((rec >= 0).sum()))
            if fetch:
                info["d2h"] = sum(s.nbytes + c.nbytes for s, c in fis.levels.values()) + rec.nbytes
            last.update(fis=fis, fi=fi, rec=rec)
            return rec

        # The job's result — itemsets + supports on the host, ready to be written — is produced once, by rank 0 (the rank
        # that writes freqItemset, Utils.saveFreqItemset); the other ranks hold identical replicas on their devices and
        # only return their users' recommendations.
        fetch = rank == 0
        for _ in range(warmup):
            step(resident, fetch)
        eng.reset_stats()
        sampler = ClockSampler(self.local).start() if (rank == 0 and with_clocks) else None
        ms_total = self.timed(lambda: step(resident, fetch), steps)
        st = eng.stats()
        clocks = sampler.stop() if sampler else None
        if self.args.skip_e2e:
            ms_e2e = float("nan")
        else:
            step(None, fetch)  # warm the host-buffer path once
            ms_e2e = self.timed(lambda: step(None, fetch), steps)
        out = {
            "value": rows_total * steps / (ms_total / 1e3), "unit": UNIT, "ms_per_step": ms_total / steps, "steps": steps, "warmup": warmup,
            "scaling": scaling, "config": dict(config_dict(wl, rows_per_gpu, users_per_gpu, world, scaling), **info),
            "e2e": {"value": rows_total * steps / (ms_e2e / 1e3), "unit": UNIT, "ms_per_step": ms_e2e / steps,
                    "h2d_bytes_per_step": int(h_off.numel() * 8 + data_host.ids.nbytes + users.offsets.nbytes + users.ids.nbytes),
                    "ids_dtype": str(data_host.ids.dtype),
                    "d2h_bytes_per_step": int(info.get("d2h", last["rec"].nbytes))},
            "gpu_launches": int(st["kernel_launches"]),
            "device_ms_per_step": {k: st[k] / steps for k in st if k.startswith("ms_")},
            "generate_s": round(t_gen, 1), "clocks": clocks, "stats": st,
        }
        if not self.args.no_digest:
            # untimed pass with the itemsets fetched on EVERY rank: all ranks must hold the same lattice and rules
            step(resident, True)
            fis, fi, rec = last["fis"], last["fi"], last["rec"]
            recs = self.gather_arrays(rec)
            dig = Utils.runDigest(fis.levels, fi, np.concatenate(recs), eng.rules_digest())
            digs = self.gather_arrays(dig["all"])
            gkey = f"{key}:{rows_total}:{users_total}"
            out["digest"] = dict(dig, ranks_equal=len(set(digs)) == 1, golden_key=gkey,
                                 matches_golden=(dig["all"] == self.golden[gkey]["all"]) if gkey in self.golden else None)
        last.clear()
        del d_off, d_ids, h_off, h_ids, data, data_host, h_ids16
        torch.cuda.empty_cache()
        return out

    # c3: the pair pass alone on device-synthesised rows (text or host CSR of 10 M x 327 items would be 13 GB)
    def run_c3(self, rows_total, n_items=16384, density=0.02, steps=2, warmup=1):
        torch = self.torch
        world, rank, eng = self.world, self.rank, self.eng
        rows = rows_total // world
        per_row = max(2, int(n_items * density))
        g = torch.Generator(device="cuda").manual_seed(3 + 7919 * rank)
        # fixed row length keeps the generator trivial; duplicates inside a row are legal (they collapse in the bitmap)
        ids = torch.randint(0, n_items, (rows * per_row,), dtype=torch.int32, device="cuda", generator=g)
        off = torch.arange(0, rows + 1, dtype=torch.int64, device="cuda") * per_row
        res = {}
        eng.attach_transactions(off.data_ptr(), ids.data_ptr(), rows, n_items)
        counts = eng.count_items()
        order = np.argsort(-counts.astype(np.int64), kind="stable").astype(np.int32)
        id_to_rank = np.empty(n_items, dtype=np.int32)
        id_to_rank[order] = np.arange(n_items, dtype=np.int32)
        eng.build_bitmaps(id_to_rank, counts[order])
        min_count = int(rows_total * density * density * 1.05)  # a little above the expected pair support

        def one():
            res["n2"] = eng.count_pairs(min_count)

        for _ in range(warmup):
            one()
        eng.reset_stats()
        ms = self.timed(one, steps)
        st = eng.stats()
        ops = 2.0 * rows_total * n_items * n_items
        nt = -(-n_items // 256)
        computed = 2.0 * rows_total * (nt * (nt + 1) // 2) * 256.0 * 256.0   # upper-triangle tiles of 256 x 256 items (k_pair_tc4)
        peak = 4500.0  # TOP/s, nominal dense int8 of B200; cuBLASLt int8 8192^3 measured on this pool: 3237 burst (profiles/int8_peak.py)
        ach = computed / world / (st["ms_pairs"] / steps * 1e-3) / 1e12
        out = {"value": rows_total * steps / (ms / 1e3), "unit": UNIT, "ms_per_step": ms / steps, "steps": steps, "warmup": warmup,
               "scaling": "strong",
               "config": {"workload": "dense 2-itemset stress (BASELINE.json configs[2]): pair pass only, rows sharded over the GPUs",
                          "rows_total": rows_total, "rows_per_gpu": rows, "items": n_items, "density": density, "nnz_per_gpu": int(ids.numel()),
                          "min_count": min_count, "frequent_pairs": int(res["n2"]), "data": "Bernoulli rows synthesised on the device"},
               "ms_pairs_kernel_per_step": st["ms_pairs"] / steps, "ms_exchange_per_step": st["ms_exchange"] / steps,
               "int8_TOPs_nominal_2NF2": ops / (ms / steps * 1e-3) / 1e12,
               "roofline": {"kernel": "k_pair_tc4 (K3: tcgen05 cta_group::2 int8 GEMM, A in TMEM, 256 x 256 items per CTA pair)", "bound": "tensor",
                            "achieved": ach, "peak": peak, "unit": "TOP/s", "frac": ach / peak, "traffic": None,
                            "peak_source": "nominal dense int8 (4500 TOP/s); the cuBLASLt int8 GEMM measured on this pool reaches 3237 (profiles/int8_peak.py), "
                                           "which this kernel exceeds, so the nominal figure is the denominator",
                            "frac_of_measured_cublaslt_int8": ach / 3237.0,
                            "note": "achieved = int8 ops this GPU actually issues (upper-triangle tiles) / K3 kernel time"},
               "bitmap_GB_per_gpu": st["bitmap_bytes"] / 1e9}
        if not self.args.no_digest:
            from fastapriori_b200 import utils as Utils
            # independent recount of 16 sampled pairs (8 frequent, 8 arbitrary) straight from the CSR of every rank, summed
            # over the ranks, against the all-reduced pair matrix the tensor-core kernel produced
            ptr, F = eng.pair_counts_device()

            class _Dev:
                def __init__(self, p, n):
                    self.__cuda_array_interface__ = {"shape": (n, n), "typestr": "<i4", "data": (p, False), "version": 3, "strides": None}

            P = torch.as_tensor(_Dev(ptr, F), device="cuda")
            rk = torch.from_numpy(id_to_rank).cuda()
            rng = np.random.default_rng(5)
            sets2_all, _ = eng.get_level(2)
            pairs = [tuple(sets2_all[i].tolist()) for i in rng.choice(len(sets2_all), size=min(8, len(sets2_all)), replace=False).tolist()]
            pairs += [tuple(sorted(rng.choice(n_items, size=2, replace=False).tolist())) for _ in range(8)]
            local = torch.zeros(len(pairs), dtype=torch.int64, device="cuda")
            step_rows = 1 << 20
            for r0 in range(0, rows, step_rows):   # in slabs: the int64 index copy of the whole CSR would be 26 GB
                r1 = min(rows, r0 + step_rows)
                rows_rk = rk[ids[r0 * per_row:r1 * per_row].long()].view(r1 - r0, per_row)
                local += torch.stack([((rows_rk == a).any(1) & (rows_rk == b).any(1)).sum() for a, b in pairs]).to(torch.int64)
            del rows_rk
            if world > 1:
                self.dist.all_reduce(local)
            got = torch.stack([P[a, b] for a, b in pairs]).to(torch.int64)
            out["recount"] = {"pairs_checked": len(pairs), "ok": bool(torch.equal(local, got)), "max_support_checked": int(local.max())}
            sets2, counts2 = sets2_all, eng.get_level(2)[1]
            dig = Utils.levelsDigest({2: (sets2, counts2)})
            digs = self.gather_arrays(dig)
            out["digest"] = {"pairs": dig, "ranks_equal": len(set(digs)) == 1,
                             "note": "rows are seeded per rank, so the digest is comparable across the ranks of one run only; "
                                     "tests/test_gpu_sizes.py checks the FULL pair matrix at 16384 items against an int32 matrix product"}
        del ids, off
        torch.cuda.empty_cache()
        return out


def roofline_k4(st, steps, peaks, l1_peak):
    ms = st["ms_count"]
    achieved = st["count_bytes"] / (ms / 1e3) / 1e9 if ms > 0 else 0.0
    loaded = st["count_load_bytes"] / (ms / 1e3) / 1e9 if ms > 0 else 0.0
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "k4_traffic.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    n = max(st["count_launches"], 1)
    dram = traffic.get("dram_bytes_per_step") if traffic else None
    r = {
        "kernel": "K4 support counting (prefix AND + popcount)", "bound": "l1",
        "achieved": loaded, "peak": l1_peak, "unit": "GB/s", "frac": loaded / l1_peak if l1_peak else None,
        "peak_source": "measured live: fa_probe_l1_bandwidth (16-byte read-only loads that all hit in L1, 2 CTAs x 8 warps per SM)",
        "what": "bytes of item rows the kernel requests (every load counted, prefix re-use already taken out) / K4 time; ncu shows "
                "the L1 data pipe as the busiest unit, the vertical database is L2-resident and HBM idles",
        "algorithmic": {"achieved": achieved, "unit": "GB/s", "bytes_per_launch": st["count_bytes"] / n,
                        "definition": "W*sum_groups(k-1+|ys|)+4*C_k per level (SURVEY.md 8d): the reference's own access pattern, "
                                      "of which prefix re-use skips a part; not a physical rate"},
        "ms_per_launch": ms / n, "launches": st["count_launches"], "traffic": dram,
        "hbm": {"peak": hbm_peak, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s",
                "dram_bytes_per_step": dram,
                "dram_gbs": (dram / (ms / steps * 1e-3) / 1e9) if dram and ms > 0 else None,
                "source": traffic.get("source") if traffic else None},
    }
    if r["hbm"]["dram_gbs"] is not None:
        r["hbm"]["frac"] = r["hbm"]["dram_gbs"] / hbm_peak
    return r


def run_ours(args):
    R = Runner(args)
    wl = args.wl
    head = R.run_mining(args.workload, wl, args.rows_per_gpu, args.users_per_gpu, "weak", args.steps, args.warmup)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    l1_peak = R.eng.probe_l1_bandwidth()
    st = head.pop("stats")
    extras = {}
    scale = args.extra_scale
    for e in args.extras:
        t0 = time.time()
        try:
            if e == "c3":
                extras[e] = R.run_c3(int(10_000_000 * scale))
            elif e in ("c4", "c5"):
                w = dict(WORKLOADS[e])
                rows = int(w["rows"] * scale) // R.world
                users = int(w["users"] * scale) // R.world
                # (configs[3] steps are short — 21 ms at eight GPUs — so it gets more of them than the 0.7 s steps of configs[4])
                n_steps, n_warm = (5, 2) if e == "c4" else (2, 1)
                x = R.run_mining(e, w, rows, users, "strong", steps=n_steps, warmup=n_warm, with_clocks=False)
                xs = x.pop("stats")
                x["roofline"] = roofline_k4(xs, n_steps, peaks, l1_peak)
                x["roofline"]["traffic"] = x["roofline"]["hbm"]["dram_bytes_per_step"] = x["roofline"]["hbm"]["dram_gbs"] = None
                extras[e] = x
            else:
                extras[e] = {"error": "unknown extra"}
        except Exception as ex:  # an extra must not take the headline down with it
            extras[e] = {"error": f"{type(ex).__name__}: {ex}"[:400]}
            R.torch.cuda.empty_cache()
        extras[e]["wall_s"] = round(time.time() - t0, 1)
    if R.rank != 0:
        if R.world > 1:
            R.dist.destroy_process_group()
        return
    line = {
        "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": R.world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32 bitmaps / int32 counts / f64 confidence", "data": "synthetic",
        "config": head["config"], "roofline": roofline_k4(st, args.steps, peaks, l1_peak), "e2e": head["e2e"],
        "gpu_launches": head["gpu_launches"], "device_ms_per_step": head["device_ms_per_step"], "clocks": head["clocks"],
        "digest": head.get("digest"), "host_threads": int(os.environ["OMP_NUM_THREADS"]), "extras": extras,
    }
    if R.world == 1 and not args.no_cpu_baseline:
        from oracle import oracle_api
        rows = args.cpu_sample_rows or 10000
        users = max(20, rows // 40)
        v, secs, cinfo = cpu_run(wl, args.rows_per_gpu, rows, users)
        line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": oracle_api.threads(), "kind": "port",
                                "sample": _sample_text(wl, args.rows_per_gpu, args.users_per_gpu, rows, users, cinfo)}
    print(json.dumps(line), flush=True)
    if R.world > 1:
        R.dist.destroy_process_group()
    R.sc.close()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
