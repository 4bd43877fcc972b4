#!/usr/bin/env python
"""Benchmark of the uvaia n-best database search hot path on B200 (BASELINE.json metric).

A "step" = one complete search of all queries against the whole device-resident packed database (K2 pair counts +
K3 replay over every round), i.e. BASELINE.json configs[1]: 1k synthetic SARS-CoV-2 queries vs a 1M-sequence
synthetic 29,903-site reference alignment, nbest=100.  `value` = queries/s with the packed DB and the query planes
already in HBM; `e2e` = the same search driven through the C ABI from HOST buffers (query text + reference text in
pinned memory -> H2D -> encode -> search -> heaps D2H inside the timed region).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--refs R] [--queries Q]
For N>1 launch with torchrun (one rank per GPU); the database is row-sharded, rounds are merged with NCCL all-gathers.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NCHAR = 29903
# result checksums (sum over every table of the six counts and the reference ordinals) of single-GPU runs, asserted at any N
KNOWN_CHECKSUMS = {(1000, 1_000_000, 29903, 100, 0.0, 200): 31652904200}
# algorithmic instruction counts per (pair, 32-site word) of the two pair-count kernels
OPS = {"pivot": {"kernel": "k_pairpivot", "lop3": 2.0, "popc": 2.0, "db_bytes_per_word": 9,      # per pair-word 2 AND + 2 POPC (+ 2 IMAD on the FMA pipe): every counted word
                                                                                                  # gets its own POPC (the first half of the round used a carry-save step:
                                                                                                  # 4 LOP3 + 1 POPC); staged: planes + counts, 9 KB per 128 references x 8 words
                 "ncu": "ncu_pairpivot_r02bb_four_refs_per_lane.json"},
       "general": {"kernel": "k_paircount", "lop3": 4.5, "popc": 1.25, "db_bytes_per_word": 12,  # 3 for valid+match, 1.5 carry-save tree; 1 + 0.25 POPC
                   "ncu": "ncu_paircount_r01e_csa.json"}}


def ncu_traffic(name):
    """DRAM bytes (read + write) of one launch of the dominant kernel from the committed `ncu --set full` capture"""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", name)))
        unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        rd, wr = d["dram__bytes_read.sum"], d["dram__bytes_write.sum"]
        grid = int(float(d["launch__grid_size"][0]))
        return {"bytes": float(rd[0]) * unit[rd[1]] + float(wr[0]) * unit[wr[1]], "launch_grid_blocks": grid, "source": "profiles/" + name}
    except Exception:
        return None


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--refs", type=int, default=1_000_000)
    ap.add_argument("--queries", type=int, default=1000)
    ap.add_argument("--nbest", type=int, default=100)
    ap.add_argument("--nchar", type=int, default=NCHAR)
    ap.add_argument("--ambiguous", type=float, default=0.0, help="config-4 stress: fraction of ambiguous query sites")
    ap.add_argument("--early-abort", type=int, default=0)
    ap.add_argument("--max-round", type=int, default=0)
    ap.add_argument("--pair-kernel", default="auto", choices=["auto", "general", "pivot"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-packed", action="store_true", help="skip the packed-cache variant of the e2e leg")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-refs", type=int, default=3000, help="references in the bounded CPU-baseline sample")
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4, 5],
                    help="BASELINE.json configs[N-1] at its stated shape (default 2 = the metric's configuration; 1 = bundled data is a golden test, tests/golden/cases/real_c1_full)")
    ap.add_argument("--clades", type=int, default=200)
    ap.add_argument("--closed-form", action="store_true", help="single GPU: pool-independent closed form + H1 flags instead of the literal --pool batches")
    a = ap.parse_args()
    given = {x.split("=")[0] for x in sys.argv[1:] if x.startswith("--")}
    preset = {2: {}, 3: {"queries": 10000, "refs": 5_000_000}, 4: {"ambiguous": 0.2},
              5: {"nchar": 5_000_000, "refs": 20000, "clades": 20}}[a.config]
    for k, v in preset.items():
        if "--" + k not in given:
            setattr(a, k, v)
    return a


# ---------------------------------------------------------------------------------------------------------------
# clocks sampler (B200_PROFILING.md "clocks DURING the timed region")
# ---------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # median over the samples taken under load (power above the idle floor)
        load = [s for s, p in zip(sm, pw) if p > 0.5 * max(pw)] or sm
        return {"sm_mhz": float(np.median(load)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "power_w_max": max(pw)}


# ---------------------------------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------------------------------
def make_model(nchar, n_clades=200):
    from uvaia_b200.synth import SynthModel
    m = SynthModel(nsites=nchar, n_clades=n_clades, seed=1)
    return np.stack(m.clades)


def make_queries(eng, clades, nq, nchar, ambiguous):
    """synthetic query rows from the device generator (seed 2002), filtered and ordered like the uvaia host does"""
    from uvaia_b200 import host
    d = eng.device_alloc(nq * nchar)
    eng.synth_rows_device(clades, 2002, 0, nq, d, ambiguous_frac=ambiguous)
    text = eng.d2h(d, nq * nchar).reshape(nq, nchar)
    eng.device_free(d)
    keep = host.filter_queries([r.tobytes() for r in text], 0.5)
    text = text[keep]
    order = host.query_order(text, 0, False)
    return np.ascontiguousarray(text[order]), [b"Q%06d" % keep[i] for i in order]


from uvaia_b200.shard import rank_tiles, rounds as shard_rounds  # noqa: E402  (round schedule of the row sharding, twin of uvg_shard_of)


# ---------------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own multithreaded search (oracle/_ref/uvaia_ref) on a bounded sample of the workload
# ---------------------------------------------------------------------------------------------------------------
def write_fasta(path, names, text):
    with open(path, "wb") as fh:
        for nm, row in zip(names, text):
            fh.write(b">" + nm + b"\n" + row.tobytes() + b"\n")


def run_reference_sample(qtext, qnames, rtext, nbest, threads, workdir):
    """-> (seconds, csv_text or None, kind).  kind 'reference' = compiled reference binary, 'port' = numpy oracle."""
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "uvaia_ref")
    rnames = [b"R%07d" % i for i in range(len(rtext))]
    if os.path.exists(ref_bin):
        import gzip
        qf, rf, pre = os.path.join(workdir, "q.fa"), os.path.join(workdir, "r.fa"), os.path.join(workdir, "out")
        write_fasta(qf, qnames, qtext)
        write_fasta(rf, rnames, rtext)
        cmd = [ref_bin, "-n", str(nbest), "-t", str(threads), "-A", "1.0", "-r", rf, "-o", pre, qf]
        # torchrun exports OMP_NUM_THREADS=1 to its workers and the reference clamps -t to omp_get_max_threads()
        # (nearest.c:183-189): give the reference all the cores explicitly
        env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_DYNAMIC="false")
        env.pop("OMP_THREAD_LIMIT", None)
        t0 = time.perf_counter()
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, cwd=workdir, env=env)
        dt = time.perf_counter() - t0
        if p.returncode != 0:
            raise RuntimeError("uvaia_ref failed: " + p.stderr.decode(errors="replace")[-400:])
        raw = open(pre + ".csv.xz", "rb").read()
        try:
            csv = gzip.decompress(raw).decode()           # zlib-only build writes gzip data under the .xz name (H4)
        except OSError:
            import lzma
            csv = lzma.decompress(raw).decode()
        return dt, csv, "reference"
    from oracle import uvaia_oracle as orc
    t0 = time.perf_counter()
    res = orc.search(list(qnames), [r.tobytes() for r in qtext], None, nbest=nbest, ambig_r=1.0,
                     refs=[(rnames, [r.tobytes() for r in rtext])])
    return time.perf_counter() - t0, res.csv(), "port"


def gpu_csv_for_sample(qtext, qnames, rtext, nbest):
    from uvaia_b200 import host
    from uvaia_b200.engine import Engine
    with Engine(qtext.shape[1], nbest=nbest) as e:
        e.set_queries(qtext)
        e.push_refs(rtext, 0)
        e.search()
        out = host.SearchOutput()
        out.qnames = list(qnames)
        out.rows = e.heaps()
        out.ref_names = [b"R%07d" % i for i in range(len(rtext))]
        return out.csv()


def reference_timing(qtext, qnames, rtext, n_total_refs, nbest, threads, reps=1, budget_s=240.0, warmup=0):
    """The reference's own multithreaded search (oracle/_ref/uvaia_ref, else the numpy port) on a bounded sample of the
    workload: the same queries x the first n2 = len(rtext) references and x the first n1 = n2 // 3.  The full-database time
    is extrapolated from the MARGINAL cost per reference, (t2 - t1) / (n2 - n1): process start, query read / filter /
    index, output compression and the expensive start-up phase of the search (heaps still filling, nothing truncates)
    are in t1 and are not multiplied.  -> (dict for the JSON line, csv of the n2 run, seconds per timed rep)"""
    n2 = len(rtext)
    n1 = max(1, n2 // 3)
    t1s, t2s, csv, kind = [], [], None, None
    with tempfile.TemporaryDirectory() as wd:
        for _ in range(warmup):
            run_reference_sample(qtext, qnames, rtext, nbest, threads, wd)
        t_start = time.perf_counter()
        for i in range(reps):
            d1, _, kind = run_reference_sample(qtext, qnames, rtext[:n1], nbest, threads, wd)
            d2, csv, kind = run_reference_sample(qtext, qnames, rtext, nbest, threads, wd)
            t1s.append(d1); t2s.append(d2)
            if (time.perf_counter() - t_start) * (i + 2) / (i + 1) > budget_s:
                break
    t1, t2 = float(np.mean(t1s)), float(np.mean(t2s))
    marginal = (t2 - t1) / max(1, n2 - n1)
    if marginal <= 0:                                     # noise on a tiny sample: plain proportional scaling
        marginal = t2 / n2
    full_s = t2 + marginal * max(0, n_total_refs - n2)
    Q, L = qtext.shape
    return ({"value": Q / full_s, "unit": "queries/s", "cores": threads if kind == "reference" else 1, "kind": kind,
             "sample": f"{Q} queries x the first {n1} and the first {n2} of {n_total_refs} references ({t1:.2f} s and {t2:.2f} s per run, {len(t2s)} runs each); "
                       f"LINEAR EXTRAPOLATION to {n_total_refs} references from the marginal cost {marginal * 1e3:.3f} ms per reference "
                       f"(fixed part {t1 - marginal * n1:.2f} s: start-up, query preparation, output); the reference's timed run includes reading its FASTA input",
             "seconds_sample": t2, "seconds_full_extrapolated": full_s, "marginal_ms_per_ref": marginal * 1e3,
             "gsite_cmp_per_s": Q * L / marginal / 1e9, "runs": len(t2s)}, csv, [a + b for a, b in zip(t1s, t2s)])


def cpu_baseline(eng, clades, qtext, qnames, args):
    """our arm: the bounded reference run beside the GPU number, with the reference's table diffed against the CUDA path"""
    L = args.nchar
    n = min(args.cpu_refs, args.refs)
    d = eng.device_alloc(n * L)
    eng.synth_rows_device(clades, 1001, 0, n, d)
    rtext = eng.d2h(d, n * L).reshape(n, L)
    eng.device_free(d)
    cb, csv, _ = reference_timing(qtext, qnames, rtext, args.refs, args.nbest, os.cpu_count() or 1)
    cb["parity_with_gpu_on_sample"] = bool(gpu_csv_for_sample(qtext, qnames, rtext, args.nbest) == csv)
    return cb


def reference_arm(args):
    """--impl reference: rank 0 alone, no GPU, no product library -- inputs from the numpy twin of the device generator
    (uvaia_b200/synth.py, same population model and seeds), the timed call is the stock reference binary"""
    from uvaia_b200 import host
    from uvaia_b200.synth import SynthModel
    L, R = args.nchar, args.refs
    m = SynthModel(nsites=L, n_clades=args.clades, seed=1)
    _, qraw = m.sample(args.queries, 2002, ambiguous_frac=args.ambiguous, prefix="Q")
    keep = host.filter_queries([r.tobytes() for r in qraw], 0.5)
    qtext = qraw[keep]
    order = host.query_order(qtext, 0, False)
    qtext, qnames = np.ascontiguousarray(qtext[order]), [b"Q%06d" % keep[i] for i in order]
    n = min(args.cpu_refs, R)
    _, rtext = m.sample(n, 1001, prefix="R")
    threads = os.cpu_count() or 1
    cb, _, per_rep = reference_timing(qtext, qnames, rtext, R, args.nbest, threads, reps=max(1, args.steps), warmup=min(1, args.warmup))
    Q = len(qtext)
    v = cb["value"]
    line = {"metric": "queries_per_s_vs_%dk_db" % (R // 1000), "unit": "queries/s", "impl": "reference", "value": v, "n_gpus": args.gpus,
            "steps": cb["runs"], "warmup": min(1, args.warmup), "ms_per_step": float(np.mean(per_rep)) * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "int32 counts (the reference's arithmetic)", "data": "synthetic (uvaia_b200/synth.py, seeds 1/1001/2002)",
            "config": {"workload": f"{Q} synthetic SARS-CoV-2 queries (of {args.queries} generated; -a 0.5 filter) x {R} references x {L} sites, nbest={args.nbest}",
                       "nbest": args.nbest, "threads": threads, "omp_num_threads_passed": threads},
            "cpu_baseline": cb, "gsite_cmp_per_s": cb["gsite_cmp_per_s"],
            "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ---------------------------------------------------------------------------------------------------------------
# main
# ---------------------------------------------------------------------------------------------------------------
def load_peaks():
    peaks = {"hbm_gbs": 6650.0, "hbm_src": "fallback"}
    try:
        m = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        peaks["hbm_gbs"], peaks["hbm_src"] = float(m["hbm_gbs"]), "measured"
    except Exception:
        pass
    try:
        u = json.load(open(os.path.join(ROOT, "profiles", "int_pipes_b200_r01.json")))
        peaks["lop3_gips"] = float(u["lop3"]["gthread_inst_per_s"])
        peaks["popc_gips"] = float(u["popc"]["gthread_inst_per_s"])
    except Exception:
        peaks["lop3_gips"], peaks["popc_gips"] = 18400.0, 4630.0
    return peaks


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return reference_arm(args) if rank == 0 else 0      # rank 0 alone runs the CPU arm; no GPU, no product library
    import torch
    dist = None
    if world > 1 and args.impl == "ours":
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))
    from uvaia_b200.engine import Engine

    L, R = args.nchar, args.refs
    # small first rounds only serve the early-abort bounds (they fill fast); without early abort start with full waves
    first_round, max_round = (1024 if args.early_abort else 16384), (args.max_round or 65536)
    clades = make_model(L, args.clades)
    dev = local_rank if torch.cuda.is_available() else 0
    tiles = rank_tiles(R, rank, world, first_round, max_round)
    n_local = sum(c for _, c in tiles)
    # --pool of the reference program on this box: 64 x omp_get_max_threads() (nearest.c:81); single GPU: the search emulates
    # those batches literally (hazard H1), sharded contexts run the pool-independent closed form and count the flags
    pool = 64 * (os.cpu_count() or 1)
    eng = Engine(L, nbest=args.nbest, device=dev, first_round=first_round, max_round=max_round, early_abort=bool(args.early_abort), pool=pool, closed_form=bool(args.closed_form),
                 rank=rank, world=world, db_capacity=n_local,
                 pair_kernel={"auto": 0, "general": 1, "pivot": 2}[args.pair_kernel])
    qtext, qnames = make_queries(eng, clades, args.queries, L, args.ambiguous)
    Q = len(qtext)
    workload = f"{Q} synthetic SARS-CoV-2 queries (of {args.queries} generated; -a 0.5 filter) x {R} references x {L} sites, nbest={args.nbest}"
    base = {"metric": "queries_per_s_vs_%dk_db" % (R // 1000), "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32 bit-planes / int32 counts",
            "data": "synthetic (device generator, seeds 1/1001/2002)",
            "config": {"workload": workload, "nbest": args.nbest, "pool": pool, "h1": "literal --pool batches" if world == 1 and not args.closed_form else "closed form + flags", "early_abort": bool(args.early_abort), "sharding": "rows x%d" % world,
                       "l2": "inputs larger than L2 (packed DB %.1f GB; the pivot form streams planes + counts of its derived database, %.1f GB of %.1f GB, once per pivot tile)"
                             % (R * ((L + 255) // 256 * 256) * 3 / 8 / 1e9, R * ((L + 255) // 256 * 256) / 32 * 9 / 1e9, R * ((L + 255) // 256 * 256) / 32 * 25 / 1e9)}}

    # ---- build the device-resident database (K1), optionally keeping the text on the host for the e2e leg ----
    want_e2e = not args.no_e2e
    host_buf = None
    if want_e2e:
        import psutil
        need = n_local * L
        if psutil.virtual_memory().available < need * 1.5 + (8 << 30):
            want_e2e = False
        else:
            host_buf = eng._lib.uvg_host_alloc_pinned(need)
            want_e2e = bool(host_buf)
    eng.set_queries(qtext)
    tile_rows = max(64, min(16384, (1 << 30) // L))            # <= 1 GiB of generated text at a time (config 5: 5 Mb rows)
    d_tile = eng.device_alloc(tile_rows * L)
    t_build = time.perf_counter()
    off = 0
    for gb, cnt in tiles:
        for s in range(0, cnt, tile_rows):
            m = min(tile_rows, cnt - s)
            eng.synth_rows_device(clades, 1001, gb + s, m, d_tile)
            eng.push_refs_device(d_tile, m, 0)
            if want_e2e:
                eng.d2h_into(host_buf + off * L, d_tile, m * L)
            off += m
    eng.device_free(d_tile)
    t_build = time.perf_counter() - t_build
    assert eng.db_count == n_local
    if world > 1:
        uid = [Engine.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        eng.nccl_init(uid[0])
        eng.set_global_count(R)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- timed: device-resident search ----
    tab_items, tab_rows = np.zeros((eng.n_query, eng.K), Engine.ITEM_DTYPE), np.zeros(eng.n_query, np.int32)      # result buffers of the serving loop
    for _ in range(args.warmup):                               # whole steps: the first table read allocates the library's pinned result buffer
        eng.reset_search(); eng.search(); eng.heaps_raw(tab_items, tab_rows)
    sampler = ClockSampler(dev)
    barrier()
    sampler.start()
    ev_ms, pair_ms, replay_ms, launches, chunks, tables_ms = 0.0, 0.0, 0.0, 0, 0, 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        eng.reset_search(); eng.search()
        t_tab = time.perf_counter()
        eng.heaps_raw(tab_items, tab_rows)                  # the step ends with the final tables on the HOST, in the caller's buffers
        tables_ms += (time.perf_counter() - t_tab) * 1e3
        st = eng.stats()
        ev_ms += st["search_ms"]; pair_ms += st["pair_kernel_ms"]; replay_ms += st["replay_ms"]
        launches += st["pair_kernel_launches"] + st["other_kernel_launches"]
        chunks += st["warp_chunks_executed"]
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    # step time = CUDA-event time of the search (first pair-count launch -> last replay / final gathers) + host time of the tables
    tt = torch.tensor([ev_ms + tables_ms, wall * 1e3, pair_ms, tables_ms], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    step_ms_total, wall_ms, pair_ms_max, tables_ms = [float(x) for x in tt.tolist()]
    ms_step = step_ms_total / args.steps
    qps = Q / (ms_step * 1e-3)
    st = eng.stats()
    tables = [(tab_items["score"][q, :k], tab_items["ref"][q, :k]) for q, k in enumerate(tab_rows.tolist())]
    checksum = int(sum(int(sc.astype(np.int64).sum()) + int(r.sum()) for sc, r in tables) & 0xffffffffffff)
    # the same workload gives the same tables on any number of GPUs: known checksums of the default workloads (N = 1 runs)
    key = (args.queries, R, L, args.nbest, args.ambiguous, args.clades)
    known = KNOWN_CHECKSUMS.get(key)
    if known is not None and checksum != known:
        raise SystemExit(f"result checksum {checksum} differs from the single-GPU value {known} for this workload: the tables are WRONG")

    # ---- e2e: same search through the C ABI from host buffers ----
    e2e = None
    if want_e2e:
        e2e_t = []
        for i in range(1 + args.e2e_steps):
            barrier()
            t1 = time.perf_counter()
            eng.clear_refs()
            eng.set_queries(qtext)
            t_sq = time.perf_counter() - t1
            o = 0
            if world == 1:
                # pipelined host loop (the reference's pool loop, nearest.c:245-338): the copy + encoder of piece i+1 run
                # under the search rounds of piece i
                for s0 in range(0, n_local, 65536):
                    m = min(65536, n_local - s0)
                    eng.push_refs_ptr(host_buf + s0 * L, m, 0)
                    eng.search_async()
                eng.wait()
            else:
                # pipelined sharded host: every rank pushes its tile of round j (copy + encoder) and enqueues the round; the
                # pair-count launches of round j run under the copies of round j+1.  uvg_set_search_limit keeps a rank from
                # searching a round whose rows have not arrived everywhere it matters: its own.
                eng.set_global_count(R)
                for S, T in shard_rounds(R, world, first_round, max_round):
                    gb = S + rank * T
                    cnt = max(0, min(gb + T, R) - gb)
                    if cnt:
                        eng.push_refs_ptr(host_buf + o * L, cnt, 0)
                        o += cnt
                    eng.set_search_limit(min(S + world * T, R))
                    eng.search_async()
                eng.set_search_limit(-1)
                eng.wait()
            t_ps = time.perf_counter() - t1 - t_sq
            heaps = eng.heaps()
            barrier()
            if i > 0:
                e2e_t.append(time.perf_counter() - t1)
                _s = eng.stats()
                e2e_parts = {"set_queries_s": t_sq, "set_queries_host_ms": _s["set_queries_host_ms"], "set_queries_device_ms": _s["set_queries_device_ms"], "push_and_search_s": t_ps, "tables_d2h_s": time.perf_counter() - t1 - t_sq - t_ps}
        t = torch.tensor([float(np.mean(e2e_t))], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
        chk2 = int(sum(int(sc.astype(np.int64).sum()) + int(r.sum()) for sc, r in heaps) & 0xffffffffffff)
        e2e = {"value": Q / e2e_s, "unit": "queries/s", "h2d_bytes_per_step": int(Q * L * world + R * L),
               "d2h_bytes_per_step": int(world * Q * ((args.nbest + 1) * 32 + 8)), "seconds_per_step": e2e_s,
               "same_result_as_resident_run": chk2 == checksum, "last_step_breakdown": e2e_parts}
        eng._lib.uvg_host_free_pinned(host_buf)
        if world == 1 and not args.no_packed:
            # same search fed from the packed reference cache (uvg_export_packed once, outside the timed region; what a
            # host keeps in a .uvg file): the host pushes 0.375 bytes per site and the encoder does not run
            import ctypes as C
            piece = 65536
            sizes = []
            for s0 in range(0, n_local, piece):
                nb = C.c_size_t(0)
                eng._ck(eng._lib.uvg_packed_bytes(eng._h, s0, min(piece, n_local - s0), C.byref(nb)), "uvg_packed_bytes")
                sizes.append(nb.value)
            blob_buf = eng._lib.uvg_host_alloc_pinned(sum(sizes))
            if blob_buf:
                o = 0
                for k, s0 in enumerate(range(0, n_local, piece)):
                    eng._ck(eng._lib.uvg_export_packed(eng._h, s0, min(piece, n_local - s0), blob_buf + o, sizes[k]), "uvg_export_packed")
                    o += sizes[k]
                pk_t = []
                for i in range(1 + args.e2e_steps):
                    barrier()
                    t1 = time.perf_counter()
                    eng.clear_refs()
                    eng.set_queries(qtext)
                    t_sq = time.perf_counter() - t1
                    o, t_imp = 0, 0.0
                    for k in range(len(sizes)):
                        t2 = time.perf_counter()
                        eng.import_packed((blob_buf + o, sizes[k]))
                        t_imp += time.perf_counter() - t2
                        eng.search_async()
                        o += sizes[k]
                    t2 = time.perf_counter()
                    eng.wait()
                    t_wait = time.perf_counter() - t2
                    heaps = eng.heaps()
                    barrier()
                    if i > 0:
                        pk_t.append(time.perf_counter() - t1)
                        pk_parts = {"set_queries_s": t_sq, "import_s": t_imp, "wait_s": t_wait, "tables_d2h_s": time.perf_counter() - t2 - t_wait}
                chk3 = int(sum(int(sc.astype(np.int64).sum()) + int(r.sum()) for sc, r in heaps) & 0xffffffffffff)
                pk_s = float(np.mean(pk_t))
                e2e["packed_cache"] = {"value": Q / pk_s, "unit": "queries/s", "h2d_bytes_per_step": int(Q * L + sum(sizes)),
                                       "seconds_per_step": pk_s, "same_result_as_resident_run": chk3 == checksum, "last_step_breakdown": pk_parts,
                                       "note": "references enter as the packed blobs of uvg_export_packed (pinned host memory) instead of text"}
                eng._lib.uvg_host_free_pinned(blob_buf)

    # ---- roofline of the dominant kernel (k_paircount) ----
    peaks = load_peaks()
    W = (L + 31) // 32
    pairwords_dense = Q * n_local * W                         # per step on this rank, no early abort
    pairwords_rank = chunks / args.steps * 4096.0             # executed: (warp, chunk) iterations x 8 words x 16 pairs x 32 lanes
    k_ms = (pair_ms_max / args.steps)
    ops = OPS["pivot" if st["pivot_fix_entries"] >= 0 else "general"]
    lop3 = ops["lop3"] * pairwords_rank / (k_ms * 1e-3) / 1e9
    popc = ops["popc"] * pairwords_rank / (k_ms * 1e-3) / 1e9
    f_l, f_p = lop3 / peaks["lop3_gips"], popc / peaks["popc_gips"]
    bound = "int-popc" if f_p >= f_l else "int-lop3"
    n_qtiles = (Q + 1 + 63) // 64                             # + the consensus pseudo-query
    db_bytes = n_local * ((W + 7) // 8 * 8) * ops["db_bytes_per_word"]        # streamed database bytes per 32-site word and reference
    tr = ncu_traffic(ops["ncu"])
    if tr:   # algorithmic bytes of the captured launch: its slice of the streamed database once + one 16-byte record per pair
        tiles_k = (Q + 1 + 59) // 60 if st["pivot_fix_entries"] >= 0 else n_qtiles      # the pivot kernel's blocks take 60 slots (pivot tiles)
        refs = tr["launch_grid_blocks"] // tiles_k * (128 if st["pivot_fix_entries"] >= 0 else 64)    # references per thread block: pivot form 4 x 32, general 2 x 32
        tr["algorithmic_bytes"] = refs * ((W + 7) // 8 * 8) * ops["db_bytes_per_word"] + refs * (Q + 1) * 16
        tr["launch_refs"] = refs
    roofline = {"kernel": ops["kernel"], "bound": bound, "achieved": popc if bound == "int-popc" else lop3,
                "peak": peaks["popc_gips"] if bound == "int-popc" else peaks["lop3_gips"], "unit": "Gthread-inst/s",
                "frac": max(f_l, f_p), "traffic": tr["bytes"] if tr else None, "traffic_detail": tr,
                "note": "INT pipes (north_star): %.2f LOP3 + %.2f POPC per EXECUTED pair x 32-site word (padding, early abort and skipped units accounted for); peaks = profiles/int_pipes_b200_r01.json (measured)"
                        % (ops["lop3"], ops["popc"]),
                "lop3": {"achieved": lop3, "peak": peaks["lop3_gips"], "frac": f_l}, "popc": {"achieved": popc, "peak": peaks["popc_gips"], "frac": f_p},
                "hbm": {"bound": "hbm", "achieved": db_bytes / (k_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                        "frac": db_bytes / (k_ms * 1e-3) / 1e9 / peaks["hbm_gbs"], "peak_src": peaks["hbm_src"],
                        "note": "algorithmic = one pass over the packed DB per step; L2-level re-reads x%d query tiles" % n_qtiles},
                "kernel_ms_per_step": k_ms, "kernel_share_of_step": k_ms / ms_step,
                "executed_pairwords_per_step": pairwords_rank, "dense_pairwords_per_step": pairwords_dense,
                "executed_fraction_of_dense": pairwords_rank / max(1, pairwords_dense),
                # the pivot form skips (query, 128-site unit) pairs where the query is complete and equal to the pivot: the same
                # search priced with the DENSE op count of the dense pivot formulation (4 LOP3 per pair-word on the LOP3 peak: what a
                # kernel that does not skip would have to issue)
                "frac_if_priced_dense": 4.0 * pairwords_dense / (k_ms * 1e-3) / 1e9 / peaks["lop3_gips"],
                # continuity with round 1, whose kernel executed 0.3588 of the dense pair-words of this workload (whole 128-site
                # units) at 4 LOP3 each: the same search priced on THAT work.  The r02 kernel gives the same tables executing 0.096 (all-N
                # units come from a counts table, the rest is counted in 64-site halves, SNP-only halves are settled by fix entries) at
                # 2 AND + 2 POPC each, so `frac` per executed pair-word is a different quantity; this key is an explanation, not the
                # roofline claim
                "frac_at_r01_work": (4.0 * 0.3588 * pairwords_dense / (k_ms * 1e-3) / 1e9 / peaks["lop3_gips"]) if st["pivot_fix_entries"] >= 0 else None,
                "pivot_fix_entries_per_query": (st["pivot_fix_entries"] / max(1, Q + 1)) if st["pivot_fix_entries"] >= 0 else None}

    cb = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cb = cpu_baseline(eng, clades, qtext, qnames, args)

    if rank == 0:
        line = dict(base, value=qps, ms_per_step=ms_step, wall_ms_per_step=wall_ms / args.steps, gsite_cmp_per_s=Q * R * L / (ms_step * 1e-3) / 1e9,
                    clocks=clocks, e2e=e2e, gpu_launches=int(launches), roofline=roofline, cpu_baseline=cb,
                    replay_ms_per_step=replay_ms / args.steps, tables_ms_per_step=tables_ms / args.steps,
                    timed_region="reset -> rounds (CUDA events: first pair-count launch -> final tables sorted on the device, gathered) -> tables on the host (one D2H into the caller's buffers)", db_build_s=t_build,
                    search_stats={k: st[k] for k in ("n_rounds", "n_admitted", "n_h1_flags", "max_mismatch")},
                    result_checksum=checksum)
        print(json.dumps(line))
    eng.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
