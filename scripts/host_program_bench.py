#!/usr/bin/env python
"""Throughput of the C host program (uvaia_b200/bin/uvaia_b200) from FASTA files on local disk: writes a synthetic
SARS-CoV-2-shaped query file and reference file (device generator, raw FASTA), runs the program, reports MB/s of
reference text and queries/s.   python scripts/host_program_bench.py [--refs 200000] [--queries 1000]"""
import argparse, json, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import bench
from uvaia_b200.engine import Engine

ap = argparse.ArgumentParser()
ap.add_argument("--refs", type=int, default=200000)
ap.add_argument("--queries", type=int, default=1000)
ap.add_argument("--dir", default=None)
ap.add_argument("--devices", default=None, help="also run with the references row-sharded over these GPUs, e.g. 0-3")
a = ap.parse_args()
L = 29903
clades = bench.make_model(L)
wd = a.dir or tempfile.mkdtemp(prefix="uvb_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
with Engine(L, nbest=100) as eng:
    d = eng.device_alloc(a.queries * L)
    eng.synth_rows_device(clades, 2002, 0, a.queries, d)
    q = eng.d2h(d, a.queries * L).reshape(a.queries, L)
    eng.device_free(d)
    bench.write_fasta(os.path.join(wd, "q.fa"), [b"Q%06d" % i for i in range(a.queries)], q)
    step = 16384
    d = eng.device_alloc(step * L)
    with open(os.path.join(wd, "r.fa"), "wb") as fh:
        for s in range(0, a.refs, step):
            m = min(step, a.refs - s)
            eng.synth_rows_device(clades, 1001, s, m, d)
            t = eng.d2h(d, m * L).reshape(m, L)
            for i in range(m):
                fh.write(b">R%07d\n" % (s + i)); fh.write(t[i].tobytes()); fh.write(b"\n")
    eng.device_free(d)
size = os.path.getsize(os.path.join(wd, "r.fa"))
import lzma
import re
BIN = os.path.join(ROOT, "uvaia_b200", "bin", "uvaia_b200")
runs = {}
variants = [("text_serial_parser", ["--serial"]), ("text", []), ("cache_write", ["--cache", os.path.join(wd, "refs.uvg")]), ("cache_read", ["--cache", os.path.join(wd, "refs.uvg")])]
if a.devices:
    variants.append(("text_devices_" + a.devices.replace(",", "_"), ["--devices", a.devices]))
for label, extra in variants:
    t0 = time.perf_counter()
    env = dict(os.environ, UVG_PARSE_TIMING="1")
    if extra == ["--serial"]:
        extra, env["UVG_PARSE_THREADS"] = [], "1"
    p = subprocess.run([BIN, "-n", "100", "-r", os.path.join(wd, "r.fa"), "-o", os.path.join(wd, "out_" + label)] + extra + [os.path.join(wd, "q.fa")],
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, env=env)
    dt = time.perf_counter() - t0
    err = p.stderr.decode(errors="replace")
    marks = {}
    for key, pat in (("search_done_s", r"too ambiguous \(excluded\)\. ([0-9.]+) secs elapsed"), ("sharded_search_done_s", r"Search over \d+ GPUs finished, ([0-9.]+) secs elapsed"), ("aln_done_s", r"Saved \d+ sequences to file .* ([0-9.]+) secs elapsed"),
                     ("csv_done_s", r"Saved distance table to file .* ([0-9.]+) secs elapsed")):
        m = re.search(pat, err)
        marks[key] = float(m.group(1)) if m else None
    runs[label] = {"returncode": p.returncode, "wall_s": dt, "ref_text_MB_per_s": size / dt / 1e6, **marks, "parser": [ln for ln in err.splitlines() if "parallel parse" in ln or "Finished reading" in ln],
                   "stderr_tail": err.strip().splitlines()[-4:]}
same = all(lzma.open(os.path.join(wd, "out_%s.csv.xz" % k)).read() == lzma.open(os.path.join(wd, "out_text.csv.xz")).read() for k in runs)
print(json.dumps({"refs": a.refs, "ref_file_bytes": size, "cache_file_bytes": os.path.getsize(os.path.join(wd, "refs.uvg")), "same_csv": same, "runs": runs}))
