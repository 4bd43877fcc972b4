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
t0 = time.perf_counter()
p = subprocess.run([os.path.join(ROOT, "uvaia_b200", "bin", "uvaia_b200"), "-n", "100", "-r", os.path.join(wd, "r.fa"), "-o", os.path.join(wd, "out"),
                    os.path.join(wd, "q.fa")], stdout=subprocess.PIPE, stderr=subprocess.PIPE)
dt = time.perf_counter() - t0
err = p.stderr.decode(errors="replace")
print(json.dumps({"returncode": p.returncode, "refs": a.refs, "ref_file_bytes": size, "wall_s": dt, "ref_text_MB_per_s": size / dt / 1e6,
                  "stderr_tail": err.strip().splitlines()[-6:]}))
