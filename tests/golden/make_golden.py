#!/usr/bin/env python
"""Generate the golden fixtures of tests/golden/ from the UNMODIFIED reference, compiled by oracle/build_ref.sh:

  * unit_vectors.json  -- outputs of the reference's own scoring / heap functions called through ctypes in
                          oracle/_ref/libuvaia_ref_score.so (biomcmc_pairwise_score_matches_truncated_idx,
                          quick_pairwise_score_acgt_and_valid, quick_count_sequence_non_N, heap_insert,
                          heap_finalise_heap_qsort) on: the README toy trio (reference README.md:307-316), the full
                          byte-pair truth table, the truncation vector, random idx/limit cases, heap tie cases.
  * cases/<name>/      -- end-to-end runs of oracle/_ref/uvaia_ref (= src/nearest.c): inputs (xz FASTA), args.json,
                          expected.csv.xz (decompressed CSV of the reference, re-compressed), expected_aln_names.txt.

Runs only in the build container (needs /root/reference for the binary build and the bundled alignment).  The GPU box
only reads the committed fixtures.  Usage: python tests/golden/make_golden.py [--only NAME ...]
"""
from __future__ import annotations

import argparse
import ctypes as C
import gzip
import json
import lzma
import os
import random
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from uvaia_b200.synth import SynthModel, write_fasta  # noqa: E402

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "uvaia_ref")
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libuvaia_ref_score.so")
BALL_BIN = os.path.join(ROOT, "oracle", "_ref", "uvaiaball_ref")
BUNDLED = "/root/reference/data/03.unique_acgt.aln.xz"
ALPHABET = "ACGTUMRWSYKVHDBNOX?-.EZ*acgtn"


# ------------------------------------------------------------------------------------------------------------------
# unit vectors through the reference's own functions
# ------------------------------------------------------------------------------------------------------------------
class QItem(C.Structure):
    _fields_ = [("score", C.c_int * 6), ("name", C.c_char_p)]


class HeapStruct(C.Structure):
    _fields_ = [("seq", C.POINTER(QItem)), ("heap_size", C.c_int), ("n", C.c_int), ("max_incompatible", C.c_int)]


def load_ref_lib():
    lib = C.CDLL(REF_LIB)
    lib.initialise_acgt()          # is_acgt[] / is_indel[] tables (src/utils.c:255-264); the programs call it at start-up
    lib.biomcmc_pairwise_score_matches_truncated_idx.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_int),
                                                                 C.POINTER(C.c_size_t)]
    lib.biomcmc_pairwise_score_matches_truncated_idx.restype = None
    lib.quick_pairwise_score_acgt_and_valid.argtypes = [C.c_char_p, C.c_char_p, C.c_size_t, C.c_int, C.POINTER(C.c_int),
                                                        C.POINTER(C.c_size_t)]
    lib.quick_pairwise_score_acgt_and_valid.restype = None
    lib.quick_count_sequence_non_N.argtypes = [C.c_char_p, C.c_size_t]
    lib.quick_count_sequence_non_N.restype = C.c_int
    lib.new_heap_t.argtypes = [C.c_int]
    lib.new_heap_t.restype = C.POINTER(HeapStruct)
    lib.heap_insert.argtypes = [C.POINTER(HeapStruct), QItem]
    lib.heap_insert.restype = C.c_ubyte
    lib.heap_finalise_heap_qsort.argtypes = [C.POINTER(HeapStruct)]
    lib.heap_finalise_heap_qsort.restype = None
    return lib


def ref_truncated_idx(lib, s1: bytes, s2: bytes, idx, limit):
    res = (C.c_int * 4)()
    arr = (C.c_size_t * max(1, len(idx)))(*idx)
    lib.biomcmc_pairwise_score_matches_truncated_idx(s1, s2, len(idx), limit, res, arr)
    return list(res)


def ref_acgt_valid(lib, s1: bytes, s2: bytes, idx, limit):
    res = (C.c_int * 2)()
    arr = (C.c_size_t * max(1, len(idx)))(*idx)
    lib.quick_pairwise_score_acgt_and_valid(s1, s2, len(idx), limit, res, arr)
    return list(res)


def ref_heap_run(lib, nbest, items):
    """items: list of 6-int tuples, inserted in order with name = str(index).  -> (admitted flags, final order of names,
    final scores)"""
    h = lib.new_heap_t(nbest)
    flags = []
    for i, sc in enumerate(items):
        it = QItem((C.c_int * 6)(*sc), str(i).encode())
        flags.append(int(lib.heap_insert(h, it)))
    n = h.contents.n
    K = h.contents.heap_size
    if n == K - 1:
        raise ValueError("hazard H2 configuration (n == K-1) crashes the reference's writer; not a golden")
    lib.heap_finalise_heap_qsort(h)
    order = [int(h.contents.seq[i].name) for i in range(n)]
    scores = [list(h.contents.seq[i].score) for i in range(n)]
    return flags, order, scores


def make_unit_vectors():
    lib = load_ref_lib()
    big = 1 << 24
    out = {"alphabet": ALPHABET}
    # byte-pair truth table (single site)
    tab, tab_acgt = {}, {}
    for a in ALPHABET:
        for b in ALPHABET:
            tab[a + b] = ref_truncated_idx(lib, a.encode(), b.encode(), [0], big)
            tab_acgt[a + b] = ref_acgt_valid(lib, a.encode(), b.encode(), [0], big)
    out["pair_table"] = tab
    out["pair_table_acgt"] = tab_acgt
    # README toy trio (reference README.md:307-316)
    trio = ["AACGTTA--", "AACG-TAM-", "MNCGTTMC-"]
    out["readme_trio"] = {"seqs": trio,
                          "results": {f"{i}{j}": ref_truncated_idx(lib, trio[i].encode(), trio[j].encode(), list(range(9)), big)
                                      for i in range(3) for j in range(3)}}
    # truncation vector
    s1, s2 = b"AAAAAAAAAA", b"ACACACACAC"
    out["truncation"] = {"s1": s1.decode(), "s2": s2.decode(),
                         "results": {str(m): ref_truncated_idx(lib, s1, s2, list(range(10)), m) for m in range(0, 8)},
                         "results_acgt": {str(m): ref_acgt_valid(lib, s1, s2, list(range(10)), m) for m in range(0, 8)}}
    # non_N counter
    strings = [ALPHABET, "NNNN", "ACGT-acgt?XxOo.", "A", "", "nNxX-?oO." * 3 + "ACGTRYKM"]
    out["non_n"] = [{"s": s, "n": int(lib.quick_count_sequence_non_N(s.encode(), len(s)))} for s in strings]
    # random cases: sequences over a SARS-like alphabet, random sorted idx subsets, random limits
    rng = random.Random(20261104)
    cases = []
    weights = [("A", 24), ("C", 15), ("G", 16), ("T", 26), ("N", 10), ("-", 2), ("R", 1), ("Y", 1), ("M", 1), ("K", 1), ("S", 1), ("W", 1),
               ("B", 1), ("D", 1), ("H", 1), ("V", 1), ("U", 1), ("X", 1), ("?", 1), ("E", 1)]
    pop = [c for c, w in weights for _ in range(w)]
    for k in range(60):
        L = rng.choice([1, 7, 33, 64, 200])
        a = "".join(rng.choice(pop) for _ in range(L))
        # second sequence: mostly a copy with a few changes, so matches and mismatches both occur
        b = "".join(ch if rng.random() < 0.8 else rng.choice(pop) for ch in a)
        idx = sorted(rng.sample(range(L), rng.randint(0, L)))
        lim = rng.choice([0, 1, 2, 3, 5, 10, 1000])
        cases.append({"s1": a, "s2": b, "idx": idx, "limit": lim,
                      "full": ref_truncated_idx(lib, a.encode(), b.encode(), idx, lim) if idx else [0, 0, 0, 0],
                      "acgt": ref_acgt_valid(lib, a.encode(), b.encode(), idx, lim) if idx else [0, 0]})
    out["random_truncated"] = cases
    # heap: admission flags + final order, with many ties (first-come tie-break, glibc qsort order)
    heaps = []
    for nbest, n_items, spread in [(5, 12, 1), (5, 40, 2), (1, 9, 3), (2, 2, 2), (10, 4, 3), (7, 100, 4), (100, 350, 6), (3, 50, 1), (16, 17, 2)]:
        items = [[rng.randrange(spread) + 10, rng.randrange(spread) + 10, rng.randrange(spread) + 10, rng.randrange(spread) + 12,
                  rng.randrange(spread), rng.randrange(2) + 20] for _ in range(n_items)]
        flags, order, scores = ref_heap_run(lib, nbest, items)
        heaps.append({"nbest": nbest, "items": items, "admitted": flags, "order": order, "scores": scores})
    out["heap"] = heaps
    with open(os.path.join(HERE, "unit_vectors.json"), "w") as fh:
        json.dump(out, fh, separators=(",", ":"))
    print("unit_vectors.json:", len(tab), "byte pairs,", len(cases), "random cases,", len(heaps), "heap runs")


# ------------------------------------------------------------------------------------------------------------------
# end-to-end cases through the reference binary
# ------------------------------------------------------------------------------------------------------------------
def xz_write(path, data: bytes):
    with lzma.open(path, "wb", preset=9 | lzma.PRESET_EXTREME) as fh:
        fh.write(data)


def fasta_bytes(names, text, width=0):
    with tempfile.NamedTemporaryFile() as tf:
        write_fasta(tf.name, names, text, width)
        return open(tf.name, "rb").read()


def run_reference(workdir, qfile, rfiles, args):
    cmd = [REF_BIN] + args + sum((["-r", r] for r in rfiles), []) + ["-o", os.path.join(workdir, "out"), qfile]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, cwd=workdir)
    if p.returncode != 0:
        raise RuntimeError(f"uvaia_ref failed ({p.returncode}): {' '.join(cmd)}\n{p.stderr.decode(errors='replace')[-800:]}")
    csv = gzip.decompress(open(os.path.join(workdir, "out.csv.xz"), "rb").read())     # zlib-only build: gzip under .xz (H4)
    aln = gzip.decompress(open(os.path.join(workdir, "out.aln.xz"), "rb").read())
    names = [ln[1:] for ln in aln.split(b"\n") if ln.startswith(b">")]
    return csv, names, p.stderr.decode(errors="replace")


def save_case(name, q_fa: bytes, r_fas, args, note, threads=8, h1_expected=False):
    d = os.path.join(HERE, "cases", name)
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d)
    with tempfile.TemporaryDirectory() as wd:
        qf = os.path.join(wd, "q.fa")
        open(qf, "wb").write(q_fa)
        rfs = []
        for i, r in enumerate(r_fas):
            rf = os.path.join(wd, f"r{i}.fa")
            open(rf, "wb").write(r)
            rfs.append(rf)
        csv, aln_names, stderr = run_reference(wd, qf, rfs, args + ["-t", str(threads)])
        # the result must not depend on threads / pool unless the case says so (hazard H1 never fires in the goldens)
        pool_dependent = False
        if "-p" not in args or h1_expected is None:
            base = [a for i, a in enumerate(args) if not (a == "-p" or (i > 0 and args[i - 1] == "-p"))]
            csv2, _, _ = run_reference(wd, qf, rfs, base + ["-t", "1", "-p", "3"])
            pool_dependent = csv2 != csv
            assert h1_expected is None or pool_dependent == h1_expected, \
                f"{name}: reference output depends on --pool = {pool_dependent} (hazard H1), expected {h1_expected}"
    xz_write(os.path.join(d, "q.fa.xz"), q_fa)
    for i, r in enumerate(r_fas):
        xz_write(os.path.join(d, f"r{i}.fa.xz"), r)
    xz_write(os.path.join(d, "expected.csv.xz"), csv)
    open(os.path.join(d, "expected_aln_names.txt"), "wb").write(b"\n".join(aln_names) + b"\n")
    idx_line = [ln for ln in stderr.split("\n") if "segregating" in ln or "idx" in ln.lower()]
    json.dump({"args": args, "n_ref_files": len(r_fas), "note": note, "h1_pool_dependent": pool_dependent, "reference_stderr_partition_line": idx_line[:2],
               "csv_rows": csv.count(b"\n") - 1}, open(os.path.join(d, "args.json"), "w"), indent=1)
    print(f"case {name}: {csv.count(10) - 1} rows, {len(aln_names)} aln names")


def case_synthetic():
    m = SynthModel(nsites=3000, n_clades=24, seed=11, clade_subs=12)
    qn, Q = m.sample(60, 12, prefix="Wales/QRY")
    rn, R = m.sample(3000, 13, prefix="England/REF")
    qf, rf = fasta_bytes(qn, Q), fasta_bytes(rn, R)
    save_case("syn_default", qf, [rf], ["-n", "10"], "60 q x 3000 refs x 3000 sites, default mode")
    save_case("syn_reversed", qf, [fasta_bytes(rn[::-1], R[::-1])], ["-n", "10"], "same DB in reverse order (order-dependent admission)")
    save_case("syn_acgt", qf, [rf], ["-n", "10", "--acgt"], "--acgt mode")
    save_case("syn_trim", qf, [rf], ["-n", "10", "--trim", "230"], "--trim 230")
    save_case("syn_trim_acgt", qf, [rf], ["-n", "7", "--trim", "100", "--acgt"], "--trim 100 --acgt")
    save_case("syn_n1", qf, [rf], ["-n", "1"], "-n 1 gives 2 rows per query (H3)")
    save_case("syn_n100", qf, [rf], ["-n", "100"], "default nbest")
    save_case("syn_pool1", qf, [rf], ["-n", "10", "-p", "1"], "-p 1")
    half = 1500
    save_case("syn_two_files", qf, [fasta_bytes(rn[:half], R[:half]), fasta_bytes(rn[half:], R[half:])], ["-n", "10"], "two -r files")
    save_case("syn_ref_filter", qf, [rf], ["-n", "10", "-A", "0.12"], "-A 0.12 drops N-rich references")
    save_case("syn_query_filter", qf, [rf], ["-n", "10", "-a", "0.15"], "-a 0.15 drops N-rich queries")
    # -x: queries are part of the DB
    mix_n = rn[:700] + qn[:30] + rn[700:1400]
    mix_t = np.concatenate([R[:700], Q[:30], R[700:1400]])
    save_case("syn_exclude_self", qf, [fasta_bytes(mix_n, mix_t)], ["-n", "5", "-x"], "-x with 30 of the queries inside the DB")
    save_case("syn_include_self", qf, [fasta_bytes(mix_n, mix_t)], ["-n", "5"], "same DB without -x: self hits rank first")
    # few references (n < K-1): rows shrink to n
    save_case("syn_few_refs", qf, [fasta_bytes(rn[:4], R[:4])], ["-n", "10"], "4 references, nbest 10")
    save_case("syn_one_ref", qf, [fasta_bytes(rn[:1], R[:1])], ["-n", "10"], "1 reference")
    # a single query: every usable site is idx_c, the whole score is the consensus stage computed with the batch-start
    # limit -> the reference's own result depends on --pool (hazard H1).  One golden per pool value.
    for pool in (1, 7, 512):
        save_case(f"h1_one_query_p{pool}", fasta_bytes(qn[:1], Q[:1]), [rf], ["-n", "10", "-p", str(pool)],
                  f"single query, -p {pool}: pool-dependent truncation of the consensus stage (H1)", h1_expected=None)


def case_ties():
    m = SynthModel(nsites=600, n_clades=3, seed=21, clade_subs=4)
    qn, Q = m.sample(6, 22, prefix="Q")
    rn, R = m.sample(40, 23, prefix="R")
    # 12 identical copies of one row scattered through the DB, plus duplicates of a query
    R[5:17] = R[5]
    R[30] = Q[2]
    R[31] = Q[2]
    save_case("ties_n5", fasta_bytes(qn, Q), [fasta_bytes(rn, R)], ["-n", "5"], "12 identical references: first-come tie-break + qsort order")
    save_case("ties_n5_acgt", fasta_bytes(qn, Q), [fasta_bytes(rn, R)], ["-n", "5", "--acgt"], "ties in --acgt mode")
    save_case("ties_n30", fasta_bytes(qn, Q), [fasta_bytes(rn, R)], ["-n", "30"], "heap almost as large as the DB")


def case_ambiguous():
    m = SynthModel(nsites=2500, n_clades=16, seed=31, clade_subs=10)
    qn, Q = m.sample(48, 32, ambiguous_frac=0.20, prefix="Q")
    rn, R = m.sample(2000, 33, ambiguous_frac=0.04, prefix="R")
    qf, rf = fasta_bytes(qn, Q), fasta_bytes(rn, R)
    save_case("amb_default", qf, [rf], ["-n", "12"], "config-4 shape: 20 % ambiguous query sites (15 % N, 5 % IUPAC), refs 4 %")
    save_case("amb_acgt", qf, [rf], ["-n", "12", "--acgt"], "same, --acgt")
    save_case("amb_strict", qf, [rf], ["-n", "12", "-a", "0.2", "-A", "0.2"], "same, stricter filters")


def case_hygiene():
    m = SynthModel(nsites=500, n_clades=4, seed=41, clade_subs=5)
    qn, Q = m.sample(5, 42, prefix="Q")
    rn, R = m.sample(30, 43, prefix="R")

    def messy(names, text, crlf):
        out = [b"# a comment line", b"; another comment", b""]
        nl = b"\r\n" if crlf else b"\n"
        for i, (nm, row) in enumerate(zip(names, text)):
            s = row.tobytes()
            if i % 3 == 0:
                s = s.lower()
            out.append(b">" + nm + (b" extra words | x=1" if i % 4 == 0 else b""))
            w = [60, 70, 33][i % 3]
            for k in range(0, len(s), w):
                ln = s[k:k + w]
                if i % 5 == 0 and len(ln) > 10:
                    ln = ln[:5] + b" " + ln[5:10] + b"\t" + ln[10:]
                out.append(ln)
            if i % 2 == 0:
                out.append(b"")
        return nl.join(out) + nl

    save_case("hygiene", messy(qn, Q, False), [messy(rn, R, True)], ["-n", "4"],
              "multi-line records, lower case, blanks, tabs, CRLF, comments, names with spaces")


def case_keep_resolved():
    """-k / --keep_resolved (nearest.c:219-224, fastaseq.c:797-841): query sets with exact duplicates, masked (less
    resolved) copies, chains of copies, copies masked at complementary places (both kept) and copies with a private SNP."""
    m = SynthModel(nsites=2000, n_clades=12, seed=51, clade_subs=10)
    qn, Q = m.sample(36, 52, prefix="Wales/QRY")
    rn, R = m.sample(1500, 53, prefix="England/REF")
    rng = np.random.default_rng(54)
    names, rows = list(qn), [q.copy() for q in Q]

    def add(tag, src, row):
        names.append(qn[src] + b"|" + tag.encode())
        rows.append(row)

    def masked(row, n_runs, run, char=ord("N")):
        out = row.copy()
        for _ in range(n_runs):
            a = int(rng.integers(200, 1800))
            out[a:a + run] = char
        return out

    for src in range(0, 8):
        add("dup", src, Q[src].copy())                                   # identical: the later one goes
    for src in range(4, 16):
        add("lessN", src, masked(Q[src], 3, 25))                         # less resolved copy
    for src in range(12, 18):
        a = masked(Q[src], 2, 30)
        add("chain1", src, a)
        add("chain2", src, masked(a, 2, 30, ord("-")))                   # chain: src > chain1 > chain2
    for src in range(18, 24):
        a, b = Q[src].copy(), Q[src].copy()
        a[300:340] = ord("N"); b[900:960] = ord("N")                     # complementary masks: neither resolves the other
        add("compA", src, a); add("compB", src, b)
    for src in range(24, 30):
        a = masked(Q[src], 2, 20)
        pos = int(rng.integers(400, 1600))
        while a[pos] not in b"ACGT":
            pos += 1
        a[pos] = b"ACGT"[(b"ACGT".index(bytes([a[pos]])) + 1) % 4]       # masked copy with a private SNP: not redundant
        add("snp", src, a)
    for src in range(30, 34):
        a = Q[src].copy()
        sel = np.where(np.isin(a, list(b"ACGT")))[0]
        a[rng.choice(sel, 12, replace=False)] = ord("R")                 # IUPAC codes: differ as text, equal-ish as sets
        add("iupac", src, a)
    perm = rng.permutation(len(names))
    names = [names[i] for i in perm]
    text = np.stack([rows[i] for i in perm])
    qf, rf = fasta_bytes(names, text), fasta_bytes(rn, R)
    save_case("keep_default", qf, [rf], ["-n", "8", "-k"], "-k: redundant queries pruned, more resolved kept (default mode)")
    save_case("keep_acgt", qf, [rf], ["-n", "8", "-k", "--acgt"], "-k with --acgt")
    save_case("keep_trim", qf, [rf], ["-n", "8", "-k", "--trim", "350"], "-k with --trim 350: masks inside the trimmed ends do not count")
    save_case("keep_off", qf, [rf], ["-n", "8"], "same inputs without -k: every query is searched")


def case_alphabet_soup():
    """every symbol class of the truth table inside end-to-end runs: U, lower case, the indel spellings (n x o ? . -),
    letters outside the IUPAC alphabet (E Z *), two- and three-base codes, in the queries and in the references"""
    m = SynthModel(nsites=1800, n_clades=10, seed=61, clade_subs=9)
    qn, Q = m.sample(30, 62, prefix="Q")
    rn, R = m.sample(1200, 63, prefix="R")
    rng = np.random.default_rng(64)
    soup = np.frombuffer(b"UuacgtnxXoO?.-EZ*MRWSYKVHDBmrwsykvhdb", np.uint8)

    def sprinkle(T, frac):
        T = T.copy()
        mask = rng.random(T.shape) < frac
        T[mask] = soup[rng.integers(0, len(soup), size=int(mask.sum()))]
        return T

    Qs, Rs = sprinkle(Q, 0.04), sprinkle(R, 0.04)
    Rs[7] = Qs[3]                                                  # a reference identical to a query, odd bytes included
    Rs[8] = np.frombuffer(Qs[3].tobytes().lower(), np.uint8)       # the same in lower case (the reader upper-cases)
    qf, rf = fasta_bytes(qn, Qs), fasta_bytes(rn, Rs)
    save_case("soup_default", qf, [rf], ["-n", "9"], "full symbol alphabet (U, lower case, indel spellings, non-IUPAC letters, IUPAC codes) at 4 % of the sites")
    save_case("soup_acgt", qf, [rf], ["-n", "9", "--acgt"], "same, --acgt")
    save_case("soup_trim_k", qf, [rf], ["-n", "9", "--trim", "77", "-k"], "same, --trim 77 -k")


def case_c2_shape():
    m = SynthModel(nsites=29903, n_clades=40, seed=1)
    qn, Q = m.sample(24, 2002, prefix="England/QRY")
    rn, R = m.sample(900, 1001, prefix="England/NORW")
    save_case("c2_shape", fasta_bytes(qn, Q), [fasta_bytes(rn, R)], ["-n", "100"], "config-2 shape: 29,903 sites, nbest 100")


def case_bundled():
    if not os.path.exists(BUNDLED):
        print("bundled alignment not present, skipping real_c1_small")
        return
    names, seqs = [], []
    cur = []
    for line in lzma.open(BUNDLED, "rb"):
        line = line.rstrip(b"\r\n")
        if line.startswith(b">"):
            if cur:
                seqs.append(b"".join(cur))
            names.append(line[1:])
            cur = []
        elif line:
            cur.append(line)
    seqs.append(b"".join(cur))
    rng = random.Random(20261104)
    pick = rng.sample(range(len(names)), 420)
    qi, ri = pick[:20], sorted(pick[20:])
    qf = b"".join(b">" + names[i] + b"\n" + seqs[i] + b"\n" for i in qi)
    rf = b"".join(b">" + names[i] + b"\n" + seqs[i] + b"\n" for i in ri)
    save_case("real_c1_small", qf, [rf], ["-n", "10"],
              "config-1 shape: 20 queries x 400 references sampled (seed 20261104) from the reference's bundled data/03.unique_acgt.aln.xz")
    # BASELINE.json configs[0] at its stated shape (SURVEY.md 8c golden iv): 100 sampled queries against the remaining
    # 9,085 sequences of the bundled alignment, nbest 10 -- and the same with the database order reversed (the n-best SET
    # depends on the order for most queries)
    random.seed(20261104)
    qi = random.sample(range(len(names)), 100)
    qset = set(qi)
    ri = [i for i in range(len(names)) if i not in qset]
    qf = b"".join(b">" + names[i] + b"\n" + seqs[i] + b"\n" for i in qi)
    rf = b"".join(b">" + names[i] + b"\n" + seqs[i] + b"\n" for i in ri)
    save_case("real_c1_full", qf, [rf], ["-n", "10"],
              "BASELINE config 1: 100 queries (random.seed(20261104); random.sample(range(9185), 100)) x the remaining 9,085 sequences of data/03.unique_acgt.aln.xz")
    rf = b"".join(b">" + names[i] + b"\n" + seqs[i] + b"\n" for i in reversed(ri))
    save_case("real_c1_full_reversed", qf, [rf], ["-n", "10"], "BASELINE config 1 with the database in reverse file order")


# ------------------------------------------------------------------------------------------------------------------
# uvaiaball (src/ball.c): which references are within the radius of some query
# ------------------------------------------------------------------------------------------------------------------
def save_ball_case(name, q_fa: bytes, r_fas, args, note):
    d = os.path.join(HERE, "ball_cases", name)
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d)
    with tempfile.TemporaryDirectory() as wd:
        qf = os.path.join(wd, "q.fa")
        open(qf, "wb").write(q_fa)
        rfs = []
        for i, r in enumerate(r_fas):
            rf = os.path.join(wd, f"r{i}.fa")
            open(rf, "wb").write(r)
            rfs.append(rf)
        cmd = [BALL_BIN] + args + sum((["-r", r] for r in rfs), []) + ["-o", os.path.join(wd, "out"), qf]
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, cwd=wd)
        if p.returncode != 0:
            raise RuntimeError(f"uvaiaball_ref failed ({p.returncode}): {p.stderr.decode(errors='replace')[-800:]}")
        aln = gzip.decompress(open(os.path.join(wd, "out.aln.xz"), "rb").read())         # zlib-only build: names are intact, sequences are not (H4)
        kept = [ln[1:] for ln in aln.split(b"\n") if ln.startswith(b">")]
        stderr = p.stderr.decode(errors="replace")
    xz_write(os.path.join(d, "q.fa.xz"), q_fa)
    for i, r in enumerate(r_fas):
        xz_write(os.path.join(d, f"r{i}.fa.xz"), r)
    open(os.path.join(d, "expected_kept_names.txt"), "wb").write(b"\n".join(kept) + b"\n")
    lines = [ln for ln in stderr.split("\n") if "Query database" in ln or "radius distance" in ln or "sequences within radius" in ln]
    json.dump({"args": args, "n_ref_files": len(r_fas), "note": note, "n_kept": len(kept), "reference_stderr": lines},
              open(os.path.join(d, "args.json"), "w"), indent=1)
    print(f"ball case {name}: {len(kept)} kept")


def case_ball():
    m = SynthModel(nsites=3000, n_clades=24, seed=11, clade_subs=12)
    qn, Q = m.sample(40, 12, prefix="Wales/QRY")
    rn, R = m.sample(3000, 13, prefix="England/REF")
    # some references are copies of queries with 0..7 extra substitutions, so that every radius keeps a different set
    rng = np.random.default_rng(5)
    for k in range(300):
        row = Q[k % len(Q)].copy()
        pos = rng.choice(3000, size=k % 8, replace=False)
        row[pos] = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=len(pos))]
        R[10 * k] = row
    qf, rf = fasta_bytes(qn, Q), fasta_bytes(rn, R)
    save_ball_case("ball_default", qf, [rf], [], "radius 5 (default), 40 queries x 3000 references")
    save_ball_case("ball_d0", qf, [rf], ["-d", "0"], "radius 0: only references identical on the compared sites")
    save_ball_case("ball_d2", qf, [rf], ["-d", "2"], "radius 2")
    save_ball_case("ball_d40", qf, [rf], ["-d", "40"], "radius 40: most references")
    save_ball_case("ball_acgt", qf, [rf], ["-d", "3", "-x"], "--acgt (-x) mode")
    save_ball_case("ball_keep", qf, [rf], ["-d", "3", "-k"], "-k: the more resolved of two equivalent queries stays")
    save_ball_case("ball_trim", qf, [rf], ["-d", "4", "--trim", "230"], "--trim 230")
    save_ball_case("ball_filters", qf, [rf], ["-d", "6", "-a", "0.2", "-A", "0.85"], "-a 0.2, -A 0.85 (the reference filter is non_n >= nchar * A)")
    save_ball_case("ball_two_files", qf, [fasta_bytes(rn[:1500], R[:1500]), fasta_bytes(rn[1500:], R[1500:])], ["-d", "5"], "two -r files")
    save_ball_case("ball_one_query", fasta_bytes(qn[:1], Q[:1]), [rf], ["-d", "5"], "a single query: every site is a consensus site")
    # the whole alphabet: U, lower case, codes, letters outside IUPAC -- the ball distance compares BYTES
    m2 = SynthModel(nsites=900, n_clades=5, seed=77, clade_subs=6)
    qn2, Q2 = m2.sample(12, 78, prefix="Q")
    rn2, R2 = m2.sample(400, 79, prefix="R")
    soup = np.frombuffer(ALPHABET.encode(), np.uint8)
    for T, frac in ((Q2, 0.01), (R2, 0.02)):
        mask = rng.random(T.shape) < frac
        T[mask] = soup[rng.integers(0, len(soup), size=int(mask.sum()))]
    save_ball_case("ball_soup", fasta_bytes(qn2, Q2), [fasta_bytes(rn2, R2)], ["-d", "40"], "every symbol class in queries and references")
    save_ball_case("ball_soup_acgt", fasta_bytes(qn2, Q2), [fasta_bytes(rn2, R2)], ["-d", "6", "-x"], "same, --acgt")


GROUPS = {"unit": make_unit_vectors, "ball": case_ball, "synthetic": case_synthetic, "ties": case_ties, "ambiguous": case_ambiguous,
          "hygiene": case_hygiene, "keep_resolved": case_keep_resolved, "alphabet_soup": case_alphabet_soup, "c2_shape": case_c2_shape, "bundled": case_bundled}

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", nargs="*", default=None, choices=list(GROUPS))
    a = ap.parse_args()
    if not (os.path.exists(REF_BIN) and os.path.exists(REF_LIB)):
        sys.exit("build the reference first: bash oracle/build_ref.sh")
    for g in (a.only or list(GROUPS)):
        GROUPS[g]()
