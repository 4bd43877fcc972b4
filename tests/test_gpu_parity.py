"""Parity tests proper: the CUDA path, called through the C ABI (ctypes -> libuvaia_b200.so), against
  * the golden CSV tables written by the compiled reference binary (tests/golden/cases),
  * the numpy oracle on seeded random inputs,
  * the reference binary itself when oracle/_ref travelled to the box.
Bit-exact: same neighbours, same six counts, same order."""
import gzip
import os
import subprocess
import tempfile

import numpy as np
import pytest

from helpers import ROOT, Case, case_names, orc
from uvaia_b200 import host
from uvaia_b200.engine import Engine, UvgError
from uvaia_b200.synth import SynthModel, write_fasta

pytestmark = pytest.mark.gpu



@pytest.fixture(autouse=True, params=["pivot", "general"])
def pair_kernel(request, monkeypatch):
    """every test of this module runs twice: with the pivot form of the pair-count kernel forced wherever the mode allows
    it (default scoring, < 65,536 columns, no early abort) and with the general 3-plane kernel; the tables must not move"""
    monkeypatch.setenv("UVG_PAIR_KERNEL", request.param)
    return request.param


REF_BIN = os.path.join(ROOT, "oracle", "_ref", "uvaia_ref")
NON_H1 = [n for n in case_names() if not Case(n).h1]
H1 = [n for n in case_names() if Case(n).h1]


def gpu_csv(case, **engine_kw):
    kw = case.options()
    out = host.search_files(case.query_file, case.ref_files, **kw, **engine_kw)
    return out, out.csv(kw["acgt"])


# ---------------------------------------------------------------------------------------------------------------
# golden tables of the reference binary
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", case_names())
def test_golden_csv_bit_exact(name):
    """ALL cases, the pool-dependent ones (hazard H1) included: the search emulates the reference's --pool batches"""
    c = Case(name)
    out, csv = gpu_csv(c)
    assert csv == c.expected_csv
    # .aln.xz content: every reference that ever entered a heap, in file order (nearest.c:303-314)
    got = [out.ref_names[i] for i in np.nonzero(out.ever_best)[0]]
    assert got == c.expected_aln_names
    if not c.h1:
        assert out.stats["n_h1_flags"] == 0


@pytest.mark.parametrize("name", H1)
@pytest.mark.parametrize("knobs", [dict(first_round=32, max_round=64), dict(batch_rows=97), dict(batch_rows=5, first_round=64, max_round=64)])
def test_h1_cases_independent_of_round_schedule(name, knobs):
    """the reference's batches are independent of how the host cuts its pushes and of the scoring rounds"""
    c = Case(name)
    _, csv = gpu_csv(c, **knobs)
    assert csv == c.expected_csv


@pytest.mark.parametrize("name", ["syn_default", "syn_acgt", "syn_n100", "ties_n5", "amb_default", "c2_shape", "syn_exclude_self"])
@pytest.mark.parametrize("knobs", [dict(first_round=32, max_round=64), dict(first_round=32, max_round=256, early_abort=True),
                                   dict(batch_rows=97), dict(batch_rows=1000, first_round=64, max_round=64, early_abort=True)])
def test_golden_csv_independent_of_round_schedule(name, knobs):
    """rounds, streaming batches and the early abort are performance knobs: the table must not move"""
    c = Case(name)
    _, csv = gpu_csv(c, **knobs)
    assert csv == c.expected_csv


@pytest.mark.parametrize("name", H1)
def test_closed_form_mode_flags_the_h1_hazard(name):
    """uvg_params.closed_form = 1: the pool-independent closed form (what sharded contexts run).  It must equal the
    oracle's closed form and FLAG that the reference would have truncated (never silently differ)."""
    c = Case(name)
    kw = c.options()
    out, csv = gpu_csv(c, closed_form=True)
    closed = orc.run_files(c.query_file, c.ref_files, **kw).csv(kw["acgt"])
    assert csv == closed
    lit = orc.run_files(c.query_file, c.ref_files, literal=True, **kw)
    assert lit.csv(kw["acgt"]) == c.expected_csv
    if lit.h1_events:
        assert out.stats["n_h1_flags"] > 0
    if csv != c.expected_csv:
        assert out.stats["n_h1_flags"] > 0


# ---------------------------------------------------------------------------------------------------------------
# kernels one by one against the oracle
# ---------------------------------------------------------------------------------------------------------------
def make_inputs(L, nq, nr, seed, amb_q=0.0, amb_r=0.0, clades=12):
    m = SynthModel(nsites=L, n_clades=clades, seed=seed, clade_subs=max(3, L // 500))
    qn, Q = m.sample(nq, seed + 1, ambiguous_frac=amb_q, prefix="Q")
    rn, R = m.sample(nr, seed + 2, ambiguous_frac=amb_r, prefix="R")
    return qn, Q, rn, R


@pytest.mark.parametrize("L,acgt", [(1000, False), (1000, True), (2017, False), (29903, False), (33, False), (7, True)])
def test_encoder_parity(L, acgt):
    qn, Q, rn, R = make_inputs(L, 5, 300, 7 + L, amb_r=0.05)
    R[3, :] = ord("N")                       # a reference the filter must drop
    R[4, : L // 2 + 1] = ord("-")
    R[5] = np.frombuffer((b"acgtnxo?.-RYKMSWBDHVUEZ*" * (L // 24 + 1))[:L], np.uint8)      # every symbol class, lower case too
    names, Qo, tr, cons, idx, idx_m, idx_c = orc.prepare_queries(qn, [q.tobytes() for q in Q], 0, acgt, 1.0)
    with Engine(L, nbest=5, acgt=acgt) as eng:
        eng.set_queries(Qo)
        min_non_n = int(L * 0.5)
        nn, acc, na = eng.push_refs(R, min_non_n)
        want_nn = (~orc.INDEL[R]).sum(1)
        assert np.array_equal(nn, want_nn)                                   # quick_count_sequence_non_N
        assert np.array_equal(acc.astype(bool), want_nn >= min_non_n)       # nearest.c:264
        assert na == int((want_nn >= min_non_n).sum()) == eng.db_count
        lut = orc.CODE.copy()
        lut[[ord(c) for c in "acgtumrwsykvhdb"]] = lut[[ord(c) for c in "ACGTUMRWSYKVHDB"]]      # reader upper-cases
        if acgt:
            lut = np.zeros(256, np.uint8)
            for k, ch in enumerate(b"ACGT"):
                lut[ch] = 1 << k
                lut[ch + 32] = 1 << k
        kept = np.nonzero(acc)[0]
        for j in list(range(min(6, len(kept)))) + [len(kept) // 2, len(kept) - 1]:
            assert np.array_equal(eng.debug_unpack_ref(j), lut[R[kept[j]]]), j


@pytest.mark.parametrize("L,nq,nr,amb_q,amb_r,trim", [(1000, 70, 200, 0.0, 0.0, 0), (1500, 33, 130, 0.2, 0.05, 0), (2000, 9, 97, 0.1, 0.1, 123),
                                                      (29903, 10, 64, 0.02, 0.01, 0), (40, 3, 33, 0.1, 0.1, 0)])
def test_pair_counts_default_mode(L, nq, nr, amb_q, amb_r, trim):
    qn, Q, rn, R = make_inputs(L, nq, nr, 17 + L, amb_q, amb_r)
    names, Qo, tr, cons, idx, idx_m, idx_c = orc.prepare_queries(qn, [q.tobytes() for q in Q], trim, False, 1.0)
    inS = np.zeros(L, bool)
    inS[tr:L - tr] = True
    qc = np.where(inS[None, :], orc.CODE[Qo], 0).astype(np.uint8)
    with Engine(L, nbest=5, trim=trim) as eng:
        info = eng.set_queries(Qo)
        assert (info.n_idx, info.n_idx_m, info.n_idx_c, info.trim) == (len(idx), len(idx_m), len(idx_c), tr)
        eng.push_refs(R, 0)
        got = eng.debug_pair_counts(0, nr)
        want = np.stack(orc.pair_counts(qc, orc.CODE[R]), -1)
        assert np.array_equal(got, want)
        # a sub-range that does not start at a group boundary, and the row-restricted variant
        if nr > 40:
            assert np.array_equal(eng.debug_pair_counts(5, 30), want[:, 5:35])
            assert np.array_equal(eng.debug_pair_counts_rows(1, 2, 7, 20), want[1:3, 7:27])
        # per-reference consensus-stage values
        nn, c0, ci = eng.debug_ref_meta(0, nr)
        rc = orc.CODE[R]
        mask_c = np.zeros(L, bool)
        mask_c[idx_c] = True
        cc = orc.CODE[cons]
        want_c0 = ((cc[None, :] == rc) & (orc.POPC4[rc] == 1) & mask_c[None, :]).sum(1)
        want_v = ((rc != 0) & (cc[None, :] != 0) & mask_c[None, :]).sum(1)
        assert np.array_equal(c0, want_c0) and np.array_equal(ci, want_v - want_c0)
        assert np.array_equal(nn, (~orc.INDEL[R]).sum(1))


@pytest.mark.parametrize("L,nq,nr,amb,trim", [(1000, 70, 200, 0.0, 0), (1500, 33, 130, 0.2, 0), (2000, 9, 97, 0.1, 77)])
def test_pair_counts_acgt_mode(L, nq, nr, amb, trim):
    qn, Q, rn, R = make_inputs(L, nq, nr, 23 + L, amb, amb / 2)
    names, Qo, tr, cons, idx, idx_m, idx_c = orc.prepare_queries(qn, [q.tobytes() for q in Q], trim, True, 1.0)
    inS = np.zeros(L, bool)
    inS[tr:L - tr] = True
    qa = orc.ACGT[Qo] & inS[None, :]
    ra = orc.ACGT[R]
    mask_i = np.zeros(L, bool)
    mask_i[idx] = True
    f = np.float32
    V = (qa.astype(f) @ ra.astype(f).T).astype(np.int32)
    eq = sum(((qa & (Qo == ch)).astype(f) @ (R == ch).astype(f).T).astype(np.int32) for ch in b"ACGT")
    eqi = sum((((qa & (Qo == ch)) & mask_i[None, :]).astype(f) @ (R == ch).astype(f).T).astype(np.int32) for ch in b"ACGT")
    Vi = ((qa & mask_i[None, :]).astype(f) @ ra.astype(f).T).astype(np.int32)
    with Engine(L, nbest=5, trim=trim, acgt=True) as eng:
        eng.set_queries(Qo)
        eng.push_refs(R, 0)
        got = eng.debug_pair_counts(0, nr)
    assert np.array_equal(got[..., 0], eq)              # matches
    assert np.array_equal(got[..., 1], V)               # valid ACGT pairs
    assert np.array_equal(got[..., 2], Vi - eqi)        # mismatches on segregating (idx) sites


# ---------------------------------------------------------------------------------------------------------------
# whole searches against the oracle on seeded random inputs
# ---------------------------------------------------------------------------------------------------------------
def run_vs_oracle(L, nq, nr, nbest, acgt=False, trim=0, amb=0.0, seed=5, ambig_r=0.5, **engine_kw):
    qn, Q, rn, R = make_inputs(L, nq, nr, seed, amb, amb / 2)
    qs = [q.tobytes() for q in Q]
    rs = [r.tobytes() for r in R]
    ref = orc.search(qn, qs, None, nbest=nbest, trim=trim, acgt=acgt, refs=[(rn, rs)], pool=64, ambig_r=ambig_r)
    names, Qo, tr, *_ = orc.prepare_queries(qn, qs, trim, acgt, 0.5)
    if nq * nr <= 60000:
        # small enough for the oracle's literal (staged, truncated, batch by batch) mode: the library's default mode
        # emulates the reference's --pool batches and must agree with it even where the closed form does not
        for pool in (64, 5):
            lit = orc.search(qn, qs, None, nbest=nbest, trim=trim, acgt=acgt, refs=[(rn, rs)], pool=pool, ambig_r=ambig_r, literal=True)
            with Engine(L, nbest=nbest, trim=trim, acgt=acgt, pool=pool, **engine_kw) as eng:
                eng.set_queries(Qo)
                eng.push_refs(R, int(L * (1 - ambig_r)))
                eng.search()
                for iq in range(len(names)):
                    sc, refs = eng.heap(iq)
                    want = lit.rows[iq]
                    assert np.array_equal(sc, np.array([w[0] for w in want], np.int32).reshape(-1, 6)), f"literal, pool {pool}, query {iq}"
                    assert np.array_equal(refs, np.array([w[2] for w in want], np.int64)), f"literal, pool {pool}, query {iq}"
                assert np.array_equal(np.nonzero(eng.ever_best(0, eng.db_count))[0], np.array(lit.ever_best, np.int64))
    with Engine(L, nbest=nbest, trim=trim, acgt=acgt, pool=64, closed_form=True, **engine_kw) as eng:
        eng.set_queries(Qo)
        nn, acc, na = eng.push_refs(R, int(L * (1 - ambig_r)))
        eng.search()
        assert na == ref.n_ref_accepted
        for iq in range(len(names)):
            sc, refs = eng.heap(iq)
            want = ref.rows[iq]
            ws = np.array([w[0] for w in want], np.int32).reshape(-1, 6)
            wr = np.array([w[2] for w in want], np.int64)
            assert np.array_equal(sc, ws) and np.array_equal(refs, wr), f"query {iq}"
        eb = eng.ever_best(0, na)
        assert np.array_equal(np.nonzero(eb)[0], np.array(ref.ever_best, np.int64))
        st = eng.stats()
    return st


@pytest.mark.parametrize("kw", [
    dict(L=1000, nq=10, nr=300, nbest=5),
    dict(L=2000, nq=70, nr=2500, nbest=10, first_round=64, max_round=512),
    dict(L=2000, nq=70, nr=2500, nbest=10, first_round=64, max_round=512, early_abort=True),
    dict(L=2000, nq=33, nr=1500, nbest=10, acgt=True),
    dict(L=2000, nq=33, nr=1500, nbest=10, trim=100),
    dict(L=2000, nq=33, nr=1500, nbest=1, amb=0.1),
    dict(L=2000, nq=40, nr=3000, nbest=20, amb=0.2, first_round=32, max_round=256),
    dict(L=2000, nq=40, nr=3000, nbest=20, amb=0.2, first_round=32, max_round=256, early_abort=True, acgt=True),
    dict(L=29903, nq=20, nr=2000, nbest=10, early_abort=True),
    dict(L=500, nq=130, nr=5000, nbest=100, ambig_r=0.2),
    dict(L=64, nq=1, nr=100, nbest=3, seed=9),
    dict(L=31, nq=65, nr=33, nbest=40, seed=10),
])
def test_search_vs_oracle(kw):
    run_vs_oracle(**kw)


def test_long_alignment_counters_do_not_wrap():
    """alignments of 65,536 columns or more take the wide-counter variant of the pair-count kernel"""
    run_vs_oracle(L=70001, nq=6, nr=96, nbest=4, amb=0.02, seed=77)


def test_five_megabase_alignment_vs_oracle():
    """BASELINE config 5's column count (5 Mb core-genome alignment) with a few queries and references: 19,532 chunks per row,
    32-bit totals in the pair-count kernel, the literal replay's prefix recount over 39,063 units"""
    run_vs_oracle(L=5_000_001, nq=3, nr=40, nbest=5, amb=0.001, seed=123)


def test_streaming_push_search_drop_equals_resident_search():
    L = 1200
    qn, Q, rn, R = make_inputs(L, 20, 4000, 41, 0.05, 0.02)
    names, Qo, *_ = orc.prepare_queries(qn, [q.tobytes() for q in Q], 0, False, 0.5)
    with Engine(L, nbest=15, first_round=64, max_round=1024) as a, Engine(L, nbest=15) as b, Engine(L, nbest=15, first_round=32, max_round=256) as p:
        a.set_queries(Qo)
        b.set_queries(Qo)
        p.set_queries(Qo)
        a.push_refs(R, 600)
        a.search()
        for s in range(0, len(R), 333):
            b.push_refs(R[s:s + 333], 600)
            b.search()
            b.drop_refs()
        # pipelined host: the next push (copy + encoder) runs under the rounds enqueued by search_async
        for s in range(0, len(R), 700):
            p.push_refs(R[s:s + 700], 600)
            p.search_async()
        p.wait()
        for iq in range(len(names)):
            sa, ra = a.heap(iq)
            sp, rp = p.heap(iq)
            assert np.array_equal(sa, sp) and np.array_equal(ra, rp)
        assert np.array_equal(a.ever_best(0, a.db_count), p.ever_best(0, a.db_count))
        for iq in range(len(names)):
            sa, ra = a.heap(iq)
            sb, rb = b.heap(iq)
            assert np.array_equal(sa, sb) and np.array_equal(ra, rb)
        na = a.db_count
        assert np.array_equal(a.ever_best(0, na), b.ever_best(0, na))
        # repeating the search over the same resident database is idempotent
        h0 = [a.heap(i) for i in range(len(names))]
        a.reset_search()
        a.search()
        for iq in range(len(names)):
            assert np.array_equal(a.heap(iq)[0], h0[iq][0]) and np.array_equal(a.heap(iq)[1], h0[iq][1])


def test_edge_cases():
    L = 300
    qn, Q, rn, R = make_inputs(L, 4, 40, 51)
    names, Qo, *_ = orc.prepare_queries(qn, [q.tobytes() for q in Q], 0, False, 1.0)
    with Engine(L, nbest=6) as eng:
        with pytest.raises(UvgError):
            eng.search()                                   # no queries yet
        eng.set_queries(Qo)
        eng.search()                                       # empty database
        assert all(len(eng.heap(i)[1]) == 0 for i in range(len(names)))
        nn, acc, na = eng.push_refs(np.full((7, L), ord("N"), np.uint8), 10)
        assert na == 0 and eng.db_count == 0               # everything filtered
        eng.search()
        assert all(len(eng.heap(i)[1]) == 0 for i in range(len(names)))
        # hazard H2: K-1 neighbours crash the reference's writer; we return the K-1 rows
        eng.push_refs(R[:5], 0)
        eng.search()
        for i in range(len(names)):
            sc, refs = eng.heap(i)
            assert len(refs) == 5 and sorted(refs.tolist()) == [0, 1, 2, 3, 4]
            assert all(tuple(sc[j]) >= tuple(sc[j + 1]) for j in range(4))
        with pytest.raises(UvgError):
            eng.heap(99)
    with pytest.raises(UvgError):
        Engine(0)
    # nbest 1 keeps two neighbours (min_heap.c:58, hazard H3)
    with Engine(L, nbest=1) as eng:
        eng.set_queries(Qo)
        eng.push_refs(R, 0)
        eng.search()
        assert all(len(eng.heap(i)[1]) == 2 for i in range(len(names)))


def test_query_identical_to_a_reference_ranks_first():
    """only for queries without ambiguity codes: an IUPAC site is a valid comparison that is never an ACGT match, so an
    exact copy of an ambiguous query has D = #ambiguous sites and the reference may refuse it (D >= max_incompatible)"""
    L = 2500
    qn, Q, rn, R = make_inputs(L, 12, 1500, 61, 0.0, 0.01)
    for q in Q:                                  # the generator adds ~1.7 two-base codes per genome: resolve them
        amb = orc.POPC4[orc.CODE[q]] > 1
        q[amb] = ord("A")
    names, Qo, *_ = orc.prepare_queries(qn, [q.tobytes() for q in Q], 0, False, 0.5)
    R = R.copy()
    planted = {}
    for iq in range(len(names)):
        pos = 100 + 97 * iq
        R[pos] = Qo[iq]
        planted[iq] = pos
    with Engine(L, nbest=10, first_round=128, max_round=512) as eng:
        eng.set_queries(Qo)
        eng.push_refs(R, 0)
        eng.search()
        for iq, pos in planted.items():
            sc, refs = eng.heap(iq)
            code = orc.CODE[Qo[iq]]
            n_valid = int((code != 0).sum())
            n_one = int((orc.POPC4[code] == 1).sum())
            j = list(refs).index(pos)
            assert sc[j, 3] == n_valid and sc[j, 0] == n_one and sc[j, 1] == n_valid and sc[j, 2] == n_valid
            assert sc[0, 0] >= sc[j, 0]


# ---------------------------------------------------------------------------------------------------------------
# the reference binary itself, when it travelled to the box (oracle/_ref is git-ignored but not gpurun-ignored)
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/uvaia_ref not present")
@pytest.mark.parametrize("seed,acgt,trim,nbest", [(201, False, 0, 25), (202, True, 0, 8), (203, False, 150, 100)])
def test_vs_live_reference_binary(seed, acgt, trim, nbest):
    m = SynthModel(nsites=4000, n_clades=10, seed=seed, clade_subs=10)
    qn, Q = m.sample(90, seed + 1, ambiguous_frac=0.03, prefix="Q")
    rn, R = m.sample(2500, seed + 2, ambiguous_frac=0.01, prefix="R")
    with tempfile.TemporaryDirectory() as wd:
        qf, rf = os.path.join(wd, "q.fa"), os.path.join(wd, "r.fa")
        write_fasta(qf, qn, Q)
        write_fasta(rf, rn, R, width=80)
        cmd = [REF_BIN, "-n", str(nbest), "-t", "4", "--trim", str(trim)] + (["--acgt"] if acgt else []) + ["-r", rf, "-o", os.path.join(wd, "o"), qf]
        p = subprocess.run(cmd, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert p.returncode == 0, p.stderr.decode()[-500:]
        want = gzip.decompress(open(os.path.join(wd, "o.csv.xz"), "rb").read()).decode()
        out = host.search_files(qf, [rf], nbest=nbest, trim=trim, acgt=acgt, first_round=256, max_round=1024)
    assert out.csv(acgt) == want


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/uvaia_ref not present")
@pytest.mark.parametrize("nq,acgt,seed", [(1, False, 301), (1, True, 302), (2, False, 303), (3, False, 304), (3, True, 305), (1, False, 306)])
@pytest.mark.parametrize("pool", [1, 7, 64, 512])
def test_pool_dependence_vs_live_reference_binary(nq, acgt, seed, pool):
    """hazard H1 against the reference program itself: one to three queries (every / most sites are consensus sites),
    N-rich references in two files, and the table the reference writes for this --pool"""
    L = 1500
    m = SynthModel(nsites=L, n_clades=6, seed=seed, clade_subs=12)
    qn, Q = m.sample(nq, seed + 1, ambiguous_frac=0.0 if seed != 306 else 0.02, prefix="Q")
    rn, R = m.sample(1300, seed + 2, ambiguous_frac=0.01, prefix="R")
    rng = np.random.default_rng(seed)
    for r in R:                                  # N runs anywhere: the consensus-stage mismatches move around
        for _ in range(int(rng.integers(0, 4))):
            a = int(rng.integers(0, L - 50))
            r[a:a + int(rng.integers(5, 200))] = ord("N")
    with tempfile.TemporaryDirectory() as wd:
        qf, rf1, rf2 = os.path.join(wd, "q.fa"), os.path.join(wd, "r1.fa"), os.path.join(wd, "r2.fa")
        write_fasta(qf, qn, Q)
        write_fasta(rf1, rn[:801], R[:801], width=80)
        write_fasta(rf2, rn[801:], R[801:])
        cmd = [REF_BIN, "-n", "12", "-t", "2", "-p", str(pool)] + (["--acgt"] if acgt else []) + ["-r", rf1, "-r", rf2, "-o", os.path.join(wd, "o"), qf]
        p = subprocess.run(cmd, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert p.returncode == 0, p.stderr.decode()[-500:]
        want = gzip.decompress(open(os.path.join(wd, "o.csv.xz"), "rb").read()).decode()
        out = host.search_files(qf, [rf1, rf2], nbest=12, acgt=acgt, pool=pool, first_round=128, max_round=512, batch_rows=333)
    assert out.csv(acgt) == want


@pytest.mark.parametrize("nbest", [2048, 2500])
def test_final_tables_sorted_on_device_and_on_host(nbest):
    """the final order of a table = heap array after the slot swap + a STABLE descending sort (glibc's qsort is a merge
    sort): up to 2,048 slots k_finalise ranks the rows on the device, larger heaps go through libc qsort on the host.
    Many identical references (ties in all six keys) on both sides of the limit, uvg_get_heap and uvg_get_all_heaps."""
    L, nq, nr = 600, 5, 3000
    qn, Q, rn, R = make_inputs(L, nq, nr, 31, 0.0, 0.0)
    R[100:900] = R[100]                                  # 800 identical rows
    R[1500:1800] = R[7]
    qs = [q.tobytes() for q in Q]
    ref = orc.search(qn, qs, None, nbest=nbest, refs=[(rn, [r.tobytes() for r in R])], pool=64)
    names, Qo, *_ = orc.prepare_queries(qn, qs, 0, False, 0.5)
    with Engine(L, nbest=nbest, pool=64, closed_form=True, first_round=256, max_round=1024) as eng:
        eng.set_queries(Qo)
        eng.push_refs(R, int(L * 0.5))
        eng.search()
        tables = eng.heaps()
        for iq in range(len(names)):
            want = ref.rows[iq]
            ws = np.array([w[0] for w in want], np.int32).reshape(-1, 6)
            wr = np.array([w[2] for w in want], np.int64)
            for sc, rf in (eng.heap(iq), tables[iq]):
                assert np.array_equal(sc, ws) and np.array_equal(rf, wr), f"query {iq}"


def test_pivot_form_resident_db_requery_and_divergent_queries(pair_kernel):
    """pivot form of the pair-count kernel: (1) references pushed BEFORE the queries get their derived planes in
    uvg_set_queries, (2) a second query set on the same resident database re-derives them, (3) queries far from any
    pivot (independent random sequences: thousands of fix entries each) stay exact when the form is forced."""
    L, nr = 3000, 700
    qn, Q, rn, R = make_inputs(L, 40, nr, 91, 0.03, 0.02)
    rng = np.random.default_rng(3)
    Qdiv = np.frombuffer(b"ACGT", np.uint8)[rng.integers(0, 4, size=(70, L))]          # no common pivot
    Qdiv[:, :50] = ord("N")
    with Engine(L, nbest=7) as eng:
        eng.push_refs(R, 0)
        for Qset in (Q, Qdiv, Q[::-1]):
            names, Qo, *_ = orc.prepare_queries([b"q%d" % i for i in range(len(Qset))], [q.tobytes() for q in Qset], 0, False, 1.0)
            eng.set_queries(Qo)
            got = eng.debug_pair_counts(0, nr)
            want = np.stack(orc.pair_counts(orc.CODE[Qo], orc.CODE[R]), -1)
            assert np.array_equal(got, want)
            st = eng.stats()
            assert (st["pivot_fix_entries"] >= 0) == (pair_kernel == "pivot")
        # automatic choice: close queries take the pivot form, divergent ones the general kernel
    with Engine(L, nbest=7, pair_kernel=0) as eng:
        eng.push_refs(R, 0)
        for Qset, want_pivot in ((Q, True), (Qdiv, False)):
            names, Qo, *_ = orc.prepare_queries([b"q%d" % i for i in range(len(Qset))], [q.tobytes() for q in Qset], 0, False, 1.0)
            eng.set_queries(Qo)
            assert (eng.stats()["pivot_fix_entries"] >= 0) == want_pivot
            assert np.array_equal(eng.debug_pair_counts(0, nr), np.stack(orc.pair_counts(orc.CODE[Qo], orc.CODE[R]), -1))
