"""The numpy oracle against the end-to-end goldens (CSV tables written by the compiled reference binary)."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from helpers import ROOT, Case, case_names, orc

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "uvaia_ref")


@pytest.mark.parametrize("name", case_names())
def test_oracle_reproduces_reference_csv(name):
    c = Case(name)
    kw = c.options()
    res = orc.run_files(c.query_file, c.ref_files, literal=c.h1, **kw)
    assert res.csv(kw["acgt"]) == c.expected_csv
    # .aln.xz: every reference that ever entered a heap, in file order (names only: H4)
    names = [res.ref_names[o] for o in res.ever_best] if res.ref_names else None
    if names is not None:
        assert names == c.expected_aln_names


@pytest.mark.parametrize("name", ["syn_default", "syn_acgt", "ties_n5", "amb_default", "hygiene"])
def test_literal_mode_equals_closed_form_when_h1_is_silent(name):
    c = Case(name)
    kw = c.options()
    a = orc.run_files(c.query_file, c.ref_files, literal=True, **kw)
    assert a.h1_events == 0
    assert a.csv(kw["acgt"]) == c.expected_csv


def test_h1_cases_are_pool_dependent_in_the_reference():
    csvs = {p: Case(f"h1_one_query_p{p}").expected_csv for p in (1, 7, 512)}
    assert len(set(csvs.values())) >= 2
    # closed form == the reference at pool 1 (every batch has one reference: nothing can be truncated late)
    c = Case("h1_one_query_p1")
    assert orc.run_files(c.query_file, c.ref_files, **c.options()).csv() == csvs[1]


@pytest.mark.skipif(not os.path.exists(REF_BIN), reason="oracle/_ref/uvaia_ref not built (bash oracle/build_ref.sh)")
@pytest.mark.parametrize("seed,acgt,trim", [(101, False, 0), (102, True, 0), (103, False, 57)])
def test_oracle_vs_live_reference_binary(seed, acgt, trim):
    import gzip
    from uvaia_b200.synth import SynthModel, write_fasta
    m = SynthModel(nsites=1200, n_clades=8, seed=seed, clade_subs=8)
    qn, Q = m.sample(17, seed + 1, ambiguous_frac=0.05, prefix="Q")
    rn, R = m.sample(700, seed + 2, ambiguous_frac=0.02, prefix="R")
    with tempfile.TemporaryDirectory() as wd:
        qf, rf = os.path.join(wd, "q.fa"), os.path.join(wd, "r.fa")
        write_fasta(qf, qn, Q)
        write_fasta(rf, rn, R, width=70)
        cmd = [REF_BIN, "-n", "8", "-t", "2", "--trim", str(trim)] + (["--acgt"] if acgt else []) + ["-r", rf, "-o", os.path.join(wd, "o"), qf]
        p = subprocess.run(cmd, cwd=wd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert p.returncode == 0, p.stderr.decode()[-500:]
        want = gzip.decompress(open(os.path.join(wd, "o.csv.xz"), "rb").read()).decode()
        got = orc.run_files(qf, [rf], nbest=8, trim=trim, acgt=acgt).csv(acgt)
    assert got == want
