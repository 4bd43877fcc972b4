"""Host-side mirror of the reference flow (uvaia_b200/host.py): FASTA hygiene, query filter, ordering, CSV layout.
No GPU: the engine is not touched here."""
import lzma

import numpy as np
import pytest

from helpers import Case, case_names, orc
from uvaia_b200 import host


@pytest.mark.parametrize("name", ["hygiene", "syn_default", "real_c1_small", "amb_default"])
def test_fasta_reader_matches_oracle_reader(name):
    c = Case(name)
    for f in [c.query_file] + c.ref_files:
        names, seqs = orc.read_fasta(f)
        got = list(host.iter_fasta(f))
        assert [g[0] for g in got] == names
        assert [g[1] for g in got] == seqs


def test_fasta_hygiene_rules(tmp_path):
    raw = (b"# comment\n; comment\n\n>seq one | x\nacg t\nNN\r\n>s2\nA\nAC-GT\n  \n x>weird header\nGGGG\n")
    p = tmp_path / "h.fa"
    p.write_bytes(raw)
    recs = list(host.iter_fasta(str(p)))
    # lines shorter than 2 bytes are dropped ("A"), '\r' acts as a line break, any line containing '>' is a header whose
    # name is everything after the first '>', sequence lines lose white space and are upper-cased
    assert recs == [(b"seq one | x", b"ACGTNN"), (b"s2", b"AC-GT"), (b"weird header", b"GGGG")]
    assert recs == list(zip(*orc.read_fasta(str(p))))


@pytest.mark.parametrize("codec", ["raw", "gz", "xz", "bz2"])
def test_codec_sniffing_by_content(tmp_path, codec):
    import bz2, gzip
    body = b">a\nACGTACGT\n>b\nACGTNNGT\n"
    data = {"raw": body, "gz": gzip.compress(body), "xz": lzma.compress(body), "bz2": bz2.compress(body)}[codec]
    p = tmp_path / "file.with.misleading.ext.xz"
    p.write_bytes(data)
    assert list(host.iter_fasta(str(p))) == [(b"a", b"ACGTACGT"), (b"b", b"ACGTNNGT")]


@pytest.mark.parametrize("name", ["syn_query_filter", "amb_strict", "amb_default", "real_c1_small"])
def test_query_filter_and_order_match_oracle(name):
    c = Case(name)
    kw = c.options()
    qn, qs = orc.read_fasta(c.query_file)
    keep = host.filter_queries(qs, min(1.0, max(0.001, kw["ambig_q"])))
    assert keep == orc.filter_queries(qn, qs, min(1.0, max(0.001, kw["ambig_q"])))
    Q = np.frombuffer(b"".join(qs[i] for i in keep), np.uint8).reshape(len(keep), -1)
    t = host.clamp_trim(kw["trim"], Q.shape[1])
    assert t == orc.clamp_trim(kw["trim"], Q.shape[1])
    assert host.query_order(Q, t, kw["acgt"]) == orc.query_order(Q, t, kw["acgt"])
    # the first CSV column of the golden lists the queries in exactly that order
    order = host.query_order(Q, t, kw["acgt"])
    want = []
    for ln in c.expected_csv.split("\n")[1:]:
        if ln and (not want or want[-1] != ln.split(",")[0]):
            want.append(ln.split(",")[0])
    assert [qn[keep[i]].decode() for i in order] == want


def test_trim_clamp():
    assert host.clamp_trim(-3, 1000) == 0
    assert host.clamp_trim(100, 1000) == 100
    assert host.clamp_trim(600, 1000) == int(1000 / 2.1)


def test_unaligned_queries_are_fatal():
    with pytest.raises(ValueError):
        host.filter_queries([b"ACGTACGT", b"ACGTACG"], 0.5)


def test_csv_layout():
    out = host.SearchOutput()
    out.qnames = [b"q1"]
    out.ref_names = [b"r0", b"r1"]
    out.rows = [(np.array([[9, 9, 9, 10, 2, 11], [8, 8, 9, 10, 1, 12]], np.int32), np.array([1, 0]))]
    assert out.csv(False) == host.HEADER_FULL + "\nq1,r1,1,9,9,9,10,2,11\nq1,r0,2,8,8,9,10,1,12\n"
    assert out.csv(True).startswith(host.HEADER_ACGT + "\n")
    assert host.HEADER_FULL == orc.HEADER_FULL and host.HEADER_ACGT == orc.HEADER_ACGT
    # headers as written by the reference binary
    assert Case("syn_default").expected_csv.split("\n")[0] == host.HEADER_FULL
    assert Case("syn_acgt").expected_csv.split("\n")[0] == host.HEADER_ACGT
