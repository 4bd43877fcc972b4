"""Packed reference cache (uvg_export_packed / uvg_import_packed): a database that went through the encoder once and comes
back as a blob gives the same tables, bit for bit, as the text it was made from."""
import numpy as np
import pytest

from helpers import orc
from uvaia_b200.engine import Engine, UvgError
from uvaia_b200.synth import SynthModel

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["pivot", "general"])
def pair_kernel(request, monkeypatch):
    monkeypatch.setenv("UVG_PAIR_KERNEL", request.param)
    return request.param


def make_inputs(L, nq, nr, seed, amb_q=0.0, amb_r=0.0):
    m = SynthModel(nsites=L, n_clades=10, seed=seed, clade_subs=max(2, L // 200))
    qn, Q = m.sample(nq, seed + 1, ambiguous_frac=amb_q, prefix="Q")
    rn, R = m.sample(nr, seed + 2, ambiguous_frac=amb_r, prefix="R")
    return qn, Q, rn, R


def tables(eng, nq):
    return [eng.heap(i) for i in range(nq)], eng.ever_best(0, eng.db_count)


def same_tables(a, b):
    (ha, ea), (hb, eb) = a, b
    return all(np.array_equal(x[0], y[0]) and np.array_equal(x[1], y[1]) for x, y in zip(ha, hb)) and np.array_equal(ea, eb)


@pytest.mark.parametrize("L,nr,acgt,amb_r,min_non_n", [(1200, 1000, False, 0.03, 500), (1200, 1024, True, 0.03, 0), (29903, 150, False, 0.01, 0),
                                                       (333, 77, False, 0.3, 0)])
def test_export_import_round_trip(L, nr, acgt, amb_r, min_non_n):
    qn, Q, rn, R = make_inputs(L, 24, nr, 61, 0.04, amb_r)
    names, Qo, *_ = orc.prepare_queries(qn, [q.tobytes() for q in Q], 0, acgt, 0.5)
    with Engine(L, nbest=9, acgt=acgt) as a, Engine(L, nbest=9, acgt=acgt) as b, Engine(L, nbest=9, acgt=acgt) as c:
        a.set_queries(Qo)
        a.push_refs(R, min_non_n)
        a.search()
        want = tables(a, len(names))
        blob = a.export_packed()
        # (1) blob into a context that already has its queries
        b.set_queries(Qo)
        assert b.import_packed(blob) == a.db_count
        b.search()
        assert same_tables(want, tables(b, len(names)))
        assert np.array_equal(b.export_packed(), blob)              # export(import(blob)) == blob
        # (2) blob first, queries afterwards; in pieces that end at multiples of 32
        n = a.db_count
        cut = (n // 2) // 32 * 32
        pieces = [a.export_packed(0, cut), a.export_packed(cut, n - cut)] if cut else [blob]
        for piece in pieces:
            c.import_packed(piece)
        c.set_queries(Qo)
        c.search()
        assert same_tables(want, tables(c, len(names)))
        assert np.array_equal(c.export_packed(), blob)
        # the reference metadata came along
        assert np.array_equal(np.stack(a.debug_ref_meta(0, n)), np.stack(c.debug_ref_meta(0, n)))


def test_blob_then_text_and_streaming():
    L = 900
    qn, Q, rn, R = make_inputs(L, 16, 640 + 123, 71, 0.02, 0.02)
    names, Qo, *_ = orc.prepare_queries(qn, [q.tobytes() for q in Q], 0, False, 0.5)
    with Engine(L, nbest=5) as a, Engine(L, nbest=5) as b, Engine(L, nbest=5) as s:
        a.set_queries(Qo)
        a.push_refs(R, 0)
        a.search()
        want = tables(a, len(names))
        blob = a.export_packed(0, 640)
        b.set_queries(Qo)
        b.import_packed(blob)
        b.push_refs(R[640:], 0)                                     # text appended behind a blob
        b.search()
        assert same_tables(want, tables(b, len(names)))
        # streaming host: import, search, drop
        s.set_queries(Qo)
        for first in range(0, a.db_count, 256):
            s.import_packed(a.export_packed(first, min(256, a.db_count - first)))
            s.search()
            s.drop_refs()
        for iq in range(len(names)):
            assert np.array_equal(s.heap(iq)[0], want[0][iq][0]) and np.array_equal(s.heap(iq)[1], want[0][iq][1])


def test_refusals():
    L = 400
    qn, Q, rn, R = make_inputs(L, 4, 100, 81)
    with Engine(L, nbest=3) as a, Engine(L + 1, nbest=3) as other, Engine(L, nbest=3, acgt=True) as mode:
        a.push_refs(R, 0)
        blob = a.export_packed()
        with pytest.raises(UvgError, match="multiple of 32"):
            a.export_packed(5, 10)
        with pytest.raises(UvgError, match="not resident"):
            a.export_packed(0, 101)
        with pytest.raises(UvgError, match="append at multiples"):
            a.import_packed(blob)                                   # 100 resident references: not a group boundary
        with pytest.raises(UvgError, match="nchar"):
            other.import_packed(blob)
        with pytest.raises(UvgError, match="acgt_mode"):
            mode.import_packed(blob)
        with pytest.raises(UvgError, match="truncated"):
            Engine(L, nbest=3).import_packed(blob[:-16])
        bad = blob.copy()
        bad[:4] = 0
        with pytest.raises(UvgError, match="magic"):
            Engine(L, nbest=3).import_packed(bad)
