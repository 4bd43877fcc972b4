"""BASELINE.json configs[1] at full size (1k queries x 1M references x 29,903 sites, nbest 100) through size-independent
properties, plus an exact re-derivation of sampled queries' tables with the oracle's heap from the raw pair counts."""
import numpy as np
import pytest

from helpers import orc
from uvaia_b200 import host
from uvaia_b200.engine import Engine
from uvaia_b200.synth import SynthModel

pytestmark = pytest.mark.gpu

L, R, NQ, NBEST = 29903, 1_000_000, 1000, 100


def replay_with_oracle_heap(counts, non_n, c0, nbest, L):
    """sequential admission of SURVEY.md 3.3 over raw counts (n,4) with the oracle's Heap; numpy only skips stretches
    that cannot be admitted because D >= max_incompatible"""
    h = orc.Heap(nbest, L)
    D = (counts[:, 3] - counts[:, 0]).astype(np.int64)
    n = len(D)
    pos = 0
    admitted = []
    while pos < n:
        cand = np.nonzero(D[pos:pos + 200000] < h.max_inc)[0]
        if len(cand) == 0:
            pos += 200000
            continue
        moved = False
        for i in cand:
            r = pos + int(i)
            s = counts[r]
            item = ((int(s[0]), int(s[1]), int(s[2]), int(s[3]), int(s[0]) - int(c0[r]), int(non_n[r])), b"", r)
            old = h.max_inc
            if h.insert(item):
                admitted.append(r)
                if h.n == h.K:
                    root = h.a[1][0]
                    h.max_inc = root[3] - root[0] + 1
                if h.max_inc != old:            # the candidate list was computed with the old limit: rescan from r+1
                    pos = r + 1
                    moved = True
                    break
        if not moved:
            pos += 200000
    return h.finalise(), admitted


@pytest.fixture(scope="module")
def big():
    clades = np.stack(SynthModel(nsites=L, n_clades=200, seed=1).clades)
    eng = Engine(L, nbest=NBEST, db_capacity=R)
    d = eng.device_alloc(NQ * L)
    eng.synth_rows_device(clades, 2002, 0, NQ, d)
    qtext = eng.d2h(d, NQ * L).reshape(NQ, L)
    eng.device_free(d)
    keep = host.filter_queries([r.tobytes() for r in qtext], 0.5)
    qtext = qtext[keep]
    qtext = np.ascontiguousarray(qtext[host.query_order(qtext, 0, False)])
    # plant exact copies of five queries as references (known answers at known ordinals)
    # (queries without IUPAC codes only: a copy of an ambiguous query has D > 0 and the reference may refuse it)
    clean = [iq for iq in range(len(qtext)) if not (orc.POPC4[orc.CODE[qtext[iq]]] > 1).any()]
    assert len(clean) >= 5
    pick = [clean[0], clean[len(clean) // 4], clean[len(clean) // 2], clean[3 * len(clean) // 4], clean[-1]]
    planted = dict(zip(pick, [12_345, 250_000, 640_001, 999_999, 31]))
    eng.set_queries(qtext)
    tile = 16384
    d_tile = eng.device_alloc(tile * L)
    non_n = np.empty(R, np.int32)
    for s in range(0, R, tile):
        m = min(tile, R - s)
        eng.synth_rows_device(clades, 1001, s, m, d_tile)
        for iq, pos in planted.items():
            if s <= pos < s + m:
                eng._ck(eng._lib.uvg_memcpy_h2d(eng._h, d_tile + (pos - s) * L, qtext[iq].ctypes.data, L), "h2d")
        nn, acc, na = eng.push_refs_device(d_tile, m, 0, want_outputs=True)
        assert na == m
        non_n[s:s + m] = nn
    eng.device_free(d_tile)
    eng.search()
    yield eng, qtext, planted, non_n
    eng.close()


def test_fullsize_shapes_and_invariants(big):
    eng, qtext, planted, non_n = big
    K = NBEST
    st = eng.stats()
    assert st["n_refs"] == R and eng.db_count == R
    for iq in range(0, eng.n_query, 37):
        sc, refs = eng.heap(iq)
        assert len(refs) == K and len(set(refs.tolist())) == K
        tup = [tuple(x) for x in sc.tolist()]
        assert tup == sorted(tup, reverse=True)                      # rows are in descending lexicographic order
        assert (sc[:, 0] <= sc[:, 1]).all() and (sc[:, 1] <= sc[:, 2]).all() and (sc[:, 2] <= sc[:, 3]).all()
        assert np.array_equal(sc[:, 5], non_n[refs])                 # valid_ref_sites column
        assert (sc[:, 3] <= sc[:, 5]).all()
    eb = eng.ever_best(0, R)
    assert int(eb.sum()) <= st["n_admitted"]
    for iq in range(0, eng.n_query, 37):
        assert eb[eng.heap(iq)[1]].all()


def test_fullsize_planted_copies(big):
    eng, qtext, planted, non_n = big
    for iq, pos in planted.items():
        sc, refs = eng.heap(iq)
        code = orc.CODE[qtext[iq]]
        n_valid, n_one = int((code != 0).sum()), int((orc.POPC4[code] == 1).sum())
        assert pos in refs.tolist(), (iq, pos)
        j = refs.tolist().index(pos)
        assert (sc[j, 0], sc[j, 1], sc[j, 2], sc[j, 3]) == (n_one, n_valid, n_valid, n_valid)


def test_fullsize_idempotent_and_abort_invariant(big):
    eng, qtext, planted, non_n = big
    before = [eng.heap(iq) for iq in range(0, eng.n_query, 11)]
    eng.reset_search()
    eng.search()
    for k, iq in enumerate(range(0, eng.n_query, 11)):
        sc, refs = eng.heap(iq)
        assert np.array_equal(sc, before[k][0]) and np.array_equal(refs, before[k][1])


@pytest.mark.parametrize("iq", [0, 481, 961])
def test_fullsize_sampled_queries_exact(big, iq):
    """the table of a query re-derived on the host: raw counts of its 1M pairs (pair-count kernel, itself checked against
    the oracle in test_gpu_parity) -> the oracle's sequential admission and heap -> must equal what the device replay kept"""
    eng, qtext, planted, non_n = big
    iq = min(iq, eng.n_query - 1)
    counts = np.empty((R, 4), np.int32)
    c0 = np.empty(R, np.int32)
    step = 32768
    for s in range(0, R, step):
        m = min(step, R - s)
        counts[s:s + m] = eng.debug_pair_counts_rows(iq, 1, s, m)[0]
        c0[s:s + m] = eng.debug_ref_meta(s, m)[1]
    rows, admitted = replay_with_oracle_heap(counts, non_n, c0, NBEST, L)
    sc, refs = eng.heap(iq)
    assert np.array_equal(refs, np.array([r[2] for r in rows], np.int64))
    assert np.array_equal(sc, np.array([r[0] for r in rows], np.int32))
    eb = eng.ever_best(0, R)
    assert eb[np.array(admitted)].all()
