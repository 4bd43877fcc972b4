"""-k / --keep_resolved: uvg_prune_queries (host threads inside the C-ABI library, no GPU needed) against the oracle's
restatement of exclude_redundant_query_sequences (src/fastaseq.c:797-841) and against the query names the reference
itself kept in the `keep_*` goldens."""
from __future__ import annotations

import numpy as np
import pytest

from helpers import Case, orc
from uvaia_b200.engine import prune_queries
from uvaia_b200.host import iter_fasta


def _ordered_queries(case):
    kw = case.options()
    recs = list(iter_fasta(case.query_file))
    qn, qs = [r[0] for r in recs], [r[1] for r in recs]
    names, Q, trim, cons, idx, idx_m, idx_c = orc.prepare_queries(qn, qs, kw["trim"], kw["acgt"], kw["ambig_q"], keep_resolved=False)
    return names, Q, kw, idx, idx_m


@pytest.mark.parametrize("name", ["keep_default", "keep_acgt", "keep_trim"])
def test_prune_matches_reference_golden(name):
    case = Case(name)
    names, Q, kw, idx, idx_m = _ordered_queries(case)
    mask = prune_queries(Q, kw["trim"], kw["acgt"], True)
    kept = [nm.decode() for nm, k in zip(names, mask) if k]
    rows = case.expected_csv.split("\n")[1:]
    expected = []
    for ln in rows:
        if ln and (not expected or expected[-1] != ln.split(",")[0]):
            expected.append(ln.split(",")[0])
    assert kept == expected                     # the reference's CSV lists the surviving queries in their order
    assert 0 < len(kept) < len(names)


def _random_set(rng, n_base, n_derived, L):
    base = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=L)
    rows = []
    for _ in range(n_base):
        r = base.copy()
        pos = rng.choice(L, 6, replace=False)
        r[pos] = rng.choice(np.frombuffer(b"ACGT", np.uint8), size=6)
        rows.append(r)
    for _ in range(n_derived):
        r = rows[int(rng.integers(0, len(rows)))].copy()
        kind = int(rng.integers(0, 5))
        if kind >= 1:
            for _ in range(int(rng.integers(1, 4))):
                a = int(rng.integers(0, L - 8))
                r[a:a + int(rng.integers(1, 8))] = rng.choice(np.frombuffer(b"N-?XRYn", np.uint8))
        if kind == 4:
            p = int(rng.integers(0, L))
            r[p] = rng.choice(np.frombuffer(b"ACGT", np.uint8))
        rows.append(r)
    Q = np.stack(rows)
    return Q[rng.permutation(len(rows))]


@pytest.mark.parametrize("acgt", [False, True])
@pytest.mark.parametrize("keep_more", [True, False])
def test_prune_matches_oracle_on_random_sets(acgt, keep_more):
    rng = np.random.default_rng(7 + 2 * acgt + keep_more)
    for trial in range(12):
        L = int(rng.integers(40, 400))
        Q = _random_set(rng, int(rng.integers(2, 8)), int(rng.integers(3, 40)), L)
        trim = int(rng.integers(0, L // 3)) if trial % 3 == 0 else 0
        t = orc.clamp_trim(trim, L)
        order = orc.query_order(Q, t, acgt)
        Q = np.ascontiguousarray(Q[order])
        cons, idx, idx_m, idx_c = orc.partition(Q, t, acgt)
        want = orc.exclude_redundant(Q, idx, idx_m, acgt, keep_more)
        got = np.where(prune_queries(Q, trim, acgt, keep_more))[0].tolist()
        assert got == want, (trial, L, trim)


def test_prune_single_query_and_all_identical():
    Q = np.frombuffer(b"ACGTNACGTA", np.uint8).reshape(1, -1)
    assert prune_queries(Q).tolist() == [True]
    same = np.repeat(Q, 300, axis=0)                       # more rows than one classification block
    assert prune_queries(same).tolist() == [True] + [False] * 299
    assert prune_queries(same, keep_more_resolved=False).tolist() == [True] + [False] * 299
