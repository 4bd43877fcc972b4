"""uvaiaball on the GPU (uvg_ball_*, csrc/ball.cu) through the C ABI: the references kept must be exactly the ones the
compiled reference program keeps (goldens of tests/golden/ball_cases) and the ones the numpy oracle keeps on random inputs."""
import numpy as np
import pytest

from helpers import orc
from test_oracle_ball import BallCase, ball_case_names
from uvaia_b200 import host
from uvaia_b200.engine import Ball
from uvaia_b200.synth import SynthModel

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ball_case_names())
def test_ball_golden_kept_set(name):
    c = BallCase(name)
    kept, n_invalid = host.ball_files(c.query_file, c.ref_files, batch_rows=777, **c.options())
    assert kept == c.expected


@pytest.mark.parametrize("L,nq,nr,dist,acgt,trim,seed", [(1500, 30, 2000, 4, False, 0, 1), (1500, 30, 2000, 4, True, 0, 2), (999, 3, 700, 9, False, 50, 3),
                                                         (4000, 120, 1500, 2, False, 0, 4), (64, 5, 300, 3, False, 0, 5), (29903, 40, 600, 6, False, 230, 6)])
def test_ball_vs_oracle(L, nq, nr, dist, acgt, trim, seed):
    m = SynthModel(nsites=L, n_clades=6, seed=seed, clade_subs=max(2, L // 300))
    qn, Q = m.sample(nq, seed + 10, prefix="Q")
    rn, R = m.sample(nr, seed + 20, ambiguous_frac=0.005, prefix="R")
    rng = np.random.default_rng(seed)
    for k in range(0, nr, 7):                                   # near copies of queries, so that the radius cuts through them
        row = Q[k % nq].copy()
        pos = rng.choice(L, size=k % 11, replace=False)
        row[pos] = np.frombuffer(b"ACGTN", np.uint8)[rng.integers(0, 5, size=len(pos))]
        R[k] = row
    qs, rs = [q.tobytes() for q in Q], [r.tobytes() for r in R]
    want, d, _ = orc.ball_search(qn, qs, [(rn, rs)], dist=dist, trim=trim, acgt=acgt)
    names, Qo, tr, cons, idx, idx_m, idx_c = orc.ball_prepare_queries(qn, qs, trim, acgt, 0.5, False)
    with Ball(L, dist=dist, trim=trim, acgt=acgt) as ball:
        assert ball.dist == d
        info = ball.set_queries(Qo)
        assert (info.n_idx, info.n_idx_m, info.n_idx_c) == (len(idx), len(idx_m), len(idx_c))
        keep, nn = ball.push_refs(R, int(L * 0.5))
        assert np.array_equal(nn, (~orc.INDEL[R]).sum(1))
        got = [nm for nm, k in zip(rn, keep) if k == 2]
    assert got == want
