"""uvg_ball_classify_device_noalloc (csrc/ball.cu): the radius classification for callers that must not allocate -- rows and
results in the caller's device buffers, the scratch between the two kernels in a file-scope device array.  Through the C ABI,
against the existing entry point uvg_ball_push_refs on the same rows: bit-exact (keep codes and valid-site counts are
integers), on two calls in a row whose row counts differ (the array keeps the first call's contents).  Collected last on
purpose; passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

from helpers import orc
from uvaia_b200.engine import Ball, UvgError
from uvaia_b200.synth import SynthModel

pytestmark = pytest.mark.gpu


def _inputs(L, nq, nr, seed):
    m = SynthModel(nsites=L, n_clades=6, seed=seed, clade_subs=max(2, L // 300))
    qn, Q = m.sample(nq, seed + 10, prefix="Q")
    rn, R = m.sample(nr, seed + 20, ambiguous_frac=0.005, prefix="R")
    rng = np.random.default_rng(seed)
    for k in range(0, nr, 5):                                   # near copies of queries: rows on both sides of the radius
        row = Q[k % nq].copy()
        pos = rng.choice(L, size=k % 13, replace=False)
        row[pos] = np.frombuffer(b"ACGTN", np.uint8)[rng.integers(0, 5, size=len(pos))]
        R[k] = row
    for k in range(3, nr, 17):                                  # N-rich rows: the filter (keep code 0)
        R[k, : L * 3 // 4] = ord("N")
    names, Qo, *_ = orc.ball_prepare_queries(qn, [q.tobytes() for q in Q], 0, False, 0.5, False)
    return Qo, np.ascontiguousarray(R)


def test_no_alloc_entry_point_matches_push_refs_on_two_calls_in_a_row():
    import torch
    L, dist = 1500, 4
    Qo, R = _inputs(L, 30, 2700, 1)
    R1, R2 = np.ascontiguousarray(R[:2000]), np.ascontiguousarray(R[2000:])      # second call: fewer, different rows over the first call's scratch
    dev = torch.device("cuda", 0)
    with Ball(L, dist=dist) as ball:
        ball.set_queries(Qo)
        min_non_n = int(L * 0.5)
        # the caller's buffers, made before the calls that must not allocate
        d_text = [torch.from_numpy(R).to(dev) for R in (R1, R2)]
        d_keep = [torch.full((len(R),), 255, dtype=torch.uint8, device=dev) for R in (R1, R2)]
        d_nn = [torch.full((len(R),), -1, dtype=torch.int32, device=dev) for R in (R1, R2)]
        want = [ball.push_refs(R, min_non_n) for R in (R1, R2)]             # the existing entry point (host rows, allocates its staging)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        for k in range(2):                                                   # two calls in a row
            ball.classify_device_noalloc(d_text[k].data_ptr(), len(d_text[k]), min_non_n, d_keep[k].data_ptr(), d_nn[k].data_ptr())
        ball.classify_device_noalloc(d_text[0].data_ptr(), len(d_text[0]), min_non_n, d_keep[0].data_ptr())    # once more, valid-site counts into the scratch
        assert torch.cuda.mem_get_info(dev)[0] == free_before                # nothing was allocated on the device
        for k in range(2):
            keep, nn = d_keep[k].cpu().numpy(), d_nn[k].cpu().numpy()
            assert np.array_equal(keep, want[k][0]), f"call {k}: keep codes"
            assert np.array_equal(nn, want[k][1]), f"call {k}: valid sites"
            assert set(np.unique(keep)) == {0, 1, 2}                         # the case exercises the filter and both sides of the radius
        # beyond the static array: a clear error, nothing launched
        with pytest.raises(UvgError, match="exceed the 65536 rows"):
            ball.classify_device_noalloc(d_text[0].data_ptr(), Ball.NOALLOC_MAX_ROWS + 1, min_non_n, d_keep[0].data_ptr())
        # n = 0 is a no-op
        ball.classify_device_noalloc(0, 0, min_non_n, 0)
