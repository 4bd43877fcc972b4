"""uvaiaball (src/ball.c): the numpy restatement (oracle.ball_search) against the goldens written by the compiled reference
program oracle/_ref/uvaiaball_ref (tests/golden/make_golden.py, group `ball`).  CPU only."""
import json
import os

import pytest

from helpers import GOLDEN, orc

BALL_DIR = os.path.join(GOLDEN, "ball_cases")


def ball_case_names():
    return sorted(d for d in os.listdir(BALL_DIR) if os.path.isdir(os.path.join(BALL_DIR, d)))


class BallCase:
    def __init__(self, name):
        self.dir = os.path.join(BALL_DIR, name)
        self.meta = json.load(open(os.path.join(self.dir, "args.json")))
        self.args = self.meta["args"]
        self.query_file = os.path.join(self.dir, "q.fa.xz")
        self.ref_files = [os.path.join(self.dir, f"r{i}.fa.xz") for i in range(self.meta["n_ref_files"])]

    @property
    def expected(self):
        return [ln for ln in open(os.path.join(self.dir, "expected_kept_names.txt"), "rb").read().split(b"\n") if ln]

    def options(self):
        kw = dict(dist=5, trim=0, acgt=False, ambig_q=0.5, ambig_r=0.5, keep_more_resolved=False)
        a, i = self.args, 0
        while i < len(a):
            if a[i] == "-d": kw["dist"] = int(a[i + 1]); i += 2
            elif a[i] == "--trim": kw["trim"] = int(a[i + 1]); i += 2
            elif a[i] == "-a": kw["ambig_q"] = float(a[i + 1]); i += 2
            elif a[i] == "-A": kw["ambig_r"] = float(a[i + 1]); i += 2
            elif a[i] == "-x": kw["acgt"] = True; i += 1
            elif a[i] == "-k": kw["keep_more_resolved"] = True; i += 1
            else: raise ValueError(a[i])
        return kw


@pytest.mark.parametrize("name", ball_case_names())
def test_oracle_reproduces_uvaiaball(name):
    c = BallCase(name)
    kept, dist, n_invalid = orc.ball_run_files(c.query_file, c.ref_files, **c.options())
    assert kept == c.expected
