"""Shared helpers of the test-suite: golden-case loading and the oracle import (tests may use oracle/, the product may not)."""
from __future__ import annotations

import json
import lzma
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
CASES_DIR = os.path.join(GOLDEN, "cases")

from oracle import uvaia_oracle as orc  # noqa: E402


def case_names():
    return sorted(d for d in os.listdir(CASES_DIR) if os.path.isdir(os.path.join(CASES_DIR, d)))


class Case:
    """one end-to-end golden: inputs, the reference's command line options, the reference's CSV"""

    def __init__(self, name):
        self.name = name
        self.dir = os.path.join(CASES_DIR, name)
        meta = json.load(open(os.path.join(self.dir, "args.json")))
        self.meta = meta
        self.args = meta["args"]
        self.n_ref_files = meta["n_ref_files"]
        self.query_file = os.path.join(self.dir, "q.fa.xz")
        self.ref_files = [os.path.join(self.dir, f"r{i}.fa.xz") for i in range(self.n_ref_files)]
        self.h1 = bool(meta.get("h1_pool_dependent"))

    @property
    def expected_csv(self) -> str:
        return lzma.open(os.path.join(self.dir, "expected.csv.xz"), "rb").read().decode()

    @property
    def expected_aln_names(self):
        data = open(os.path.join(self.dir, "expected_aln_names.txt"), "rb").read()
        return [ln for ln in data.split(b"\n") if ln]

    def options(self):
        """the reference's CLI options as keyword arguments of oracle.search / host.search_files"""
        kw = dict(nbest=100, trim=0, acgt=False, ambig_q=0.5, ambig_r=0.5, pool=512, exclude_self=False, keep_resolved=False)
        a = self.args
        i = 0
        while i < len(a):
            if a[i] == "-n": kw["nbest"] = int(a[i + 1]); i += 2
            elif a[i] == "--trim": kw["trim"] = int(a[i + 1]); i += 2
            elif a[i] == "-a": kw["ambig_q"] = float(a[i + 1]); i += 2
            elif a[i] == "-A": kw["ambig_r"] = float(a[i + 1]); i += 2
            elif a[i] == "-p": kw["pool"] = int(a[i + 1]); i += 2
            elif a[i] == "--acgt": kw["acgt"] = True; i += 1
            elif a[i] == "-x": kw["exclude_self"] = True; i += 1
            elif a[i] == "-k": kw["keep_resolved"] = True; i += 1
            else: raise ValueError(a[i])
        return kw
