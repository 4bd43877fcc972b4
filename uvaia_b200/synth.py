"""Seeded synthetic SARS-CoV-2-shaped alignments (host side, numpy) for parity tests and small benches.

Shape statistics follow the bundled alignment data/03.unique_acgt.aln.xz as measured in SURVEY.md 8(d):
a root genome, a shallow clade tree, a handful of private substitutions per genome, N-masking at both
ends plus internal N runs, rare gaps and rare IUPAC two-base codes.  The full-size (1M x 29,903) database
of the benchmark is produced by the device generator in csrc/ (same model, counter-based RNG); this module
is the small-scale, file-writing twin used where the reference binary has to read the same input.
"""
from __future__ import annotations

import numpy as np

BASES = np.frombuffer(b"TACG", np.uint8)
BASE_P = np.array([0.32, 0.30, 0.20, 0.18])          # T A G C order mapped below
IUPAC2 = np.frombuffer(b"MRWSYK", np.uint8)


def _mutate(rng, g, pos):
    """substitute g[pos] by a different base"""
    cur = g[pos]
    alt = rng.integers(0, 3, size=len(pos))
    table = {ord("A"): b"CGT", ord("C"): b"AGT", ord("G"): b"ACT", ord("T"): b"ACG"}
    out = np.array([table[int(c)][int(a)] for c, a in zip(cur, alt)], np.uint8) if len(pos) else np.zeros(0, np.uint8)
    g[pos] = out


class SynthModel:
    """Holds the root/clade genomes so queries and database rows come from the same population."""

    def __init__(self, nsites=29903, n_clades=200, seed=1, hot_frac=0.27, clade_subs=30, n_frac_mean=0.197,
                 clean=False):
        rng = np.random.default_rng(seed)
        self.L = L = int(nsites)
        self.root = rng.choice(np.frombuffer(b"TAGC", np.uint8), size=L, p=BASE_P)
        self.hot = np.sort(rng.choice(L, size=max(1, int(L * hot_frac)), replace=False))
        self.clades = []
        for _ in range(n_clades):
            g = self.root.copy()
            k = max(0, int(rng.normal(clade_subs, clade_subs / 2)))
            k = min(k, len(self.hot))
            pos = rng.choice(self.hot, size=k, replace=False)
            _mutate(rng, g, pos)
            self.clades.append(g)
        self.n_frac_mean = 0.01 if clean else n_frac_mean
        self.clean = clean

    def sample(self, n, seed, private_subs=8, ambiguous_frac=0.0, prefix="England/SYN"):
        """-> (names list[bytes], text (n, L) uint8)"""
        rng = np.random.default_rng(seed)
        L = self.L
        out = np.empty((n, L), np.uint8)
        names = []
        scale = L / 29903.0
        for i in range(n):
            g = self.clades[int(rng.integers(0, len(self.clades)))].copy()
            k = int(rng.poisson(private_subs))
            if k:
                _mutate(rng, g, rng.integers(0, L, size=k))
            # N masking: both ends + internal runs; per-genome N fraction ~ Beta with the requested mean
            lead = int(rng.integers(30, 151) * scale) if L > 400 else int(rng.integers(0, 3))
            trail = int(rng.integers(30, 121) * scale) if L > 400 else int(rng.integers(0, 3))
            g[:lead] = ord("N")
            if trail:
                g[L - trail:] = ord("N")
            m = self.n_frac_mean
            f = float(rng.beta(1.2, 1.2 * (1 - m) / m))
            nruns = int(rng.poisson(20 if not self.clean else 2))
            if nruns:
                mean_len = max(1.0, L * f / nruns)
                starts = rng.integers(0, L, size=nruns)
                lens = rng.geometric(1.0 / mean_len, size=nruns)
                for s, ln in zip(starts, lens):
                    g[s:s + int(ln)] = ord("N")
            # gaps 0.08 % as runs of 1-9
            ngap = int(rng.poisson(max(0.2, 0.0008 * L / 5)))
            for s in rng.integers(0, L, size=ngap):
                g[s:s + int(rng.integers(1, 10))] = ord("-")
            # IUPAC two-base codes
            nia = int(rng.poisson(1.7))
            if nia:
                g[rng.integers(0, L, size=nia)] = rng.choice(IUPAC2, size=nia)
            if ambiguous_frac > 0:          # config-4 style stress: 3/4 N, 1/4 IUPAC (incl. 3-base codes)
                pos = np.nonzero(rng.random(L) < ambiguous_frac)[0]
                kind = rng.random(len(pos))
                g[pos[kind < 0.75]] = ord("N")
                amb = pos[kind >= 0.75]
                g[amb] = rng.choice(np.frombuffer(b"MRWSYKVHDB", np.uint8), size=len(amb))
            out[i] = g
            names.append(b"%s-%07X/2022" % (prefix.encode(), int(rng.integers(0, 1 << 28))))
        return names, out


def write_fasta(path, names, text, width=0):
    with open(path, "wb") as fh:
        for nm, row in zip(names, text):
            fh.write(b">" + nm + b"\n")
            s = row.tobytes() if isinstance(row, np.ndarray) else bytes(row)
            if width and width > 0:
                for k in range(0, len(s), width):
                    fh.write(s[k:k + width] + b"\n")
            else:
                fh.write(s + b"\n")
