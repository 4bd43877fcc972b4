"""Host-side mirror of the `uvaia` program flow (src/nearest.c main, 159-353) on top of the C ABI.

Everything that scores or keeps neighbours runs on the GPU through uvaia_b200.engine.Engine; this module only does
what the reference's host code does around the seam: FASTA hygiene, the query quality filter and ordering, the
reference stream with `-x`, and the CSV table.  The native C host program (src/ of this repo) follows the same steps.
"""
from __future__ import annotations

import bz2
import gzip
import lzma

import numpy as np

from .engine import Engine, prune_queries

HEADER_FULL = ("query,reference,rank,ACGT_matches,text_matches,partial_matches,valid_pair_comparisons,"
               "ACGT_matches_unique,valid_ref_sites")          # src/nearest.c:525
HEADER_ACGT = ("query,reference,rank,ACGT_matches,valid_ACGT_comparisons,ACGT_matches_unique,valid_ref_sites,"
               "dist_consensus,dist_unique")                   # src/nearest.c:523

_SPACE = b" \t\n\v\f\r"
_UPPER = bytes(range(256)).upper()

# byte classes used by the host-side query filter (biomcmc alignment.c:886-904 via char2bit degeneracy)
_DEGEN = np.zeros(256, np.uint8)
for _ch, _d in dict(A=1, C=1, G=1, T=1, U=1, M=2, R=2, W=2, S=2, Y=2, K=2, V=3, H=3, D=3, B=3).items():
    _DEGEN[ord(_ch)] = _d
_VALIDTXT = np.ones(256, bool)
for _ch in "NnXx-?Oo.":
    _VALIDTXT[ord(_ch)] = False
_ISACGT = np.zeros(256, bool)
for _ch in "ACGTacgt":
    _ISACGT[ord(_ch)] = True


def open_any(path):
    """codec sniffing by content like biomcmc_open_compress (file_compression.c:21-47)"""
    with open(path, "rb") as fh:
        magic = fh.read(6)
    if magic.startswith(b"\xfd7zXZ\x00"):
        return lzma.open(path, "rb")
    if magic.startswith(b"BZh"):
        return bz2.open(path, "rb")
    if magic.startswith(b"\x1f\x8b"):
        return gzip.open(path, "rb")
    return open(path, "rb")


def iter_fasta(path):
    """yields (name, sequence) with the reference's hygiene rules (H5): lines shorter than 2 bytes and lines starting
    with # or ; are dropped (nexus_common.c:133-145), any line containing '>' starts a record named by everything after
    the first '>' (fastaseq.c:432-442), '\\r' is a line break (file_compression.c:216), whitespace and '>' are removed
    from sequence lines which are upper-cased (fastaseq.c:451-452)."""
    name, chunks = None, []
    with open_any(path) as fh:
        data = fh.read()
    for line in data.replace(b"\r", b"\n").split(b"\n"):
        if len(line) < 2:
            continue
        s = line.lstrip(_SPACE)
        if not s or s[:1] in (b"#", b";"):
            continue
        k = line.find(b">")
        if k >= 0:
            if name is not None:
                yield name, b"".join(chunks)
            name, chunks = line[k + 1:], []
        elif name is not None:
            chunks.append(line.translate(_UPPER, _SPACE + b">"))
    if name is not None:
        yield name, b"".join(chunks)


def filter_queries(seqs, ambiguity):
    """uvaia_keep_only_valid_sequences (src/utils.c:10-48): indices of the queries that stay."""
    keep = []
    for i, s in enumerate(seqs):
        n = len(s)
        if n < 5:
            continue
        d = _DEGEN[np.frombuffer(s, np.uint8)]
        if float((d == 0).sum()) / n > ambiguity:
            continue
        if float((d == 1).sum()) / n < 1.0 - 1.1 * ambiguity:
            continue
        keep.append(i)
    if len({len(seqs[i]) for i in keep}) > 1:
        raise ValueError("query sequences are not aligned")          # utils.c:38-42
    return keep


def clamp_trim(trim, nchar):
    """new_query_structure_from_fasta (src/fastaseq.c:708-712)"""
    trim = max(0, int(trim))
    if trim > nchar / 2.1:
        trim = int(nchar / 2.1)
    return trim


def query_order(Q, trim, acgt):
    """reorder_query_structure (src/fastaseq.c:779-795): ascending usable sites in the trimmed range, ties by index."""
    L = Q.shape[1]
    sub = Q[:, trim:L - trim]
    cnt = (_ISACGT[sub] if acgt else _VALIDTXT[sub]).sum(1)
    return sorted(range(Q.shape[0]), key=lambda i: (int(cnt[i]), i))


class SearchOutput:
    def __init__(self):
        self.qnames = []
        self.rows = []          # per query: (scores[n,6], ref ordinals[n])
        self.ref_names = []     # accepted references by ordinal
        self.ever_best = None
        self.info = None
        self.stats = None
        self.n_invalid = 0

    def csv(self, acgt=False):
        out = [HEADER_ACGT if acgt else HEADER_FULL]
        for qn, (sc, refs) in zip(self.qnames, self.rows):
            for j in range(len(refs)):
                out.append("%s,%s,%d," % (qn.decode(), self.ref_names[int(refs[j])].decode(), j + 1)
                           + ",".join(str(int(x)) for x in sc[j]))
        return "\n".join(out) + "\n"


def ball_files(query_file, ref_files, dist=5, trim=0, acgt=False, ambig_q=0.5, ambig_r=0.5, keep_more_resolved=False, batch_rows=4096, device=0):
    """End-to-end `uvaiaball` run (ball.c:138-291) with the distances on the GPU.  -> (names of the references kept, in file
    order; number too ambiguous)"""
    from .engine import Ball
    ambig_q = min(1.0, max(0.001, ambig_q))                     # ball.c:154-157
    ambig_r = min(1.0, max(0.001, ambig_r))
    recs = list(iter_fasta(query_file))
    qn, qs = [r[0] for r in recs], [r[1] for r in recs]
    keep = filter_queries(qs, ambig_q)
    if not keep:
        raise ValueError("no valid query sequences")
    Q = np.frombuffer(b"".join(qs[i] for i in keep), np.uint8).reshape(len(keep), -1)
    L = Q.shape[1]
    t = clamp_trim(trim, L)
    Q = np.ascontiguousarray(Q[query_order(Q, t, acgt)])
    Q = np.ascontiguousarray(Q[prune_queries(Q, trim, acgt, keep_more_resolved)])       # ball.c:190: always, -k picks the side
    min_non_n = int(L * ambig_r)                                # ball.c:201 (A, not 1 - A)
    kept, n_invalid = [], 0
    with Ball(L, dist=dist, trim=trim, acgt=acgt, device=device) as ball:
        ball.set_queries(Q)
        for rf in ref_files:
            names, rows = [], []

            def flush():
                nonlocal n_invalid
                if not rows:
                    return
                text = np.frombuffer(b"".join(rows), np.uint8).reshape(len(rows), L)
                k, _ = ball.push_refs(text, min_non_n)
                kept.extend(nm for nm, x in zip(names, k) if x == 2)
                n_invalid += int((k == 0).sum())
                names.clear(); rows.clear()

            for nm, sq in iter_fasta(rf):
                if len(sq) != L:
                    # ball.c:216-231: a poor-quality reference of the wrong length is skipped, a good one is fatal
                    if int(_VALIDTXT[np.frombuffer(sq, np.uint8)].sum()) < min_non_n:
                        n_invalid += 1
                        continue
                    raise ValueError("all sequences must be aligned")
                names.append(nm); rows.append(sq)
                if len(rows) >= batch_rows:
                    flush()
            flush()
    return kept, n_invalid


def search_files(query_file, ref_files, nbest=100, trim=0, acgt=False, ambig_q=0.5, ambig_r=0.5, pool=512, exclude_self=False,
                 keep_resolved=False, batch_rows=4096, device=0, **engine_kw):
    """End-to-end `uvaia` run (nearest.c:159-353) with scoring and heaps on the GPU.  Returns a SearchOutput."""
    ambig_q = min(1.0, max(0.001, ambig_q))                     # nearest.c:175-178
    ambig_r = min(1.0, max(0.001, ambig_r))
    nbest = max(1, nbest)
    recs = list(iter_fasta(query_file))
    qn, qs = [r[0] for r in recs], [r[1] for r in recs]
    keep = filter_queries(qs, ambig_q)
    if not keep:
        raise ValueError("no valid query sequences")
    Q = np.frombuffer(b"".join(qs[i] for i in keep), np.uint8).reshape(len(keep), -1)
    names = [qn[i] for i in keep]
    L = Q.shape[1]
    t = clamp_trim(trim, L)
    order = query_order(Q, t, acgt)
    Q = np.ascontiguousarray(Q[order])
    names = [names[i] for i in order]
    out = SearchOutput()
    min_non_n = int(L * (1.0 - ambig_r))                        # nearest.c:237
    with Engine(L, nbest=nbest, trim=trim, acgt=acgt, pool=pool, device=device, **engine_kw) as eng:
        if keep_resolved:                                        # nearest.c:219-224
            mask = prune_queries(Q, trim, acgt, True)
            Q = np.ascontiguousarray(Q[mask])
            names = [nm for nm, k in zip(names, mask) if k]
        out.qnames = names
        qset = set(names) if exclude_self else None             # nearest.c:226-230 (after the pruning)
        out.info = eng.set_queries(Q)
        for rf in ref_files:
            buf_names, buf_rows = [], []

            def flush():
                if not buf_rows:
                    return
                text = np.frombuffer(b"".join(buf_rows), np.uint8).reshape(len(buf_rows), L)
                nn, acc, na = eng.push_refs(text, min_non_n)
                out.ref_names.extend(nm for nm, a in zip(buf_names, acc) if a)
                out.n_invalid += int(len(buf_rows) - na)
                eng.search()
                eng.drop_refs()
                buf_names.clear()
                buf_rows.clear()

            for nm, sq in iter_fasta(rf):
                if qset is not None and nm in qset:                 # nearest.c:255-261
                    continue
                if len(sq) != L:
                    # nearest.c:263-278: a poor-quality reference of the wrong length is skipped, a good one is fatal (H7)
                    nn = int(_VALIDTXT[np.frombuffer(sq, np.uint8)].sum())
                    if nn < min_non_n:
                        out.n_invalid += 1
                        continue
                    raise ValueError("all sequences must be aligned: reference %r has %d sites, queries have %d"
                                     % (nm.decode(errors="replace"), len(sq), L))
                buf_names.append(nm)
                buf_rows.append(sq)
                if len(buf_rows) >= batch_rows:
                    flush()
            flush()
            eng.mark_file_end()                                  # the reference's batch ends with the file (nearest.c:245-286)
        out.rows = eng.heaps()
        out.ever_best = eng.ever_best(0, len(out.ref_names))
        out.stats = eng.stats()
    return out
