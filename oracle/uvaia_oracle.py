"""CPU oracle for the uvaia n-best search path (reference: src/nearest.c -> binary `uvaia`).

TEST INFRASTRUCTURE ONLY.  Nothing under uvaia_b200/ may import this module; only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline leg may.  It is a numpy restatement of the
reference algorithm, written from the reference's behaviour (citations are file:line under
/root/reference) and PINNED against the compiled reference binary (oracle/_ref/uvaia_ref, built by
oracle/build_ref.sh) through the golden fixtures in tests/golden/ (see tests/golden/make_golden.py).

Two execution modes of the search:
  * closed form  (default): per-pair totals over the whole trimmed range, then the sequential
    admission.  Equivalent to the reference whenever hazard H1 does not fire.
  * literal=True: emulates the staged, truncated computation batch by batch exactly as nearest.c
    does it (consensus stage with the batch-wide limit, idx_m stage, idx stage), including H1.
"""
from __future__ import annotations

import functools
import gzip
import lzma
import bz2
from dataclasses import dataclass, field

import numpy as np

# ---------------------------------------------------------------------------------------------
# byte tables
# ---------------------------------------------------------------------------------------------
# char2bit[][0] restricted to degeneracy 1..3 (biomcmc alignment.c:906-931 + the skip rule at :870):
# N, O, X, ?, - and every unknown byte are INVALID sites -> code 0.
CODE = np.zeros(256, np.uint8)
for _ch, _v in dict(A=1, B=14, C=2, D=13, G=4, H=11, K=12, M=3, R=5, S=6, T=8, U=8, V=7, W=9, Y=10).items():
    CODE[ord(_ch)] = _v
# is_indel[] (src/utils.c:263): bytes that are NOT "valid text"
INDEL = np.zeros(256, bool)
for _ch in "NnXx-?Oo.":
    INDEL[ord(_ch)] = True
# is_acgt[] (src/utils.c:262)
ACGT = np.zeros(256, bool)
for _ch in "ACGTacgt":
    ACGT[ord(_ch)] = True
POPC4 = np.array([bin(i).count("1") for i in range(16)], np.uint8)

HEADER_FULL = ("query,reference,rank,ACGT_matches,text_matches,partial_matches,valid_pair_comparisons,"
               "ACGT_matches_unique,valid_ref_sites")          # src/nearest.c:525
HEADER_ACGT = ("query,reference,rank,ACGT_matches,valid_ACGT_comparisons,ACGT_matches_unique,valid_ref_sites,"
               "dist_consensus,dist_unique")                   # src/nearest.c:523


# ---------------------------------------------------------------------------------------------
# FASTA reading (hygiene rules H5: nexus_common.c:133-145, fastaseq.c:422-474, alignment.c:80-115)
# ---------------------------------------------------------------------------------------------
def _open_any(path):
    with open(path, "rb") as fh:
        magic = fh.read(6)
    if magic.startswith(b"\xfd7zXZ\x00"):
        return lzma.open(path, "rb")
    if magic.startswith(b"BZh"):
        return bz2.open(path, "rb")
    if magic.startswith(b"\x1f\x8b"):
        return gzip.open(path, "rb")
    return open(path, "rb")


_SPACE = b" \t\n\v\f\r"


def _nonempty_fasta_line(line: bytes) -> bool:
    if len(line) < 2:
        return False
    s = line.lstrip(_SPACE)
    if not s:
        return False
    return s[:1] not in (b"#", b";")


def read_fasta(path):
    """-> (names: list[bytes], seqs: list[bytes]) with sequence text space-stripped and upper-cased."""
    names, chunks = [], []
    with _open_any(path) as fh:
        data = fh.read()
    data = data.replace(b"\r", b"\n")                      # file_compression.c:216
    for line in data.split(b"\n"):
        if not _nonempty_fasta_line(line):
            continue
        k = line.find(b">")
        if k >= 0:
            names.append(line[k + 1:])
            chunks.append([])
        else:
            if not chunks:
                continue
            s = line.translate(None, _SPACE + b">").upper()
            chunks[-1].append(s)
    return names, [b"".join(c) for c in chunks]


# ---------------------------------------------------------------------------------------------
# query preparation
# ---------------------------------------------------------------------------------------------
def filter_queries(names, seqs, ambiguity):
    """uvaia_keep_only_valid_sequences (src/utils.c:10-48). Returns indices kept."""
    keep = []
    for i, s in enumerate(seqs):
        n = len(s)
        if n < 5:
            continue
        c = CODE[np.frombuffer(s, np.uint8)]
        d = POPC4[c]
        n_acgt = int((d == 1).sum())
        n_bad = int((c == 0).sum())
        if n_bad / float(n) > ambiguity:
            continue
        if n_acgt / float(n) < 1.0 - 1.1 * ambiguity:
            continue
        keep.append(i)
    lens = {len(seqs[i]) for i in keep}
    if len(lens) > 1:
        raise ValueError("query sequences are not aligned")
    return keep


def clamp_trim(trim, nchar):
    """new_query_structure_from_fasta (src/fastaseq.c:708-712)."""
    if trim < 0:
        trim = 0
    if trim > nchar / 2.1:
        trim = int(nchar / 2.1)
    return trim


def partition(Q, trim, acgt):
    """create_query_indices (src/fastaseq.c:732-777). Q: (n, L) uint8 text.
    -> consensus bytes (L,), idx, idx_m, idx_c (int arrays)."""
    n, L = Q.shape
    usable = ACGT[Q] if acgt else ~INDEL[Q]
    inrange = np.zeros(L, bool)
    inrange[trim:L - trim] = True
    anyu = usable.any(0)
    first = np.where(usable, Q, 0).max(0)                      # some usable byte of the column
    same = np.where(usable, Q == first, True).all(0)
    allu = usable.all(0)
    cons = np.full(L, ord("N"), np.uint8)
    sel = inrange & anyu
    cons[sel & same] = first[sel & same]
    cons[sel & ~same] = ord("#")
    idx = np.where(sel & ~same)[0]
    idx_c = np.where(sel & same & allu)[0]
    idx_m = np.where(sel & same & ~allu)[0]
    return cons, idx, idx_m, idx_c


def query_order(Q, trim, acgt):
    """reorder_query_structure (src/fastaseq.c:779-795): ascending count of usable sites in the trimmed
    range, ties by original index (biomcmc empirical_frequency.c:38-45)."""
    L = Q.shape[1]
    sub = Q[:, trim:L - trim]
    cnt = (ACGT[sub] if acgt else ~INDEL[sub]).sum(1)
    return sorted(range(Q.shape[0]), key=lambda i: (int(cnt[i]), i))


def _resolved(v1, v2):
    """left_is_resolved_right[_acgt] (src/fastaseq.c:598-641) on the usable-site flags of two sequences over one index
    list: -1 if only the left sequence has usable sites the right lacks, +1 for the opposite, 0 if none, 0xff if both."""
    left, right = bool((v1 & ~v2).any()), bool((~v1 & v2).any())
    if left and right:
        return 0xff
    return -1 if left else (1 if right else 0)


def exclude_redundant(Q, idx, idx_m, acgt, keep_more_resolved):
    """exclude_redundant_query_sequences (src/fastaseq.c:797-841) on the ORDERED query text Q (n, L) with the index
    lists of create_query_indices.  Returns the indices of the queries that stay.  Pairs (i < j) are visited in
    lexicographic order and only while both are still kept; a pair is a candidate when no site of idx holds two usable,
    different bytes (quick_pairwise_score_truncated_idx_indelcheck / quick_pairwise_score_acgt with maxdist 1,
    fastaseq.c:562-583)."""
    n = Q.shape[0]
    use = ACGT if acgt else ~INDEL
    A, Am = Q[:, idx], Q[:, idx_m]
    U, Um = use[A], use[Am]
    valid = np.ones(n, bool)
    for i in range(n - 1):
        if not valid[i]:
            continue
        # candidates of row i in one sweep (the classification of a pair does not depend on `valid`)
        differ = (U[i][None, :] & U[i + 1:] & (A[i][None, :] != A[i + 1:])).any(1)
        for j in (np.where(~differ)[0] + i + 1):
            if not (valid[i] and valid[j]):
                continue
            red1 = _resolved(U[i], U[j])
            if red1 > 1:
                continue
            red2 = _resolved(Um[i], Um[j])
            if red2 > 1:
                continue
            if red1 == 0 and red2 == 0:
                valid[j] = False
            red1 += red2
            if red1 == 0:
                continue
            if keep_more_resolved:
                if red1 > 0: valid[i] = False
                else: valid[j] = False
            else:
                if red1 > 0: valid[j] = False
                else: valid[i] = False
    return [i for i in range(n) if valid[i]]


# ---------------------------------------------------------------------------------------------
# heap (src/min_heap.c)
# ---------------------------------------------------------------------------------------------
def _cmp(a, b):
    """compare_q_item_score (min_heap.c:41-47): first nonzero b[i]-a[i]."""
    for x, y in zip(a[0], b[0]):
        if x != y:
            return y - x
    return 0


class Heap:
    def __init__(self, nbest, max_inc):
        self.K = max(2, nbest)              # min_heap.c:58
        self.a = [None] * (self.K + 1)
        self.n = 0
        self.max_inc = max_inc              # nearest.c:375,387

    def _down(self, p):
        while True:
            c, m = 2 * p, p
            for i in (0, 1):
                if c + i <= self.n and _cmp(self.a[m], self.a[c + i]) < 0:
                    m = c + i
            if m == p:
                return
            self.a[p], self.a[m] = self.a[m], self.a[p]
            p = m

    def _up(self, i):
        while i > 1:
            p = i // 2
            if _cmp(self.a[p], self.a[i]) < 0:
                self.a[p], self.a[i] = self.a[i], self.a[p]
                i = p
            else:
                return

    def insert(self, item):
        """heap_insert (min_heap.c:93-117)."""
        if self.n == self.K:
            if _cmp(item, self.a[1]) < 0:
                self.a[1] = item
                self._down(1)
                return True
            return False
        self.n += 1
        self.a[self.n] = item
        self._up(self.n)
        return True

    def finalise(self):
        """heap_finalise_heap_qsort (min_heap.c:149-158): swap slot0<->slot n, stable sort (glibc qsort is a
        mergesort for these sizes).  Returns the n rows (H2: we never emit the NULL-name rows)."""
        if self.n == 0:
            return []
        arr = [self.a[self.n]] + self.a[1:self.n]
        arr.sort(key=functools.cmp_to_key(_cmp))
        return arr


# ---------------------------------------------------------------------------------------------
# pair scoring
# ---------------------------------------------------------------------------------------------
def pair_counts(qc, rc):
    """Closed form of biomcmc_pairwise_score_matches_truncated_idx without truncation
    (biomcmc alignment.c:855-884) for code arrays qc (Q,L) and rc (R,L): four (Q,R) int32 matrices
    [acgt, text, partial, valid], computed with exact small-integer float32 matrix products."""
    Qn, L = qc.shape
    assert L < (1 << 24)
    f = np.float32
    qv, rv = (qc != 0).astype(f), (rc != 0).astype(f)
    valid = qv @ rv.T
    text = np.zeros_like(valid)
    acgt = np.zeros_like(valid)
    for c in range(1, 15):
        m = (qc == c).astype(f) @ (rc == c).astype(f).T
        text += m
        if c in (1, 2, 4, 8):
            acgt += m
    # partial = #sites with (q & r) != 0, by inclusion-exclusion over the non-empty base subsets T
    partial = np.zeros_like(valid)
    for T in range(1, 16):
        sign = 1.0 if POPC4[T] % 2 == 1 else -1.0
        partial += sign * (((qc & T) == T).astype(f) @ ((rc & T) == T).astype(f).T)
    return [x.astype(np.int32) for x in (acgt, text, partial, valid)]


def truncated_idx(s1c, s2c, idx, max_incompatible):
    """Literal biomcmc_pairwise_score_matches_truncated_idx (alignment.c:855-884) on CODE arrays."""
    if len(idx) == 0:
        return [0, 0, 0, 0]
    if max_incompatible < 1:
        max_incompatible = 1
    a, b = s1c[idx], s2c[idx]
    v = (a != 0) & (b != 0)
    ex = v & (a == b)
    ac = ex & (POPC4[a] == 1)
    pa = v & ((a & b) != 0)
    mism = np.cumsum(v.astype(np.int64) - ac.astype(np.int64))
    hit = np.nonzero(mism >= max_incompatible)[0]
    end = len(idx) if len(hit) == 0 else int(hit[0]) + 1      # the site that reaches the limit is counted
    return [int(ac[:end].sum()), int(ex[:end].sum()), int(pa[:end].sum()), int(v[:end].sum())]


def truncated_acgt(s1, s2, idx, maxdist):
    """Literal quick_pairwise_score_acgt_and_valid (src/fastaseq.c:585-596) on raw bytes.
    NB: no clamp of maxdist here, unlike the IUPAC variant."""
    if len(idx) == 0 or maxdist <= 0:
        return [0, 0]
    a, b = s1[idx], s2[idx]
    v = ACGT[a] & ACGT[b]
    d = v & (a != b)
    cs = np.cumsum(d.astype(np.int64))
    hit = np.nonzero(cs >= maxdist)[0]
    end = len(idx) if len(hit) == 0 else int(hit[0]) + 1
    return [int(d[:end].sum()), int(v[:end].sum())]


@dataclass
class SearchResult:
    qnames: list
    rows: list            # per query: list of (score6 tuple, ref name, ref ordinal)
    ever_best: list       # ordinals (into the accepted-ref sequence) of refs ever admitted, file order
    n_idx: tuple          # (idx, idx_m, idx_c) sizes
    n_ref_accepted: int = 0
    n_ref_invalid: int = 0
    h1_events: int = 0
    ref_names: list = field(default_factory=list)

    def csv(self, acgt=False):
        out = [HEADER_ACGT if acgt else HEADER_FULL]
        for qn, rows in zip(self.qnames, self.rows):
            for j, (sc, nm, _o) in enumerate(rows):
                out.append("%s,%s,%d," % (qn.decode(), nm.decode(), j + 1) + ",".join(str(int(x)) for x in sc))
        return "\n".join(out) + "\n"


def prepare_queries(qnames, qseqs, trim=0, acgt=False, ambiguity=0.5, keep_resolved=False):
    """filter -> trim clamp -> partition -> reorder -> with -k: prune and partition again (nearest.c:203-224)."""
    keep = filter_queries(qnames, qseqs, min(1.0, max(0.001, ambiguity)))
    if not keep:
        raise ValueError("no valid query sequences")
    Q = np.frombuffer(b"".join(qseqs[i] for i in keep), np.uint8).reshape(len(keep), -1).copy()
    names = [qnames[i] for i in keep]
    trim = clamp_trim(trim, Q.shape[1])
    cons, idx, idx_m, idx_c = partition(Q, trim, acgt)
    order = query_order(Q, trim, acgt)
    Q = Q[order]
    names = [names[i] for i in order]
    if keep_resolved:                                          # nearest.c:219-224: hires->count is 1, so "keep more resolved"
        kept = exclude_redundant(Q, idx, idx_m, acgt, True)
        Q = Q[kept]
        names = [names[i] for i in kept]
        cons, idx, idx_m, idx_c = partition(Q, trim, acgt)
    return names, Q, trim, cons, idx, idx_m, idx_c


def search(qnames, qseqs, ref_files, nbest=100, trim=0, acgt=False, ambig_q=0.5, ambig_r=0.5, pool=512,
           exclude_self=False, literal=False, refs=None, keep_resolved=False):
    """End-to-end restatement of nearest.c main (159-353).
    ref_files: list of FASTA paths, or pass refs=[(names, seqs), ...] directly (one entry per file)."""
    ambig_r = min(1.0, max(0.001, ambig_r))
    nbest = max(1, nbest)
    names, Q, trim, cons, idx, idx_m, idx_c = prepare_queries(qnames, qseqs, trim, acgt, ambig_q, keep_resolved)
    nq, L = Q.shape
    qset = set(names) if exclude_self else None
    non_n_ref = int(L * (1.0 - ambig_r))                       # nearest.c:237
    heaps = [Heap(nbest, L) for _ in range(nq)]
    K = heaps[0].K
    inS = np.zeros(L, bool)
    inS[trim:L - trim] = True
    qc = CODE[Q]
    qc_S = np.where(inS[None, :], qc, 0).astype(np.uint8)
    cons_code = CODE[cons]
    mask_c = np.zeros(L, bool); mask_c[idx_c] = True
    mask_i = np.zeros(L, bool); mask_i[idx] = True
    q_acgt_S = ACGT[Q] & inS[None, :]
    res = SearchResult(names, [], [], (len(idx), len(idx_m), len(idx_c)))
    ordinal = 0
    files = refs if refs is not None else [read_fasta(f) for f in ref_files]
    for rnames, rseqs in files:
        batch = []                                             # accepted refs of the current batch
        stream = list(zip(rnames, rseqs))
        pos = 0
        while True:
            batch = []
            while len(batch) < pool and pos < len(stream):
                nm, sq = stream[pos]; pos += 1
                if qset is not None and nm in qset:
                    continue
                r = np.frombuffer(sq, np.uint8)
                nn = int((~INDEL[r]).sum())                    # fastaseq.c:642-648 (whole sequence)
                if nn < non_n_ref:
                    res.n_ref_invalid += 1
                    continue
                if len(r) != L:
                    raise ValueError("all sequences must be aligned")      # nearest.c:272-278
                batch.append((nm, r, nn))
            if batch:
                _process_batch(batch, heaps, K, Q, qc, qc_S, q_acgt_S, cons, cons_code, idx, idx_m, idx_c,
                               mask_c, mask_i, inS, acgt, literal, res, ordinal)
                ordinal += len(batch)
            if pos >= len(stream):
                break
    res.n_ref_accepted = ordinal
    res.ever_best = sorted(set(res.ever_best))
    for h in heaps:
        res.rows.append(h.finalise())
    return res


def _process_batch(batch, heaps, K, Q, qc, qc_S, q_acgt_S, cons, cons_code, idx, idx_m, idx_c, mask_c, mask_i,
                   inS, acgt, literal, res, ordinal0):
    nq, L = Q.shape
    R = np.stack([b[1] for b in batch])
    rc = CODE[R]
    cq_max = max(h.max_inc for h in heaps)                     # nearest.c:290-291
    if literal:
        for c, (nm, r, nn) in enumerate(batch):
            if acgt:
                cres = truncated_acgt(r, cons, idx_c, cq_max)                  # nearest.c:431
            else:
                cres = truncated_idx(rc[c], cons_code, idx_c, cq_max)          # nearest.c:432
            full_ci = None
            for iq, h in enumerate(heaps):
                item = _literal_pair(acgt, h, cres, r, rc[c], Q[iq], qc[iq], idx, idx_m, nn)
                if item is None:
                    continue
                # H1 bookkeeping: was the consensus stage truncated for an accepted pair?
                if full_ci is None:
                    full_ci = (truncated_acgt(r, cons, idx_c, 1 << 30)[0] if acgt else
                               (lambda t: t[3] - t[0])(truncated_idx(rc[c], cons_code, idx_c, 1 << 30)))
                if full_ci >= cq_max:
                    res.h1_events += 1
                _admit(h, K, (item, nm, ordinal0 + c), acgt, res)
        return
    # closed form
    if acgt:
        ra = ACGT[R]
        f = np.float32
        V = (q_acgt_S.astype(f) @ ra.astype(f).T).astype(np.int32)
        eq = np.zeros((nq, len(batch)), np.int32)
        eqi = np.zeros((nq, len(batch)), np.int32)
        Vi = (((q_acgt_S & mask_i[None, :]).astype(f)) @ ra.astype(f).T).astype(np.int32)
        for ch in b"ACGT":
            qm = (q_acgt_S & (Q == ch)).astype(f)
            rm = (R == ch).astype(f)
            eq += (qm @ rm.T).astype(np.int32)
            eqi += ((qm * mask_i[None, :].astype(f)) @ rm.T).astype(np.int32)
        M = V - eq
        Mi = Vi - eqi
        cmatch = ((R == cons[None, :]) & ra & mask_c[None, :]).sum(1)
        for c, (nm, r, nn) in enumerate(batch):
            for iq, h in enumerate(heaps):
                if M[iq, c] >= h.max_inc:
                    continue
                sc = (int(V[iq, c] - M[iq, c]), int(V[iq, c]), int(V[iq, c] - M[iq, c] - cmatch[c]), nn,
                      int(M[iq, c] - Mi[iq, c]), int(Mi[iq, c]))
                _admit(h, K, (sc, nm, ordinal0 + c), acgt, res)
    else:
        s0, s1, s2, s3 = pair_counts(qc_S, rc)
        c0 = ((cons_code[None, :] == rc) & (POPC4[rc] == 1) & mask_c[None, :]).sum(1)
        for c, (nm, r, nn) in enumerate(batch):
            for iq, h in enumerate(heaps):
                if s3[iq, c] - s0[iq, c] >= h.max_inc:
                    continue
                sc = (int(s0[iq, c]), int(s1[iq, c]), int(s2[iq, c]), int(s3[iq, c]), int(s0[iq, c] - c0[c]), nn)
                _admit(h, K, (sc, nm, ordinal0 + c), acgt, res)


def _admit(h, K, item, acgt, res):
    if h.insert(item):
        res.ever_best.append(item[2])
        if h.n == K:
            root = h.a[1][0]
            h.max_inc = (root[1] - root[0] + 1) if acgt else (root[3] - root[0] + 1)   # nearest.c:475 / :508


def _literal_pair(acgt, h, cres, r, rcode, q, qcode, idx, idx_m, nn):
    """queue_update_min_heaps_full / _acgt (nearest.c:479-510 / 442-477), stage by stage."""
    if acgt:
        ci = cres[0]
        if ci >= h.max_inc:
            return None
        tol = h.max_inc - ci
        r1 = truncated_acgt(r, q, idx_m, tol)
        if r1[0] >= tol:
            return None
        tol -= r1[0]
        r1[0] += cres[0]; r1[1] += cres[1]
        r2 = truncated_acgt(r, q, idx, tol)
        if r2[0] >= tol:
            return None
        s0 = r1[1] + r2[1] - r1[0] - r2[0]
        return (s0, r1[1] + r2[1], s0 - (cres[1] - cres[0]), nn, r1[0], r2[0])
    ci = cres[3] - cres[0]
    if ci >= h.max_inc:
        return None
    tol = h.max_inc - ci
    r1 = truncated_idx(rcode, qcode, idx_m, tol)
    if r1[3] - r1[0] >= tol:
        return None
    tol = tol - r1[3] + r1[0]
    r2 = truncated_idx(rcode, qcode, idx, tol)
    if r2[3] - r2[0] >= tol:
        return None
    s = [r1[i] + r2[i] + cres[i] for i in range(4)]
    return (s[0], s[1], s[2], s[3], r1[0] + r2[0], nn)


# ---------------------------------------------------------------------------------------------
# uvaiaball (src/ball.c): radius search
# ---------------------------------------------------------------------------------------------
def ball_clamp_dist(dist, nchar, trim):
    """new_query_structure_from_fasta (src/fastaseq.c:713-715); trim already clamped"""
    dist = max(0, int(dist))
    if dist > (nchar - 2 * trim) / 10:
        dist = int((nchar - 2 * trim) / 10)
    return dist


def ball_prepare_queries(qnames, qseqs, trim=0, acgt=False, ambiguity=0.5, keep_more_resolved=False):
    """ball.c:172-194: filter -> partition -> reorder -> ALWAYS prune (-k keeps the more resolved) -> partition again"""
    keep = filter_queries(qnames, qseqs, min(1.0, max(0.001, ambiguity)))
    if not keep:
        raise ValueError("no valid query sequences")
    Q = np.frombuffer(b"".join(qseqs[i] for i in keep), np.uint8).reshape(len(keep), -1).copy()
    names = [qnames[i] for i in keep]
    trim = clamp_trim(trim, Q.shape[1])
    cons, idx, idx_m, idx_c = partition(Q, trim, acgt)
    order = query_order(Q, trim, acgt)
    Q = Q[order]
    names = [names[i] for i in order]
    kept = exclude_redundant(Q, idx, idx_m, acgt, keep_more_resolved)
    Q = Q[kept]
    names = [names[i] for i in kept]
    cons, idx, idx_m, idx_c = partition(Q, trim, acgt)
    return names, Q, trim, cons, idx, idx_m, idx_c


def _ball_count(s1, s2, idx, maxdist, acgt):
    """quick_pairwise_score_truncated_idx_indelcheck / quick_pairwise_score_acgt (src/fastaseq.c:562-583): sites of idx
    where both bytes are usable and differ, the loop stops when the count reaches maxdist"""
    if len(idx) == 0 or maxdist <= 0:
        return 0
    a, b = s1[idx], s2[idx]
    use = (ACGT[a] & ACGT[b]) if acgt else (~INDEL[a] & ~INDEL[b])
    d = int((use & (a != b)).sum())
    return min(d, maxdist)


def ball_min_dist(r, Q, cons, idx, idx_m, idx_c, radius, acgt):
    """seq_ball_against_query_structure (src/fastaseq.c:660-696), literally: `radius` = dist + 1"""
    md = _ball_count(r, cons, idx_c, radius, acgt)
    if md >= radius:
        return md
    c = md
    md = _ball_count(r, cons, idx_m, radius, acgt) + c
    if md >= radius:
        return md
    c = md
    i = 0
    while i < Q.shape[0] and (md + c) >= radius:               # NB the first test compares 2c with the radius
        md = _ball_count(r, Q[i], idx, radius - c, acgt)
        i += 1
    return md + c


def ball_search(qnames, qseqs, refs, dist=5, trim=0, acgt=False, ambig_q=0.5, ambig_r=0.5, keep_more_resolved=False):
    """End-to-end restatement of uvaiaball (src/ball.c:138-291).  refs = [(names, seqs), ...] one entry per -r file.
    -> (names of the references kept, in file order; clamped dist; number too ambiguous)"""
    ambig_r = min(1.0, max(0.001, ambig_r))
    names, Q, trim, cons, idx, idx_m, idx_c = ball_prepare_queries(qnames, qseqs, trim, acgt, ambig_q, keep_more_resolved)
    L = Q.shape[1]
    dist = ball_clamp_dist(dist, L, trim)
    non_n_ref = int(L * ambig_r)                               # ball.c:201 -- A, not 1 - A
    kept, n_invalid = [], 0
    for rnames, rseqs in refs:
        for nm, sq in zip(rnames, rseqs):
            r = np.frombuffer(sq, np.uint8)
            if int((~INDEL[r]).sum()) < non_n_ref:
                n_invalid += 1
                continue
            if len(r) != L:
                raise ValueError("all sequences must be aligned")
            if ball_min_dist(r, Q, cons, idx, idx_m, idx_c, dist + 1, acgt) <= dist:
                kept.append(nm)
    return kept, dist, n_invalid


def ball_run_files(query_file, ref_files, **kw):
    qn, qs = read_fasta(query_file)
    return ball_search(qn, qs, [read_fasta(f) for f in ref_files], **kw)


def run_files(query_file, ref_files, **kw):
    qn, qs = read_fasta(query_file)
    return search(qn, qs, ref_files, **kw)


if __name__ == "__main__":
    import argparse, sys
    ap = argparse.ArgumentParser(description="uvaia n-best search oracle (numpy restatement; test infrastructure)")
    ap.add_argument("query"); ap.add_argument("-r", "--reference", action="append", required=True)
    ap.add_argument("-n", "--nbest", type=int, default=100); ap.add_argument("--trim", type=int, default=0)
    ap.add_argument("-a", type=float, default=0.5); ap.add_argument("-A", type=float, default=0.5)
    ap.add_argument("-p", "--pool", type=int, default=512); ap.add_argument("--acgt", action="store_true")
    ap.add_argument("-x", action="store_true"); ap.add_argument("--literal", action="store_true")
    a = ap.parse_args()
    r = run_files(a.query, a.reference, nbest=a.nbest, trim=a.trim, acgt=a.acgt, ambig_q=a.a, ambig_r=a.A,
                  pool=a.pool, exclude_self=a.x, literal=a.literal)
    sys.stdout.write(r.csv(a.acgt))
