"""Developer check (GPU box): GPU search vs the numpy oracle on small synthetic inputs.  Not part of the product."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from uvaia_b200.synth import SynthModel
from uvaia_b200.engine import Engine
from oracle import uvaia_oracle as orc

def run_case(L, nq, nr, nbest, acgt=False, trim=0, amb=0.0, first_round=0, max_round=0, early_abort=False, seed=5):
    m = SynthModel(nsites=L, n_clades=12, seed=seed, clade_subs=max(3, L // 500))
    qn, Q = m.sample(nq, seed + 1, ambiguous_frac=amb, prefix="Q")
    rn, R = m.sample(nr, seed + 2, ambiguous_frac=amb / 2, prefix="R")
    qs = [q.tobytes() for q in Q]; rs = [r.tobytes() for r in R]
    t0 = time.time()
    ref = orc.search(qn, qs, None, nbest=nbest, trim=trim, acgt=acgt, refs=[(rn, rs)], pool=64)
    t1 = time.time()
    names, Qo, tr, cons, idx, idx_m, idx_c = orc.prepare_queries(qn, qs, trim, acgt, 0.5)
    with Engine(L, nbest=nbest, trim=trim, acgt=acgt, pool=64, first_round=first_round, max_round=max_round,
                early_abort=early_abort) as eng:
        info = eng.set_queries(Qo)
        assert (info.n_idx, info.n_idx_m, info.n_idx_c) == (len(idx), len(idx_m), len(idx_c)), (info.n_idx, info.n_idx_m, info.n_idx_c, len(idx), len(idx_m), len(idx_c))
        nn, acc, na = eng.push_refs(R, int(L * 0.5))
        # encoder parity
        lut = orc.CODE.copy()
        if acgt:
            lut = np.zeros(256, np.uint8)
            for k, ch in enumerate(b"ACGT"): lut[ch] = 1 << k
        kept = np.nonzero(acc)[0]
        assert np.array_equal(nn, (~orc.INDEL[R]).sum(1)), "non_n mismatch"
        for j in (0, len(kept) // 2, len(kept) - 1):
            assert np.array_equal(eng.debug_unpack_ref(j), lut[R[kept[j]]]), "encode mismatch"
        # pair counts parity on a slice
        if not acgt:
            inS = np.zeros(L, bool); inS[tr:L - tr] = True
            qc = np.where(inS[None, :], orc.CODE[Qo], 0).astype(np.uint8)
            n = min(100, na)
            want = np.stack(orc.pair_counts(qc, orc.CODE[R[kept[:n]]]), -1)
            got = eng.debug_pair_counts(0, n)
            assert np.array_equal(got, want), ("pair counts mismatch", np.argwhere(got != want)[:5])
        eng.search()
        st = eng.stats()
        bad = 0
        for iq in range(len(names)):
            sc, refs = eng.heap(iq)
            want = ref.rows[iq]
            ws = np.array([w[0] for w in want], np.int32).reshape(-1, 6); wr = np.array([w[2] for w in want], np.int64)
            if not (np.array_equal(sc, ws) and np.array_equal(refs, wr)):
                bad += 1
                if bad == 1:
                    print("  MISMATCH query", iq, "\n   got", sc[:3], refs[:3], "\n  want", ws[:3], wr[:3], len(refs), len(wr))
        eb = eng.ever_best(0, na)
        eb_ok = np.array_equal(np.nonzero(eb)[0], np.array(ref.ever_best, np.int64))
    print(f"L={L} nq={nq} nr={nr} K={nbest} acgt={acgt} trim={trim} amb={amb} fr={first_round} mr={max_round} ea={early_abort}: "
          f"{'OK' if bad == 0 and eb_ok else 'FAIL'} bad={bad} ever_best_ok={eb_ok} oracle={t1-t0:.1f}s rounds={st['n_rounds']} "
          f"adm={st['n_admitted']} h1={st['n_h1_flags']} pair_ms={st['pair_kernel_ms']:.2f}")
    return bad == 0 and eb_ok

ok = True
ok &= run_case(1000, 10, 300, 5)
ok &= run_case(2000, 70, 2500, 10, first_round=64, max_round=512)
ok &= run_case(2000, 70, 2500, 10, first_round=64, max_round=512, early_abort=True)
ok &= run_case(2000, 33, 1500, 10, acgt=True)
ok &= run_case(2000, 33, 1500, 10, trim=100)
ok &= run_case(2000, 33, 1500, 1, amb=0.1)
ok &= run_case(2000, 40, 3000, 20, amb=0.2, first_round=32, max_round=256)
ok &= run_case(2000, 40, 3000, 20, amb=0.2, first_round=32, max_round=256, early_abort=True, acgt=True)
ok &= run_case(29903, 20, 2000, 10, early_abort=True)
print("ALL OK" if ok else "SOME FAILED")
sys.exit(0 if ok else 1)
