"""One rank of a sharded (row-partitioned) search on real GPUs; launched by torchrun from test_gpu_sharded.py / by hand:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/sharded_worker.py
Every rank pushes the tiles uvg_shard_of assigns to it, the ranks search together (score tiles to the query owners over
peer memory, NCCL for the barrier / final gathers), and every rank checks ALL tables against the numpy oracle and
rank 0 also against a single-GPU engine holding the whole database."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch
import torch.distributed as dist

from helpers import orc
from uvaia_b200.engine import Engine
from uvaia_b200.shard import rank_tiles
from uvaia_b200.synth import SynthModel


def run_case(rank, world, L, nq, nr, nbest, acgt=False, amb=0.0, first_round=64, max_round=256, early_abort=False, seed=5, check_oracle=True):
    m = SynthModel(nsites=L, n_clades=12, seed=seed, clade_subs=max(3, L // 500))
    qn, Q = m.sample(nq, seed + 1, ambiguous_frac=amb, prefix="Q")
    rn, R = m.sample(nr, seed + 2, ambiguous_frac=amb / 2, prefix="R")
    qs = [q.tobytes() for q in Q]
    names, Qo, *_ = orc.prepare_queries(qn, qs, 0, acgt, 0.5)
    min_non_n = int(L * 0.5)
    keep = np.nonzero((~orc.INDEL[R]).sum(1) >= min_non_n)[0]          # the host applies the reference filter before dealing rows out
    Racc = np.ascontiguousarray(R[keep])
    total = len(Racc)
    dev = int(os.environ.get("LOCAL_RANK", rank))
    with Engine(L, nbest=nbest, acgt=acgt, pool=64, device=dev, first_round=first_round, max_round=max_round, early_abort=early_abort,
                rank=rank, world=world) as eng:
        eng.set_queries(Qo)
        for gb, cnt in rank_tiles(total, rank, world, first_round, max_round):
            nn, acc, na = eng.push_refs(Racc[gb:gb + cnt], min_non_n)
            assert na == cnt
        uid = [Engine.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        eng.nccl_init(uid[0])
        eng.set_global_count(total)
        eng.search()
        got = [eng.heap(iq) for iq in range(len(names))]
        eb = eng.ever_best(0, total)
        st = eng.stats()
        # a second query set on the resident shards: the communicator exists now, so rank 0 alone prepares the queries and
        # the other ranks receive the arrays by broadcast (uvg_set_queries)
        names2, Qo2, *_ = orc.prepare_queries(qn[::2], qs[::2], 0, acgt, 0.5)
        info2 = eng.set_queries(Qo2)
        eng.search()
        got2 = [eng.heap(iq) for iq in range(len(names2))]
        eb2 = eng.ever_best(0, total)
        idx2 = (info2.n_idx, info2.n_idx_m, info2.n_idx_c)
    bad = 0
    if check_oracle:
        ref = orc.search(qn, qs, None, nbest=nbest, acgt=acgt, refs=[(rn, [r.tobytes() for r in R])], pool=64)
        for iq in range(len(names)):
            ws = np.array([w[0] for w in ref.rows[iq]], np.int32).reshape(-1, 6)
            wr = np.array([w[2] for w in ref.rows[iq]], np.int64)
            if not (np.array_equal(got[iq][0], ws) and np.array_equal(got[iq][1], wr)):
                bad += 1
        if not np.array_equal(np.nonzero(eb)[0], np.array(ref.ever_best, np.int64)):
            bad += 1000
    if rank == 0:
        # sharded contexts run the pool-independent closed form (uvg_params.closed_form) unless the query set fits one tile
        # of 64 slots, where they emulate the reference's batches literally like a single GPU does: compare like with like
        def single_gpu(Qset, want, want_eb):
            bad = 0
            with Engine(L, nbest=nbest, acgt=acgt, pool=64, device=dev, first_round=first_round, max_round=max_round, closed_form=len(Qset) >= 64) as one:
                info = one.set_queries(Qset)
                one.push_refs(Racc, min_non_n)
                one.search()
                for iq in range(len(Qset)):
                    sc, rf = one.heap(iq)
                    if not (np.array_equal(want[iq][0], sc) and np.array_equal(want[iq][1], rf)):
                        bad += 1
                if not np.array_equal(want_eb, one.ever_best(0, total)):
                    bad += 1000
            return bad, (info.n_idx, info.n_idx_m, info.n_idx_c)
        bad += single_gpu(Qo, got, eb)[0]
        b2, i2 = single_gpu(Qo2, got2, eb2)
        bad += b2
        if idx2 != i2:
            bad += 100000
    if check_oracle:                                                    # every rank got the same site-set sizes through the broadcast
        c2, x2, xm2, xc2 = orc.partition(Qo2, 0, acgt)
        if idx2 != (len(x2), len(xm2), len(xc2)):
            bad += 100000
    t = torch.tensor([bad], device=f"cuda:{dev}")
    dist.all_reduce(t)
    if rank == 0:
        print(f"sharded x{world}: L={L} nq={nq} nr={nr} K={nbest} acgt={acgt} amb={amb} rounds {first_round}..{max_round} ea={early_abort}: "
              f"{'OK' if int(t.item()) == 0 else 'FAIL ' + str(int(t.item()))}  rounds={st['n_rounds']} admitted(rank0)={st['n_admitted']}", flush=True)
    return int(t.item()) == 0


def run_literal_case(rank, world, nq, pool, acgt=False, seed=301, first_round=64, max_round=256):
    """hazard H1 on shards: a query set of one tile (<= 63 queries) runs the literal emulation of the reference's --pool
    batches on rank 0, with the packed planes of truncated references read from the rank that holds them; results = the
    oracle's literal mode = a single-GPU context in its default (literal) mode"""
    L, nbest = 1500, 12
    m = SynthModel(nsites=L, n_clades=6, seed=seed, clade_subs=12)
    qn, Q = m.sample(nq, seed + 1, ambiguous_frac=0.0 if seed != 306 else 0.02, prefix="Q")
    rn, R = m.sample(1300, seed + 2, ambiguous_frac=0.01, prefix="R")
    rng = np.random.default_rng(seed)
    for r in R:                                  # N runs anywhere: the consensus-stage mismatches move around
        for _ in range(int(rng.integers(0, 4))):
            a = int(rng.integers(0, L - 50))
            r[a:a + int(rng.integers(5, 200))] = ord("N")
    qs = [q.tobytes() for q in Q]
    names, Qo, *_ = orc.prepare_queries(qn, qs, 0, acgt, 0.5)
    min_non_n = int(L * 0.5)
    keep = np.nonzero((~orc.INDEL[R]).sum(1) >= min_non_n)[0]
    Racc = np.ascontiguousarray(R[keep])
    total = len(Racc)
    file_end = int((keep < 801).sum())                                   # two -r files: a batch ends with the first file
    dev = int(os.environ.get("LOCAL_RANK", rank))
    with Engine(L, nbest=nbest, acgt=acgt, pool=pool, device=dev, first_round=first_round, max_round=max_round, rank=rank, world=world) as eng:
        eng.set_queries(Qo)
        for gb, cnt in rank_tiles(total, rank, world, first_round, max_round):
            nn, acc, na = eng.push_refs(Racc[gb:gb + cnt], min_non_n)
            assert na == cnt
        eng.mark_file_end(file_end)
        uid = [Engine.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        eng.nccl_init(uid[0])
        eng.set_global_count(total)
        eng.search()
        got = [eng.heap(iq) for iq in range(len(names))]
        eb = eng.ever_best(0, total)
        st = eng.stats()
    ref = orc.search(qn, qs, None, nbest=nbest, acgt=acgt, refs=[(rn[:801], [r.tobytes() for r in R[:801]]), (rn[801:], [r.tobytes() for r in R[801:]])],
                     pool=pool, literal=True)
    bad = 0
    for iq in range(len(names)):
        ws = np.array([w[0] for w in ref.rows[iq]], np.int32).reshape(-1, 6)
        wr = np.array([w[2] for w in ref.rows[iq]], np.int64)
        if not (np.array_equal(got[iq][0], ws) and np.array_equal(got[iq][1], wr)):
            bad += 1
    if not np.array_equal(np.nonzero(eb)[0], np.array(ref.ever_best, np.int64)):
        bad += 1000
    if rank == 0:
        with Engine(L, nbest=nbest, acgt=acgt, pool=pool, device=dev, first_round=first_round, max_round=max_round) as one:
            one.set_queries(Qo)
            one.push_refs(Racc, min_non_n)
            one.mark_file_end(file_end)
            one.search()
            for iq in range(len(names)):
                sc, rf = one.heap(iq)
                if not (np.array_equal(got[iq][0], sc) and np.array_equal(got[iq][1], rf)):
                    bad += 1
    t = torch.tensor([bad], device=f"cuda:{dev}")
    dist.all_reduce(t)
    if rank == 0:
        print(f"sharded x{world} literal: nq={nq} pool={pool} acgt={acgt} seed={seed}: {'OK' if int(t.item()) == 0 else 'FAIL ' + str(int(t.item()))}"
              f"  oracle h1 events={ref.h1_events} flags(rank0)={st['n_h1_flags']}", flush=True)
    return int(t.item()) == 0


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local))
    ok = True
    ok &= run_case(rank, world, 2000, 70, 2500, 10)
    ok &= run_case(rank, world, 2000, 200, 3000, 20, amb=0.1, first_round=32, max_round=128)
    ok &= run_case(rank, world, 2000, 130, 2500, 10, acgt=True)
    ok &= run_case(rank, world, 1500, 65, 999, 100, first_round=32, max_round=64, early_abort=True)
    ok &= run_case(rank, world, 29903, 300, 6000, 100, first_round=256, max_round=1024, check_oracle=False)
    # many queries and full-size rounds: the score tiles exceed the 3 GiB a single-GPU context caps them at, and the round
    # schedule of the shards must still be the one uvg_shard_of deals the rows with (first_round, max_round only)
    ok &= run_case(rank, world, 400, 4200, 140000, 10, first_round=16384, max_round=65536, check_oracle=False, seed=11)
    for nq, pool, acgt, seed in ((1, 1, False, 301), (1, 7, False, 302), (1, 64, True, 303), (2, 7, False, 304), (3, 5, True, 305), (3, 64, False, 306),
                                 (1, 512, False, 307), (2, 1, True, 308)):
        ok &= run_literal_case(rank, world, nq, pool, acgt=acgt, seed=seed)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("SHARDED ALL OK" if ok else "SHARDED FAILED", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
