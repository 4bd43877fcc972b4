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
        # sharded contexts run the pool-independent closed form (uvg_params.closed_form): compare like with like
        with Engine(L, nbest=nbest, acgt=acgt, pool=64, device=dev, first_round=first_round, max_round=max_round, closed_form=True) as one:
            one.set_queries(Qo)
            one.push_refs(Racc, min_non_n)
            one.search()
            for iq in range(len(names)):
                sc, rf = one.heap(iq)
                if not (np.array_equal(got[iq][0], sc) and np.array_equal(got[iq][1], rf)):
                    bad += 1
            if not np.array_equal(eb, one.ever_best(0, total)):
                bad += 1000
            i2 = one.set_queries(Qo2)
            one.search()
            for iq in range(len(names2)):
                sc, rf = one.heap(iq)
                if not (np.array_equal(got2[iq][0], sc) and np.array_equal(got2[iq][1], rf)):
                    bad += 1
            if not np.array_equal(eb2, one.ever_best(0, total)):
                bad += 1000
            if idx2 != (i2.n_idx, i2.n_idx_m, i2.n_idx_c):
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
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("SHARDED ALL OK" if ok else "SHARDED FAILED", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
