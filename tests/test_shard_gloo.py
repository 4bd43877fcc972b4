"""The N>1 protocol of the row-sharded search, on CPU with world_size-2 gloo processes (no CUDA): every rank scores only
the tiles uvaia_b200.shard deals to it (scores from the ORACLE's pair counts -- this test is about the distribution and
the order of the replay, not about kernels), hands each query row to the rank that owns the query (all_to_all), the
owners replay their queries over the segments of all ranks in file order with the oracle's heap, and the gathered
tables must equal the single-process oracle search."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import ROOT, orc
from uvaia_b200 import shard
from uvaia_b200.synth import SynthModel

WORLD = 2


def _inputs():
    L = 600
    m = SynthModel(nsites=L, n_clades=6, seed=3, clade_subs=5)
    qn, Q = m.sample(150, 4, ambiguous_frac=0.03, prefix="Q")       # 150 queries -> 3 tiles of 64 (+ consensus): both ranks own some
    rn, R = m.sample(700, 5, ambiguous_frac=0.02, prefix="R")
    return L, qn, Q, rn, R


def _worker(rank, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    L, qn, Q, rn, R = _inputs()
    nbest, first_round, max_round = 7, 32, 128
    names, Qo, trim, cons, idx, idx_m, idx_c = orc.prepare_queries(qn, [q.tobytes() for q in Q], 0, False, 0.5)
    nq = len(names)
    keep = np.nonzero((~orc.INDEL[R]).sum(1) >= L // 2)[0]
    Racc = R[keep]
    total = len(Racc)
    qc = orc.CODE[Qo]
    cons_code = orc.CODE[cons]
    mask_c = np.zeros(L, bool)
    mask_c[idx_c] = True
    owned = shard.owned_queries(nq, rank, WORLD)
    heaps = {q: orc.Heap(nbest, L) for q in owned}
    K = max(2, nbest)
    ever = np.zeros(total, np.uint8)
    for S, T in shard.rounds(total, WORLD, first_round, max_round):
        # --- this rank's tile of the round ---
        gb = S + rank * T
        ge = min(gb + T, total)
        cnt = max(0, ge - gb)
        if cnt:
            rc = orc.CODE[Racc[gb:ge]]
            s = np.stack(orc.pair_counts(qc, rc), -1).astype(np.int32)                       # (nq, cnt, 4)
            c0 = ((cons_code[None, :] == rc) & (orc.POPC4[rc] == 1) & mask_c[None, :]).sum(1).astype(np.int32)
            nn = (~orc.INDEL[Racc[gb:ge]]).sum(1).astype(np.int32)
        else:
            s, c0, nn = np.zeros((nq, 0, 4), np.int32), np.zeros(0, np.int32), np.zeros(0, np.int32)
        # --- every query row goes to its owner; the per-reference records go to everybody (fixed-size padded buffers) ---
        send = []
        for dst in range(WORLD):
            rows = shard.owned_queries(nq, dst, WORLD)
            buf = np.zeros((len(rows) + 2, T, 4), np.int32)
            buf[:len(rows), :cnt] = s[rows]
            buf[len(rows), :cnt, 0] = c0
            buf[len(rows) + 1, :cnt, 0] = nn
            send.append(torch.from_numpy(buf))
        recv = [torch.zeros_like(send[rank]) for _ in range(WORLD)]
        # gloo has no all_to_all: one gather per destination does the same exchange
        for dst in range(WORLD):
            dist.gather(send[dst], recv if rank == dst else None, dst=dst)
        # --- replay of the owned queries over the segments of all ranks, in rank (= file) order ---
        for g in range(WORLD):
            seg_first = S + g * T
            seg_cnt = max(0, min(T, total - seg_first))
            seg = recv[g].numpy()
            for k, q in enumerate(owned):
                h = heaps[q]
                for i in range(seg_cnt):
                    a, t, p, v = (int(x) for x in seg[k, i])
                    if v - a >= h.max_inc:
                        continue
                    item = ((a, t, p, v, a - int(seg[len(owned), i, 0]), int(seg[len(owned) + 1, i, 0])), b"", seg_first + i)
                    if h.insert(item):
                        ever[seg_first + i] = 1
                        if h.n == K:
                            root = h.a[1][0]
                            h.max_inc = root[3] - root[0] + 1
    # --- final gathers: tables of all queries + OR of the ever-best flags ---
    mine = {q: [(r[0], r[2]) for r in heaps[q].finalise()] for q in owned}
    allrows = [None] * WORLD
    dist.all_gather_object(allrows, mine)
    ev = torch.from_numpy(ever)
    dist.all_reduce(ev, op=dist.ReduceOp.MAX)
    if rank == 0:
        merged = {}
        for d in allrows:
            merged.update(d)
        np.save(os.path.join(out_dir, "ever.npy"), ev.numpy())
        import pickle
        pickle.dump(merged, open(os.path.join(out_dir, "rows.pkl"), "wb"))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_protocol_is_exact_world2(tmp_path):
    port = 29000 + os.getpid() % 2000
    mp.spawn(_worker, args=(port, str(tmp_path)), nprocs=WORLD, join=True)
    import pickle
    merged = pickle.load(open(tmp_path / "rows.pkl", "rb"))
    ever = np.load(tmp_path / "ever.npy")
    L, qn, Q, rn, R = _inputs()
    ref = orc.search(qn, [q.tobytes() for q in Q], None, nbest=7, refs=[(rn, [r.tobytes() for r in R])])
    assert len(merged) == len(ref.qnames)
    for iq, rows in enumerate(ref.rows):
        assert [(r[0], r[2]) for r in rows] == merged[iq], iq
    assert np.nonzero(ever)[0].tolist() == ref.ever_best


def test_ownership_and_tiles_are_partitions():
    for world in (1, 2, 4, 8):
        for nq in (1, 63, 64, 65, 962, 10000):
            seen = sorted(q for r in range(world) for q in shard.owned_queries(nq, r, world))
            assert seen == list(range(nq))
        for total in (0, 1, 31, 1000, 300000):
            cover = sorted(t for r in range(world) for t in shard.rank_tiles(total, r, world, 64, 4096))
            pos = 0
            for gb, cnt in cover:
                assert gb == pos
                pos += cnt
            assert pos == total
            for o in (0, total // 2, total - 1):
                if total:
                    r, l = shard.shard_of(o, world, 64, 4096, total)
                    tiles = shard.rank_tiles(total, r, world, 64, 4096)
                    assert any(gb <= o < gb + cnt for gb, cnt in tiles)
