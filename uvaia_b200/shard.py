"""Row sharding of the reference database over the GPUs of one box (host-side mirror of uvg_shard_of, engine.cu).

The reference is a single process; the partition is ours.  The per-query admission of the reference is
order dependent, so shards are NOT independent top-k problems: references are dealt to ranks in rounds of contiguous
tiles in file order -- round j covers global ordinals [S_j, S_j + world*T_j), rank g scores [S_j + g*T_j, S_j + (g+1)*T_j),
T_1 = first_round, T_{j+1} = min(4*T_j, max_round); towards the end of the database the rounds taper off (round_size_at:
half of what is left per rank, never below 4,096, rounded up to 64), so that every rank gets the same share of the tail
and the last replay is short -- and the scores of every round are handed to the rank that owns
the query (query tile qt belongs to rank qt % world), which replays them in global file order.
"""
from __future__ import annotations

GROUP = 32          # references interleaved per 16-byte column word (csrc/common.cuh)
QTILE = 64          # queries per thread block of the pair-count kernel


def round_sizes(first_round=0, max_round=0):
    """(first, max) after the library's rounding (round_size_first / round_size_max, engine.cu)"""
    first = first_round if first_round > 0 else 1024
    first = (first + GROUP - 1) // GROUP * GROUP
    mx = max_round if max_round > 0 else 65536
    mx = (mx + GROUP - 1) // GROUP * GROUP
    return first, max(mx, first)


ROUND_TAPER_MIN = 4096


def round_size_at(S, T, total, world):
    """size per rank of the round that starts at global ordinal S with nominal size T (twin of round_size_at, engine.cu):
    once less than two nominal rounds remain per rank, every round takes half of what is left (the last replay, which no
    pair counting hides, becomes short; the ranks share the tail evenly)"""
    if total is None or total < 0:
        return T
    rem = max(0, total - S)
    share = (rem + world - 1) // world
    if share >= 2 * T:
        return T
    v = share if share <= ROUND_TAPER_MIN else max(ROUND_TAPER_MIN, (share + 1) // 2)
    return min(T, max(64, (v + 63) // 64 * 64))


def rounds(total, world, first_round=0, max_round=0):
    """[(S_j, T_j)] for every round needed to cover `total` accepted references"""
    T, mx = round_sizes(first_round, max_round)
    out, S = [], 0
    while S < total:
        Te = round_size_at(S, T, total, world)
        out.append((S, Te))
        S += world * Te
        T = min(4 * T, mx)
    return out


def rank_tiles(total, rank, world, first_round=0, max_round=0):
    """[(global_first, count)] of the reference ordinals rank `rank` holds, round by round"""
    out = []
    for S, T in rounds(total, world, first_round, max_round):
        gb = S + rank * T
        ge = min(gb + T, total)
        if ge > gb:
            out.append((gb, ge - gb))
    return out


def shard_of(ordinal, world, first_round=0, max_round=0, total=None):
    """(rank, local index) of a global ordinal: python twin of uvg_shard_of"""
    T, mx = round_sizes(first_round, max_round)
    S = Ls = 0
    while True:
        Te = round_size_at(S, T, total, world)
        if ordinal < S + world * Te:
            return (ordinal - S) // Te, Ls + (ordinal - S) % Te
        S += world * Te
        Ls += Te
        T = min(4 * T, mx)


def query_owner(iq, world):
    """rank that replays query iq (ownership goes by tiles of QTILE queries, round-robin)"""
    return (iq // QTILE) % world


def owned_queries(n_query, rank, world):
    return [q for q in range(n_query) if query_owner(q, world) == rank]
