"""The C-ABI shared library: it loads, exports every symbol include/uvaia_b200.h declares, refuses to run without a
CUDA device (no CPU fallback), and its host-only helpers (shard schedule) behave.  No compute calls here."""
import ctypes as C
import os
import re

import pytest

from helpers import ROOT
from uvaia_b200 import _capi

HEADER = os.path.join(ROOT, "include", "uvaia_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(uvg_[a-z0-9_]+)\s*\(", src)))


def test_library_is_built_in_tree():
    assert os.path.exists(_capi.LIB_PATH), "build with `make -C uvaia_b200/csrc` or __graft_entry__.build()"
    assert os.path.realpath(_capi.LIB_PATH).startswith(os.path.realpath(ROOT))


def test_every_declared_symbol_is_exported_and_bound():
    lib = C.CDLL(_capi.LIB_PATH)
    decl = declared_symbols()
    assert len(decl) >= 25
    for name in decl:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert sorted(_capi.SYMBOLS) == decl, "ctypes binding table and header disagree"


def test_no_torch_types_in_the_abi():
    src = open(HEADER).read()
    assert "torch" not in src.lower().replace("no torch", "") and "at::" not in src and "std::" not in src


def test_abi_version_and_struct_sizes():
    lib = _capi.lib()
    assert lib.uvg_abi_version() == _capi.UVG_ABI_VERSION
    assert C.sizeof(_capi.uvg_item) == 32


def test_create_rejects_bad_arguments_and_missing_device():
    import torch
    lib = _capi.lib()
    h = C.c_void_p()
    p = _capi.uvg_params()
    p.abi_version = 999
    assert lib.uvg_create(C.byref(p), C.byref(h)) == -1            # UVG_EINVAL
    p.abi_version = _capi.UVG_ABI_VERSION
    p.nchar, p.nbest = 0, 10
    assert lib.uvg_create(C.byref(p), C.byref(h)) == -1
    if not torch.cuda.is_available():
        p.nchar = 100
        rc = lib.uvg_create(C.byref(p), C.byref(h))
        assert rc == -2, "without a CUDA device uvg_create must fail with UVG_ENODEV (there is no CPU fallback)"
        assert b"no CPU fallback" in lib.uvg_last_error(None)
        from uvaia_b200.engine import Engine, UvgError
        with pytest.raises(UvgError):
            Engine(100)


def test_shard_schedule_partitions_ordinals_in_file_order():
    from uvaia_b200.engine import Engine
    from uvaia_b200.shard import rank_tiles
    # incl. totals below one round, below the taper's floor of 4,096 per rank, and totals that end exactly on a round boundary
    for world, first, mx, total in [(1, 32, 128, 1000), (2, 32, 128, 1000), (4, 64, 256, 5000), (8, 1024, 65536, 300000), (3, 32, 32, 333),
                                    (2, 1024, 65536, 1), (8, 1024, 65536, 63), (2, 1024, 65536, 2 * (1024 + 4096 + 16384)), (4, 1024, 65536, 1_000_000),
                                    (8, 1024, 65536, 1_000_000), (2, 16384, 65536, 140_000), (5, 64, 64, 64 * 5 * 7)]:
        seen = {}
        for rank in range(world):
            local = 0
            for gb, cnt in rank_tiles(total, rank, world, first, mx):
                for o in (gb, gb + cnt - 1):
                    r, l = Engine.shard_of(o, world, first, mx, total)
                    assert r == rank
                seen[rank] = seen.get(rank, []) + [(gb, cnt, local)]
                # local index of the tile start as the library computes it
                r, l = Engine.shard_of(gb, world, first, mx, total)
                assert l >= local
                local = l + cnt
        # tiles of all ranks cover [0, total) exactly once
        cover = sorted((gb, cnt) for v in seen.values() for gb, cnt, _ in v)
        pos = 0
        for gb, cnt in cover:
            assert gb == pos
            pos += cnt
        assert pos == total
        # the tail round is shared evenly: no rank holds more than 64 + one round's rounding above the mean
        per = [sum(cnt for _, cnt, _ in seen.get(r, [])) for r in range(world)]
        assert max(per) - min(per) <= 64 + first, per
