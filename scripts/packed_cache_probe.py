#!/usr/bin/env python
"""Times uvg_import_packed alone (no search): pinned and pageable blobs, pivot form on and off."""
import ctypes as C
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from uvaia_b200.engine import Engine
from uvaia_b200.synth import SynthModel

L, N, PIECE = 29903, 196608, 65536
m = SynthModel(nsites=L, n_clades=40, seed=1)
qn, Q = m.sample(962, 2002, prefix="Q")
rn, R = m.sample(4096, 1001, prefix="R")
R = np.ascontiguousarray(np.tile(R, (N // 4096, 1)))
for kernel in (2, 1):
    with Engine(L, nbest=100, pair_kernel=kernel, db_capacity=N) as eng:
        eng.set_queries(Q)
        t = time.perf_counter(); eng.push_refs(R, 0); t_text = time.perf_counter() - t
        sizes = []
        for s0 in range(0, N, PIECE):
            nb = C.c_size_t(0)
            eng._ck(eng._lib.uvg_packed_bytes(eng._h, s0, PIECE, C.byref(nb)), "bytes")
            sizes.append(nb.value)
        buf = eng._lib.uvg_host_alloc_pinned(sum(sizes))
        o = 0
        t = time.perf_counter()
        for k, s0 in enumerate(range(0, N, PIECE)):
            eng._ck(eng._lib.uvg_export_packed(eng._h, s0, PIECE, buf + o, sizes[k]), "export"); o += sizes[k]
        t_exp = time.perf_counter() - t
        pageable = np.ctypeslib.as_array((C.c_uint8 * sum(sizes)).from_address(buf)).copy()
        for label, base in (("pinned", buf), ("pageable", pageable.ctypes.data)):
            for rep in range(2):
                eng.clear_refs()
                o = 0
                ts = []
                for k in range(len(sizes)):
                    t = time.perf_counter(); eng.import_packed((base + o, sizes[k])); ts.append(time.perf_counter() - t); o += sizes[k]
            print(f"kernel {kernel} {label}: import {[round(x * 1e3, 1) for x in ts]} ms per {sizes[0] / 1e6:.0f} MB piece "
                  f"-> {sum(sizes) / sum(ts) / 1e9:.1f} GB/s; stream time of last piece {eng.stats()['encode_ms']:.1f} ms; "
                  f"text push of the same rows {t_text * 1e3:.0f} ms, export {t_exp * 1e3:.0f} ms")
        # the pipelined host loop: import a piece, enqueue its rounds, import the next one under them
        for rep in range(2):
            eng.clear_refs()
            eng.set_queries(Q)
            o, ts = 0, []
            t0 = time.perf_counter()
            for k in range(len(sizes)):
                t = time.perf_counter(); eng.import_packed((buf + o, sizes[k])); t1 = time.perf_counter(); eng.search_async(); t2 = time.perf_counter()
                ts.append((round((t1 - t) * 1e3, 1), round((t2 - t1) * 1e3, 1))); o += sizes[k]
            t = time.perf_counter(); eng.wait(); tw = time.perf_counter() - t
            print(f"kernel {kernel} pipelined: (import, search_async) ms {ts}, wait {tw * 1e3:.1f} ms, total {(time.perf_counter() - t0) * 1e3:.1f} ms; search_ms {eng.stats()['search_ms']:.1f}")
        eng._lib.uvg_host_free_pinned(buf)
