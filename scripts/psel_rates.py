#!/usr/bin/env python3
"""Hand-over rate, passes and list sizes of the selection engine's planning on synthetic stacks, through the host
replay (tests/helpers/libpsel_host.so, built by tests/test_psel_host.py).  Usage: psel_rates.py [K]"""
import ctypes as C, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_sorted_host import make_stack
from oracle import oracle as O
L = C.CDLL(os.path.join(ROOT, "tests", "helpers", "libpsel_host.so"))
L.psel_host_pixel.restype = C.c_int
L.psel_host_pixel.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_void_p,
                              C.POINTER(C.c_int), C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_int]
K = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
SIG = {O.SIGMA: (3, 3), O.PERCENTILE: (0.2, 0.3), O.WINSORIZED: (3, 3)}
NAME = {O.SIGMA: "sigma", O.PERCENTILE: "percentile", O.WINSORIZED: "winsorized"}

def benchlike(rng, n):
    st = (0.1 + 0.01 * rng.standard_normal(n)).astype(np.float32)
    m = rng.random(n) < 1e-3
    st[m] += rng.uniform(0.2, 0.6, m.sum()).astype(np.float32)
    return st

rng = np.random.default_rng(1)
for algo, n, T in ((O.SIGMA, 300, 4), (O.SIGMA, 100, 2), (O.SIGMA, 64, 1), (O.PERCENTILE, 300, 4)):
    for kind in ("clean", "hot", "benchlike"):
        ho = 0; passes = []; ms = []
        for _ in range(K):
            st = benchlike(rng, n) if kind == "benchlike" else make_stack(rng, n, np.float32, kind)
            rej = (C.c_int * 2)(); res = C.c_double(); diag = (C.c_int * 6)()
            s = L.psel_host_pixel(1, st.ctypes.data_as(C.c_void_p), n, T, 1, algo, SIG[algo][0], SIG[algo][1], None, rej, C.byref(res), diag, 0)
            ho += s != 0; passes.append(diag[1]); ms.append(diag[2] + diag[3] + diag[4])
        ms = np.array(ms)
        print(f"{NAME[algo]:10s} n={n:3d} {kind:9s} handover {ho / K:.4f} passes {np.mean(passes):.3f} m {ms.mean():.1f} +- {ms.std():.1f}  P(m>48) {np.mean(ms > 48):.4f} P(m>64) {np.mean(ms > 64):.4f}")
