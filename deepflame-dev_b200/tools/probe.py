"""Quick device probe: FP64 peak, and host-API solve timing on a synthetic batch.
usage: python deepflame-dev_b200/tools/probe.py [mech] [cells] [rtol] [atol]"""
import ctypes
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "deepflame-dev_b200"))
import dfchem_host  # noqa: E402
from dfchem_host import synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "gri30"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
rtol = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-6
atol = float(sys.argv[4]) if len(sys.argv) > 4 else 1e-10
lib = dfchem_host.load_library()
pk = ctypes.c_double()
print("fp64 peak rc", lib.dfchem_bench_fp64_peak(0, ctypes.byref(pk)), "TFLOP/s", pk.value, flush=True)
h = dfchem_host.Handle(synth.load_blob(name), 0)
h.set_tolerances(rtol, atol)
b = synth.make_batch(name, h.species, h.W, n)
cost = np.zeros(n)
status = np.zeros(n, dtype=np.int32)
for rep in range(3):
    t0 = time.time()
    RR, Q = h.solve(b["T"], b["p"], b["rho"], 1e-6, b["Y"], cost=cost, status=status)
    dtm = time.time() - t0
    c = h.counters()
    print(f"{name} n={n} rtol={rtol:g} atol={atol:g}: {dtm*1e3:.1f} ms -> {n/dtm:.3e} cells/s; "
          f"steps acc {c['accepted']} rej {c['rejected']} failed {c['failed']}; cost mean {cost.mean():.2f} "
          f"max {cost.max():.0f}; {(c['accepted']+c['rejected'])/dtm:.3e} step attempts/s", flush=True)
print("cost histogram", np.histogram(cost, bins=[0, 1.5, 2.5, 4.5, 8.5, 16.5, 32.5, 64.5, 128.5, 1e9])[0])
print("finite", np.isfinite(RR).all(), np.isfinite(Q).all())
