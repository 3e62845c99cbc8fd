"""Development helper: device-resident replay of a synthetic batch, prints kernel GCUPS (CUDA events)."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fgdp_b200
from fgdp_b200 import synth

which = sys.argv[1] if len(sys.argv) > 1 else "c2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 400
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
exact = "--exact" in sys.argv
lg = int(os.environ.get("LANE_GROUP", "0"))
t = time.time()
b = {"c1": lambda: synth.config1(), "c2": lambda: synth.config2(n), "c3": lambda: synth.config3(),
     "c4": lambda: synth.config4(), "c5": lambda: synth.config5(n)}[which]()
cells = b.cells()
print("gen %.1fs  %s pairs=%d cells=%.3g" % (time.time() - t, b.name, b.n_pairs, cells), flush=True)
e = fgdp_b200.Engine(exact_fp32=exact, use_double=bool(int(os.environ.get("USE_DOUBLE", "0"))), lane_group=lg, max_ctas_per_sm=int(os.environ.get("MAX_CTAS", "0")))
t = time.time(); e.resident_load(b); print("load %.2fs" % (time.time() - t), flush=True)
res = []
for i in range(reps):
    ms, msf = e.resident_run()
    res.append((ms, msf))
    print("run %d: total %.3f ms  float %.3f ms  -> %.1f GCUPS (float pass %.1f)" % (i, ms, msf, cells / ms / 1e6, cells / msf / 1e6), flush=True)
out, resc = e.resident_fetch()
ms, msf = min(res)
peak = 148 * 128 * 1.965e9
print(json.dumps({"workload": b.name, "cells": cells, "best_ms": ms, "gcups": cells / ms / 1e6, "float_gcups": cells / msf / 1e6,
                  "fma_frac_vs_1965MHz": cells / (msf * 1e-3) * 8 / peak, "rescued_frac": float(resc.mean())}))
t = time.time(); o2, r2 = e.compute_batch(b); dt = time.time() - t
print("e2e compute_batch %.3fs -> %.1f GCUPS; same=%s" % (dt, cells / dt / 1e9, np.array_equal(o2, out)))
print(e.stats())
