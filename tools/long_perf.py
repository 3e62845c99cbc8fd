"""Development helper: throughput of the row-striped kernel for reads above 384 bases (csrc/long.cu) on a batch of long
reads only, device-resident, next to the same cells as 150-base reads through the packed kernels."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fgdp_b200
from fgdp_b200 import synth
R = int(sys.argv[1]) if len(sys.argv) > 1 else 600
for name, b in (("reads of %d bases x 8 haplotypes of ~%d" % (R, 2 * R), synth.make_batch(0xB200 + 41, [64] * 40, [np.full(8, 2 * R)] * 40, R, "clean", foreign_frac=0.05)),
                ("reads of 150 bases, same haplotypes", synth.make_batch(0xB200 + 41, [64 * R // 150] * 40, [np.full(8, 2 * R)] * 40, 150, "clean", foreign_frac=0.05))):
    e = fgdp_b200.Engine()
    e.resident_load(b)
    best = min(e.resident_run() for _ in range(5))
    out, resc = e.resident_fetch()
    e.close()
    print(json.dumps({"workload": name, "pairs": b.n_pairs, "cells": b.cells(), "ms": round(best[0], 3), "gcups": round(b.cells() / best[0] / 1e6, 1),
                      "rescued_frac": round(float(resc.mean()), 4)}), flush=True)
