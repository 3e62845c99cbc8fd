"""Development helper: the packed FP32 pass with and without prefix sharing on device-resident
batches. Reports for each workload the GCUPS of DP cells actually swept (the honest kernel rate, what the roofline
fraction is computed from) and the EFFECTIVE GCUPS = algorithmic cells R x H of every pair / time."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fgdp_b200
from fgdp_b200 import synth

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
work = [("snv-family 600 regions x 200 reads x 12 haps, 150 x 300, SNV-only variants", lambda: synth.snv_family(600, 200, 12, 150, 300, dup_frac=0.0)),
        ("snv-family with 30 % indel haplotypes (other lengths never share)", lambda: synth.snv_family(600, 200, 12, 150, 300, dup_frac=0.0, indel_frac=0.3)),
        ("BASELINE configs[1] (600 regions; every haplotype has an indel, lengths differ by up to 10)", lambda: synth.config2(600))]
for name, make in work:
    b = make()
    row = {"workload": name, "pairs": b.n_pairs}
    outs = []
    for pfx in (False, True):
        e = fgdp_b200.Engine(prefix_sharing=pfx, dedup_reads=False)
        e.resident_load(b)
        best = min(e.resident_run() for _ in range(reps))
        outs.append(e.resident_fetch())
        st = e.stats()
        e.close()
        cells, swept = st["cells"] // reps, st["cells_swept"] // reps
        k = "shared" if pfx else "per_pair"
        row[k] = {"float_ms": round(best[1], 3), "swept_frac": round(swept / cells, 4),
                  "kernel_gcups_swept_cells": round(swept / best[1] / 1e6, 1), "effective_gcups": round(cells / best[1] / 1e6, 1)}
    row["identical_bits"] = bool(np.array_equal(outs[0][0].view(np.uint64), outs[1][0].view(np.uint64)) and np.array_equal(outs[0][1], outs[1][1]))
    print(json.dumps(row), flush=True)
