"""Development helper: whole-path time of C3 / a region-sized batch against the underflowed pairs per FP64 unit (config.reserved[6])."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fgdp_b200
from fgdp_b200 import synth
for name, b in (("c3", synth.config3()), ("c3-500reads", synth.config3(500, 32)), ("c2-1region", synth.config2(1)), ("c2-40regions", synth.config2(40))):
    cells = b.cells()
    for mp in (0, 1, 4, 32):
        e = fgdp_b200.Engine(rescue_pairs_per_unit=mp)
        e.resident_load(b)
        best = min(e.resident_run() for _ in range(8))
        e.close()
        print("%s uniq_reads~%d pairs/unit=%2d: total %.3f ms (%.0f GCUPS), after-float %.3f ms" % (name, b.n_reads, mp, best[0], cells / best[0] / 1e6, best[0] - best[1]), flush=True)
