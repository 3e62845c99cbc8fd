"""Development helper: FP32-pass GCUPS of C3 / C4 against the haplotypes-per-unit knob (config.reserved[5])."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fgdp_b200
from fgdp_b200 import synth
for name, b in (("c3", synth.config3()), ("c4", synth.config4())):
    cells = b.cells()
    for uh in (0, 1, 2, 3, 4, 6, 8, 16):
        for lg in (0, 32):
            e = fgdp_b200.Engine(unit_haps=uh, lane_group=lg)
            e.resident_load(b)
            best = min(e.resident_run() for _ in range(6))
            e.close()
            print("%s unit_haps=%2d lane_group=%2d: total %.3f ms (%.0f GCUPS), float %.3f ms (%.0f GCUPS)" % (
                name, uh, lg, best[0], cells / best[0] / 1e6, best[1], cells / best[1] / 1e6), flush=True)
