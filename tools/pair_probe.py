"""Development helper: FP64 pass with two reads per warp (config.reserved[7]) against one read per warp, device-resident replay.
usage: pair_probe.py [c3|c2|c4] [regions]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fgdp_b200
from fgdp_b200 import synth
which = sys.argv[1] if len(sys.argv) > 1 else "c3"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 2000
b = {"c3": lambda: synth.config3(), "c2": lambda: synth.config2(n), "c4": lambda: synth.config4(), "c5": lambda: synth.config5(n),
     "c3x": lambda: synth.concat([synth.config3(2000, 32)] * 1 + [synth.make_batch(7 + k, [2000], [[600] * 32], 150, profile="lowq") for k in range(n)])}[which]()
ref = None
for mode, per in ((1, 0), (0, 0), (2, 1), (2, 2), (2, 3), (2, 4), (2, 5), (2, 6), (2, 8), (2, 32), (1, 0)):
    e = fgdp_b200.Engine(rescue_read_pairing=mode, rescue_pairs_per_unit=per)
    e.resident_load(b)
    r = [e.resident_run() for _ in range(10)]
    out, resc = e.resident_fetch()
    if ref is None:
        ref = (out, resc)
    same = np.array_equal(out, ref[0]) and np.array_equal(resc, ref[1])
    ms = min(x[0] for x in r); msf = min(x[1] for x in r)
    print("%s pairing=%d per=%2d: %.3f ms (float %.3f, rest %.3f) -> %.0f GCUPS  launches/run %d  same=%s" %
          (b.name, mode, per, ms, msf, ms - msf, b.cells() / ms / 1e6, e.stats()["kernel_launches"] // 10, same), flush=True)
    e.close()
