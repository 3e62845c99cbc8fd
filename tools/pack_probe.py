"""Development helper: host packing phases (PAIRHMM_CUDA_TRACE=1) of one C2 batch at several pack-thread counts."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fgdp_b200
from fgdp_b200 import synth
b = synth.config2(int(sys.argv[1]) if len(sys.argv) > 1 else 3000)
for th in [int(x) for x in (sys.argv[2:] or ["4", "16"])]:
    e = fgdp_b200.Engine(pack_threads=th)
    out = np.zeros(b.n_pairs); rs = np.zeros(b.n_pairs, np.uint8)
    for i in range(2):
        e.stats_reset(); t = time.perf_counter(); e.compute_batch(b, out, rs); dt = time.perf_counter() - t
        st = e.stats()
        print("e2e threads=%d %.1f ms (%.0f GCUPS) | pack %.1f scatter %.1f gpu %.1f" % (
            th, dt * 1e3, b.cells() / dt / 1e9, st["host_ms_pack"], st["host_ms_scatter"], st["gpu_ms_total"]), file=sys.stderr, flush=True)
    e.close()
