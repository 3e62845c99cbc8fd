"""Development helper: host packing phases of one configs[1] batch at a given number of pack threads (PAIRHMM_CUDA_TRACE=1)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fgdp_b200
from fgdp_b200 import synth
b = synth.config2(int(sys.argv[1]) if len(sys.argv) > 1 else 10000)
for th in (4, 16):
    e = fgdp_b200.Engine(pack_threads=th)
    out = np.zeros(b.n_pairs); rs = np.zeros(b.n_pairs, np.uint8)
    for i in range(3):
        e.stats_reset()
        print("---- %d pack threads, call %d" % (th, i), file=sys.stderr, flush=True)
        t = time.perf_counter(); e.compute_batch(b, out, rs); dt = time.perf_counter() - t
        st = e.stats()
        print("threads %d: e2e %.1f ms -> %.0f GCUPS | gpu_total %.1f pack %.1f scatter %.1f host_total %.1f" % (
            th, dt * 1e3, b.cells() / dt / 1e9, st["gpu_ms_total"], st["host_ms_pack"], st["host_ms_scatter"], st["host_ms_total"]), flush=True)
    e.close()
