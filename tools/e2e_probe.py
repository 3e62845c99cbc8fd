"""Development helper: end-to-end (host arrays -> host results) timing of one C2 batch, with the engine's chunk trace."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fgdp_b200
from fgdp_b200 import synth
b = synth.config2(int(sys.argv[1]) if len(sys.argv) > 1 else 10000)
e = fgdp_b200.Engine(chunk_cells=int(float(sys.argv[2])) if len(sys.argv) > 2 else 0, pack_threads=int(sys.argv[3]) if len(sys.argv) > 3 else 0)
out = np.zeros(b.n_pairs); rs = np.zeros(b.n_pairs, np.uint8)
for i in range(4):
    e.stats_reset()
    print("---- call", i, file=sys.stderr, flush=True)
    t = time.perf_counter(); e.compute_batch(b, out, rs); dt = time.perf_counter() - t
    st = e.stats()
    print("e2e %.1f ms -> %.0f GCUPS | gpu_total %.1f pack %.1f scatter %.1f host_total %.1f launches %d" % (
        dt*1e3, b.cells()/dt/1e9, st["gpu_ms_total"], st["host_ms_pack"], st["host_ms_scatter"], st["host_ms_total"], st["kernel_launches"]), flush=True)
