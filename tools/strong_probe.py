"""Development helper: BASELINE configs[4] through ONE multi-device handle (config.device_mask over all visible GPUs), e2e only.
usage: strong_probe.py [regions=25000] [reps=3]   (PAIRHMM_CUDA_TRACE=1 prints every device's chunk timeline to stderr)"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import fgdp_b200
from fgdp_b200 import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 25000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
nd = fgdp_b200.device_count()
job = synth.generate_parallel("c5", n)
cells = job.cells()
out = np.zeros(job.n_pairs); rs = np.zeros(job.n_pairs, np.uint8)
for mask in ([1, (1 << nd) - 1] if nd > 1 else [1]):
    e = fgdp_b200.Engine(device_mask=mask)
    e.compute_batch(job, out, rs)
    for r in range(reps):
        e.stats_reset()
        t0 = time.perf_counter(); e.compute_batch(job, out, rs); dt = time.perf_counter() - t0
        st = e.stats()
        print("mask %#x rep %d: e2e %.1f ms = %.0f GCUPS | gpu max %.1f ms | pack max %.1f | scatter max %.1f | host_total %.1f" %
              (mask, r, dt * 1e3, cells / dt / 1e9, st["gpu_ms_total"], st["host_ms_pack"], st["host_ms_scatter"], st["host_ms_total"]), flush=True)
    e.close()
