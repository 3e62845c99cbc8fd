"""Development helper: end-to-end time of the raw-read entry point (quality pre-pass on the device, 2 bytes per base over
PCIe) against the plain one (5 bytes per base, qualities prepared on the host beforehand), same reads."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "oracle"))
import numpy as np
import fgdp_b200
from fgdp_b200 import synth
b = synth.config2(int(sys.argv[1]) if len(sys.argv) > 1 else 2000)
rng = np.random.default_rng(1)
mapq = np.full(b.n_reads, 60, np.int32)
e = fgdp_b200.Engine()
for i in range(3):
    e.stats_reset(); t = time.perf_counter(); plain, _ = e.compute_batch(b); t_plain = time.perf_counter() - t; s0 = e.stats()
    e.stats_reset(); t = time.perf_counter(); raw, _ = e.compute_batch_raw(b, b.base_q, mapq); t_raw = time.perf_counter() - t; s1 = e.stats()
    print("plain %.1f ms (%.0f GCUPS, H2D %.0f MB, pack %.0f ms) | raw %.1f ms (%.0f GCUPS, H2D %.0f MB, pack %.0f ms)" % (
        t_plain * 1e3, b.cells() / t_plain / 1e9, s0["h2d_bytes"] / 1e6, s0["host_ms_pack"],
        t_raw * 1e3, b.cells() / t_raw / 1e9, s1["h2d_bytes"] / 1e6, s1["host_ms_pack"]), flush=True)
# host cost of the same pre-pass with the C oracle on one thread, on a sample
import oracle_py
o = oracle_py.load_oracle()
n = min(b.n_reads, 20000)
t = time.perf_counter()
o.prepare_reads(b.read_bases, b.base_q, mapq[:n], b.read_off[:n], b.read_len[:n])
dt = time.perf_counter() - t
print("host pre-pass (C oracle, 1 thread, incl. python loop): %.1f us per read -> %.0f ms for the batch's %d reads" % (
    dt / n * 1e6, dt / n * b.n_reads * 1e3, b.n_reads))
