"""Development helper: end-to-end cost of the fused per-read epilogue (device) against the same epilogue done by the
oracle on the host after a plain call."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "oracle"))
import numpy as np
import fgdp_b200
from fgdp_b200 import synth
b = synth.config2(int(sys.argv[1]) if len(sys.argv) > 1 else 2000)
e = fgdp_b200.Engine()
for i in range(3):
    t = time.perf_counter(); raw, _ = e.compute_batch(b); t_plain = time.perf_counter() - t
    t = time.perf_counter(); got, _, summ = e.compute_batch_normalized(b); t_norm = time.perf_counter() - t
    print("plain %.1f ms (%.0f GCUPS) | fused epilogue %.1f ms (%.0f GCUPS) | poorly modelled %.2f%%" % (
        t_plain * 1e3, b.cells() / t_plain / 1e9, t_norm * 1e3, b.cells() / t_norm / 1e9, 100 * summ["poorly_modeled"].mean()), flush=True)
import oracle_py
o = oracle_py.load_oracle()
t = time.perf_counter()
offs = b.region_out_offsets()
for g in range(b.n_regions):
    r0, r1 = int(b.reg_read_start[g]), int(b.reg_read_start[g + 1]); nh = int(b.reg_hap_start[g + 1] - b.reg_hap_start[g])
    o.epilogue(raw[int(offs[g]):int(offs[g]) + (r1 - r0) * nh].reshape(r1 - r0, nh), b.read_len[r0:r1])
print("host epilogue (C oracle, 1 thread, incl. python loop) %.1f ms" % ((time.perf_counter() - t) * 1e3))
