"""Development helper: where the time of ONE blocking per-region call goes (engine counters, 300 regions of configs[1])."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import numpy as np
import fgdp_b200
from fgdp_b200 import synth
from fgdp_b200.pairhmm import PhmmRead, PhmmHap
b = synth.config2(300)
lib = fgdp_b200.load_library()
lib.pairhmm_cuda_compute.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
u8 = lambda a, o: a.ctypes.data + int(o)
regs = []
for g in range(b.n_regions):
    r0, r1 = int(b.reg_read_start[g]), int(b.reg_read_start[g + 1]); h0, h1 = int(b.reg_hap_start[g]), int(b.reg_hap_start[g + 1])
    ra = (PhmmRead * (r1 - r0))(); ha = (PhmmHap * (h1 - h0))()
    for i, r in enumerate(range(r0, r1)):
        o, n = int(b.read_off[r]), int(b.read_len[r])
        ra[i] = PhmmRead(u8(b.read_bases, o), u8(b.base_q, o), u8(b.ins_q, o), u8(b.del_q, o), u8(b.gcp, o), n)
    for i, h in enumerate(range(h0, h1)):
        ha[i] = PhmmHap(u8(b.hap_bases, int(b.hap_off[h])), int(b.hap_len[h]))
    regs.append((ra, r1 - r0, ha, h1 - h0, np.zeros((r1 - r0) * (h1 - h0))))
e = fgdp_b200.Engine(pack_threads=1)
e.initialize(150, 512)
for rep in range(2):
    e.stats_reset()
    t0 = time.perf_counter()
    for ra, nr, ha, nh, out in regs:
        lib.pairhmm_cuda_compute(e.h, ra, nr, ha, nh, out.ctypes.data, None)
    dt = time.perf_counter() - t0
st = e.stats()
n = len(regs)
print("per call: wall %.1f us | host_total %.1f | pack %.1f | scatter %.1f | gpu kernels (events) %.1f (float %.1f, after-float %.1f) | launches %.1f | cells %.3g"
      % (dt / n * 1e6, st["host_ms_total"] / n * 1e3, st["host_ms_pack"] / n * 1e3, st["host_ms_scatter"] / n * 1e3,
         st["gpu_ms_total"] / n * 1e3, st["gpu_ms_float"] / n * 1e3, st["gpu_ms_double"] / n * 1e3, st["kernel_launches"] / n, st["cells"] / n))
