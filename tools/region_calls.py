"""How Mutect2Cpp would drive the engine unchanged: T host threads (one PairHMM = one handle each, M2/main.cpp:199), every
thread calls pairhmm_cuda_compute ONCE PER REGION (VectorLoglessPairHMM::computeLog10Likelihoods, PHLCE.cpp:88-91), blocking.
Regions of BASELINE configs[1]; the pointer structs are built before the timed loop (a C++ caller has them for free), the
ctypes call releases the GIL. Prints aggregate GCUPS and the mean call latency for each T; `--submit` overlaps a thread's own
next region with the current one through pairhmm_cuda_submit / _wait."""
import sys, os, time, threading, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import numpy as np
import fgdp_b200
from fgdp_b200 import synth
from fgdp_b200.pairhmm import PhmmRead, PhmmHap

n_regions = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 1200
use_submit = "--submit" in sys.argv
b = synth.config2(n_regions)
lib = fgdp_b200.load_library()
regs = []
u8 = lambda a, o, n: a.ctypes.data + int(o)
for g in range(b.n_regions):
    r0, r1 = int(b.reg_read_start[g]), int(b.reg_read_start[g + 1])
    h0, h1 = int(b.reg_hap_start[g]), int(b.reg_hap_start[g + 1])
    ra = (PhmmRead * (r1 - r0))(); ha = (PhmmHap * (h1 - h0))()
    cells = 0
    for i, r in enumerate(range(r0, r1)):
        o, n = int(b.read_off[r]), int(b.read_len[r])
        ra[i] = PhmmRead(u8(b.read_bases, o, n), u8(b.base_q, o, n), u8(b.ins_q, o, n), u8(b.del_q, o, n), u8(b.gcp, o, n), n)
    for i, h in enumerate(range(h0, h1)):
        ha[i] = PhmmHap(u8(b.hap_bases, int(b.hap_off[h]), 0), int(b.hap_len[h]))
    cells = int(b.read_len[r0:r1].astype(np.int64).sum()) * int(b.hap_len[h0:h1].astype(np.int64).sum())
    regs.append((ra, r1 - r0, ha, h1 - h0, np.zeros((r1 - r0) * (h1 - h0)), cells))
total_cells = sum(r[5] for r in regs)
ref_out, _ = fgdp_b200.Engine().compute_batch(b)
lib.pairhmm_cuda_compute.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
lib.pairhmm_cuda_submit.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]
lib.pairhmm_cuda_wait.argtypes = [C.c_void_p, C.c_int64]


def worker(eng, mine, errs):
    h = eng.h
    if not use_submit:
        for ra, nr, ha, nh, out, _ in mine:
            st = lib.pairhmm_cuda_compute(h, ra, nr, ha, nh, out.ctypes.data, None)
            if st: errs.append(st)
    else:
        prev = None
        for ra, nr, ha, nh, out, _ in mine:
            t = C.c_int64()
            st = lib.pairhmm_cuda_submit(h, ra, nr, ha, nh, out.ctypes.data, None, C.byref(t))
            if st: errs.append(st)
            if prev is not None and lib.pairhmm_cuda_wait(h, prev): errs.append(-99)
            prev = t.value
        if prev is not None and lib.pairhmm_cuda_wait(h, prev): errs.append(-99)


for T in (1, 2, 4, 8, 16, 32):
    engs = [fgdp_b200.Engine(pack_threads=1) for _ in range(T)]
    for e in engs: e.initialize(150, 512)
    errs = []
    for rep in range(2):                                       # first pass warms the handles' buffers
        for r in regs: r[4][:] = 0
        ths = [threading.Thread(target=worker, args=(engs[t], regs[t::T], errs)) for t in range(T)]
        t0 = time.perf_counter()
        for th in ths: th.start()
        for th in ths: th.join()
        dt = time.perf_counter() - t0
    got = np.concatenate([r[4] for r in regs])
    print(json.dumps({"threads": T, "mode": "submit/wait" if use_submit else "compute", "regions": len(regs), "seconds": round(dt, 4),
                      "gcups": round(total_cells / dt / 1e9, 1), "mean_call_ms": round(dt * T / len(regs) * 1e3, 3),
                      "errors": len(errs), "identical_to_batch_call": bool(np.array_equal(got, ref_out))}), flush=True)
    for e in engs: e.close()
