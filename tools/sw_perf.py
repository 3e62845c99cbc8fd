"""Development helper: throughput of the batched CUDA Smith-Waterman against the reference's own aligner on the host
cores (oracle/_ref/libref_sw.so), on read-vs-haplotype pairs like the post-PairHMM realignment (150 bp x ~400 bp)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import numpy as np
import oracle_py
from fgdp_b200 import smithwaterman as sw
n = int(sys.argv[1]) if len(sys.argv) > 1 else 200000
rng = np.random.default_rng(5)
acgt = np.frombuffer(b"ACGT", np.uint8)
haps = [rng.choice(acgt, int(h)) for h in rng.integers(300, 501, 2000)]
refs, alts = [], []
for k in range(n):
    hap = haps[k % len(haps)]
    st = int(rng.integers(0, len(hap) - 150))
    read = hap[st:st + 150].copy()
    m = rng.random(150) < 0.01
    read[m] = rng.choice(acgt, int(m.sum()))
    if k % 50 == 0:
        read = np.concatenate([read[:70], read[75:], rng.choice(acgt, 5)])    # a 5-base deletion
    refs.append(bytes(hap)); alts.append(bytes(read))
cells = sum(len(r) * len(a) for r, a in zip(refs, alts))
for i in range(3):
    t = time.perf_counter(); c, o, ms = sw.align_batch(refs, alts); dt = time.perf_counter() - t
    print("GPU: %d pairs, %.3g cells: kernel %.2f ms (%.1f GCUPS), call %.1f ms (incl. python list packing)" % (
        n, cells, ms, cells / ms / 1e6, dt * 1e3), flush=True)
ns = min(n, 20000)
p = oracle_py.SWPairs(refs[:ns], alts[:ns])
th = os.cpu_count() or 1
rc, ro, secs = oracle_py.reference_sw(p, threads=th)
print("reference on %d host threads: %d pairs in %.3f s = %.2f GCUPS; identical on the sample: %s" % (
    th, ns, secs, p.cells() / secs / 1e9, rc == c[:ns] and np.array_equal(ro, o[:ns])))
