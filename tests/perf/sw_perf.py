"""Development helper: throughput of the batched CUDA Smith-Waterman against the reference's own aligner on the host
cores (oracle/_ref/libref_sw.so), on read-vs-haplotype pairs like the post-PairHMM realignment (150 bp x ~400 bp)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
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
# the C call on flat arrays (what a C++ caller pays): host arrays in, CIGAR text + offsets out, H2D / D2H inside
rl = np.array([len(x) for x in refs], np.int32); al = np.array([len(x) for x in alts], np.int32)
ro = np.concatenate([[0], np.cumsum(rl)[:-1]]).astype(np.int64); ao = np.concatenate([[0], np.cumsum(al)[:-1]]).astype(np.int64)
rb = np.frombuffer(b"".join(refs), np.uint8); ab = np.frombuffer(b"".join(alts), np.uint8)
for i in range(3):
    t = time.perf_counter(); cig, off, ms = sw.align_flat(rb, ro, rl, ab, ao, al); dt = time.perf_counter() - t
    print("C call, flat host arrays: %d pairs end to end %.1f ms = %.1f GCUPS (kernel %.2f ms)" % (n, dt * 1e3, cells / dt / 1e9, ms), flush=True)
assert np.array_equal(off, o)
for i in range(2):    # a caller that knows its reads are short passes a short CIGAR buffer: 64 B instead of 2 * max(len)
    t = time.perf_counter(); cig, off, ms = sw.align_flat(rb, ro, rl, ab, ao, al, cigar_stride=64); dt = time.perf_counter() - t
    print("C call, flat host arrays, cigar_stride 64: end to end %.1f ms = %.1f GCUPS (kernel %.2f ms)" % (dt * 1e3, cells / dt / 1e9, ms), flush=True)
assert np.array_equal(off, o) and [bytes(r).split(b"\0")[0].decode() for r in cig[:2000]] == c[:2000]
# per-region use: 200 alignments per call (one call per active region, after the PairHMM)
m = 200
calls = min(500, n // m)
sw.align_flat(rb, ro[:m], rl[:m], ab, ao[:m], al[:m], cigar_stride=64)
t = time.perf_counter()
for k in range(calls):
    q = slice(k * m, (k + 1) * m)
    sw.align_flat(rb, ro[q], rl[q], ab, ao[q], al[q], cigar_stride=64)
dt = time.perf_counter() - t
print("per-region calls: %d calls x %d pairs: %.3f ms per call (%.1f GCUPS)" % (calls, m, dt / calls * 1e3, float((rl[:calls * m].astype(np.int64) * al[:calls * m]).sum()) / dt / 1e9), flush=True)
ns = min(n, 20000)
p = oracle_py.SWPairs(refs[:ns], alts[:ns])
th = os.cpu_count() or 1
rc, ro, secs = oracle_py.reference_sw(p, threads=th)
print("reference on %d host threads: %d pairs in %.3f s = %.2f GCUPS; identical on the sample: %s" % (
    th, ns, secs, p.cells() / secs / 1e9, rc == c[:ns] and np.array_equal(ro, o[:ns])))
