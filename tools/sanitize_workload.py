"""Small pass over every kernel of the library for compute-sanitizer (tools/sanitize.sh): packed / scalar / exact FP32
kernels, FP64 rescue (one and two reads per warp), prefix sharing, raw-read pre-pass, fused epilogue, both Smith-Waterman kernels."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import fgdp_b200
from fgdp_b200 import synth, smithwaterman as sw

outs = {}
for name, kw in (("packed", {}), ("scalar_ug", {"scalar_uniform": True}), ("general", {"force_general": True}),
                 ("exact", {"exact_fp32": True}), ("double", {"use_double": True}), ("w8", {"lane_group": 8}), ("w32", {"lane_group": 32}),
                 ("prefix", {"prefix_sharing": 2, "unit_haps": 32}), ("rescue_pairs", {"rescue_read_pairing": 2}),
                 ("rescue_pairs_1", {"rescue_read_pairing": 2, "rescue_pairs_per_unit": 1})):
    e = fgdp_b200.Engine(**kw)
    for b in (synth.config1("lowq"), synth.edge_cases(), synth.snv_family(3, 20, 9, 150, 203, indel_frac=0.2)):
        out, resc = e.compute_batch(b)
        outs.setdefault(b.name, []).append((name, out, resc))
    e.close()
for bname, rs in outs.items():
    base = [r for r in rs if r[0] == "packed"][0]
    for name, out, resc in rs:
        if name in ("double",):
            continue
        fin = np.isfinite(base[1])
        assert np.array_equal(np.isfinite(out), fin) and np.abs(out[fin] - base[1][fin]).max() < 1e-4, (bname, name)
    print(bname, "ok:", [r[0] for r in rs], flush=True)
e = fgdp_b200.Engine()
b = synth.config1("lowq")
e.compute_batch_normalized(b)                                                             # fused per-read epilogue
rq = np.random.default_rng(9)
e.compute_batch_raw(b, rq.integers(2, 42, b.read_bases.shape[0]).astype(np.uint8), rq.integers(0, 61, b.n_reads).astype(np.int32))   # quality pre-pass
e.close()
print("epilogue + raw reads ok", flush=True)
rng = np.random.default_rng(5)
acgt = np.frombuffer(b"ACGT", np.uint8)
refs = [bytes(rng.choice(acgt, int(n))) for n in rng.integers(30, 500, 300)]
alts = [bytes(rng.choice(acgt, int(n))) for n in rng.integers(10, 300, 300)]
for strat in (sw.SOFTCLIP, sw.INDEL, sw.LEADING_INDEL, sw.IGNORE):
    sw.align_batch(refs, alts, strategy=strat)
os.environ["SW_CUDA_KERNEL"] = "diag"
sw.align_batch(refs[:100], alts[:100])
print("smith-waterman ok", flush=True)
