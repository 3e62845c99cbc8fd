#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the REFERENCE's own code (oracle/_ref/libref_pairhmm.so, built by
oracle/Makefile from the unmodified sources under /root/reference). Run in the container that has
/root/reference:   python tests/golden/make_golden.py

Each fixture stores the input bytes and the reference's outputs (log10 likelihoods as float64, rescue flags), plus
the build facts (ISA path the reference's CPUID test selected, compiler, flags). The reference ships no golden
vectors of its own for this path, so these are what pins the oracle and the CUDA engine.
"""
import os
import subprocess
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle_py            # noqa: E402
import fgdp_b200            # noqa: E402
from fgdp_b200 import synth  # noqa: E402

FIELDS = ("read_bases", "base_q", "ins_q", "del_q", "gcp", "read_off", "read_len", "hap_bases", "hap_off", "hap_len",
          "reg_read_start", "reg_hap_start")


def sw_pairs(seed, n):
    """(haplotype-like ref, read-like alt) pairs: reads cut from the ref with substitutions and indels, some unrelated,
    some longer than the ref, some single bases"""
    rng = np.random.default_rng(seed)
    acgt = np.frombuffer(b"ACGT", np.uint8)
    refs, alts = [], []
    for k in range(n):
        H, R = int(rng.integers(1, 520)), int(rng.integers(1, 260))
        hap = rng.choice(acgt, H)
        if k % 6 == 0:
            read = rng.choice(acgt, R)
        else:
            st = int(rng.integers(0, max(1, H - R)))
            src, out = hap[st:st + R], []
            for c in src:
                r = rng.random()
                if r < 0.03:
                    out.append(int(rng.choice(acgt)))
                elif r < 0.04:
                    continue
                elif r < 0.05:
                    out.append(int(c)); out.extend(int(x) for x in rng.choice(acgt, int(rng.integers(1, 6))))
                else:
                    out.append(int(c))
            read = np.array(out if out else [65], np.uint8)
        refs.append(bytes(hap)); alts.append(bytes(read))
    return refs, alts


def main():
    ref = oracle_py.load_reference()
    gxx = subprocess.run(["g++", "--version"], capture_output=True, text=True).stdout.splitlines()[0]
    meta = "reference ISA path: AVX-%d; %s; -O3, default -ffp-contract (see oracle/Makefile)" % (ref.isa(), gxx)
    cases = {
        "c1_clean": synth.config1("clean"),
        "c1_lowq": synth.config1("lowq"),
        "edge": synth.edge_cases(),
        "c3_small": synth.config3(48, 6),
        "c4_small": synth.config4(2, 6, 3),
        "c2_small": synth.config2(3),
        "long_reads": synth.long_reads(),                    # reads of 385..2049 bases (row stripes in the reference too)
        "long_reads_gcp": synth.long_reads(uniform_gcp=False, profile="lowq"),
    }
    only = [a for a in sys.argv[1:] if a in cases]
    if only:                                                 # python make_golden.py long_reads ...: just these fixtures
        for name in only:
            b = cases[name]
            out, resc, _ = ref.compute(b.as_flat(), threads=8)
            out_d, _, _ = ref.compute(b.as_flat(), use_double=True, threads=8)
            np.savez_compressed(os.path.join(HERE, name + ".npz"), log10=out, rescued=resc, log10_use_double=out_d,
                                meta=np.array(meta), **{f: getattr(b, f) for f in FIELDS})
            print(name, b.n_pairs, "pairs, rescued", int(resc.sum()))
        return
    for name, b in cases.items():
        out, resc, _ = ref.compute(b.as_flat(), threads=8)
        out_d, _, _ = ref.compute(b.as_flat(), use_double=True, threads=8)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), log10=out, rescued=resc, log10_use_double=out_d,
                            meta=np.array(meta), **{f: getattr(b, f) for f in FIELDS})
        print(name, b.n_pairs, "pairs, rescued", int(resc.sum()))
    # per-read epilogue: the reference's AlleleLikelihoods code on the reference's own matrices and
    # on synthetic rows with ties, -inf and single-haplotype regions
    rep = oracle_py.ReferenceEpilogue()
    rng = np.random.default_rng(0xB200)
    ep = {}
    k = 0
    for name in ("c1_clean", "c1_lowq", "c3_small"):
        b = cases[name]
        out, _, _ = ref.compute(b.as_flat(), threads=8)
        for r in range(b.n_regions):
            r0, r1 = int(b.reg_read_start[r]), int(b.reg_read_start[r + 1])
            nh = int(b.reg_hap_start[r + 1] - b.reg_hap_start[r])
            o0 = int(b.region_out_offsets()[r])
            m = out[o0:o0 + (r1 - r0) * nh].reshape(r1 - r0, nh)
            for cap, rate, nl in ((-4.5, 0.02, 0), (-4.5, 0.02, 1), (-np.inf, 0.02, 0), (-1.0, 0.0, 0)):
                mo, removed = rep.epilogue(m, b.read_len[r0:r1], cap, rate, nl)
                ep["m%d" % k] = m; ep["len%d" % k] = b.read_len[r0:r1]; ep["par%d" % k] = np.array([cap, rate, nl])
                ep["out%d" % k] = mo; ep["rem%d" % k] = removed
                k += 1
    for t in range(40):
        n, nh = int(rng.integers(1, 30)), int(rng.integers(1, 12))
        m = -rng.random((n, nh)) * float(rng.choice([1, 10, 100]))
        if t % 3 == 0:
            m = np.round(m)
        if t % 7 == 0:
            m[int(rng.integers(0, n))] = -np.inf
        rl = rng.integers(1, 300, n).astype(np.int32)
        cap, rate, nl = float(rng.choice([-4.5, -1.0, -np.inf])), float(rng.choice([0.02, 0.0, 0.01])), t % 2
        mo, removed = rep.epilogue(m, rl, cap, rate, nl)
        ep["m%d" % k] = m; ep["len%d" % k] = rl; ep["par%d" % k] = np.array([cap, rate, nl])
        ep["out%d" % k] = mo; ep["rem%d" % k] = removed
        k += 1
    np.savez_compressed(os.path.join(HERE, "epilogue.npz"), n_cases=np.array(k), meta=np.array(
        "reference AlleleLikelihoods.h via oracle/ref_epilogue_driver.cpp; " + gxx), **ep)
    print("epilogue", k, "cases")
    # Smith-Waterman: the reference's own SmithWaterman_align (oracle/_ref/libref_sw.so) on
    # read-vs-haplotype-like pairs, every overhang strategy, the three parameter sets Mutect2Cpp uses
    refs, alts = sw_pairs(0xB200 + 33, 160)
    sp = oracle_py.SWPairs(refs, alts)
    sw = {"ref_bases": sp.ref_bases, "ref_off": sp.ref_off, "ref_len": sp.ref_len, "alt_bases": sp.alt_bases,
          "alt_off": sp.alt_off, "alt_len": sp.alt_len}
    k = 0
    for params in ((10, -15, -30, -5), (200, -150, -260, -11), (25, -50, -110, -6)):
        for strat in (9, 10, 11, 12):
            cig, off, _ = oracle_py.reference_sw(sp, params, strat, threads=4)
            sw["par%d" % k] = np.array(list(params) + [strat]); sw["cig%d" % k] = np.array(cig); sw["off%d" % k] = off
            k += 1
    np.savez_compressed(os.path.join(HERE, "sw.npz"), n_cases=np.array(k), meta=np.array(
        "reference SmithWaterman_align (M2/intel/smithwaterman) via oracle/ref_sw_driver.cpp; " + gxx), **sw)
    print("sw", k, "cases x", sp.n, "pairs")
    pf, mf, pd, md = ref.tables()
    np.savez_compressed(os.path.join(HERE, "tables.npz"), ph2pr_f=pf, mm_f=mf[:8256], ph2pr_d=pd, mm_d=md[:8256], meta=np.array(meta))


if __name__ == "__main__":
    main()
