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
sys.path.insert(0, os.path.join(ROOT, "tests"))
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


def m2_region_cases():
    """active regions with RAW reads (as the BAM holds them) for the reference's whole caller path
    (PairHMMLikelihoodCalculationEngine::computeReadLikelihoods): name -> dict of inputs for ReferenceM2.region"""
    from util import raw_quals
    cases = {}

    def case(name, b, seed, with_indel=False, n_samples=1, dup=0, mapq_lo=0, synth_q=True, **kw):
        assert b.n_regions == 1
        bq, mapq, iq, dq = raw_quals(b, seed, with_indel)
        if synth_q:       # the generator's own base qualities as the BAM's (mostly high: some reads survive filterPoorlyModeledEvidence)
            bq = b.base_q.copy()
        rng = np.random.default_rng(seed + 1)
        mapq = np.maximum(mapq, mapq_lo).astype(np.int32)
        rb, ro, rl = b.read_bases.copy(), b.read_off.copy(), b.read_len.copy()
        bqs = bq.copy()
        if dup:    # exact duplicates of earlier reads (bases, qualities, mapping quality): the de-dup path of VLPH.cpp:83-103
            for k in range(dup):
                src, dst = 2 * k, 2 * k + 1
                if rl[src] != rl[dst]:
                    continue
                rb[ro[dst]:ro[dst] + rl[dst]] = rb[ro[src]:ro[src] + rl[src]]
                bqs[ro[dst]:ro[dst] + rl[dst]] = bqs[ro[src]:ro[src] + rl[src]]
                mapq[dst] = mapq[src]
                if iq is not None:
                    iq[ro[dst]:ro[dst] + rl[dst]] = iq[ro[src]:ro[src] + rl[src]]
                    dq[ro[dst]:ro[dst] + rl[dst]] = dq[ro[src]:ro[src] + rl[src]]
        has = None
        if with_indel == "mixed":
            has = (np.arange(b.n_reads) % 2).astype(np.uint8)
        sample = rng.integers(0, n_samples, b.n_reads).astype(np.int32)
        d = dict(read_bases=rb, base_q=bqs, mapq=mapq, read_off=ro, read_len=rl, hap_bases=b.hap_bases, hap_off=b.hap_off,
                 hap_len=b.hap_len, read_sample=sample, n_samples=n_samples)
        if iq is not None:
            d.update(ins_q=iq, del_q=dq)
        if has is not None:
            d.update(has_indel_q=has)
        d.update(kw)
        cases[name] = d

    case("c1", synth.config1("clean"), 1, mapq_lo=40)
    case("c1_lowmapq", synth.config1("clean"), 2, synth_q=False)                      # mapping quality 0..60 caps the base qualities
    case("c1_two_samples_dups", synth.config1("clean"), 3, n_samples=2, dup=12, mapq_lo=30)
    case("c1_indel_q", synth.config1("clean"), 4, with_indel=True, mapq_lo=30)
    case("c1_indel_q_mixed", synth.config1("clean"), 5, with_indel="mixed", n_samples=3, mapq_lo=20)
    case("c1_aggressive_pcr", synth.config1("clean"), 6, pcr_rate_factor=2, base_q_threshold=20, mapq_lo=30)
    # use_double on a region the reference does NOT send down its trie path (6 haplotypes of >= 2 lengths): with use_double
    # the reference's trie dispatch writes nothing at all (IntelPairHmm.cc:698-699, a bug not to be replicated)
    case("c3_use_double", synth.config3(40, 6), 7, use_double=True, mapq_lo=30)
    case("c3_small", synth.config3(48, 6), 8, n_samples=2, dup=6, mapq_lo=10)   # rescue-heavy
    e = synth.edge_cases()
    for r in range(e.n_regions):
        one = e.subset([r])
        if one.n_reads and one.n_haps == len(set(bytes(one.hap_bases[int(o):int(o) + int(n)]) for o, n in zip(one.hap_off, one.hap_len))) \
                and (one.hap_len > 0).all() and (one.read_len > 0).all():
            case("edge%d" % r, one, 20 + r, n_samples=1 + r % 2, mapq_lo=15)
    return cases


def main():
    ref = oracle_py.load_reference()
    gxx = subprocess.run(["g++", "--version"], capture_output=True, text=True).stdout.splitlines()[0]
    if not sys.argv[1:] or "prepass" in sys.argv[1:] or "m2_region" in sys.argv[1:]:
        # the caller side, from the reference's own PairHMMLikelihoodCalculationEngine.cpp & co (oracle/_ref/libref_m2.so)
        from util import raw_quals, repeat_rich_batch
        m2 = oracle_py.ReferenceM2()
        meta2 = "reference PairHMMLikelihoodCalculationEngine.cpp / GATKVariantContextUtils.cpp / VectorLoglessPairHMM.cpp ... via oracle/ref_m2_driver.cpp; " + gxx
        # N1: modifyReadQualities + buildGapContinuationPenalties. Repeat-rich reads with N of 1..600 bases, WES-like reads;
        # no / all / every-other read with BI + BD arrays; MAPQ 0..60; the three PCR models; thresholds 18 / 6 / 25; gcp 10 / 12
        pp, k = {}, 0
        for bi, b in enumerate((repeat_rich_batch(max_len=600), synth.config2(1))):
            for with_indel in (False, True, "mixed"):
                bq, mapq, iq, dq = raw_quals(b, 100 + bi, with_indel)
                has = (np.arange(b.n_reads) % 2).astype(np.uint8) if with_indel == "mixed" else None
                tag = "in%d" % (k // 1000)
                for pcr, thr, gcp in ((3, 18, 10), (2, 18, 10), (1, 18, 10), (3, 6, 12), (2, 25, 10))[:5 if bi == 0 else 2]:
                    out = m2.prepare_reads(b.read_bases, bq, mapq, b.read_off, b.read_len, iq, dq, has, pcr, thr, gcp)
                    pp["par%d" % k] = np.array([pcr, thr, gcp, bi, {False: 0, True: 1, "mixed": 2}[with_indel]])
                    for nm, arr in zip(("bq", "iq", "dq", "gcp"), out):
                        pp["%s%d" % (nm, k)] = arr
                    k += 1
                key = "b%d_i%d_" % (bi, {False: 0, True: 1, "mixed": 2}[with_indel])
                pp[key + "bases"] = b.read_bases; pp[key + "base_q"] = bq; pp[key + "mapq"] = mapq
                pp[key + "read_off"] = b.read_off; pp[key + "read_len"] = b.read_len
                if iq is not None:
                    pp[key + "ins_q"] = iq; pp[key + "del_q"] = dq
                if has is not None:
                    pp[key + "has"] = has
        pp["lmt_len"] = np.arange(0, 400, dtype=np.int32)
        pp["lmt"] = np.array([m2.log10_min_true_likelihood(int(n), 0.02) for n in pp["lmt_len"]])
        np.savez_compressed(os.path.join(HERE, "prepass.npz"), n_cases=np.array(k), meta=np.array(meta2), **pp)
        print("prepass", k, "cases")
        # a13 / a14 end to end: whole regions through the reference's computeReadLikelihoods (stage 0: raw matrix in the
        # SampleMatrix, stage 1: + normalizeLikelihoods + filterPoorlyModeledEvidence), results mapped back to input order
        rg, names = {}, []
        for name, d in m2_region_cases().items():
            try:
                for stage in (0, 1):
                    out, removed, order = m2.region(stage=stage, **d)
                    rg["%s.out%d" % (name, stage)] = out; rg["%s.removed%d" % (name, stage)] = removed
                    rg["%s.order" % name] = order
            except RuntimeError as ex:      # e.g. Allele.cpp rejects haplotype bytes outside ACGTN ("Unexpected base in allele bases")
                print("m2_region", name, "skipped: the reference's caller refuses this region:", ex)
                for kk in [q for q in rg if q.startswith(name + ".")]:
                    del rg[kk]
                continue
            for kk, v in d.items():
                rg["%s.in.%s" % (name, kk)] = np.asarray(v)
            names.append(name)
            print("m2_region", name, out.shape, "removed", int(removed.sum()), "allele order", order.tolist())
        np.savez_compressed(os.path.join(HERE, "m2_region.npz"), names=np.array(names), meta=np.array(meta2), **rg)
        if sys.argv[1:]:
            return
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
