"""CPU tests (run with -m "not gpu"): the oracle against the golden vectors made from the reference's own code,
against the live reference build when oracle/_ref is present, and the lookup tables."""
import numpy as np
import pytest

from fgdp_b200 import synth
from util import assert_parity, load_golden, GOLDEN, TOL


@pytest.mark.parametrize("name", GOLDEN)
def test_oracle_matches_golden(oracle, name):
    """fixtures hold outputs of the reference (AVX-512 path, FMA-contracted); the oracle is the unfused scalar
    restatement: <= 1e-4 in log10 (observed <= 4e-6 = one float ulp at |log10| ~ 40), identical rescue set"""
    b, want, wresc, want_d = load_golden(name)
    got, resc, _ = oracle.compute(b.as_flat(), threads=4)
    assert_parity(got, resc, want, wresc, what=name)
    fin = np.isfinite(want)
    assert np.abs(got[fin] - want[fin]).max() < 1e-5
    got_d, resc_d, _ = oracle.compute(b.as_flat(), use_double=True, threads=4)
    assert resc_d.all()
    fin = np.isfinite(want_d)
    assert np.array_equal(np.isfinite(got_d), fin)
    assert np.abs(got_d[fin] - want_d[fin]).max() < 1e-9      # FP64 route: only FMA-vs-no-FMA rounding


def test_golden_covers_both_routes():
    tot = resc = 0
    for name in GOLDEN:
        _, want, wresc, _ = load_golden(name)
        tot += wresc.size; resc += int(wresc.sum())
    assert resc > 500 and tot - resc > 5000


def test_tables_match_golden_bitwise(oracle):
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "tables.npz"))
    pf, mf, pd, md = oracle.tables()
    for got, key in ((pf, "ph2pr_f"), (mf[:8256], "mm_f"), (pd, "ph2pr_d"), (md[:8256], "mm_d")):
        assert np.array_equal(got.view(np.uint8), z[key].view(np.uint8)), key


def test_oracle_vs_live_reference(oracle, reference):
    """when the reference build is present: fresh random inputs, not only the committed fixtures"""
    for b in (synth.config2(2, seed_extra=5), synth.config3(40, 5), synth.edge_cases(seed=99)):
        flat = b.as_flat()
        want, wresc, _ = reference.compute(flat, threads=8)
        got, resc, _ = oracle.compute(flat, threads=8)
        assert_parity(got, resc, want, wresc, what=b.name)
    to, tr = oracle.tables(), reference.tables()
    for a, c in zip(to, tr):
        assert np.array_equal(a.view(np.uint8), c.view(np.uint8))


def test_oracle_edge_semantics(oracle):
    """N matches everything, non-ACGTN bytes behave as 'A', quals are masked with &127"""
    A = np.frombuffer(b"ACGTACGTAC", np.uint8)

    def one(read, hap, q=30, i=45, d=45, c=10):
        n = len(read)
        b = synth.RegionBatch(np.asarray(read, np.uint8), np.full(n, q, np.uint8), np.full(n, i, np.uint8),
                              np.full(n, d, np.uint8), np.full(n, c, np.uint8), np.zeros(1, np.int64),
                              np.array([n], np.int32), np.asarray(hap, np.uint8), np.zeros(1, np.int64),
                              np.array([len(hap)], np.int32), np.array([0, 1], np.int64), np.array([0, 1], np.int64))
        return oracle.compute(b.as_flat())[0][0]

    base = one(A, A)
    n_read = A.copy(); n_read[3] = ord("N")
    n_hap = A.copy(); n_hap[3] = ord("N")
    assert one(n_read, A) == base and one(A, n_hap) == base          # N never costs a mismatch
    lower = A.copy(); lower[0] = ord("a")
    junk = A.copy(); junk[4] = ord("x")                               # position 4 is 'A'
    assert one(lower, A) == base and one(junk, A) == base
    assert one(A, A, q=30 | 0x80, i=45 | 0x80, d=45 | 0x80, c=10 | 0x80) == base
    mism = A.copy(); mism[5] = ord("A")
    assert one(mism, A) < base - 1.0


def _check_derived(m_in, mo, best, second, bh, pm, rl, cap, rate, nl):
    """best / second_best / best_hap / poorly_modeled against plain numpy on the same rows"""
    n, nh = m_in.shape
    scale = np.log(10.0) if nl else 1.0
    for r in range(n):
        row = m_in[r]
        assert bh[r] == int(np.argmax(row))                       # first maximum
        assert best[r] == row.max() * scale
        if nh > 1:
            srt = np.sort(mo[r])[::-1]
            assert second[r] == srt[1]                            # runner-up on the normalised row
        else:
            assert second[r] == -np.inf
        assert pm[r] == (row.max() < min(2.0, np.ceil(rl[r] * rate)) * -4.0)


def test_epilogue_oracle_matches_reference_golden(oracle):
    """oracle_epilogue() vs tests/golden/epilogue.npz = outputs of the reference's own AlleleLikelihoods.h
    (normalizeLikelihoods, filterPoorlyModeledEvidence, switchToNaturalLog): bit for bit, removal set identical."""
    import os
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "epilogue.npz"))
    n = int(z["n_cases"])
    assert n >= 50
    for k in range(n):
        m, rl, (cap, rate, nl) = z["m%d" % k], z["len%d" % k], z["par%d" % k]
        mo, best, second, bh, pm = oracle.epilogue(m, rl, cap, rate, bool(nl))
        assert np.array_equal(mo, z["out%d" % k]), k
        assert np.array_equal(pm, z["rem%d" % k]), k
        _check_derived(m, mo, best, second, bh, pm, rl, cap, rate, nl)


def test_epilogue_oracle_vs_live_reference(oracle):
    """the same against oracle/_ref/libref_epilogue.so run here, on fresh random matrices"""
    import oracle_py
    if not oracle_py.reference_epilogue_available():
        pytest.skip("oracle/_ref/libref_epilogue.so not built (no /root/reference on this machine)")
    ref = oracle_py.ReferenceEpilogue()
    rng = np.random.default_rng(7)
    for t in range(200):
        n, nh = int(rng.integers(1, 50)), int(rng.integers(1, 33))
        m = -rng.random((n, nh)) * float(rng.choice([1, 10, 100]))
        if t % 3 == 0:
            m = np.round(m)
        if t % 5 == 0:
            m[int(rng.integers(0, n))] = -np.inf
        rl = rng.integers(1, 385, n).astype(np.int32)
        cap, rate, nl = float(rng.choice([-4.5, -0.5, -np.inf])), float(rng.choice([0.02, 0.0, 0.005])), t % 2
        mo, best, second, bh, pm = oracle.epilogue(m, rl, cap, rate, bool(nl))
        mr, removed = ref.epilogue(m, rl, cap, rate, bool(nl))
        assert np.array_equal(mo, mr) and np.array_equal(pm, removed), t


def _prepass_cases():
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "prepass.npz"))
    for k in range(int(z["n_cases"])):
        pcr, thr, gcp, bi, mode = (int(x) for x in z["par%d" % k])
        key = "b%d_i%d_" % (bi, mode)
        inp = {f: z[key + f] for f in ("bases", "base_q", "mapq", "read_off", "read_len")}
        inp["ins_q"] = z[key + "ins_q"] if mode else None
        inp["del_q"] = z[key + "del_q"] if mode else None
        inp["has"] = z[key + "has"] if mode == 2 else None
        yield k, (pcr, thr, gcp), mode, inp, tuple(z["%s%d" % (n, k)] for n in ("bq", "iq", "dq", "gcp"))


def test_read_prepass_oracle_matches_reference_golden(oracle):
    """oracle_prepare_read against tests/golden/prepass.npz = outputs of the reference's OWN modifyReadQualities /
    applyPCRErrorModel / findTandemRepeatUnits / capMinimumReadQualities / buildGapContinuationPenalties
    (PairHMMLikelihoodCalculationEngine.cpp:123-292 + GATKVariantContextUtils.cpp:59-105, compiled unmodified into
    oracle/_ref/libref_m2.so): every byte equal. Repeat-rich reads with N of 1..600 bases and WES-like reads; no / all /
    every-other read with BI + BD arrays; MAPQ 0..60; the three PCR models; thresholds 18 / 6 / 25; gcp 10 / 12."""
    n = 0
    for k, (pcr, thr, gcp), mode, i, want in _prepass_cases():
        got = oracle.prepare_reads(i["bases"], i["base_q"], i["mapq"], i["read_off"], i["read_len"], i["ins_q"], i["del_q"],
                                   pcr, thr, gcp, has_indel_q=i["has"])
        for g, w, name in zip(got, want, ("base_q", "ins_q", "del_q", "gcp")):
            assert np.array_equal(g, w), (k, name, np.nonzero(g != w)[0][:5])
        n += 1
    assert n >= 20


def test_log10_min_true_likelihood_matches_reference_golden(oracle):
    """the filterPoorlyModeledEvidence threshold inside oracle_epilogue against the reference's own log10MinTrueLikelihood
    (PairHMMLikelihoodCalculationEngine.cpp:294-299, via libref_m2.so) for read lengths 0..399"""
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "prepass.npz"))
    lens, lmt = z["lmt_len"], z["lmt"]
    assert np.array_equal(lmt, np.minimum(2.0, np.ceil(lens * 0.02)) * -4.0)
    # a row whose best value sits exactly on the threshold is kept, just below it is flagged
    m = np.stack([lmt, lmt - 1.0], axis=1)
    _, _, _, _, pm = oracle.epilogue(m, lens)
    assert not pm.any()
    _, _, _, _, pm = oracle.epilogue(np.nextafter(m, -np.inf), lens)
    assert pm.all()


def _m2_region_cases():
    import os
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "m2_region.npz"))
    for name in z["names"]:
        name = str(name)
        d = {k[len(name) + 4:]: z[k] for k in z.files if k.startswith(name + ".in.")}
        for k in ("n_samples", "pcr_rate_factor", "base_q_threshold", "use_double"):
            if k in d:
                d[k] = int(d[k])
        yield name, d, z[name + ".out0"], z[name + ".out1"], z[name + ".removed1"], z[name + ".order"]


def test_reference_caller_golden_equals_oracle_pipeline(oracle):
    """tests/golden/m2_region.npz = whole regions through the reference's own PairHMMLikelihoodCalculationEngine::
    computeReadLikelihoods (raw reads -> modifyReadQualities -> VectorLoglessPairHMM (per-pair and trie paths, de-dup,
    allele-order scatter into the SampleMatrix) -> normalizeLikelihoods -> filterPoorlyModeledEvidence). The oracle's
    three restatements chained (prepare_read -> forward -> epilogue) must give the same matrices: stage 0 within 1e-4
    (observed 4e-6: FMA vs no FMA), stage 1 the same kept / removed reads."""
    n = 0
    for name, d, out0, out1, removed1, order in _m2_region_cases():
        pb, pi, pd, pg = oracle.prepare_reads(d["read_bases"], d["base_q"], d["mapq"], d["read_off"], d["read_len"],
                                              d.get("ins_q"), d.get("del_q"), d.get("pcr_rate_factor", 3),
                                              d.get("base_q_threshold", 18), 10, has_indel_q=d.get("has_indel_q"))
        nr, nh = out0.shape
        b = synth.RegionBatch(d["read_bases"], pb, pi, pd, pg, d["read_off"], d["read_len"], d["hap_bases"], d["hap_off"],
                              d["hap_len"], np.array([0, nr], np.int64), np.array([0, nh], np.int64), name)
        got, _, _ = oracle.compute(b.as_flat(), use_double=bool(d.get("use_double", 0)), threads=4)
        got = got.reshape(nr, nh)
        assert np.isfinite(out0).all() and np.abs(got - out0).max() <= TOL, (name, np.abs(got - out0).max())
        m, best, second, bh, pm = oracle.epilogue(got, d["read_len"])
        # reads whose best value is within 1e-4 of the threshold may legitimately fall on either side
        thr = np.minimum(2.0, np.ceil(d["read_len"] * 0.02)) * -4.0
        sure = np.abs(got.max(axis=1) - thr) > TOL
        assert np.array_equal(pm[sure], removed1[sure]), name
        kept = removed1 == 0
        assert np.isnan(out1[~kept]).all()
        if (kept & sure).any():
            assert np.abs(m[kept & sure] - out1[kept & sure]).max() <= TOL, name
        assert sorted(order.tolist()) == list(range(nh))
        n += 1
    assert n >= 20


def test_oracle_and_goldens_vs_live_reference_caller(oracle):
    """when oracle/_ref/libref_m2.so is present (the container with /root/reference): fresh random reads through the
    reference's own pre-pass, and the committed goldens reproduced by it"""
    import oracle_py
    if not oracle_py.reference_m2_available():
        pytest.skip("oracle/_ref/libref_m2.so not built (needs /root/reference at build time)")
    from util import raw_quals, repeat_rich_batch
    m2 = oracle_py.ReferenceM2()
    for seed in (41, 42):
        b = repeat_rich_batch(seed=seed, max_len=600)
        for with_indel in (False, True):
            bq, mapq, iq, dq = raw_quals(b, seed, with_indel)
            for pcr in (1, 2, 3):
                want = m2.prepare_reads(b.read_bases, bq, mapq, b.read_off, b.read_len, iq, dq, None, pcr, 18, 10)
                got = oracle.prepare_reads(b.read_bases, bq, mapq, b.read_off, b.read_len, iq, dq, pcr, 18, 10)
                for g, w in zip(got, want):
                    assert np.array_equal(g, w)
    for name, d, out0, out1, removed1, order in list(_m2_region_cases())[:6]:
        o0, _, od = m2.region(stage=0, **d)
        assert np.array_equal(o0, out0) and np.array_equal(od, order), name


def test_read_prepass_oracle_known_answers(oracle):
    """oracle_prepare_read / oracle_tandem_repeat_length (restating PairHMMLikelihoodCalculationEngine.cpp:123-267):
    hand-derived known answers, kept beside the reference-generated goldens above (which are what pins the row)."""
    s = b"ACGTTTTTTGCACACACACAGGATTCTTC"
    # offset 2 ('G' before T x 6): BW unit G x1, FW unit T x6, FW unit does not repeat backwards -> 6;
    # offset 8 (last T): BW unit T x6 but FW unit G x1 differs and does not repeat backwards -> 1 (GATK's behaviour);
    # offsets 9..15 (GCACACAC): alternating 5 / 4 repeats of CA / AC
    want = [1, 1, 6, 6, 6, 6, 6, 6, 1, 5, 4, 5, 4, 5, 4, 5, 1, 1, 1, 2, 2, 1, 2, 2, 1, 2, 2, 1, 2]
    assert [oracle.tandem_repeat_length(s, i) for i in range(len(s))] == want
    assert oracle.tandem_repeat_length(b"A" * 50, 10) == 20            # MAX_REPEAT_LENGTH
    assert oracle.tandem_repeat_length(b"A", 0) == 1
    # getErrorModelAdjustedQual(i, 3) = max(10, round(40 - exp(i / (3 pi))) + 1)
    import ctypes as C
    t = (C.c_uint8 * 21)(); oracle.lib.oracle_pcr_table(3, t)
    assert list(t) == [max(10, int(40.0 - np.exp(i / (3 * np.pi)) + 0.5) + 1) for i in range(21)]
    assert t[0] == 40 and t[20] == 33
    # one read: base quality capped by MAPQ 25 then < 18 -> 6; indel default 45 lowered by the PCR table except at
    # the last base (applyPCRErrorModel stops at len-2); gcp constant
    bq = np.full(len(s), 30, np.uint8); bq[3] = 10; bq[5] = 17
    b, i, d, g = oracle.prepare_reads(np.frombuffer(s, np.uint8), bq, [25], [0], [len(s)])
    assert list(b) == [25, 25, 25, 6, 25, 6] + [25] * 23
    assert list(i) == [t[w] for w in want[:-1]] + [45] and list(d) == list(i) and set(g) == {10}
    # existing indel qualities below 6 are raised to 6, above the table value lowered to it; no PCR model = untouched
    iq = np.full(len(s), 3, np.uint8); dq = np.full(len(s), 60, np.uint8)
    b, i, d, g = oracle.prepare_reads(np.frombuffer(s, np.uint8), bq, [60], [0], [len(s)], ins_q=iq, del_q=dq)
    assert set(i) == {6} and list(d) == [t[w] for w in want[:-1]] + [60] and b[0] == 30
    b, i, d, g = oracle.prepare_reads(np.frombuffer(s, np.uint8), bq, [60], [0], [len(s)], ins_q=iq, del_q=dq, pcr_rate_factor=0)
    assert set(d) == {60}
