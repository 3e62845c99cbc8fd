"""-m gpu: the CUDA path, called through the C ABI, against the CPU oracle on identical bytes."""
import numpy as np
import pytest

import fgdp_b200
from fgdp_b200 import synth
from util import assert_parity, load_golden, GOLDEN, TOL, raw_quals, repeat_rich_batch

pytestmark = pytest.mark.gpu


def _region_lists(b, g=0):
    r0, r1 = int(b.reg_read_start[g]), int(b.reg_read_start[g + 1])
    h0, h1 = int(b.reg_hap_start[g]), int(b.reg_hap_start[g + 1])
    reads = []
    for r in range(r0, r1):
        o, n = int(b.read_off[r]), int(b.read_len[r])
        reads.append(tuple(a[o:o + n] for a in (b.read_bases, b.base_q, b.ins_q, b.del_q, b.gcp)))
    haps = [b.hap_bases[int(b.hap_off[h]):int(b.hap_off[h]) + int(b.hap_len[h])] for h in range(h0, h1)]
    return reads, haps


@pytest.mark.parametrize("profile", ["clean", "lowq"])
def test_config1_region_call(engine, oracle, profile):
    """BASELINE configs[0]: 64 reads x 8 haps, through pairhmm_cuda_compute as the adapter calls it."""
    b = synth.config1(profile)
    want, wresc, _ = oracle.compute(b.as_flat(), threads=4)
    reads, haps = _region_lists(b)
    engine.initialize(150, 300)
    got, gresc = engine.compute(reads, haps)
    assert got.shape == (64, 8)
    assert_parity(got, gresc, want, wresc, what="C1-" + profile)


@pytest.mark.parametrize("name", GOLDEN)
def test_golden_vectors_from_the_reference(engine, name):
    """committed outputs of the reference's own code (tests/golden/make_golden.py)"""
    b, want, wresc, want_d = load_golden(name)
    got, gresc = engine.compute_batch(b)
    assert_parity(got, gresc, want, wresc, what="golden " + name)
    e = fgdp_b200.Engine(use_double=True)
    try:
        got_d, resc_d = e.compute_batch(b)
    finally:
        e.close()
    fin = np.isfinite(want_d)
    assert resc_d.all() and np.array_equal(np.isfinite(got_d), fin)
    assert np.abs(got_d[fin] - want_d[fin]).max() < 1e-9


def test_edge_cases(engine, oracle):
    b = synth.edge_cases()
    want, wresc, _ = oracle.compute(b.as_flat(), threads=4)
    got, gresc = engine.compute_batch(b)
    assert_parity(got, gresc, want, wresc, what="edge")
    assert wresc.sum() > 50 and (wresc == 0).sum() > 50     # both routes are exercised


def test_exact_mode_is_bit_identical_to_the_unfused_oracle(oracle):
    """exact_fp32=1 disables FMA contraction: same IEEE operations in the same order as the scalar oracle
    (and as the reference's AVX-256 build), so the raw FP32 forward sums must agree to the last bit.
    (The final log10 is not compared bitwise: glibc's log10f is not correctly rounded, the device's is.)"""
    e = fgdp_b200.Engine(exact_fp32=True, dedup_reads=False)
    try:
        for b in (synth.config1("clean"), synth.config1("lowq"), synth.edge_cases(), synth.config3(96, 6),
                  synth.config4(2, 6, 3)):
            flat = b.as_flat()
            want_raw = oracle.forward_float(flat)
            want, wresc, _ = oracle.compute(flat, threads=4)
            e.resident_load(b)
            e.resident_run()
            raw = e.resident_fetch_raw()
            got, gresc = e.resident_fetch()
            assert np.array_equal(raw.view(np.uint32), want_raw.view(np.uint32)), \
                "%s: %d of %d FP32 sums differ in bits" % (b.name, (raw.view(np.uint32) != want_raw.view(np.uint32)).sum(), raw.size)
            assert_parity(got, gresc, want, wresc, what=b.name)
    finally:
        e.close()


def test_uniform_gcp_kernels_equal_general_kernels():
    """the packed (two reads per lane, FFMA2) and the scalar constant-bank uniform-gcp kernels do the same IEEE
    operations as the per-row-register kernels: raw FP32 sums identical to the last bit at every lane-group width"""
    b = synth.config2(5)
    assert b.gcp.min() == b.gcp.max() == 10
    raws = []
    for force_general, lane_group, scalar_uniform in ((True, 32, False), (False, 32, True), (False, 32, False),
                                                      (False, 16, False), (False, 8, False), (False, 0, False)):
        e = fgdp_b200.Engine(force_general=force_general, dedup_reads=False, lane_group=lane_group,
                             scalar_uniform=scalar_uniform)
        try:
            e.resident_load(b); e.resident_run()
            raws.append(e.resident_fetch_raw())
        finally:
            e.close()
    for r in raws[1:]:
        assert np.array_equal(raws[0].view(np.uint32), r.view(np.uint32))


@pytest.mark.parametrize("lane_group", [0, 8, 16, 32])
def test_every_bin(oracle, lane_group):
    """one region per read length 1..384: every (lane-group width, rows-per-lane, last-row slot) kernel
    instantiation runs, on the uniform-gcp path (gcp == 10 everywhere) and, with gcp perturbed, on the
    general path"""
    rng = np.random.default_rng(77)
    engine = fgdp_b200.Engine(lane_group=lane_group)
    for perturb in ((False, True) if lane_group in (0, 32) else (False,)):
        lens = np.arange(1, 385)
        hls = [np.array([max(1, int(R * 1.3) + 7), 40]) for R in lens]
        parts = [synth.make_batch(1000 + int(R), [5], [hl], int(R), "clean", foreign_frac=0.4) for R, hl in zip(lens, hls)]
        b = parts[0].__class__(*[np.concatenate([getattr(p, f) for p in parts]) for f in
                                 ("read_bases", "base_q", "ins_q", "del_q", "gcp")],
                               None, np.concatenate([p.read_len for p in parts]),
                               np.concatenate([p.hap_bases for p in parts]), None,
                               np.concatenate([p.hap_len for p in parts]),
                               np.arange(0, 5 * len(parts) + 1, 5, dtype=np.int64),
                               np.arange(0, 2 * len(parts) + 1, 2, dtype=np.int64), "allbins")
        b.read_off = synth._offsets(b.read_len); b.hap_off = synth._offsets(b.hap_len)
        if perturb:
            b.gcp = b.gcp.copy(); b.gcp[rng.random(b.gcp.shape[0]) < 0.3] = 7
        want, wresc, _ = oracle.compute(b.as_flat(), threads=8)
        got, gresc = engine.compute_batch(b)
        assert_parity(got, gresc, want, wresc, what="all bins, perturb=%s" % perturb)
    engine.close()


def test_config3_rescue_heavy(engine, oracle):
    b = synth.config3(200, 8)
    want, wresc, _ = oracle.compute(b.as_flat(), threads=4)
    got, gresc = engine.compute_batch(b)
    assert_parity(got, gresc, want, wresc, what="C3")
    assert wresc.mean() > 0.2


def test_rescue_with_two_reads_per_warp_equals_one_read_per_warp(oracle):
    """FP64 pass, W = 16 instance (config.reserved[7]): reads that underflow are paired up on the device (which two meet depends
    on timing) and sweep the union of the haplotypes either needs. Forced on for small inputs: the bits of every result and
    the rescued flags equal those of the one-read-per-warp pass, for every size of the haplotype runs, reads of many
    lengths (bins of the W = 16 instance, and reads it does not take), duplicate reads and ragged regions; and the oracle
    agrees. use_double handles take the W = 16 instance for every read that fits."""
    rng = np.random.default_rng(77)
    mixed = synth.make_batch(991, [9, 40, 1, 33], [[260, 300, 310], [90] * 7, [400], [333, 350] * 9], 150, profile="lowq", name="mixed-150")
    lens = synth.concat([synth.make_batch(1000 + R, [6], [[200, 220, 231, 250, 180]], R, profile="lowq", foreign_frac=0.4) for R in
                         (1, 15, 16, 17, 33, 97, 128, 129, 144, 145, 159, 160, 161, 200)], name="many-read-lengths")
    off = fgdp_b200.Engine(rescue_read_pairing=1)
    engines = [fgdp_b200.Engine(rescue_read_pairing=2, rescue_pairs_per_unit=k) for k in (0, 1, 3, 32)]
    try:
        for b in (synth.config3(300, 12), mixed, lens, synth.edge_cases(), synth.config2(30, profile="lowq")):
            want, wresc = off.compute_batch(b)
            for e in engines:
                for rep in range(2):
                    got, gresc = e.compute_batch(b)
                    assert np.array_equal(got, want) and np.array_equal(gresc, wresc), b.name
            ow, owr, _ = oracle.compute(b.as_flat(), threads=8)
            assert_parity(want, wresc, ow, owr, what=b.name)
        st = engines[0].stats()
        assert st["rescued_pairs"] > 0
        # the default decides by the chunk's size: a large launch group pairs, and gives the same bits
        big = synth.config3(1500, 16)
        want, wresc = off.compute_batch(big)
        dflt = fgdp_b200.Engine()
        try:
            got, gresc = dflt.compute_batch(big)
        finally:
            dflt.close()
        assert np.array_equal(got, want) and np.array_equal(gresc, wresc) and wresc.mean() > 0.2
    finally:
        off.close()
        for e in engines:
            e.close()
    # use_double: W = 16 for the reads that fit, W = 32 for the others, one chunk
    e = fgdp_b200.Engine(use_double=True)
    try:
        for b in (lens, mixed):
            got, gresc = e.compute_batch(b)
            ow = oracle.compute(b.as_flat(), threads=8, use_double=True)[0]
            fin = np.isfinite(ow)
            assert gresc.all() and np.array_equal(np.isfinite(got), fin) and np.abs(got[fin] - ow[fin]).max() < 1e-9, b.name
    finally:
        e.close()


def test_config4_long_haps(engine, oracle):
    b = synth.config4(3, 12, 4)
    want, wresc, _ = oracle.compute(b.as_flat(), threads=4)
    got, gresc = engine.compute_batch(b)
    assert_parity(got, gresc, want, wresc, what="C4")


def test_config2_many_regions_chunked(oracle):
    """several regions, forced through several chunks and both staging slots"""
    b = synth.config2(12)
    want, wresc, _ = oracle.compute(b.as_flat(), threads=8)
    e = fgdp_b200.Engine(chunk_cells=int(b.cells() // 5))
    try:
        got, gresc = e.compute_batch(b)
        st = e.stats()
    finally:
        e.close()
    assert_parity(got, gresc, want, wresc, what="C2")
    assert st["pairs"] == b.n_pairs and st["cells"] == b.cells() and st["regions"] == 12
    assert st["kernel_launches"] >= 5 * 3


def test_use_double_everything_rescued(oracle):
    b = synth.config1("clean")
    want, _, _ = oracle.compute(b.as_flat(), use_double=True, threads=4)
    e = fgdp_b200.Engine(use_double=True)
    try:
        got, gresc = e.compute_batch(b)
    finally:
        e.close()
    assert gresc.all()
    assert np.abs(got - want).max() < 1e-9


def test_dedup_and_duplicates(engine, oracle):
    """identical reads share one DP (VectorLoglessPairHMM.cpp:83-103) and still get their own rows"""
    b = synth.config1("clean")
    reads, haps = _region_lists(b)
    reads = reads[:10] + reads[:10] + reads[3:5]
    engine.stats_reset()
    got, _ = engine.compute(reads, haps)
    st = engine.stats()
    assert st["reads"] == 22 and st["unique_reads"] == 10
    assert np.array_equal(got[:10], got[10:20]) and np.array_equal(got[3:5], got[20:22])
    want, _, _ = oracle.compute(b.as_flat(), threads=2)
    assert np.abs(got[:10].reshape(-1) - want[:80]).max() <= TOL


def test_submit_wait_overlap(engine, oracle):
    b = synth.config2(3)
    want, wresc, _ = oracle.compute(b.as_flat(), threads=4)
    off = b.region_out_offsets()
    tickets = []
    for g in range(3):
        reads, haps = _region_lists(b, g)
        tickets.append(engine.submit(reads, haps))
    for g, (t, out, resc) in enumerate(tickets):
        engine.wait(t)
        assert_parity(out, resc, want[off[g]:off[g + 1]], wresc[off[g]:off[g + 1]], what="submit %d" % g)


def test_pair_list(engine, oracle):
    b = synth.config1("lowq")
    reads, haps = _region_lists(b)
    rng = np.random.default_rng(5)
    ri = rng.integers(0, 64, size=300).astype(np.uint32); hi = rng.integers(0, 8, size=300).astype(np.uint32)
    want, _, _ = oracle.compute(b.as_flat(), threads=2)
    got, _ = engine.compute_pairs(reads, haps, ri, hi)
    assert np.abs(got - want.reshape(64, 8)[ri, hi]).max() <= TOL


def test_empty_and_errors(engine):
    out, resc = engine.compute([], [np.frombuffer(b"ACGT", np.uint8)])
    assert out.shape == (0, 1)
    with pytest.raises(fgdp_b200.PairHMMError):
        engine.initialize(10, 0)
    r = tuple(np.full(385, v, np.uint8) for v in (65, 30, 45, 45, 10))
    out, resc = engine.compute([r], [np.frombuffer(b"ACGT", np.uint8)])   # > 384 rows: the row-striped kernel
    assert np.isfinite(out).all()
    with pytest.raises(fgdp_b200.PairHMMError):
        engine.initialize(0, 10)
    z = tuple(np.zeros(0, np.uint8) for _ in range(5))
    out, resc = engine.compute([z], [np.frombuffer(b"ACGT", np.uint8)])
    assert np.isneginf(out).all() and resc.all()


def test_null_outputs_and_resident_state_are_refused(engine, oracle):
    """a null result array is PAIRHMM_ERR_ARG (not a segfault after the GPU work); a batch staged by
    _resident_load is gone once another call re-packs the slot (ERR_STATE instead of replaying the wrong batch)"""
    import ctypes as C
    b = synth.config1("clean")
    reads = [tuple(a[int(o):int(o) + int(n)] for a in (b.read_bases, b.base_q, b.ins_q, b.del_q, b.gcp))
             for o, n in zip(b.read_off, b.read_len)]
    haps = [b.hap_bases[int(o):int(o) + int(n)] for o, n in zip(b.hap_off, b.hap_len)]
    ra, ha, keep = engine._structs(reads, haps)
    st = engine.lib.pairhmm_cuda_compute(engine.h, ra, len(reads), ha, len(haps), None, None)
    assert st == -1 and b"null output" in engine.lib.pairhmm_cuda_last_error(engine.h)
    t = C.c_int64()
    assert engine.lib.pairhmm_cuda_submit(engine.h, ra, len(reads), ha, len(haps), None, None, C.byref(t)) == -1
    reg = (fgdp_b200.pairhmm.PhmmRegion * 1)()
    reg[0] = fgdp_b200.pairhmm.PhmmRegion(ra, len(reads), ha, len(haps), None, None)
    assert engine.lib.pairhmm_cuda_compute_batch(engine.h, reg, 1) == -1
    # resident state
    engine.resident_load(b)
    engine.resident_run()
    other = synth.config2(2)
    engine.compute_batch(other)
    with pytest.raises(fgdp_b200.PairHMMError) as ei:
        engine.resident_run()
    assert ei.value.status == -5
    with pytest.raises(fgdp_b200.PairHMMError):
        engine.resident_fetch()
    engine.resident_load(b)
    engine.resident_run()
    got, gresc = engine.resident_fetch()
    want, wresc, _ = oracle.compute(b.as_flat(), threads=2)
    assert_parity(got, gresc, want, wresc, what="resident after reload")


def test_raw_batch_with_an_empty_leading_region(engine, oracle):
    """the read kind of a chunk comes from its first region WITH reads: an empty first region (reads == NULL) in a
    raw batch is valid input"""
    import ctypes as C
    P = fgdp_b200.pairhmm
    b = synth.config1("clean")
    bq, mapq, _, _ = raw_quals(b, 4, False)

    class RawRegion(C.Structure):
        _fields_ = [("reads", C.c_void_p), ("n_reads", C.c_int32), ("haps", C.c_void_p), ("n_haps", C.c_int32),
                    ("log10_out", C.c_void_p), ("rescued_out", C.c_void_p)]
    rr = (P.PhmmRawRead * b.n_reads)()
    for i, (o, n) in enumerate(zip(b.read_off, b.read_len)):
        rr[i] = P.PhmmRawRead(b.read_bases.ctypes.data + int(o), bq.ctypes.data + int(o), None, None, int(n), int(mapq[i]))
    ha = (P.PhmmHap * b.n_haps)()
    for i, (o, n) in enumerate(zip(b.hap_off, b.hap_len)):
        ha[i] = P.PhmmHap(b.hap_bases.ctypes.data + int(o), int(n))
    out = np.zeros(b.n_pairs); resc = np.zeros(b.n_pairs, np.uint8)
    regs = (RawRegion * 3)()
    regs[0] = RawRegion(None, 0, C.cast(ha, C.c_void_p), b.n_haps, None, None)
    regs[1] = RawRegion(C.cast(rr, C.c_void_p), b.n_reads, C.cast(ha, C.c_void_p), b.n_haps, out.ctypes.data, resc.ctypes.data)
    regs[2] = RawRegion(None, 0, None, 0, None, None)
    m = P.ReadModel()
    engine.lib.pairhmm_cuda_compute_batch_raw.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(P.ReadModel)]
    st = engine.lib.pairhmm_cuda_compute_batch_raw(engine.h, C.cast(regs, C.c_void_p), 3, C.byref(m))
    assert st == 0, engine.lib.pairhmm_cuda_last_error(engine.h)
    want, wresc = engine.compute_batch_raw(b, bq, mapq)
    assert np.array_equal(out, want) and np.array_equal(resc, wresc)


def test_fp64_underflow_window(oracle):
    """the reference runs with MXCSR flush-to-zero (IntelPairHmm.cc:230), which also flushes DOUBLE denormals; a
    2^1020-scaled FP64 sum below DBL_MIN is -inf there. The engine flushes the final sum (forward.cuh: flush_denormal);
    intermediate FP64 denormals cannot be flushed on the GPU, so inside the window (true log10 about -631 .. -600) a pair may
    be finite here and -inf in the reference. Checked on all-mismatch reads with very high qualities, through the wide FP64
    kernel (uniform gcp), the general FP64 kernel (per-base gcp) and the long-read kernel: identical wherever the reference
    is finite and above the window; where the reference says -inf the engine says -inf or a value below -600."""
    def batch(R, H, q, jitter):
        read = np.full(R, ord("A"), np.uint8); hap = np.full(H, ord("C"), np.uint8)
        g = np.full(R, q, np.uint8)
        if jitter:
            g[::2] -= 1
        return synth.RegionBatch(read, np.full(R, q, np.uint8), np.full(R, q, np.uint8), np.full(R, q, np.uint8), g,
                                 np.array([0], np.int64), np.array([R], np.int32), hap, np.array([0], np.int64),
                                 np.array([H], np.int32), np.array([0, 1], np.int64), np.array([0, 1], np.int64), "w")
    e = fgdp_b200.Engine()
    seen_inf = seen_finite = 0
    try:
        for jitter in (False, True):
            for R in list(range(120, 161, 2)) + [400, 450, 460, 470, 480, 490, 500, 520]:    # > 384 rows: long.cu
                q = 45 if R < 200 else 14
                b = batch(R, 200 if R < 200 else 500, q, jitter)
                want, wresc, _ = oracle.compute(b.as_flat())
                got, gresc = e.compute_batch(b)
                assert np.array_equal(gresc, wresc)
                if np.isfinite(want[0]):
                    assert np.isfinite(got[0]) and abs(got[0] - want[0]) <= TOL, (R, jitter, got, want)
                    seen_finite += 1
                else:
                    assert np.isneginf(got[0]) or got[0] < -600.0, (R, jitter, got, want)
                    seen_inf += 1
    finally:
        e.close()
    assert seen_inf > 4 and seen_finite > 4


def test_resident_replay_matches(engine, oracle):
    b = synth.config2(6)
    want, wresc, _ = oracle.compute(b.as_flat(), threads=8)
    engine.resident_load(b)
    for _ in range(2):
        ms, ms_f = engine.resident_run()
        assert ms > 0 and ms_f > 0
    got, gresc = engine.resident_fetch()
    assert_parity(got, gresc, want, wresc, what="resident")


def test_long_reads_double_and_exact_paths(oracle):
    """reads of 257..384 bases (rows-per-lane 9..12 at one read per warp): FP32 exact, FP64-only and rescue paths"""
    rng = np.random.default_rng(5)
    lens = [257, 288, 300, 301, 320, 351, 352, 383, 384]
    parts = [synth.make_batch(4000 + R, [6], [np.array([R + 40, R - 30, 500])], R, "lowq" if i % 2 else "clean", foreign_frac=0.5)
             for i, R in enumerate(lens)]
    for p in parts:
        flat = p.as_flat()
        want, wresc, _ = oracle.compute(flat, threads=4)
        for kw in ({}, {"exact_fp32": True}):
            e = fgdp_b200.Engine(**kw)
            try:
                got, gresc = e.compute_batch(p)
            finally:
                e.close()
            assert_parity(got, gresc, want, wresc, what="long reads %s %s" % (p.name, kw))
        e = fgdp_b200.Engine(use_double=True)
        try:
            got_d, resc_d = e.compute_batch(p)
        finally:
            e.close()
        assert resc_d.all()
        m = wresc.astype(bool)                       # where the reference also took the double path: same numbers
        if m.any():
            assert np.abs(got_d[m] - want[m]).max() < 1e-9


def _rows(b, flat, g):
    """region g of a flat result array as [n_reads, n_haps]"""
    r0, r1 = int(b.reg_read_start[g]), int(b.reg_read_start[g + 1])
    nh = int(b.reg_hap_start[g + 1] - b.reg_hap_start[g])
    o0 = int(b.region_out_offsets()[g])
    return flat[o0:o0 + (r1 - r0) * nh].reshape(r1 - r0, nh), r0, r1


@pytest.mark.parametrize("cap,rate,nl", [(-4.5, 0.02, False), (-4.5, 0.02, True), (-np.inf, 0.02, False), (-1.0, 0.0, False)])
def test_fused_epilogue_equals_oracle_on_the_same_matrix(engine, oracle, cap, rate, nl):
    """pairhmm_cuda_compute_flat_normalized = PairHMM + normalizeLikelihoods + filterPoorlyModeledEvidence's predicate
    [+ switchToNaturalLog] on the device. Applied to the engine's own raw matrix, the oracle's epilogue (pinned against
    the reference's AlleleLikelihoods.h, tests/test_oracle.py) must give the same bits; the raw matrix itself is held to
    the usual parity bar by the other tests."""
    for b in (synth.config1("lowq"), synth.config3(96, 7), synth.config2(4), synth.edge_cases()):
        raw, resc = engine.compute_batch(b)
        ep = fgdp_b200.pairhmm.Epilogue(cap, rate, nl)
        got, gresc, summ = engine.compute_batch_normalized(b, ep)
        assert np.array_equal(gresc, resc)
        for g in range(b.n_regions):
            m, r0, r1 = _rows(b, raw, g)
            want, best, second, bh, pm = oracle.epilogue(m, b.read_len[r0:r1], cap, rate, nl)
            gm, _, _ = _rows(b, got, g)
            assert np.array_equal(gm, want), (b.name, g)
            s = summ[r0:r1]
            assert np.array_equal(s["best"], best) and np.array_equal(s["second_best"], second)
            assert np.array_equal(s["best_hap"], bh) and np.array_equal(s["poorly_modeled"], pm)


def test_fused_epilogue_region_call_duplicates_and_empty_reads(engine, oracle):
    """single-region entry point; duplicate reads share their unique read's row, a zero-length read reports -inf"""
    b = synth.config1("lowq")
    reads, haps = _region_lists(b)
    reads = reads[:10] + [reads[3]] + [tuple(a[:0] for a in reads[0])] + reads[10:20]
    raw, _ = engine.compute(reads, haps)
    got, _, summ = engine.compute_normalized(reads, haps)
    lens = np.array([len(r[0]) for r in reads], np.int32)
    want, best, second, bh, pm = oracle.epilogue(raw, lens)
    assert np.array_equal(got, want)
    assert np.array_equal(summ["best"], best) and np.array_equal(summ["best_hap"], bh)
    assert np.array_equal(summ["poorly_modeled"], pm) and np.array_equal(summ["second_best"], second)
    assert np.array_equal(got[10], got[3]) and summ[11]["poorly_modeled"] == 1 and np.isneginf(got[11]).all()
    with pytest.raises(fgdp_b200.PairHMMError):
        engine.compute_normalized(reads, haps, fgdp_b200.pairhmm.Epilogue(likelihood_cap=1.0))
    # the plain entry points are unaffected by an earlier normalised call
    again, _ = engine.compute(reads, haps)
    assert np.array_equal(again, raw)


_raw_quals, _repeat_rich_batch = raw_quals, repeat_rich_batch


@pytest.mark.parametrize("with_indel", [False, True])
@pytest.mark.parametrize("pcr", [0, 1, 2, 3])
def test_device_read_prepass_equals_oracle(engine, oracle, with_indel, pcr):
    """prep_reads_kernel (applyPCRErrorModel + capMinimumReadQualities + constant gcp on the device) against the oracle's
    restatement of PairHMMLikelihoodCalculationEngine.cpp:123-267: every byte equal"""
    for b in (_repeat_rich_batch(), synth.config2(2), synth.long_reads()):
        bq, mapq, iq, dq = _raw_quals(b, 11 + pcr, with_indel)
        model = fgdp_b200.pairhmm.ReadModel(pcr, 18, 10)
        got = engine.prepare_reads(b, bq, mapq, iq, dq, model)
        want = oracle.prepare_reads(b.read_bases, bq, mapq, b.read_off, b.read_len, iq, dq, pcr, 18, 10)
        for g, w, name in zip(got, want, ("base_q", "ins_q", "del_q", "gcp")):
            assert np.array_equal(g, w), (b.name, name, np.nonzero(g != w)[0][:5])


def test_device_read_prepass_equals_reference_golden(engine):
    """prep_reads_kernel against tests/golden/prepass.npz = outputs of the reference's OWN modifyReadQualities &
    buildGapContinuationPenalties (PairHMMLikelihoodCalculationEngine.cpp:123-292, GATKVariantContextUtils.cpp:59-105;
    oracle/_ref/libref_m2.so): every byte equal -- row N1 pinned on the reference, not on a restatement"""
    from test_oracle import _prepass_cases
    n = 0
    for k, (pcr, thr, gcp), mode, i, want in _prepass_cases():
        if mode == 2:
            continue        # BI / BD on every other read: the flat pre-pass takes them for all reads or none; see the region test below
        b = synth.RegionBatch(i["bases"], i["base_q"], i["base_q"], i["base_q"], i["base_q"], i["read_off"], i["read_len"],
                              np.zeros(0, np.uint8), np.zeros(0, np.int64), np.zeros(0, np.int32), np.array([0, len(i["read_len"])], np.int64),
                              np.array([0, 0], np.int64), "prepass-golden")
        got = engine.prepare_reads(b, i["base_q"], i["mapq"], i["ins_q"], i["del_q"], fgdp_b200.pairhmm.ReadModel(pcr, thr, gcp))
        for g, w, name in zip(got, want, ("base_q", "ins_q", "del_q", "gcp")):
            assert np.array_equal(g, w), (k, name, np.nonzero(g != w)[0][:5])
        n += 1
    assert n >= 12


def test_raw_regions_equal_the_reference_callers_matrices(engine):
    """pairhmm_cuda_compute_raw / _compute_normalized against tests/golden/m2_region.npz = whole regions through the
    reference's own PairHMMLikelihoodCalculationEngine::computeReadLikelihoods (pre-pass, VectorLoglessPairHMM incl. its trie
    path and de-dup, normalizeLikelihoods, filterPoorlyModeledEvidence): raw matrix within 1e-4; fused epilogue: same reads
    flagged, same normalised values"""
    from test_oracle import _m2_region_cases
    n = 0
    for name, d, out0, out1, removed1, order in _m2_region_cases():
        ro, rl = d["read_off"], d["read_len"]
        cut = lambda a, i: a[int(ro[i]):int(ro[i]) + int(rl[i])]
        has = d.get("has_indel_q")
        reads = []
        for i in range(len(rl)):
            own = "ins_q" in d and (has is None or has[i])
            reads.append((cut(d["read_bases"], i), cut(d["base_q"], i), cut(d["ins_q"], i) if own else None,
                          cut(d["del_q"], i) if own else None, int(d["mapq"][i])))
        haps = [d["hap_bases"][int(o):int(o) + int(m)] for o, m in zip(d["hap_off"], d["hap_len"])]
        model = fgdp_b200.pairhmm.ReadModel(d.get("pcr_rate_factor", 3), d.get("base_q_threshold", 18), 10)
        e = fgdp_b200.Engine(use_double=True) if d.get("use_double") else engine
        got, _ = e.compute_raw(reads, haps, model)
        assert np.abs(got - out0).max() <= TOL, (name, np.abs(got - out0).max())
        if e is not engine:
            e.close()
            continue
        # the fused epilogue on the device-prepared reads: same rows removed, same normalised values as the reference's stage 1
        b = synth.RegionBatch(d["read_bases"], d["base_q"], d["base_q"], d["base_q"], d["base_q"], ro, rl, d["hap_bases"],
                              d["hap_off"], d["hap_len"], np.array([0, len(rl)], np.int64), np.array([0, len(haps)], np.int64), name)
        if has is None:
            pb, pi, pd, pg = engine.prepare_reads(b, d["base_q"], d["mapq"], d.get("ins_q"), d.get("del_q"), model)
            preads = [(cut(d["read_bases"], i), cut(pb, i), cut(pi, i), cut(pd, i), cut(pg, i)) for i in range(len(rl))]
            norm, _, summ = engine.compute_normalized(preads, haps)
            thr = np.minimum(2.0, np.ceil(rl * 0.02)) * -4.0
            sure = np.abs(out0.max(axis=1) - thr) > TOL
            assert np.array_equal(summ["poorly_modeled"][sure], removed1[sure]), name
            kept = (removed1 == 0) & sure
            if kept.any():
                assert np.abs(norm[kept] - out1[kept]).max() <= TOL, name
        n += 1
    assert n >= 18


@pytest.mark.parametrize("with_indel", [False, True])
def test_raw_read_entry_point_equals_prepared_reads(engine, oracle, with_indel):
    """pairhmm_cuda_compute_flat_raw = device pre-pass + PairHMM: identical to feeding the oracle-prepared qualities to the
    plain entry point, and within the parity bar of the CPU oracle on them"""
    for b in (_repeat_rich_batch(), synth.config2(3), synth.long_reads()):
        bq, mapq, iq, dq = _raw_quals(b, 3, with_indel)
        got, gresc = engine.compute_batch_raw(b, bq, mapq, iq, dq)
        pb, pi, pd, pg = oracle.prepare_reads(b.read_bases, bq, mapq, b.read_off, b.read_len, iq, dq)
        pbatch = synth.RegionBatch(b.read_bases, pb, pi, pd, pg, b.read_off, b.read_len, b.hap_bases, b.hap_off, b.hap_len,
                                   b.reg_read_start, b.reg_hap_start, b.name + "-prepared")
        same, sresc = engine.compute_batch(pbatch)
        assert np.array_equal(got, same) and np.array_equal(gresc, sresc)
        want, wresc, _ = oracle.compute(pbatch.as_flat(), threads=4)
        assert_parity(got, gresc, want, wresc, what="raw " + b.name)
    st = engine.stats()
    assert st["kernel_launches"] > 0
    with pytest.raises(fgdp_b200.PairHMMError):
        engine.compute_batch_raw(b, bq, mapq, iq, dq, fgdp_b200.pairhmm.ReadModel(7, 18, 10))


def test_raw_read_entry_point_large_batch_multithreaded_pack(engine):
    """> 65536 reads: the flat raw entry point builds its views on several host threads and the call spans several chunks"""
    b = synth.config2(400)
    assert b.n_reads > 70000
    bq, mapq, _, _ = _raw_quals(b, 9, False)
    got, gresc = engine.compute_batch_raw(b, bq, mapq)
    pb, pi, pd, pg = engine.prepare_reads(b, bq, mapq)
    pbatch = synth.RegionBatch(b.read_bases, pb, pi, pd, pg, b.read_off, b.read_len, b.hap_bases, b.hap_off, b.hap_len,
                               b.reg_read_start, b.reg_hap_start, b.name + "-prepared")
    same, sresc = engine.compute_batch(pbatch)
    assert np.array_equal(got, same) and np.array_equal(gresc, sresc)


def test_handles_on_concurrent_host_threads(oracle):
    """SURVEY.md 8(b) threading: one handle per host thread (as one PairHMM per Mutect2Cpp worker, M2/main.cpp:199), calls
    from different threads fully concurrent (ctypes drops the GIL inside the library)."""
    import threading
    batches = [synth.config2(6, seed_extra=k) for k in range(4)]
    wants = [oracle.compute(b.as_flat(), threads=4)[:2] for b in batches]
    results, errors = [None] * 4, []

    def worker(k):
        try:
            e = fgdp_b200.Engine()
            try:
                for _ in range(3):
                    results[k] = e.compute_batch(batches[k])
            finally:
                e.close()
        except Exception as ex:   # noqa: BLE001
            errors.append(ex)

    th = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errors, errors
    for k in range(4):
        assert_parity(results[k][0], results[k][1], wants[k][0], wants[k][1], what="thread %d" % k)


def test_long_reads_row_stripes(oracle):
    """reads above 384 bases go through the row-striped kernel (long.cu), alone and mixed with ordinary reads, under every
    engine mode; the per-region call and the batch call agree"""
    for b in (synth.long_reads(), synth.long_reads(uniform_gcp=False, profile="lowq"), synth.long_reads(seed=99, profile="lowq")):
        want, wresc, _ = oracle.compute(b.as_flat(), threads=4)
        assert 0 < wresc.sum() < wresc.size
        for kw in ({}, {"prefix_sharing": 2, "unit_haps": 32}, {"force_general": True}, {"exact_fp32": True}, {"dedup_reads": False}):
            e = fgdp_b200.Engine(**kw)
            try:
                got, gresc = e.compute_batch(b)
                assert_parity(got, gresc, want, wresc, what="%s %s" % (b.name, kw))
                e.resident_load(b); e.resident_run()
                got2, gresc2 = e.resident_fetch()
                assert np.array_equal(got, got2) and np.array_equal(gresc, gresc2)
            finally:
                e.close()
        e = fgdp_b200.Engine(use_double=True)
        try:
            got_d, resc_d = e.compute_batch(b)
        finally:
            e.close()
        want_d = oracle.compute(b.as_flat(), threads=4, use_double=True)[0]
        fin = np.isfinite(want_d)
        assert resc_d.all() and np.array_equal(np.isfinite(got_d), fin) and np.abs(got_d[fin] - want_d[fin]).max() < 1e-9
    # one region through pairhmm_cuda_compute
    b = synth.long_reads()
    e = fgdp_b200.Engine()
    try:
        off = b.region_out_offsets()
        want, wresc, _ = oracle.compute(b.as_flat(), threads=4)
        for g in range(b.n_regions):
            reads, haps = _region_lists(b, g)
            got, gresc = e.compute(reads, haps)
            assert_parity(got, gresc, want[off[g]:off[g + 1]], wresc[off[g]:off[g + 1]], what="long region %d" % g)
    finally:
        e.close()
