import numpy as np

TOL = 1e-4   # BASELINE.json north_star: |dlog10| <= 1e-4 against the reference on every pair


def assert_parity(got, got_resc, want, want_resc, tol=TOL, what=""):
    got = np.asarray(got).reshape(-1); want = np.asarray(want).reshape(-1)
    assert got.shape == want.shape, what
    fin = np.isfinite(want)
    assert np.array_equal(np.isfinite(got), fin), what + ": non-finite pattern differs"
    assert np.array_equal(got[~fin], want[~fin]) or np.all(np.isnan(got[~fin]) == np.isnan(want[~fin])), what
    if fin.any():
        d = np.abs(got[fin] - want[fin]).max()
        assert d <= tol, "%s: max |dlog10| = %g > %g" % (what, d, tol)
    if want_resc is not None:
        bad = np.nonzero(np.asarray(got_resc).reshape(-1) != np.asarray(want_resc).reshape(-1))[0]
        assert bad.size == 0, "%s: rescue set differs at %d pairs, first %s" % (what, bad.size, bad[:5])


def load_golden(name):
    """tests/golden/<name>.npz -> (RegionBatch, log10, rescued, log10_use_double)"""
    import os
    from fgdp_b200 import synth
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz"))
    b = synth.RegionBatch(*[z[f] for f in ("read_bases", "base_q", "ins_q", "del_q", "gcp", "read_off", "read_len",
                                            "hap_bases", "hap_off", "hap_len", "reg_read_start", "reg_hap_start")], name)
    return b, z["log10"], z["rescued"], z["log10_use_double"]


GOLDEN = ["c1_clean", "c1_lowq", "edge", "c3_small", "c4_small", "c2_small", "long_reads", "long_reads_gcp"]


def raw_quals(b, seed, with_indel):
    """raw (BAM-like) qualities for the reads of batch b: base qualities 2..41, mapping quality 0..60 per read;
    with_indel: False = no BI / BD arrays, True = every read has them, "mixed" = every other read (has_indel_q says which)"""
    rng = np.random.default_rng(seed)
    n = b.read_bases.shape[0]
    bq = rng.integers(2, 42, n).astype(np.uint8)
    mapq = rng.integers(0, 61, b.n_reads).astype(np.int32)
    iq = rng.integers(0, 61, n).astype(np.uint8) if with_indel else None
    dq = rng.integers(0, 61, n).astype(np.uint8) if with_indel else None
    return bq, mapq, iq, dq


def repeat_rich_batch(seed=5, max_len=384):
    """reads full of homopolymers and short tandem repeats (what the PCR error model keys on) and N, lengths 1..max_len"""
    from fgdp_b200 import synth
    rng = np.random.default_rng(seed)
    lens = [1, 2, 3, 8, 9, 31, 32, 33, 150, 151, 250, 384] + [int(x) for x in rng.integers(10, 300, 60)]
    if max_len > 384:
        lens += [385, 450, 599, max_len]
    reads = []
    for n in lens:
        s = []
        while len(s) < n:
            unit = rng.choice(list(b"ACGT"), int(rng.integers(1, 9)))
            s.extend(list(unit) * int(rng.integers(1, 12)))
            if rng.random() < 0.3:
                s.append(ord("N"))
        reads.append(np.array(s[:n], np.uint8))
    b = synth.config1("clean")
    return synth.RegionBatch(np.concatenate(reads), np.zeros(sum(lens), np.uint8), np.zeros(sum(lens), np.uint8),
                             np.zeros(sum(lens), np.uint8), np.zeros(sum(lens), np.uint8),
                             np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64), np.array(lens, np.int32),
                             b.hap_bases, b.hap_off, b.hap_len, np.array([0, len(lens)], np.int64),
                             np.array([0, b.n_haps], np.int64), "repeat-rich")
