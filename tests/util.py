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
