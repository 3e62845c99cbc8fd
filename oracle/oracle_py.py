"""TEST INFRASTRUCTURE ONLY: ctypes loaders for the CPU checkers.

  * ``load_oracle()``    -> oracle/libpairhmm_oracle.so (our scalar C restatement, pairhmm_oracle.c)
  * ``load_reference()`` -> oracle/_ref/libref_pairhmm.so (the reference's own sources compiled by
    oracle/Makefile; present only if it was built in the container that has /root/reference)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg import this.
The product package never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MM_COUNT = 32640

_u8p = np.ctypeslib.ndpointer(np.uint8, flags="C")
_i64p = np.ctypeslib.ndpointer(np.int64, flags="C")
_i32p = np.ctypeslib.ndpointer(np.int32, flags="C")
_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C")
_f64p = np.ctypeslib.ndpointer(np.float64, flags="C")

_COMPUTE_ARGS = [_u8p, _u8p, _u8p, _u8p, _u8p, _i64p, _i32p, C.c_int, _u8p, _i64p, _i32p, C.c_int,
                 _u32p, _u32p, C.c_int64, C.c_int, C.c_int, C.c_int, _f64p, C.c_void_p]


def build(ref=True):
    """make the oracle (+ the reference .so when /root/reference is here)."""
    subprocess.run(["make", "-s", "-C", HERE, "libpairhmm_oracle.so"] + (["ref"] if ref else []), check=True)


class _Lib:
    def __init__(self, path, prefix):
        self.path = path
        self.lib = C.CDLL(path)
        self.prefix = prefix
        self._compute = getattr(self.lib, prefix + "_compute")
        self._compute.argtypes = _COMPUTE_ARGS
        self._compute.restype = C.c_double
        self._tables = getattr(self.lib, prefix + "_tables")
        self._tables.argtypes = [C.c_void_p] * 4 + [C.c_int]
        self._tables.restype = None

    def tables(self):
        pf = np.zeros(128, np.float32); mf = np.zeros(MM_COUNT, np.float32)
        pd = np.zeros(128, np.float64); md = np.zeros(MM_COUNT, np.float64)
        self._tables(pf.ctypes.data, mf.ctypes.data, pd.ctypes.data, md.ctypes.data, MM_COUNT)
        return pf, mf, pd, md

    def compute(self, b, use_double=False, threads=1, tables_in_timed_region=False, want_rescued=True):
        """b: a FlatBatch (see synth.py). Returns (log10[n_pairs] f64, rescued[n_pairs] u8, seconds)."""
        n = int(b.pair_read.shape[0])
        out = np.zeros(n, np.float64)
        resc = np.zeros(n, np.uint8)
        if n == 0:
            return out, resc, 0.0
        secs = self._compute(b.read_bases, b.base_q, b.ins_q, b.del_q, b.gcp, b.read_off, b.read_len,
                             int(b.read_len.shape[0]), b.hap_bases, b.hap_off, b.hap_len,
                             int(b.hap_len.shape[0]), b.pair_read, b.pair_hap, n, int(use_double),
                             int(threads), int(tables_in_timed_region), out,
                             resc.ctypes.data if want_rescued else None)
        return out, resc, secs


class Oracle(_Lib):
    def __init__(self):
        super().__init__(os.path.join(HERE, "libpairhmm_oracle.so"), "oracle")
        self.lib.oracle_forward_float_pairs.argtypes = [_u8p, _u8p, _u8p, _u8p, _u8p, _i64p, _i32p, _u8p, _i64p,
                                                        _i32p, _u32p, _u32p, C.c_int64, C.c_void_p]
        self.lib.oracle_forward_float_pairs.restype = None
        self.lib.oracle_epilogue.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_double, C.c_double, C.c_int,
                                             C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.oracle_epilogue.restype = None
        self.lib.oracle_prepare_read.argtypes = [C.c_void_p] * 4 + [C.c_int] * 5 + [C.c_void_p] * 4
        self.lib.oracle_prepare_read.restype = None
        self.lib.oracle_tandem_repeat_length.argtypes = [C.c_void_p, C.c_int, C.c_int]

    def epilogue(self, m, read_len, likelihood_cap=-4.5, max_error_per_base=0.02, natural_log=False):
        """m[n_reads, n_haps] (copied) -> (normalised m, best, second_best, best_hap, poorly_modeled)"""
        m = np.array(m, dtype=np.float64, order="C", copy=True)
        n, nh = m.shape
        rl = np.ascontiguousarray(read_len, np.int32)
        best = np.zeros(n); second = np.zeros(n); bh = np.zeros(n, np.int32); pm = np.zeros(n, np.uint8)
        self.lib.oracle_epilogue(m.ctypes.data, n, nh, rl.ctypes.data, float(likelihood_cap), float(max_error_per_base),
                                 int(natural_log), best.ctypes.data, second.ctypes.data, bh.ctypes.data, pm.ctypes.data)
        return m, best, second, bh, pm

    def prepare_reads(self, read_bases, base_q, mapq, read_off, read_len, ins_q=None, del_q=None, pcr_rate_factor=3,
                      base_q_threshold=18, constant_gcp=10, has_indel_q=None):
        """the caller's quality pre-pass (modifyReadQualities) on flat read arrays -> (base_q, ins_q, del_q, gcp);
        has_indel_q[i] == 0: read i has no BI / BD arrays even though ins_q / del_q are given"""
        bases = np.ascontiguousarray(read_bases, np.uint8); bq = np.ascontiguousarray(base_q, np.uint8)
        outs = [np.zeros(bq.shape[0], np.uint8) for _ in range(4)]
        iq = None if ins_q is None else np.ascontiguousarray(ins_q, np.uint8)
        dq = None if del_q is None else np.ascontiguousarray(del_q, np.uint8)
        for i in range(len(read_len)):
            o, n = int(read_off[i]), int(read_len[i])
            own = iq is not None and (has_indel_q is None or has_indel_q[i])
            self.lib.oracle_prepare_read(bases.ctypes.data + o, bq.ctypes.data + o,
                                         iq.ctypes.data + o if own else None,
                                         dq.ctypes.data + o if own else None, n, int(mapq[i]), pcr_rate_factor,
                                         base_q_threshold, constant_gcp, *[x.ctypes.data + o for x in outs])
        return tuple(outs)

    def tandem_repeat_length(self, bases, offset):
        b = np.frombuffer(bytes(bases), np.uint8)
        return int(self.lib.oracle_tandem_repeat_length(b.ctypes.data, len(b), int(offset)))

    def forward_float(self, b):
        """raw 2^120-scaled FP32 sums (FTZ on, no FMA) for every pair of a FlatPairs"""
        n = int(b.pair_read.shape[0])
        out = np.zeros(n, np.float32)
        self.lib.oracle_forward_float_pairs(b.read_bases, b.base_q, b.ins_q, b.del_q, b.gcp, b.read_off, b.read_len,
                                            b.hap_bases, b.hap_off, b.hap_len, b.pair_read, b.pair_hap, n,
                                            out.ctypes.data)
        return out


class Reference(_Lib):
    """oracle/_ref/libref_pairhmm.so (canonical: default -ffp-contract, as the reference's CMake builds it) or, with
    nofma=True, libref_pairhmm_nofma.so (the same sources with -ffp-contract=off): the two disagree on `< 1e-28f` only
    for threshold-ambiguous pairs"""

    def __init__(self, nofma=False):
        super().__init__(os.path.join(HERE, "_ref", "libref_pairhmm_nofma.so" if nofma else "libref_pairhmm.so"), "ref_pairhmm")
        self.lib.ref_pairhmm_isa.restype = C.c_int

    def isa(self):
        return int(self.lib.ref_pairhmm_isa())


class ReferenceEpilogue:
    """oracle/_ref/libref_epilogue.so: the reference's own AlleleLikelihoods<E, A> (normalizeLikelihoods,
    filterPoorlyModeledEvidence, switchToNaturalLog) behind oracle/ref_epilogue_driver.cpp"""

    def __init__(self):
        self.lib = C.CDLL(os.path.join(HERE, "_ref", "libref_epilogue.so"))
        self.lib.ref_epilogue.argtypes = [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_double, C.c_double, C.c_int,
                                          C.c_void_p]

    def epilogue(self, m, read_len, likelihood_cap=-4.5, max_error_per_base=0.02, natural_log=False):
        """-> (matrix after normalisation [and log conversion], removed[n_reads])"""
        m = np.array(m, dtype=np.float64, order="C", copy=True)
        n, nh = m.shape
        rl = np.ascontiguousarray(read_len, np.int32)
        removed = np.zeros(n, np.uint8)
        self.lib.ref_epilogue(m.ctypes.data, n, nh, rl.ctypes.data, float(likelihood_cap), float(max_error_per_base),
                              int(natural_log), removed.ctypes.data)
        return m, removed


def reference_epilogue_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libref_epilogue.so"))


def load_oracle():
    p = os.path.join(HERE, "libpairhmm_oracle.so")
    if not os.path.exists(p):
        build(ref=False)
    return Oracle()


def reference_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libref_pairhmm.so"))


def load_reference():
    return Reference()


def reference_nofma_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libref_pairhmm_nofma.so"))


def load_reference_nofma():
    return Reference(nofma=True)


# ---- Smith-Waterman ----
SW_SOFTCLIP, SW_INDEL, SW_LEADING_INDEL, SW_IGNORE = 9, 10, 11, 12


class SWPairs:
    """flat list of (ref, alt) sequence pairs"""

    def __init__(self, refs, alts):
        cat = lambda xs: np.ascontiguousarray(np.concatenate(xs) if len(xs) else np.zeros(0, np.uint8), np.uint8)
        self.n = len(refs)
        self.ref_len = np.array([len(x) for x in refs], np.int32); self.alt_len = np.array([len(x) for x in alts], np.int32)
        self.ref_off = np.concatenate([[0], np.cumsum(self.ref_len)[:-1]]).astype(np.int64) if self.n else np.zeros(0, np.int64)
        self.alt_off = np.concatenate([[0], np.cumsum(self.alt_len)[:-1]]).astype(np.int64) if self.n else np.zeros(0, np.int64)
        self.ref_bases = cat([np.frombuffer(bytes(x), np.uint8) for x in refs])
        self.alt_bases = cat([np.frombuffer(bytes(x), np.uint8) for x in alts])
        self.cigar_stride = int(2 * max([1] + list(self.ref_len) + list(self.alt_len)))   # IntelSmithWaterman.cpp:12

    def cells(self):
        return int((self.ref_len.astype(np.int64) * self.alt_len.astype(np.int64)).sum())


def _sw_call(fn, p, params, strategy, extra=()):
    cig = np.zeros((p.n, p.cigar_stride), np.uint8); off = np.zeros(p.n, np.int32)
    r = fn(p.ref_bases.ctypes.data, p.ref_off.ctypes.data, p.ref_len.ctypes.data, p.alt_bases.ctypes.data,
           p.alt_off.ctypes.data, p.alt_len.ctypes.data, p.n, *[int(x) for x in params], int(strategy), cig.ctypes.data,
           p.cigar_stride, off.ctypes.data, *extra)
    cigars = [bytes(row).split(b"\0")[0].decode() for row in cig]
    return cigars, off, r


_SW_ARGS = [C.c_void_p] * 6 + [C.c_int64] + [C.c_int] * 5 + [C.c_void_p, C.c_int, C.c_void_p]


def oracle_sw(p, params=(10, -15, -30, -5), strategy=SW_SOFTCLIP):
    lib = load_oracle().lib
    lib.oracle_sw_align_batch.argtypes = _SW_ARGS; lib.oracle_sw_align_batch.restype = None
    c, o, _ = _sw_call(lib.oracle_sw_align_batch, p, params, strategy)
    return c, o


def reference_sw_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libref_sw.so"))


def reference_sw(p, params=(10, -15, -30, -5), strategy=SW_SOFTCLIP, threads=1):
    """the reference's own SmithWaterman_align on every pair -> (cigars, offsets, seconds)"""
    lib = C.CDLL(os.path.join(HERE, "_ref", "libref_sw.so"))
    lib.ref_sw_align_batch.argtypes = _SW_ARGS + [C.c_int]; lib.ref_sw_align_batch.restype = C.c_double
    return _sw_call(lib.ref_sw_align_batch, p, params, strategy, (int(threads),))


# ---- the reference's caller side (oracle/_ref/libref_m2.so) ----
class ReferenceM2:
    """PairHMMLikelihoodCalculationEngine / VectorLoglessPairHMM / SAMRecord / Haplotype / AlleleLikelihoods ... of the
    reference, compiled unmodified (oracle/Makefile, ref_m2_driver.cpp). `replay=True` loads integration/replay/libm2_replay.so
    instead: the same harness with integration/CudaPairHMM.cpp linked in (use_cuda=True swaps the engine's pairHMM member)."""

    def __init__(self, replay=False):
        path = (os.path.join(HERE, "..", "integration", "replay", "libm2_replay.so") if replay
                else os.path.join(HERE, "_ref", "libref_m2.so"))
        self.path = os.path.abspath(path)
        self.lib = C.CDLL(self.path)
        self.lib.ref_m2_last_error.restype = C.c_char_p
        self.lib.ref_m2_log10_min_true_likelihood.argtypes = [C.c_int, C.c_double]
        self.lib.ref_m2_log10_min_true_likelihood.restype = C.c_double
        self.lib.ref_m2_prepass.argtypes = [C.c_void_p] * 8 + [C.c_int64, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 4
        self.lib.ref_m2_region.argtypes = ([C.c_void_p] * 9 + [C.c_int64, C.c_int] + [C.c_void_p] * 3 + [C.c_int, C.c_int] +
                                           [C.c_int] * 3 + [C.c_double] + [C.c_int] * 3 + [C.c_void_p] * 3)

    @staticmethod
    def _reads(read_bases, base_q, ins_q, del_q, has_indel_q, mapq, read_off, read_len):
        n = len(read_len)
        bases = np.ascontiguousarray(read_bases, np.uint8); bq = np.ascontiguousarray(base_q, np.uint8)
        iq = None if ins_q is None else np.ascontiguousarray(ins_q, np.uint8)
        dq = None if del_q is None else np.ascontiguousarray(del_q, np.uint8)
        if has_indel_q is None:
            has = np.full(n, 0 if iq is None else 1, np.uint8)
        else:
            has = np.ascontiguousarray(has_indel_q, np.uint8)
        mq = np.ascontiguousarray(mapq, np.int32); ro = np.ascontiguousarray(read_off, np.int64); rl = np.ascontiguousarray(read_len, np.int32)
        ptr = lambda a: None if a is None else a.ctypes.data
        return (bases, bq, iq, dq, has, mq, ro, rl), [ptr(bases), ptr(bq), ptr(iq), ptr(dq), ptr(has), ptr(mq)], ptr(ro), ptr(rl), n

    def log10_min_true_likelihood(self, read_len, max_error_per_base=0.02):
        return float(self.lib.ref_m2_log10_min_true_likelihood(int(read_len), float(max_error_per_base)))

    def prepare_reads(self, read_bases, base_q, mapq, read_off, read_len, ins_q=None, del_q=None, has_indel_q=None,
                      pcr_rate_factor=3, base_q_threshold=18, constant_gcp=10):
        """modifyReadQualities + buildGapContinuationPenalties -> (base_q, ins_q, del_q, gcp), laid out like base_q"""
        keep, ptrs, ro, rl, n = self._reads(read_bases, base_q, ins_q, del_q, has_indel_q, mapq, read_off, read_len)
        outs = [np.zeros(keep[1].shape[0], np.uint8) for _ in range(4)]
        st = self.lib.ref_m2_prepass(*ptrs, ro, rl, n, int(pcr_rate_factor), int(base_q_threshold), int(constant_gcp),
                                     *[x.ctypes.data for x in outs])
        if st != 0:
            raise RuntimeError("ref_m2_prepass: " + self.lib.ref_m2_last_error().decode())
        return tuple(outs)

    def region(self, read_bases, base_q, mapq, read_off, read_len, hap_bases, hap_off, hap_len, read_sample=None, n_samples=1,
               ins_q=None, del_q=None, has_indel_q=None, hap_ref=0, pcr_rate_factor=3, base_q_threshold=18, constant_gcp=10,
               log10_mismapping=-4.5, use_double=False, stage=0, use_cuda=False):
        """one active region through PairHMMLikelihoodCalculationEngine -> (matrix[n_reads, n_haps] in INPUT haplotype order,
        removed[n_reads], allele_order[n_haps])"""
        keep, ptrs, ro, rl, n = self._reads(read_bases, base_q, ins_q, del_q, has_indel_q, mapq, read_off, read_len)
        rs = np.zeros(n, np.int32) if read_sample is None else np.ascontiguousarray(read_sample, np.int32)
        hb = np.ascontiguousarray(hap_bases, np.uint8); ho = np.ascontiguousarray(hap_off, np.int64); hl = np.ascontiguousarray(hap_len, np.int32)
        nh = len(hl)
        out = np.zeros((n, nh), np.float64); removed = np.zeros(n, np.uint8); order = np.zeros(nh, np.int32)
        st = self.lib.ref_m2_region(*ptrs, rs.ctypes.data, ro, rl, n, int(n_samples), hb.ctypes.data, ho.ctypes.data, hl.ctypes.data, nh,
                                    int(hap_ref), int(pcr_rate_factor), int(base_q_threshold), int(constant_gcp),
                                    float(log10_mismapping), int(use_double), int(stage), int(use_cuda), out.ctypes.data,
                                    removed.ctypes.data, order.ctypes.data)
        if st != 0:
            raise RuntimeError("ref_m2_region: " + self.lib.ref_m2_last_error().decode())
        return out, removed, order


def reference_m2_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libref_m2.so"))


def replay_available():
    return os.path.exists(os.path.join(HERE, "..", "integration", "replay", "libm2_replay.so"))
