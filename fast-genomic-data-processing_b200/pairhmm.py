"""ctypes binding of libpairhmm_cuda + a host-side mirror of the reference's PairHMM plugin interface.

Reference interface mirrored (M2 = deepmutect/Mutect2Cpp-master/src):
  PairHMM::initialize(readMaxLength, haplotypeMaxLength)        M2/utils/pairhmm/PairHMM.h:37, PairHMM.cpp:14-27
  VectorLoglessPairHMM::initialize(haplotypes, reads, ...)      M2/utils/pairhmm/VectorLoglessPairHMM.cpp:18-41
  VectorLoglessPairHMM::computeLog10Likelihoods(matrix, reads, gcp)   VectorLoglessPairHMM.cpp:43-148
      result convention: read-major mLogLikelihoodArray_1D[r*nHaps+h], then matrix.set(h, r, v)
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

EXPORTED_SYMBOLS = [
    "pairhmm_cuda_abi_version", "pairhmm_cuda_device_count", "pairhmm_cuda_default_config", "pairhmm_cuda_create",
    "pairhmm_cuda_destroy", "pairhmm_cuda_initialize", "pairhmm_cuda_compute", "pairhmm_cuda_submit",
    "pairhmm_cuda_wait", "pairhmm_cuda_compute_batch", "pairhmm_cuda_compute_flat", "pairhmm_cuda_compute_pairs",
    "pairhmm_cuda_resident_load", "pairhmm_cuda_resident_run", "pairhmm_cuda_resident_fetch",
    "pairhmm_cuda_resident_fetch_raw",
    "pairhmm_cuda_compute_normalized", "pairhmm_cuda_compute_batch_normalized", "pairhmm_cuda_compute_flat_normalized",
    "pairhmm_cuda_default_read_model", "pairhmm_cuda_compute_raw", "pairhmm_cuda_compute_batch_raw",
    "pairhmm_cuda_compute_flat_raw", "pairhmm_cuda_prepare_reads_flat",
    "pairhmm_cuda_stats_get", "pairhmm_cuda_stats_reset", "pairhmm_cuda_last_error", "pairhmm_cuda_tables",
    "pairhmm_cuda_prefix_plan", "pairhmm_cuda_device_plan", "pairhmm_cuda_forward_pairs_noalloc",
]
MM_COUNT = 8256


class PairHMMError(RuntimeError):
    """Non-zero status from the C ABI (the C++ adapter throws std::runtime_error for the same cases, so
    that Mutect2Cpp's catch(std::exception&) at M2/main.cpp:303-310 keeps its behaviour)."""

    def __init__(self, status, msg):
        super().__init__("libpairhmm_cuda status %d: %s" % (status, msg))
        self.status = status


class Config(C.Structure):
    _fields_ = [("device", C.c_int32), ("use_double", C.c_int32), ("exact_fp32", C.c_int32),
                ("dedup_reads", C.c_int32), ("pack_threads", C.c_int32), ("chunk_cells", C.c_int64),
                ("reserved", C.c_int32 * 8),
                # ABI version 2
                ("device_mask", C.c_uint32), ("batch_window_us", C.c_int32), ("reserved2", C.c_int32 * 6)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("calls", "regions", "reads", "unique_reads", "pairs", "unique_pairs",
                                          "cells", "rescued_pairs", "h2d_bytes", "d2h_bytes", "kernel_launches")] + \
               [(n, C.c_double) for n in ("gpu_ms_float", "gpu_ms_double", "gpu_ms_total", "host_ms_pack",
                                          "host_ms_scatter", "host_ms_total")] + [("cells_swept", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class PhmmRead(C.Structure):
    _fields_ = [("bases", C.c_void_p), ("base_q", C.c_void_p), ("ins_q", C.c_void_p), ("del_q", C.c_void_p),
                ("gcp", C.c_void_p), ("len", C.c_int32)]


class PhmmHap(C.Structure):
    _fields_ = [("bases", C.c_void_p), ("len", C.c_int32)]


class PhmmRawRead(C.Structure):
    """phmm_raw_read: a read as the BAM record holds it (ins_q / del_q NULL = default 45)"""
    _fields_ = [("bases", C.c_void_p), ("base_q", C.c_void_p), ("ins_q", C.c_void_p), ("del_q", C.c_void_p),
                ("len", C.c_int32), ("mapq", C.c_int32)]


class PhmmRegion(C.Structure):
    _fields_ = [("reads", C.POINTER(PhmmRead)), ("n_reads", C.c_int32), ("haps", C.POINTER(PhmmHap)),
                ("n_haps", C.c_int32), ("log10_out", C.c_void_p), ("rescued_out", C.c_void_p)]


class Epilogue(C.Structure):
    """pairhmm_cuda_epilogue: normalizeLikelihoods cap, filterPoorlyModeledEvidence rate, switchToNaturalLog"""
    _fields_ = [("likelihood_cap", C.c_double), ("max_error_per_base", C.c_double), ("natural_log", C.c_int32),
                ("reserved", C.c_int32)]

    def __init__(self, likelihood_cap=-4.5, max_error_per_base=0.02, natural_log=False):
        super().__init__(float(likelihood_cap), float(max_error_per_base), 1 if natural_log else 0, 0)


class ReadModel(C.Structure):
    """pairhmm_cuda_read_model: PCR error model (rate factor), base quality threshold, constant gap-continuation penalty"""
    _fields_ = [("pcr_rate_factor", C.c_int32), ("base_quality_threshold", C.c_int32), ("constant_gcp", C.c_int32),
                ("reserved", C.c_int32)]

    def __init__(self, pcr_rate_factor=3, base_quality_threshold=18, constant_gcp=10):
        super().__init__(int(pcr_rate_factor), int(base_quality_threshold), int(constant_gcp), 0)


# phmm_read_summary as a numpy record
SUMMARY_DTYPE = np.dtype([("best", np.float64), ("second_best", np.float64), ("best_hap", np.int32),
                          ("poorly_modeled", np.uint8), ("pad", np.uint8, (3,))])
assert SUMMARY_DTYPE.itemsize == 24


def lib_path():
    # PAIRHMM_CUDA_LIB selects another build of the same library (kernel-variant experiments)
    return os.environ.get("PAIRHMM_CUDA_LIB") or os.path.join(_HERE, "libpairhmm_cuda.so")


_LIB = None


def load_library():
    """dlopen the in-tree library; raises (never falls back) if it has not been built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    p = lib_path()
    if not os.path.exists(p):
        raise PairHMMError(-3, "%s not built: run `make -C %s/csrc` (or __graft_entry__.build())" % (p, _HERE))
    lib = C.CDLL(p)
    vp, i32, i64 = C.c_void_p, C.c_int, C.c_int64
    lib.pairhmm_cuda_abi_version.restype = C.c_int
    lib.pairhmm_cuda_device_count.restype = C.c_int
    lib.pairhmm_cuda_default_config.argtypes = [C.POINTER(Config)]
    lib.pairhmm_cuda_default_config.restype = None
    lib.pairhmm_cuda_create.argtypes = [C.POINTER(Config), C.POINTER(vp)]
    lib.pairhmm_cuda_destroy.argtypes = [vp]
    lib.pairhmm_cuda_initialize.argtypes = [vp, i32, i32]
    lib.pairhmm_cuda_compute.argtypes = [vp, C.POINTER(PhmmRead), i32, C.POINTER(PhmmHap), i32, vp, vp]
    lib.pairhmm_cuda_submit.argtypes = [vp, C.POINTER(PhmmRead), i32, C.POINTER(PhmmHap), i32, vp, vp, C.POINTER(i64)]
    lib.pairhmm_cuda_wait.argtypes = [vp, i64]
    lib.pairhmm_cuda_compute_batch.argtypes = [vp, C.POINTER(PhmmRegion), i64]
    flat = [vp] * 5 + [vp, vp, i64, vp, vp, vp, i64, vp, vp, i64]
    lib.pairhmm_cuda_compute_flat.argtypes = [vp] + flat + [vp, vp]
    lib.pairhmm_cuda_compute_pairs.argtypes = [vp, C.POINTER(PhmmRead), i32, C.POINTER(PhmmHap), i32, vp, vp,
                                               C.c_size_t, vp, vp]
    lib.pairhmm_cuda_compute_normalized.argtypes = [vp, C.POINTER(PhmmRead), i32, C.POINTER(PhmmHap), i32,
                                                    C.POINTER(Epilogue), vp, vp, vp]
    lib.pairhmm_cuda_compute_batch_normalized.argtypes = [vp, C.POINTER(PhmmRegion), i64, C.POINTER(Epilogue), vp]
    lib.pairhmm_cuda_compute_flat_normalized.argtypes = [vp] + flat + [C.POINTER(Epilogue), vp, vp, vp]
    lib.pairhmm_cuda_default_read_model.argtypes = [C.POINTER(ReadModel)]
    lib.pairhmm_cuda_default_read_model.restype = None
    lib.pairhmm_cuda_compute_flat_raw.argtypes = [vp] + flat + [C.POINTER(ReadModel), vp, vp]
    lib.pairhmm_cuda_compute_raw.argtypes = [vp, C.POINTER(PhmmRawRead), i32, C.POINTER(PhmmHap), i32, C.POINTER(ReadModel), vp, vp]
    lib.pairhmm_cuda_prepare_reads_flat.argtypes = [vp] * 6 + [vp, vp, i64, C.POINTER(ReadModel), vp, vp, vp, vp]
    lib.pairhmm_cuda_resident_load.argtypes = [vp] + flat
    lib.pairhmm_cuda_resident_run.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.pairhmm_cuda_resident_fetch.argtypes = [vp, vp, vp]
    lib.pairhmm_cuda_resident_fetch_raw.argtypes = [vp, vp, i64]
    lib.pairhmm_cuda_stats_get.argtypes = [vp, C.POINTER(Stats)]
    lib.pairhmm_cuda_stats_reset.argtypes = [vp]
    lib.pairhmm_cuda_last_error.argtypes = [vp]
    lib.pairhmm_cuda_last_error.restype = C.c_char_p
    lib.pairhmm_cuda_tables.argtypes = [vp, vp, vp, vp]
    _LIB = lib
    return lib


def device_count():
    return int(load_library().pairhmm_cuda_device_count())


def prefix_plan(haps, unit_haps=0):
    """the engine's prefix-sharing plan for one region's haplotypes (pairhmm_cuda_prefix_plan; host logic, no device):
    dict of order, t_start, restore_slot, t_save, save_slot (arrays in sweep order) and shared_cols"""
    lib = load_library()
    haps = [np.ascontiguousarray(h, np.uint8) for h in haps]
    n = len(haps)
    arr = (PhmmHap * max(1, n))()
    for i, hp in enumerate(haps):
        arr[i].bases = hp.ctypes.data if hp.size else None
        arr[i].len = int(hp.size)
    order, ts, tv, ss = (np.zeros(n, np.uint32) for _ in range(4))
    rs = np.zeros(n, np.int32)
    shared = C.c_uint64(0)
    lib.pairhmm_cuda_prefix_plan.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_void_p] * 5 + [C.POINTER(C.c_uint64)]
    st = lib.pairhmm_cuda_prefix_plan(C.cast(arr, C.c_void_p), n, int(unit_haps), order.ctypes.data, ts.ctypes.data,
                                      rs.ctypes.data, tv.ctypes.data, ss.ctypes.data, C.byref(shared))
    if st != 0:
        raise PairHMMError(st, "pairhmm_cuda_prefix_plan")
    return {"order": order, "t_start": ts, "restore_slot": rs, "t_save": tv, "save_slot": ss, "shared_cols": int(shared.value)}


def device_plan(read_len_sum, hap_len_sum, n_reads, n_devices):
    """the multi-device LPT plan of a batch (pairhmm_cuda_device_plan; host logic, no device): arrays part_region,
    part_first_read, part_n_reads, part_device in (region, first read) order"""
    lib = load_library()
    rl = np.ascontiguousarray(read_len_sum, np.uint64); hl = np.ascontiguousarray(hap_len_sum, np.uint64)
    nr = np.ascontiguousarray(n_reads, np.int32)
    n = len(nr)
    cap = n + 8 * int(n_devices) * 4 + 16
    pr = np.zeros(cap, np.int64); pf = np.zeros(cap, np.int32); pn = np.zeros(cap, np.int32); pd = np.zeros(cap, np.int32)
    got = C.c_int64(0)
    lib.pairhmm_cuda_device_plan.argtypes = [C.c_void_p] * 3 + [C.c_int64, C.c_int, C.c_int64] + [C.c_void_p] * 4 + [C.POINTER(C.c_int64)]
    st = lib.pairhmm_cuda_device_plan(rl.ctypes.data, hl.ctypes.data, nr.ctypes.data, n, int(n_devices), cap, pr.ctypes.data,
                                      pf.ctypes.data, pn.ctypes.data, pd.ctypes.data, C.byref(got))
    if st != 0:
        raise PairHMMError(st, "pairhmm_cuda_device_plan")
    k = int(got.value)
    return pr[:k], pf[:k], pn[:k], pd[:k]


def host_tables():
    """the lookup tables the engine uploads: (ph2pr_f32[128], mm_f32[8256], ph2pr_f64[128], mm_f64[8256])"""
    lib = load_library()
    pf = np.zeros(128, np.float32); mf = np.zeros(MM_COUNT, np.float32)
    pd = np.zeros(128, np.float64); md = np.zeros(MM_COUNT, np.float64)
    lib.pairhmm_cuda_tables(pf.ctypes.data, mf.ctypes.data, pd.ctypes.data, md.ctypes.data)
    return pf, mf, pd, md


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a


class Engine:
    """One libpairhmm_cuda handle (one per host thread, like one PairHMM per Mutect2Cpp worker)."""

    def __init__(self, device=-1, use_double=False, exact_fp32=False, dedup_reads=True, pack_threads=0,
                 chunk_cells=0, force_general=False, lane_group=0, scalar_uniform=False, max_ctas_per_sm=0,
                 prefix_sharing=False, unit_haps=0, rescue_pairs_per_unit=0, device_mask=0, batch_window_us=0,
                 rescue_read_pairing=0):
        self.lib = load_library()
        cfg = Config()
        self.lib.pairhmm_cuda_default_config(C.byref(cfg))
        cfg.device = device; cfg.use_double = int(use_double); cfg.exact_fp32 = int(exact_fp32)
        cfg.dedup_reads = int(dedup_reads); cfg.pack_threads = int(pack_threads); cfg.chunk_cells = int(chunk_cells)
        cfg.reserved[0] = int(force_general)   # never pick the uniform-gcp kernels (testing)
        cfg.reserved[1] = int(lane_group)      # 8 | 16 | 32: force that lane-group width where it fits
        cfg.reserved[3] = int(max_ctas_per_sm) # tuning: cap on resident CTAs per SM
        cfg.reserved[2] = int(scalar_uniform)  # scalar instead of packed (FFMA2) uniform-gcp kernels (testing)
        # same-length haplotypes sharing a prefix are swept from a checkpoint (the trie analogue): True / 1 = when it pays,
        # 2 = always (testing)
        cfg.reserved[4] = int(prefix_sharing)
        cfg.reserved[5] = int(unit_haps)       # haplotypes per work unit (tuning / testing)
        cfg.reserved[6] = int(rescue_pairs_per_unit)   # underflowed pairs per FP64 unit (tuning; 0 = by chunk size)
        cfg.reserved[7] = int(rescue_read_pairing)     # FP64 pass with two reads per warp: 0 = by chunk size, 1 = never, 2 = always (testing)
        cfg.device_mask = int(device_mask)     # != 0: one handle over these CUDA ordinals (multi-GPU inside the library)
        cfg.batch_window_us = int(batch_window_us)   # > 0: blocking region calls of all such handles share one batcher
        self.h = C.c_void_p()
        st = self.lib.pairhmm_cuda_create(C.byref(cfg), C.byref(self.h))
        if st != 0:
            self.h = None
            raise PairHMMError(st, "pairhmm_cuda_create failed (no sm_100 device? this engine has no CPU fallback)")

    def close(self):
        if getattr(self, "h", None):
            self.lib.pairhmm_cuda_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, st):
        if st != 0:
            raise PairHMMError(st, (self.lib.pairhmm_cuda_last_error(self.h) or b"").decode())

    def initialize(self, read_max_len, hap_max_len):
        self._check(self.lib.pairhmm_cuda_initialize(self.h, int(read_max_len), int(hap_max_len)))

    # ---- region form (pointer structs, exactly what the C++ adapter passes) ----
    @staticmethod
    def _structs(reads, haps):
        """reads: list of 5-tuples of uint8 arrays (bases, base_q, ins_q, del_q, gcp); haps: list of arrays."""
        keep = []
        ra = (PhmmRead * max(1, len(reads)))()
        for i, r in enumerate(reads):
            arrs = [_u8(x) for x in r]
            keep.append(arrs)
            n = len(arrs[0])
            if any(len(x) != n for x in arrs):
                raise ValueError("read %d: the five arrays differ in length" % i)
            ra[i] = PhmmRead(*[x.ctypes.data for x in arrs], n)
        ha = (PhmmHap * max(1, len(haps)))()
        for i, hp in enumerate(haps):
            a = _u8(hp); keep.append(a)
            ha[i] = PhmmHap(a.ctypes.data, len(a))
        return ra, ha, keep

    def compute(self, reads, haps):
        """one region -> (log10[n_reads, n_haps] float64, rescued[n_reads, n_haps] uint8)"""
        ra, ha, keep = self._structs(reads, haps)
        out = np.zeros((len(reads), len(haps)), np.float64)
        resc = np.zeros((len(reads), len(haps)), np.uint8)
        self._check(self.lib.pairhmm_cuda_compute(self.h, ra, len(reads), ha, len(haps), out.ctypes.data, resc.ctypes.data))
        return out, resc

    def submit(self, reads, haps):
        ra, ha, keep = self._structs(reads, haps)
        out = np.zeros((len(reads), len(haps)), np.float64)
        resc = np.zeros((len(reads), len(haps)), np.uint8)
        t = C.c_int64()
        self._check(self.lib.pairhmm_cuda_submit(self.h, ra, len(reads), ha, len(haps), out.ctypes.data,
                                                 resc.ctypes.data, C.byref(t)))
        return t.value, out, resc

    def wait(self, ticket):
        self._check(self.lib.pairhmm_cuda_wait(self.h, int(ticket)))

    def compute_pairs(self, reads, haps, read_idx, hap_idx):
        ra, ha, keep = self._structs(reads, haps)
        ri = np.ascontiguousarray(read_idx, np.uint32); hi = np.ascontiguousarray(hap_idx, np.uint32)
        out = np.zeros(len(ri), np.float64); resc = np.zeros(len(ri), np.uint8)
        self._check(self.lib.pairhmm_cuda_compute_pairs(self.h, ra, len(reads), ha, len(haps), ri.ctypes.data,
                                                        hi.ctypes.data, len(ri), out.ctypes.data, resc.ctypes.data))
        return out, resc

    # ---- flat batch form (many regions in one call) ----
    @staticmethod
    def _flat_args(b):
        return [b.read_bases.ctypes.data, b.base_q.ctypes.data, b.ins_q.ctypes.data, b.del_q.ctypes.data,
                b.gcp.ctypes.data, b.read_off.ctypes.data, b.read_len.ctypes.data, b.n_reads,
                b.hap_bases.ctypes.data, b.hap_off.ctypes.data, b.hap_len.ctypes.data, b.n_haps,
                b.reg_read_start.ctypes.data, b.reg_hap_start.ctypes.data, b.n_regions]

    def compute_batch(self, b, out=None, resc=None):
        """b: synth.RegionBatch -> (log10[n_pairs], rescued[n_pairs]) region-major, read-major inside."""
        n = b.n_pairs
        out = np.zeros(n, np.float64) if out is None else out
        resc = np.zeros(n, np.uint8) if resc is None else resc
        self._check(self.lib.pairhmm_cuda_compute_flat(self.h, *self._flat_args(b), out.ctypes.data, resc.ctypes.data))
        return out, resc

    def compute_normalized(self, reads, haps, epilogue=None):
        """one region through the fused epilogue -> (normalised matrix, rescued, per-read summary records)"""
        ep = epilogue or Epilogue()
        ra, ha, keep = self._structs(reads, haps)
        out = np.zeros((len(reads), len(haps)), np.float64)
        resc = np.zeros((len(reads), len(haps)), np.uint8)
        summ = np.zeros(len(reads), SUMMARY_DTYPE)
        self._check(self.lib.pairhmm_cuda_compute_normalized(self.h, ra, len(reads), ha, len(haps), C.byref(ep),
                                                             out.ctypes.data, resc.ctypes.data, summ.ctypes.data))
        return out, resc, summ

    def compute_batch_normalized(self, b, epilogue=None):
        """b: synth.RegionBatch -> (normalised[n_pairs], rescued[n_pairs], summary[n_reads])"""
        ep = epilogue or Epilogue()
        out = np.zeros(b.n_pairs, np.float64); resc = np.zeros(b.n_pairs, np.uint8)
        summ = np.zeros(b.n_reads, SUMMARY_DTYPE)
        self._check(self.lib.pairhmm_cuda_compute_flat_normalized(self.h, *self._flat_args(b), C.byref(ep),
                                                                  out.ctypes.data, resc.ctypes.data, summ.ctypes.data))
        return out, resc, summ

    # ---- raw reads: quality pre-pass on the device ----
    def prepare_reads(self, b, base_q_raw, mapq, ins_q=None, del_q=None, model=None):
        """the pre-pass alone on the reads of RegionBatch b (its bases / offsets) with raw qualities ->
        (base_q, ins_q, del_q, gcp) as the kernels will see them"""
        m = model or ReadModel()
        bq = _u8(base_q_raw); mq = np.ascontiguousarray(mapq, np.int32)
        iq = None if ins_q is None else _u8(ins_q); dq = None if del_q is None else _u8(del_q)
        outs = [np.zeros(bq.shape[0], np.uint8) for _ in range(4)]
        self._check(self.lib.pairhmm_cuda_prepare_reads_flat(
            self.h, b.read_bases.ctypes.data, bq.ctypes.data, None if iq is None else iq.ctypes.data,
            None if dq is None else dq.ctypes.data, mq.ctypes.data, b.read_off.ctypes.data, b.read_len.ctypes.data,
            b.n_reads, C.byref(m), *[o.ctypes.data for o in outs]))
        return tuple(outs)

    def compute_raw(self, reads, haps, model=None):
        """one region of raw reads: reads = list of (bases, base_q, ins_q | None, del_q | None, mapq) ->
        (log10[n_reads, n_haps], rescued[n_reads, n_haps]); the quality pre-pass runs on the device"""
        m = model or ReadModel()
        keep = []
        ra = (PhmmRawRead * max(1, len(reads)))()
        for i, (bases, bq, iq, dq, mapq) in enumerate(reads):
            arrs = [_u8(bases), _u8(bq), None if iq is None else _u8(iq), None if dq is None else _u8(dq)]
            keep.append(arrs)
            ra[i] = PhmmRawRead(*[None if x is None else x.ctypes.data for x in arrs], len(arrs[0]), int(mapq))
        ha = (PhmmHap * max(1, len(haps)))()
        for i, hp in enumerate(haps):
            a = _u8(hp); keep.append(a)
            ha[i] = PhmmHap(a.ctypes.data, len(a))
        out = np.zeros((len(reads), len(haps)), np.float64)
        resc = np.zeros((len(reads), len(haps)), np.uint8)
        self._check(self.lib.pairhmm_cuda_compute_raw(self.h, ra, len(reads), ha, len(haps), C.byref(m), out.ctypes.data,
                                                      resc.ctypes.data))
        return out, resc

    def compute_batch_raw(self, b, base_q_raw, mapq, ins_q=None, del_q=None, model=None):
        """RegionBatch b (bases, offsets, haplotypes, regions) with raw qualities -> (log10[n_pairs], rescued[n_pairs])"""
        m = model or ReadModel()
        bq = _u8(base_q_raw); mq = np.ascontiguousarray(mapq, np.int32)
        iq = None if ins_q is None else _u8(ins_q); dq = None if del_q is None else _u8(del_q)
        out = np.zeros(b.n_pairs, np.float64); resc = np.zeros(b.n_pairs, np.uint8)
        self._check(self.lib.pairhmm_cuda_compute_flat_raw(
            self.h, b.read_bases.ctypes.data, bq.ctypes.data, None if iq is None else iq.ctypes.data,
            None if dq is None else dq.ctypes.data, mq.ctypes.data, b.read_off.ctypes.data, b.read_len.ctypes.data,
            b.n_reads, b.hap_bases.ctypes.data, b.hap_off.ctypes.data, b.hap_len.ctypes.data, b.n_haps,
            b.reg_read_start.ctypes.data, b.reg_hap_start.ctypes.data, b.n_regions, C.byref(m), out.ctypes.data,
            resc.ctypes.data))
        return out, resc

    def resident_load(self, b):
        self._check(self.lib.pairhmm_cuda_resident_load(self.h, *self._flat_args(b)))
        self._resident_pairs = b.n_pairs

    def resident_run(self):
        """-> (ms of float+finalize+rescue kernels, ms of the float kernels + finalize) by CUDA events"""
        a, f = C.c_float(), C.c_float()
        self._check(self.lib.pairhmm_cuda_resident_run(self.h, C.byref(a), C.byref(f)))
        return a.value, f.value

    def resident_fetch(self):
        out = np.zeros(self._resident_pairs, np.float64); resc = np.zeros(self._resident_pairs, np.uint8)
        self._check(self.lib.pairhmm_cuda_resident_fetch(self.h, out.ctypes.data, resc.ctypes.data))
        return out, resc

    def resident_fetch_raw(self):
        """raw 2^120-scaled FP32 sums per unique pair (test hook)"""
        out = np.zeros(self._resident_pairs, np.float32)
        self._check(self.lib.pairhmm_cuda_resident_fetch_raw(self.h, out.ctypes.data, out.shape[0]))
        return out

    def stats(self):
        s = Stats()
        self._check(self.lib.pairhmm_cuda_stats_get(self.h, C.byref(s)))
        return s.as_dict()

    def stats_reset(self):
        self._check(self.lib.pairhmm_cuda_stats_reset(self.h))


class CudaPairHMM:
    """Host-side mirror of the reference plugin: same call order and result convention as
    VectorLoglessPairHMM (initialize once per region, then one compute per sample)."""

    def __init__(self, use_double=False, **kw):
        self.engine = Engine(use_double=use_double, **kw)
        self.initialized = False
        self.haplotypes = None
        self.is_use_trietree_optimize = False   # PairHMM.h:56 -- the GPU engine never asks for the trie path

    def initialize(self, haplotypes, read_max_length, haplotype_max_length):
        if haplotype_max_length <= 0 or read_max_length <= 0:
            raise PairHMMError(-1, "haplotypeMaxLength and readMaxLength must be > 0 (PairHMM.cpp:16)")
        self.engine.initialize(read_max_length, haplotype_max_length)
        self.haplotypes = [np.ascontiguousarray(h, np.uint8) for h in haplotypes]
        self.initialized = True

    def compute_log10_likelihoods(self, processed_reads, gcp=None):
        """processed_reads: list of (bases, base_q, ins_q, del_q[, gcp]); gcp: optional list parallel to the
        reads (the reference passes it as a map read -> char[], VectorLoglessPairHMM.cpp:77).
        Returns the [n_haps][n_reads] matrix the reference fills through SampleMatrix::set(h, r, v)."""
        if not self.initialized:
            raise PairHMMError(-5, "computeLog10Likelihoods before initialize")
        if len(processed_reads) == 0:
            return np.zeros((len(self.haplotypes), 0), np.float64)   # VectorLoglessPairHMM.cpp:46-47
        reads = []
        for i, r in enumerate(processed_reads):
            g = r[4] if len(r) > 4 else (gcp[i] if gcp is not None else np.full(len(r[0]), 10, np.uint8))
            reads.append((r[0], r[1], r[2], r[3], g))
        out, _ = self.engine.compute(reads, self.haplotypes)
        return np.ascontiguousarray(out.T)

    # the reference's two trie entry points compute the same numbers
    compute_log10_likelihoods_trie = compute_log10_likelihoods
    compute_log10_likelihoods_trie_unique = compute_log10_likelihoods
