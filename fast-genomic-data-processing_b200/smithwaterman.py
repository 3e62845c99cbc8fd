"""Python binding of the batched CUDA Smith-Waterman (include/smithwaterman_cuda.h), used by the tests and tools.
Mirrors the reference's native aligner interface (M2/smithwaterman/IntelSmithWaterman.cpp:8-19: align(ref, alt,
parameters, overhangStrategy) -> (cigar, alignment offset)). No CPU fallback."""
import ctypes as C

import numpy as np

from .pairhmm import load_library, PairHMMError

SOFTCLIP, INDEL, LEADING_INDEL, IGNORE = 9, 10, 11, 12
# SWParameters used by Mutect2Cpp (M2/read/CigarUtils.cpp:14-15, SmithWatermanAligner::STANDARD_NGS)
ALIGNMENT_TO_BEST_HAPLOTYPE = (10, -15, -30, -5)
NEW_SW_PARAMETERS = (200, -150, -260, -11)
STANDARD_NGS = (25, -50, -110, -6)
SW_SYMBOLS = ["smithwaterman_cuda_align_batch", "smithwaterman_cuda_align", "smithwaterman_cuda_last_error",
              "smithwaterman_cuda_release"]


def _lib():
    lib = load_library()
    vp = C.c_void_p
    lib.smithwaterman_cuda_align_batch.argtypes = [C.c_int, vp, vp, vp, vp, vp, vp, C.c_int64] + [C.c_int] * 5 + \
                                                  [vp, C.c_int, vp, C.POINTER(C.c_float)]
    lib.smithwaterman_cuda_align.argtypes = [vp, C.c_int, vp, C.c_int, vp, C.c_int] + [C.c_int] * 4 + [C.c_uint8]
    lib.smithwaterman_cuda_last_error.restype = C.c_char_p
    return lib


def align_flat(ref_bases, ref_off, ref_len, alt_bases, alt_off, alt_len, params=ALIGNMENT_TO_BEST_HAPLOTYPE, strategy=SOFTCLIP,
               device=-1, cigar_stride=None):
    """the C call on flat numpy arrays (no Python list packing): -> (cigar bytes uint8[n, stride], offsets int32[n],
    kernel milliseconds)"""
    lib = _lib()
    rb = np.ascontiguousarray(ref_bases, np.uint8); ab = np.ascontiguousarray(alt_bases, np.uint8)
    ro = np.ascontiguousarray(ref_off, np.int64); ao = np.ascontiguousarray(alt_off, np.int64)
    rl = np.ascontiguousarray(ref_len, np.int32); al = np.ascontiguousarray(alt_len, np.int32)
    n = int(rl.shape[0])
    stride = int(cigar_stride or 2 * max(1, int(rl.max(initial=1)), int(al.max(initial=1))))
    cig = np.zeros((n, stride), np.uint8); off = np.zeros(n, np.int32)
    ms = C.c_float(0)
    st = lib.smithwaterman_cuda_align_batch(device, rb.ctypes.data, ro.ctypes.data, rl.ctypes.data, ab.ctypes.data,
                                            ao.ctypes.data, al.ctypes.data, n, *[int(x) for x in params], int(strategy),
                                            cig.ctypes.data, stride, off.ctypes.data, C.byref(ms))
    if st != 0:
        raise PairHMMError(st, (lib.smithwaterman_cuda_last_error() or b"").decode())
    return cig, off, float(ms.value)


def release():
    """frees the calling thread's cached stream and device buffers (smithwaterman_cuda_release)"""
    _lib().smithwaterman_cuda_release()


def align_batch(refs, alts, params=ALIGNMENT_TO_BEST_HAPLOTYPE, strategy=SOFTCLIP, device=-1, cigar_stride=None):
    """refs, alts: lists of byte strings -> (list of CIGAR strings, offsets int32[n], kernel milliseconds)"""
    lib = _lib()
    n = len(refs)
    rl = np.array([len(x) for x in refs], np.int32); al = np.array([len(x) for x in alts], np.int32)
    ro = np.concatenate([[0], np.cumsum(rl)[:-1]]).astype(np.int64) if n else np.zeros(0, np.int64)
    ao = np.concatenate([[0], np.cumsum(al)[:-1]]).astype(np.int64) if n else np.zeros(0, np.int64)
    rb = np.frombuffer(b"".join(bytes(x) for x in refs), np.uint8) if n else np.zeros(0, np.uint8)
    ab = np.frombuffer(b"".join(bytes(x) for x in alts), np.uint8) if n else np.zeros(0, np.uint8)
    stride = int(cigar_stride or 2 * max([1] + list(rl) + list(al)))        # IntelSmithWaterman.cpp:12
    cig = np.zeros((n, stride), np.uint8); off = np.zeros(n, np.int32)
    ms = C.c_float(0)
    st = lib.smithwaterman_cuda_align_batch(device, rb.ctypes.data, ro.ctypes.data, rl.ctypes.data, ab.ctypes.data,
                                            ao.ctypes.data, al.ctypes.data, n, *[int(x) for x in params], int(strategy),
                                            cig.ctypes.data, stride, off.ctypes.data, C.byref(ms))
    if st != 0:
        raise PairHMMError(st, (lib.smithwaterman_cuda_last_error() or b"").decode())
    return [bytes(row).split(b"\0")[0].decode() for row in cig], off, float(ms.value)
