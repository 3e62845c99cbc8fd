// Copies whose destination is not read again by this CPU soon: the packer's read planes (the next reader is the copy engine) and
// a large chunk's result matrix (the caller's arrays). Streaming stores skip the read-for-ownership of the destination lines
// and leave the caches to the source: tools/ubench/hostcopy.cpp on the GPU box's host, 16 threads, runs of 30 KB into a
// page-locked arena: memcpy 54 GB/s, streaming stores 81 GB/s.
#pragma once
#include <cstddef>

namespace phmm {
// as memcpy; small copies (and hosts without AVX2) take memcpy itself
void copy_stream(void *dst, const void *src, std::size_t n);
// orders the streaming stores of the calling thread before whatever it does next (hand the buffer to the device, wake a waiter)
void copy_stream_fence();
}  // namespace phmm
