// internal.h -- declarations shared by the translation units of libb200prim.so (not installed).
#pragma once
#include <atomic>
#include <cstdint>

#include <cuda_runtime.h>

#include "../../include/b200prim.h"

extern "C" int b200_sort_reserve(size_t N);
extern "C" int b200_scan_reserve(size_t n);
extern "C" int b200_release_scratch_of(int which);

namespace b200 {
extern std::atomic<unsigned long long> g_launches;
// scan.cu: exclusive scan of one 16-byte aligned array of n % 4 == 0 elements whose 8192-element tiles already know their
// exclusive prefixes (d_tile_prefix, produced by the kernel in front on the stream; read behind griddepcontrol.wait)
int scan_tiled(uint32_t *d_dst, const uint32_t *d_src, size_t n, const uint32_t *d_tile_prefix, cudaStream_t stream);
// Programmatic-launch guard (context.cu).  reduce_kernel, histogram_kernel and scan_tile_sums_kernel release their
// dependents at their very top, and histogram_kernel / the scans launched with the programmatic attribute read their INPUT
// before griddepcontrol.wait.  That is only sound when the input is not something a still-running predecessor writes
// (a histogram of the buffer a reduce has just zero-filled, a scan of fresh histogram bins).  Launchers of the early
// releasers note what their kernel writes; a launcher about to set the attribute asks whether its early reads overlap
// any recent note, and launches the ordinary way (full stream order) when they do.  Per device, last 8 notes, never
// cleared: a stale note can only cost one overlap, never correctness.
void pdl_note_output(int device, const void *p, size_t bytes);
bool pdl_early_read_ok(int device, const void *p, size_t bytes);
inline void count_launch(unsigned n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }
} // namespace b200
