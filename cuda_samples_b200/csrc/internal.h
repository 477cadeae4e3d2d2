// internal.h -- declarations shared by the translation units of libb200prim.so (not installed).
#pragma once
#include <atomic>

#include "../../include/b200prim.h"

extern "C" int b200_sort_reserve(size_t N);
extern "C" int b200_scan_reserve(size_t n);
extern "C" int b200_release_scratch_of(int which);

namespace b200 {
extern std::atomic<unsigned long long> g_launches;
inline void count_launch(unsigned n = 1) { g_launches.fetch_add(n, std::memory_order_relaxed); }
} // namespace b200
