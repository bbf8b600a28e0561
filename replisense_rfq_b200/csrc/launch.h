// Kernel-launch accounting (bench.py reports gpu_launches from this counter) and stage profiling.
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

namespace rfq {

extern std::atomic<uint64_t> g_launch_count;

#define RFQ_LAUNCH(kern, grid, block, smem, stream, ...)         \
  do {                                                           \
    kern<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);    \
    ++::rfq::g_launch_count;                                     \
  } while (0)

// Optional per-stage timing with CUDA events on the ctx stream.
struct Prof {
  bool on = false;
  std::vector<cudaEvent_t> ev;
  std::vector<std::string> names;
  std::vector<float> ms;
  size_t used = 0;
  void begin() { used = 0; names.clear(); ms.clear(); }
  void mark(const char* name, cudaStream_t s) {
    if (!on) return;
    if (used == ev.size()) { cudaEvent_t e; cudaEventCreate(&e); ev.push_back(e); }
    cudaEventRecord(ev[used++], s);
    names.push_back(name);
  }
  // call after the stream is synchronised: ms[i] = time from mark i to mark i+1
  void finish() {
    ms.clear();
    if (!on) return;
    for (size_t i = 0; i + 1 < used; ++i) {
      float t = 0.f; cudaEventElapsedTime(&t, ev[i], ev[i + 1]); ms.push_back(t);
    }
    if (!names.empty()) names.pop_back();  // last mark is only an end marker
  }
  ~Prof() { for (auto e : ev) cudaEventDestroy(e); }
};

// profiler of the run in progress on this thread (set by the C-ABI layer), so kernel files can
// mark individual launches
extern thread_local Prof* g_prof;
inline void prof_mark(const char* name, cudaStream_t s) { if (g_prof) g_prof->mark(name, s); }

}  // namespace rfq
