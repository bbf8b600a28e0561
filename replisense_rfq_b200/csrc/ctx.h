// The context behind the opaque rfq_ctx handle: one device, one stream, grow-only buffers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

#include "launch.h"
#include "pipeline.h"
#include "rfq_common.cuh"

#define RFQ_CK(c, call)                                                            \
  do {                                                                             \
    cudaError_t e_ = (call);                                                       \
    if (e_ != cudaSuccess) {                                                       \
      (c)->err = std::string(#call) + ": " + cudaGetErrorString(e_);               \
      cudaGetLastError();                                                          \
      return RFQ_ERR_CUDA;                                                         \
    }                                                                              \
  } while (0)

namespace rfq {

// grow-only device buffer
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

// grow-only pinned host buffer
struct HostBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr; cap = 0;
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocDefault);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

// Per-run scratch: bump allocation out of chunks; chunks are merged into one at reset() so a
// steady-state run performs no cudaMalloc at all.
struct DevPool {
  struct Chunk { uint8_t* p; size_t cap; };
  std::vector<Chunk> chunks;
  size_t used = 0;       // in the last chunk
  size_t total_used = 0; // across the run
  bool failed = false;
  void reset() {
    failed = false;
    if (chunks.size() > 1 || (chunks.size() == 1 && total_used > chunks[0].cap)) {
      size_t want = 0;
      for (auto& c : chunks) { want += c.cap; cudaFree(c.p); }
      chunks.clear();
      uint8_t* p = nullptr;
      if (cudaMalloc((void**)&p, want) == cudaSuccess) chunks.push_back({p, want});
      else cudaGetLastError();
    }
    used = 0; total_used = 0;
  }
  void* alloc_bytes(size_t bytes) {
    bytes = (bytes + 255) & ~(size_t)255;
    if (bytes == 0) bytes = 256;
    if (chunks.empty() || used + bytes > chunks.back().cap) {
      size_t want = bytes * 2 > (size_t)(8u << 20) ? bytes * 2 : (size_t)(8u << 20);
      uint8_t* p = nullptr;
      if (cudaMalloc((void**)&p, want) != cudaSuccess) { cudaGetLastError(); failed = true; return nullptr; }
      chunks.push_back({p, want});
      used = 0;
    }
    void* r = chunks.back().p + used;
    used += bytes; total_used += bytes;
    return r;
  }
  template <class T> T* alloc(size_t n) { return (T*)alloc_bytes(n * sizeof(T)); }
  void release() { for (auto& c : chunks) cudaFree(c.p); chunks.clear(); used = 0; }
};

}  // namespace rfq

struct rfq_ctx {
  int device = 0;
  cudaStream_t stream = nullptr, own_stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  std::string err;
  rfq::Prof prof;
  rfq::DevPool pool;
  // staged batch
  bool staged = false;
  uint32_t n_docs = 0, n_tiles = 0;
  uint64_t arena_bytes = 0;
  const uint8_t* h_arena = nullptr;  // host copy of the packed arena (caller's or h_stage)
  std::vector<uint32_t> h_doc_off, h_doc_len, h_tile_base;
  rfq::DevBuf d_arena, d_doctab, d_out, d_arena2;
  rfq::HostBuf h_stage, h_doctab, h_out, h_tabs, h_small;
  uint32_t *d_doc_off = nullptr, *d_doc_len = nullptr, *d_tile_base = nullptr;
  // last run
  bool have_output = false, timed = false, fetched = false;
  uint64_t out_total = 0;
  unsigned long long *d_out_off = nullptr, *d_out_len = nullptr;
  int32_t* d_status = nullptr;
  uint32_t *d_rows = nullptr, *d_cols = nullptr;
  uint8_t* d_attempt = nullptr;
  uint64_t* h_out_off = nullptr; uint64_t* h_out_len = nullptr;
  int32_t* h_status = nullptr; uint32_t* h_rows = nullptr; uint32_t* h_cols = nullptr;
  uint8_t* h_attempt = nullptr;
  // CSV/table runs keep their per-document tables on the host and their text in segments
  struct OutSeg { const uint8_t* d; uint64_t bytes; uint64_t host_base; bool copied = false; };
  cudaStream_t copy_stream = nullptr;   // D2H of finished output slices while later slices are still being emitted
  cudaEvent_t ev_slice[8] = {};
  bool overlap_fetch = false;
  std::vector<OutSeg> segs;
  bool host_tables = false;
  std::vector<uint64_t> r_off, r_len, r_chars;   // r_chars: code points of the text (~0 = not known)
  std::vector<int32_t> r_status;
  std::vector<uint32_t> r_rows, r_cols;
  std::vector<uint8_t> r_attempt;
  rfq::HostBuf h_up, h_down;   // per-attempt upload / download staging
  // fused parse -> prompt runs: what the parse stage produced per document (the text itself stays on the device)
  std::vector<uint64_t> raw_bytes, raw_chars;
  std::vector<int32_t> status_override;   // applied to the fetched status table (parse failures of a fused run)
  // row pipeline: sizes seen by the last CSV run, used to size the next run's buffers without a sync
  unsigned long long csv2_out_hint = 0;      // output bytes of an attempt
  uint32_t csv2_in_hint = 0, csv2_item_out_hint = 0;   // largest input span / output range of a work item
  // rfq_mixed_run drives the text / CSV / sheet runs over subsets of one staged batch: a nested run
  // keeps the scratch pool and the timing events of the outer call
  // device-side inflate: its output lives outside the per-run pool so that it can become the staged batch
  rfq::DevBuf d_inflated;
  std::vector<uint64_t> inf_off, inf_len;
  uint64_t inf_total = 0;
  bool inflated_valid = false;
  uint32_t out_n = 0;   // documents of the last run's output when it differs from the staged count (0 = n_docs)
  bool nested = false;
  // rfq_mixed_run gives the sheet documents to a second context of its own (own stream, scratch pool and pinned
  // buffers; it borrows the staged arena) and runs it on a host thread beside the CSV run
  rfq_ctx* helper = nullptr;
  cudaEvent_t ev_staged = nullptr, ev_helper = nullptr;
  rfq::DevBuf d_viewtab;
  rfq::HostBuf h_viewtab;
  float last_ms = 0.f;
  uint64_t h2d_bytes = 0, d2h_bytes = 0;

  unsigned long long* h_scalar() {
    if (!h_small.p) h_small.reserve(4096);
    return (unsigned long long*)h_small.p;
  }
  void release_all() {
    pool.release();
    d_arena.release(); d_doctab.release(); d_out.release(); d_arena2.release(); d_viewtab.release(); h_viewtab.release(); d_inflated.release();
    h_stage.release(); h_doctab.release(); h_out.release(); h_tabs.release(); h_small.release(); h_up.release(); h_down.release();
  }
};
