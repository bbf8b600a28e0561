// rfq_text_normalise_noalloc: the text path (UTF-8 decode with errors='ignore', universal
// newlines, str.strip()) for callers that are not allowed to allocate GPU memory.
//
// Who that is: code running inside somebody else's memory budget — a plugin of a serving process
// whose allocator owns every byte of the device and treats a foreign cudaMalloc as a fault, a
// latency-critical request handler where an allocation (an implicit device-wide synchronisation)
// is off the table, or an embedder that has frozen its memory map after start-up.  The regular
// entry points (rfq_text_run and friends) carve their per-run scratch out of a growing pool
// (cudaMalloc on first use / growth); this one never does.  Its scratch is the fixed array
// below, which is part of the library image: every GPU that loads the library holds these
// RFQ_NOALLOC_SCRATCH_BYTES until the process exits.  The price is a size limit
// (RFQ_NOALLOC_MAX_BYTES per document) and one document per call.
//
// It runs the very same kernels as rfq_text_run (txt_count / scan / txt_extent / txt_emit);
// only where their scratch lives differs, so the two paths cannot drift apart.
#include <cuda_runtime.h>
#include <stdio.h>
#include <mutex>

#include "../../include/rfq_parse.h"
#include "launch.h"
#include "pipeline.h"
#include "rfq_common.cuh"

namespace {

constexpr uint32_t kMaxTiles = RFQ_NOALLOC_MAX_BYTES / RFQ_TILE;  // 8192
constexpr size_t kScratchBytes = 96u << 10;
static_assert(kScratchBytes == RFQ_NOALLOC_SCRATCH_BYTES, "header and implementation disagree");

// The scratch itself.  In host code this name is only a stand-in for the device object (taking
// its address on the host does NOT give a device pointer); the device address is obtained once
// per device with cudaGetSymbolAddress below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

// carve-up of the scratch (byte offsets, all 256-byte aligned)
constexpr size_t kOffDesc = 0;                                   // doc_off, doc_len, tile_base[2]
constexpr size_t kOffTileCnt = 256;                              // uint32[kMaxTiles + 1]
constexpr size_t kOffTilePre = kOffTileCnt + ((4 * (kMaxTiles + 1) + 255) & ~255u);
constexpr size_t kOffSmall = kOffTilePre + ((4 * (kMaxTiles + 1) + 255) & ~255u);
// small block: first_key, last_key (u64), skip (uint32), 256 B apart, out_off, out_total (u64)
constexpr size_t kOffScan32 = kOffSmall + 6 * 256;               // scan partials (uint32[8])
constexpr size_t kOffScan64 = kOffScan32 + 256;                  // scan partials (u64[8])
constexpr size_t kUsed = kOffScan64 + 256;
static_assert(kUsed <= kScratchBytes, "scratch too small for RFQ_NOALLOC_MAX_BYTES");

__global__ void noalloc_desc_kernel(uint32_t* desc, uint32_t len, uint32_t n_tiles) {
  desc[0] = 0;        // doc_off
  desc[1] = len;      // doc_len
  desc[2] = 0;        // tile_base[0]
  desc[3] = n_tiles;  // tile_base[1]
}

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_scratch_dev[kMaxDevices];  // device address of the scratch, per device

}  // namespace

extern "C" int rfq_text_normalise_noalloc(int device, const uint8_t* d_in, uint64_t len,
                                          uint8_t* d_out, uint64_t* d_out_len, void* cuda_stream) {
  if (len > RFQ_NOALLOC_MAX_BYTES) {
    fprintf(stderr,
            "rfq_text_normalise_noalloc: document of %llu bytes exceeds the fixed scratch "
            "(limit RFQ_NOALLOC_MAX_BYTES = %u bytes); use rfq_text_parse for larger input\n",
            (unsigned long long)len, (unsigned)RFQ_NOALLOC_MAX_BYTES);
    return RFQ_ERR_TOO_LARGE;
  }
  if (!d_in || !d_out || !d_out_len || device < 0 || device >= kMaxDevices ||
      ((uintptr_t)d_in & 15) || ((uintptr_t)d_out_len & 7))
    return RFQ_ERR_ARG;
  int prev = -1;
  if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); return RFQ_ERR_CUDA; }
  if (prev != device && cudaSetDevice(device) != cudaSuccess) { cudaGetLastError(); return RFQ_ERR_CUDA; }
  int rc = RFQ_OK;
  unsigned char* base = nullptr;
  {
    // set-up, once per device, with that device current: resolve the device address of the array
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_dev[device]) {
      void* p = nullptr;
      cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
      if (e != cudaSuccess || !p) {
        fprintf(stderr, "rfq_text_normalise_noalloc: cudaGetSymbolAddress failed: %s\n",
                cudaGetErrorString(e));
        cudaGetLastError();
        rc = RFQ_ERR_CUDA;
      } else {
        g_scratch_dev[device] = (unsigned char*)p;
      }
    }
    base = g_scratch_dev[device];
  }
  if (rc == RFQ_OK) {
    cudaStream_t s = (cudaStream_t)cuda_stream;
    const uint32_t n_tiles = (uint32_t)((len + RFQ_TILE - 1) / RFQ_TILE);
    uint32_t* desc = (uint32_t*)(base + kOffDesc);
    RFQ_LAUNCH(noalloc_desc_kernel, 1, 1, 0, s, desc, (uint32_t)len, n_tiles);
    rfq::TxtArgs a{};
    a.arena = d_in;
    a.doc_off = desc; a.doc_len = desc + 1; a.tile_base = desc + 2;
    a.n_docs = 1; a.n_tiles = n_tiles;
    a.tile_cnt = (uint32_t*)(base + kOffTileCnt);
    a.tile_pre = (uint32_t*)(base + kOffTilePre);
    a.first_key = (unsigned long long*)(base + kOffSmall);
    a.last_key = (unsigned long long*)(base + kOffSmall + 256);
    a.skip = (uint32_t*)(base + kOffSmall + 512);
    a.out_off = (unsigned long long*)(base + kOffSmall + 768);
    a.out_total = (unsigned long long*)(base + kOffSmall + 1024);
    a.out_len = (unsigned long long*)d_out_len;
    a.scan_tmp32 = (uint32_t*)(base + kOffScan32);
    a.scan_tmp64 = (unsigned long long*)(base + kOffScan64);
    // The array keeps whatever the previous call left.  What must start from a known value is
    // reset here on the same stream: first_key / last_key (inside txt_pipeline) — everything
    // else is fully overwritten before it is read.
    rfq::txt_pipeline(a, s);
    rfq::txt_emit(a, d_out, s);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
      fprintf(stderr, "rfq_text_normalise_noalloc: launch failed: %s\n", cudaGetErrorString(e));
      rc = RFQ_ERR_CUDA;
    }
  }
  if (prev != device) cudaSetDevice(prev);
  return rc;
}
