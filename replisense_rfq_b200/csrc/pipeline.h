// Internal launch interfaces between the C-ABI layer (rfq_api.cu) and the kernel files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rfq {

// scratch elements an exclusive_scan over n items needs (scan.cuh: 4096 items per CTA)
inline uint32_t scan_partials_needed(uint64_t n) { return (uint32_t)((n + 4095) / 4096) + 1; }

struct TxtArgs {
  const uint8_t* arena;       // device, packed documents (16-byte aligned starts)
  const uint32_t* doc_off;    // [n_docs] byte offset of each document in arena
  const uint32_t* doc_len;    // [n_docs]
  const uint32_t* tile_base;  // [n_docs+1] exclusive prefix of ceil(len/512)
  uint32_t n_docs, n_tiles;
  uint32_t* tile_cnt;         // [n_tiles]
  uint32_t* tile_pre;         // [n_tiles]
  uint32_t* first_pos;        // [n_docs]
  uint32_t* last_end;         // [n_docs]
  unsigned long long* out_len;  // [n_docs]
  unsigned long long* out_off;  // [n_docs]
  unsigned long long* out_total;  // [1]
  uint32_t* skip;             // [n_docs]
  uint32_t* scan_tmp32;       // scan partials
  unsigned long long* scan_tmp64;
};

void txt_pipeline(const TxtArgs& a, cudaStream_t s);   // sizes: fills out_len/out_off/out_total
void txt_emit(const TxtArgs& a, uint8_t* out, cudaStream_t s);

}  // namespace rfq
