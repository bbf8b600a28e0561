// latin-1 -> UTF-8 stream transcode: the second encoding of the reference's trial loop
// (app/file_parser.py:298).  pandas decodes a non-UTF-8 file to str and hands the C tokenizer
// its UTF-8 encoding, so on the device a latin-1 attempt is "expand every byte >= 0x80 to two
// bytes, then run the same pipeline".  (cp1252 can never succeed where latin-1 failed: the
// ASCII structure is identical and latin-1 cannot raise a decode error.)
// One CTA per document (only documents whose UTF-8 attempts all failed get here).
#include <cuda_runtime.h>
#include "launch.h"
#include "pipeline.h"
#include "scan.cuh"

namespace rfq {

__global__ void __launch_bounds__(256) latin1_count_kernel(const uint8_t* __restrict__ arena,
                                                           const uint32_t* __restrict__ doc_off,
                                                           const uint32_t* __restrict__ doc_len,
                                                           uint32_t* __restrict__ hi_count) {
  uint32_t d = blockIdx.x;
  const uint8_t* doc = arena + doc_off[d];
  uint32_t len = doc_len[d];
  uint32_t cnt = 0;
  for (uint32_t base = threadIdx.x * 16; base < len; base += 256 * 16) {
    uint4 v = __ldg(reinterpret_cast<const uint4*>(doc + base));
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 16; ++k)
      if (base + k < len && ((w[k >> 2] >> ((k & 3) * 8)) & 0x80u)) ++cnt;
  }
  uint32_t tot;
  block_excl_scan<SumU32>(cnt, &tot);
  if (threadIdx.x == 0) hi_count[d] = tot;
}

__global__ void __launch_bounds__(256) latin1_expand_kernel(const uint8_t* __restrict__ arena,
                                                            const uint32_t* __restrict__ doc_off,
                                                            const uint32_t* __restrict__ doc_len,
                                                            const uint32_t* __restrict__ new_off,
                                                            uint8_t* __restrict__ out) {
  uint32_t d = blockIdx.x;
  const uint8_t* doc = arena + doc_off[d];
  uint32_t len = doc_len[d];
  uint8_t* dst = out + new_off[d];
  uint32_t carry = 0;
  for (uint32_t chunk = 0; chunk < len; chunk += 256 * 16) {
    uint32_t base = chunk + threadIdx.x * 16;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (base < len) v = __ldg(reinterpret_cast<const uint4*>(doc + base));
    uint32_t w[4] = {v.x, v.y, v.z, v.w};
    uint32_t cnt = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k)
      if (base + k < len && ((w[k >> 2] >> ((k & 3) * 8)) & 0x80u)) ++cnt;
    uint32_t tot;
    uint32_t ex = block_excl_scan<SumU32>(cnt, &tot);
    uint32_t o = base + carry + ex;
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      if (base + k >= len) break;
      uint32_t b = (w[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
      if (b < 0x80u) dst[o++] = (uint8_t)b;
      else { dst[o++] = (uint8_t)(0xC0u | (b >> 6)); dst[o++] = (uint8_t)(0x80u | (b & 0x3Fu)); }
    }
    carry += tot;
  }
}

void latin1_count(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len,
                  uint32_t n_docs, uint32_t* hi_count, cudaStream_t s) {
  if (n_docs) RFQ_LAUNCH(latin1_count_kernel, n_docs, 256, 0, s, arena, doc_off, doc_len, hi_count);
}
void latin1_expand(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len,
                   const uint32_t* new_off, uint32_t n_docs, uint8_t* out, cudaStream_t s) {
  if (n_docs) RFQ_LAUNCH(latin1_expand_kernel, n_docs, 256, 0, s, arena, doc_off, doc_len, new_off, out);
}

}  // namespace rfq
