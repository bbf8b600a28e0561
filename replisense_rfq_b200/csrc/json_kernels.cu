// Where the JSON sits inside an LLM reply: the byte work of
// RFQFieldGenerator._extract_json_from_string (app/rfq_agent.py:208-236) for a whole batch of replies:
//   strategy 2   re.search(r'\{.*\}', s, re.DOTALL)  = first '{' .. last '}'
//   strategy 3   s.strip(); drop a leading ```json and a trailing ```; .strip() again
// (Unicode whitespace, str.strip() semantics).  Output per reply: four u32 — first '{', last '}'
// (0xFFFFFFFF when there is none), and the [begin, end) of the cleaned text.  The JSON grammar itself
// (json.loads) stays with the host: it is CPython's parser in the reference too.
// One warp per reply, 16-byte loads; HBM-bound byte work, no tensor cores.
#include <cuda_runtime.h>
#include "launch.h"
#include "pipeline.h"
#include "utf8.cuh"

namespace rfq {

namespace {
struct Bytes {
  const uint8_t* p; uint32_t n;
  __device__ __forceinline__ uint32_t operator()(int64_t i) const { return (i >= 0 && i < (int64_t)n) ? p[i] : 0u; }
};
// first non-space character start and end of the last non-space character inside [a, b)
__device__ void strip_extent(const Bytes& B, uint32_t a, uint32_t b, int lane, uint32_t* s0, uint32_t* s1) {
  uint32_t lo = 0xFFFFFFFFu, hi = 0;
  for (uint32_t i = a + (uint32_t)lane; i < b; i += 32) {
    const uint32_t c = B.p[i];
    if ((c & 0xC0u) == 0x80u) continue;                        // continuation byte: not a character start
    if (u8_is_space_at(B, (int64_t)i)) continue;
    const uint32_t len = c < 0x80u ? 1u : (uint32_t)max(u8_lead_len(c), 1);
    lo = min(lo, i);
    hi = max(hi, min(i + len, b));
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if (lo == 0xFFFFFFFFu) { *s0 = a; *s1 = a; } else { *s0 = lo; *s1 = hi; }
}
}  // namespace

__global__ void __launch_bounds__(256) json_spans_kernel(const uint8_t* __restrict__ arena, const uint32_t* __restrict__ doc_off,
                                                         const uint32_t* __restrict__ doc_len, uint32_t n_docs,
                                                         uint32_t* __restrict__ out, unsigned long long* __restrict__ out_off,
                                                         unsigned long long* __restrict__ out_len) {
  const uint32_t d = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (d >= n_docs) return;
  const uint8_t* doc = arena + doc_off[d];
  const uint32_t len = doc_len[d];
  const Bytes B{doc, len};
  uint32_t first = 0xFFFFFFFFu, last = 0;
  bool any_close = false;
  for (uint32_t base = (uint32_t)lane * 16; base < len; base += 512) {
    const uint4 v = __ldg(reinterpret_cast<const uint4*>(doc + base));   // documents start 16-byte aligned, the arena has slack
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const uint32_t c = (w[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
      const uint32_t pos = base + k;
      if (pos < len) {
        if (c == '{' && pos < first) first = pos;
        if (c == '}') { last = pos; any_close = true; }
      }
    }
  }
  const unsigned closes = __ballot_sync(0xffffffffu, any_close);
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    first = min(first, __shfl_xor_sync(0xffffffffu, first, o));
    last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
  }
  uint32_t s0, s1;
  strip_extent(B, 0, len, lane, &s0, &s1);
  // cleaned.startswith('```json') -> [7:];  then .endswith('```') -> [:-3]
  uint32_t a = s0, b = s1;
  if (b - a >= 7 && doc[a] == '`' && doc[a + 1] == '`' && doc[a + 2] == '`' && doc[a + 3] == 'j' && doc[a + 4] == 's' &&
      doc[a + 5] == 'o' && doc[a + 6] == 'n') a += 7;
  if (b - a >= 3 && doc[b - 1] == '`' && doc[b - 2] == '`' && doc[b - 3] == '`') b -= 3;
  uint32_t c0, c1;
  strip_extent(B, a, b, lane, &c0, &c1);
  if (lane == 0) {
    out[4 * d + 0] = first;
    out[4 * d + 1] = closes ? last : 0xFFFFFFFFu;
    out[4 * d + 2] = c0;
    out[4 * d + 3] = c1;
    out_off[d] = 16ull * d;
    out_len[d] = 16ull;
  }
}

void json_spans_batch(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, uint32_t n_docs,
                      uint32_t* out, unsigned long long* out_off, unsigned long long* out_len, cudaStream_t s) {
  if (n_docs == 0) return;
  prof_mark("json_spans_kernel", s);
  RFQ_LAUNCH(json_spans_kernel, (n_docs + 7) / 8, 256, 0, s, arena, doc_off, doc_len, n_docs, out, out_off, out_len);
}

}  // namespace rfq
