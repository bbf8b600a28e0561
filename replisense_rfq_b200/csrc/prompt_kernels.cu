// Prompt assembly: what RFQFieldGenerator._generate_sync does with raw_text
// before the LLM call (app/rfq_agent.py:142-151):
//   * `not raw_text or not raw_text.strip()`  -> no prompt (status RFQ_DOC_EMPTY_TEXT)
//   * `len(raw_text) > 8000` (code points)    -> raw_text[:8000] + "... [truncated]"
//   * prompt = f"{TEMPLATE}\n\"\"\"\n{raw_text}\n\"\"\""
// head / tail / mark are passed in by the host (the template lives with the host mirror).
//
// One CTA per document.  Measure: 16 bytes per thread per sweep of 4 KiB, a block scan of the
// character-start counts finds the byte where character `max_chars` starts (the cut); the sweep
// stops as soon as the cut and a non-space character are both known, so a 100 MB document costs
// the same as a 32 KB one.  Emit: destination-aligned 16-byte stores, source bytes gathered from
// the four pieces.  HBM-bound byte work: no tensor cores.
#include <cuda_runtime.h>
#include "../../include/rfq_parse.h"
#include "launch.h"
#include "pipeline.h"
#include "scan.cuh"
#include "utf8.cuh"

namespace rfq {

struct DocBytes {
  const uint8_t* p; uint32_t n;
  __device__ __forceinline__ uint32_t operator()(int64_t i) const { return (i >= 0 && i < (int64_t)n) ? p[i] : 0u; }
};

__global__ void __launch_bounds__(256) prompt_measure_kernel(PromptArgs a) {
  const uint32_t d = blockIdx.x;
  const uint8_t* doc = a.arena + a.doc_off[d];
  const uint32_t len = a.doc_len[d];
  __shared__ uint32_t s_cut, s_nonspace;
  if (threadIdx.x == 0) { s_cut = 0xFFFFFFFFu; s_nonspace = 0; }
  __syncthreads();
  DocBytes B{doc, len};
  uint32_t seen = 0;   // character starts before this sweep
  for (uint32_t base = 0; base < len; base += 4096) {
    const uint32_t p0 = base + threadIdx.x * 16;
    uint32_t starts = 0, nonspace = 0;
    if (p0 < len) {
      uint4 v = __ldg(reinterpret_cast<const uint4*>(doc + p0));   // documents start 16-byte aligned
      uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        uint32_t b = (w[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
        if (p0 + k < len && (b & 0xC0u) != 0x80u) {
          starts |= 1u << k;
          if (!u8_is_space_at(B, (int64_t)p0 + k)) nonspace = 1;
        }
      }
    }
    uint32_t cnt = __popc(starts), tot;
    uint32_t ex = block_excl_scan<SumU32>(cnt, &tot);
    if (nonspace && !s_nonspace) s_nonspace = 1;
    // the character with index max_chars (0-based) starts the part that is cut off
    uint32_t want = a.max_chars - seen;   // index within this sweep, valid when seen <= max_chars
    if (seen <= a.max_chars && want >= ex && want < ex + cnt) {
      uint32_t m = starts;
      for (uint32_t q = want - ex; q; --q) m &= m - 1;
      s_cut = p0 + (uint32_t)(__ffs(m) - 1);
    }
    seen += tot;
    __syncthreads();
    if (s_cut != 0xFFFFFFFFu && s_nonspace) break;   // uniform: both were written before the barrier
  }
  if (threadIdx.x == 0) {
    bool trunc = s_cut != 0xFFFFFFFFu;
    uint32_t keep = trunc ? s_cut : len;
    bool empty = s_nonspace == 0;
    a.cut[d] = keep;
    a.trunc[d] = trunc ? 1u : 0u;
    a.status[d] = empty ? RFQ_DOC_EMPTY_TEXT : 0;
    a.out_len[d] = empty ? 0ull : (unsigned long long)a.head_len + keep + (trunc ? a.mark_len : 0u) + a.tail_len;
  }
}

__global__ void __launch_bounds__(256) prompt_emit_kernel(PromptArgs a, uint8_t* __restrict__ out) {
  const uint32_t d = blockIdx.x;
  const unsigned long long total = a.out_len[d];
  if (total == 0) return;
  const uint8_t* doc = a.arena + a.doc_off[d];
  const uint32_t e0 = a.head_len, e1 = e0 + a.cut[d], e2 = e1 + (a.trunc[d] ? a.mark_len : 0u);
  uint8_t* g = out + a.out_off[d];
  const uint32_t mis = (uint32_t)((uintptr_t)g & 15u);
  const uint32_t nvec = (uint32_t)((mis + total + 15) >> 4);
  for (uint32_t v = threadIdx.x; v < nvec; v += blockDim.x) {
    const long long p0 = (long long)v * 16 - mis;   // position of this vector's first byte in the prompt
    uint32_t w[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      long long p = p0 + k;
      uint32_t b = 0;
      if (p >= 0 && p < (long long)total) {
        uint32_t q = (uint32_t)p;
        b = q < e0 ? a.head[q] : q < e1 ? doc[q - e0] : q < e2 ? a.mark[q - e1] : a.tail[q - e2];
      }
      w[k >> 2] |= b << ((k & 3) * 8);
    }
    if (p0 >= 0 && p0 + 16 <= (long long)total) {
      *reinterpret_cast<uint4*>(g + p0) = make_uint4(w[0], w[1], w[2], w[3]);
    } else {
#pragma unroll
      for (int k = 0; k < 16; ++k) {
        long long p = p0 + k;
        if (p >= 0 && p < (long long)total) g[p] = (uint8_t)(w[k >> 2] >> ((k & 3) * 8));
      }
    }
  }
}

// Fused parse -> prompt runs (rfq_csv_prompt_run): the parsed text stays on the device; only the part
// a prompt can use (max_chars code points are at most 4 * max_chars bytes) is gathered next to the
// other documents' so that the prompt kernels see one arena.
__global__ void __launch_bounds__(256) prompt_gather_kernel(const unsigned long long* __restrict__ src_ptr,
                                                            const uint32_t* __restrict__ take,
                                                            const uint32_t* __restrict__ dst_off, uint8_t* __restrict__ dst) {
  const uint32_t d = blockIdx.x;
  const uint8_t* src = reinterpret_cast<const uint8_t*>(src_ptr[d]);
  const uint32_t n = take[d], padded = (n + 15u) & ~15u;
  uint8_t* o = dst + dst_off[d];
  // blockIdx.y slices the document: a 32 KB head is sixteen CTAs, not one
  for (uint32_t i = blockIdx.y * blockDim.x + threadIdx.x; i < padded; i += gridDim.y * blockDim.x) o[i] = i < n ? src[i] : (uint8_t)0;
}
void prompt_gather(const unsigned long long* src_ptr, const uint32_t* take, const uint32_t* dst_off, uint8_t* dst,
                   uint32_t n_docs, cudaStream_t s) {
  if (n_docs == 0) return;
  prof_mark("prompt_gather_kernel", s);
  RFQ_LAUNCH(prompt_gather_kernel, dim3(n_docs, n_docs < 4096 ? 16 : 1), 256, 0, s, src_ptr, take, dst_off, dst);
}

void prompt_measure(const PromptArgs& a, cudaStream_t s) {
  if (a.n_docs == 0) return;
  prof_mark("prompt_measure_kernel", s);
  RFQ_LAUNCH(prompt_measure_kernel, a.n_docs, 256, 0, s, a);
  prof_mark("scan(prompt sizes)", s);
  exclusive_scan<SumU64>(a.out_len, a.out_off, a.n_docs, a.scan_tmp64, a.out_total, s);
}

void prompt_emit(const PromptArgs& a, uint8_t* out, cudaStream_t s) {
  if (a.n_docs == 0) return;
  prof_mark("prompt_emit_kernel", s);
  RFQ_LAUNCH(prompt_emit_kernel, a.n_docs, 256, 0, s, a, out);
}

}  // namespace rfq
