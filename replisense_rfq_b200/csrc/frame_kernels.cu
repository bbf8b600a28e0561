// Text framing: output documents assembled from pieces the way
// FileParser._parse_pdf_sync frames what the PDF engine extracted (app/file_parser.py:166-212):
//   text_parts.append(f"=== Page {n} ===\n{page_text.strip()}")     if page_text.strip()
//   text_parts.append(table_text) / text_parts.append(images note)
//   "\n\n".join(text_parts)   or   "PDF appears to be empty or contains no extractable text"
// Three passes: (1) a warp per piece finds what str.strip() keeps of it; (2) a thread per output
// document walks its parts in order — which are printed, where every piece goes; (3) a warp per piece
// copies it (and the "\n\n" before the part it opens).  Byte work, HBM-bound: no tensor cores.
#include <cuda_runtime.h>
#include "../../include/rfq_parse.h"
#include "launch.h"
#include "pipeline.h"
#include "scan.cuh"
#include "utf8.cuh"

namespace rfq {

namespace {
struct PBytes {
  const uint8_t* p; uint32_t n;
  __device__ __forceinline__ uint32_t operator()(int64_t i) const { return (i >= 0 && i < (int64_t)n) ? p[i] : 0u; }
};
#define FRAME_NOWHERE 0xFFFFFFFFFFFFFFFFull
#define FRAME_SEP_BIT (1ull << 63)     // piece_dst: the part this piece opens is preceded by "\n\n"
}

__global__ void __launch_bounds__(256) frame_strip_kernel(FrameArgs a) {
  const uint32_t i = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= a.n_pieces) return;
  const uint32_t len = a.doc_len[i];
  uint32_t s0 = 0, s1 = len;
  if (a.piece_strip[i]) {
    const PBytes B{a.arena + a.doc_off[i], len};
    uint32_t lo = 0xFFFFFFFFu, hi = 0;
    for (uint32_t k = (uint32_t)lane; k < len; k += 32) {
      const uint32_t c = B.p[k];
      if ((c & 0xC0u) == 0x80u || u8_is_space_at(B, (int64_t)k)) continue;
      const uint32_t cl = c < 0x80u ? 1u : (uint32_t)max(u8_lead_len(c), 1);
      lo = min(lo, k); hi = max(hi, min(k + cl, len));
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
      lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
      hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if (lo == 0xFFFFFFFFu) { s0 = 0; s1 = 0; } else { s0 = lo; s1 = hi; }
  }
  if (lane == 0) { a.piece_a[i] = s0; a.piece_n[i] = s1 - s0; }
}

__global__ void __launch_bounds__(128) frame_plan_kernel(FrameArgs a) {
  const uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= a.n_out) return;
  const uint32_t p0 = a.doc_first_part[d], p1 = a.doc_first_part[d + 1];
  unsigned long long pos = 0;
  uint32_t printed_parts = 0;
  for (int pass = 0; pass < 2; ++pass) {          // pass 1 places the fallback part of a document with nothing else
    for (uint32_t p = p0; p < p1; ++p) {
      const uint32_t q0 = a.part_first_piece[p], q1 = a.part_first_piece[p + 1];
      const uint32_t kind = a.part_kind[p];
      if ((kind == RFQ_PART_FALLBACK) != (pass == 1)) continue;
      bool printed = q1 > q0;                       // a part without pieces prints nothing
      if (kind == RFQ_PART_FALLBACK) printed = printed && printed_parts == 0;
      else if (kind == RFQ_PART_IF_TEXT) printed = printed && a.piece_n[q1 - 1] != 0;
      if (!printed) { for (uint32_t q = q0; q < q1; ++q) a.piece_dst[q] = FRAME_NOWHERE; continue; }
      const bool sep = printed_parts > 0;           // "\n\n".join(text_parts)
      if (sep) pos += 2;
      for (uint32_t q = q0; q < q1; ++q) {
        a.piece_dst[q] = pos | ((sep && q == q0) ? FRAME_SEP_BIT : 0ull);
        pos += a.piece_n[q];
      }
      ++printed_parts;
    }
  }
  a.out_len[d] = pos;
}

__global__ void __launch_bounds__(256) frame_copy_kernel(FrameArgs a, uint8_t* __restrict__ out) {
  const uint32_t i = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (i >= a.n_pieces) return;
  const unsigned long long dst = a.piece_dst[i];
  if (dst == FRAME_NOWHERE) return;
  const uint8_t* src = a.arena + a.doc_off[i] + a.piece_a[i];
  uint8_t* o = out + (dst & ~FRAME_SEP_BIT);
  if ((dst & FRAME_SEP_BIT) && lane < 2) o[(int)lane - 2] = '\n';
  const uint32_t n = a.piece_n[i];
  for (uint32_t k = (uint32_t)lane; k < n; k += 32) o[k] = src[k];
}

// piece positions become absolute: + the offset of their document
__global__ void __launch_bounds__(128) frame_base_kernel(FrameArgs a) {
  const uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= a.n_out) return;
  const unsigned long long base = a.out_off[d];
  const uint32_t q0 = a.part_first_piece[a.doc_first_part[d]], q1 = a.part_first_piece[a.doc_first_part[d + 1]];
  for (uint32_t q = q0; q < q1; ++q)
    if (a.piece_dst[q] != FRAME_NOWHERE) a.piece_dst[q] += base;   // bit 63 (separator flag) is untouched: positions stay below 2^63
}

void frame_measure(const FrameArgs& a, cudaStream_t s) {
  if (a.n_out == 0) return;
  prof_mark("frame_strip+plan", s);
  if (a.n_pieces) RFQ_LAUNCH(frame_strip_kernel, (a.n_pieces + 7) / 8, 256, 0, s, a);
  RFQ_LAUNCH(frame_plan_kernel, (a.n_out + 127) / 128, 128, 0, s, a);
  exclusive_scan<SumU64>(a.out_len, a.out_off, a.n_out, a.scan_tmp64, a.out_total, s);
  RFQ_LAUNCH(frame_base_kernel, (a.n_out + 127) / 128, 128, 0, s, a);
}

void frame_emit(const FrameArgs& a, uint8_t* out, cudaStream_t s) {
  if (a.n_pieces == 0) return;
  prof_mark("frame_copy_kernel", s);
  RFQ_LAUNCH(frame_copy_kernel, (a.n_pieces + 7) / 8, 256, 0, s, a, out);
}

}  // namespace rfq
