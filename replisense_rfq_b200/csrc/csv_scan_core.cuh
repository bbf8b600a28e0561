// Pieces of the parallel tokenizer shared by the three-pass scan (csv_scan.cu: sheets' fallback and
// documents the row pipeline hands back) and the one-pass row scan (csv2_scan.cu): byte classes of a
// lane's 16 bytes, the lane's transition function, document lookup.
#pragma once
#include <cuda_runtime.h>
#include "csv_dfa.cuh"
#include "rfq_common.cuh"

namespace rfq {

struct FnCompose {
  typedef uint32_t T;
  __device__ static T identity() { return RFQ_FN_IDENTITY; }
  __device__ static T op(T a, T b) { return fn_compose(a, b); }  // a happens before b
};

__device__ __forceinline__ uint32_t scan_find_doc(const uint32_t* __restrict__ tile_base,
                                                  uint32_t n, uint32_t t) {
  uint32_t lo = 0, hi = n;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (__ldg(tile_base + mid) <= t) lo = mid; else hi = mid;
  }
  return lo;
}

// classes of this lane's 16 bytes, 4 bits each, in a 64-bit word
struct LaneBytes {
  unsigned long long cls;  // nibble k = class of byte k
};

// classes of 4 bytes at once: byte-wise compares (__vcmpeq4 gives 0xFF per equal byte) select
// the class code of each byte; then the 4 class bytes are squeezed into 4 nibbles
__device__ __forceinline__ uint32_t classes_of_word(uint32_t w, uint32_t sep4) {
  uint32_t mD = __vcmpeq4(w, sep4);
  uint32_t mQ = __vcmpeq4(w, 0x22222222u);
  uint32_t mLF = __vcmpeq4(w, 0x0A0A0A0Au);
  uint32_t mCR = __vcmpeq4(w, 0x0D0D0D0Du);
  uint32_t mWS = (__vcmpeq4(w, 0x20202020u) | __vcmpeq4(w, 0x09090909u)) & ~mD;
  uint32_t c = (mD & 0x01010101u) | (mQ & ~mD & 0x02020202u) | (mLF & ~mD & 0x03030303u) |
               (mCR & ~mD & 0x04040404u) | (mWS & 0x05050505u);
  uint32_t p = c | (c >> 4);                         // byte0 = n0 | n1 << 4, byte2 = n2 | n3 << 4
  return (p & 0xFFu) | ((p >> 8) & 0xFF00u);         // 4 nibbles
}

__device__ __forceinline__ LaneBytes load_lane_classes(const uint8_t* __restrict__ doc,
                                                       uint32_t len, uint32_t base, uint32_t sep,
                                                       bool bom) {
  uint4 v = make_uint4(0, 0, 0, 0);
  if (base < len) v = __ldg(reinterpret_cast<const uint4*>(doc + base));
  if (base + 16 <= len && !bom) {   // whole vector inside the document: no positional special cases
    uint32_t sep4 = sep * 0x01010101u;
    unsigned long long lo = classes_of_word(v.x, sep4) | (classes_of_word(v.y, sep4) << 16);
    unsigned long long hi = classes_of_word(v.z, sep4) | (classes_of_word(v.w, sep4) << 16);
    return LaneBytes{lo | (hi << 32)};
  }
  uint32_t w[4] = {v.x, v.y, v.z, v.w};
  unsigned long long cls = 0;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    uint32_t pos = base + k;
    uint32_t b = (w[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
    int c;
    if (pos < len) c = (bom && pos < 3) ? CL_PAD : csv_class(b, sep);
    else c = (pos == len) ? CL_LF : CL_PAD;
    cls |= (unsigned long long)c << (4 * k);
  }
  return LaneBytes{cls};
}

// lane function by reverse composition: G <- G o (t_2k then t_2k+1), running G kept as a byte
// table; pair_fn[c0 | c1 << 4] is the function of two consecutive bytes of classes c0, c1, so the
// 16 bytes cost 8 look-ups and 16 PRMTs
__device__ __forceinline__ uint32_t lane_function(unsigned long long cls,
                                                  const uint32_t* __restrict__ pair_fn) {
  uint32_t glo = 0x03020100u, ghi = 0x07060504u;
#pragma unroll
  for (int k = 7; k >= 0; --k) {
    uint32_t sel = pair_fn[(cls >> (8 * k)) & 0x7Fu];
    uint32_t nlo = __byte_perm(glo, ghi, sel & 0xFFFFu);
    uint32_t nhi = __byte_perm(glo, ghi, sel >> 16);
    glo = nlo; ghi = nhi;
  }
  uint32_t p0 = glo | (glo >> 4), p1 = ghi | (ghi >> 4);
  uint32_t lo = (p0 & 0xFFu) | ((p0 >> 8) & 0xFF00u);
  uint32_t hi = (p1 & 0xFFu) | ((p1 >> 8) & 0xFF00u);
  return lo | (hi << 16);
}

__device__ __forceinline__ bool doc_has_bom(const uint8_t* __restrict__ doc, uint32_t len) {
  return len >= 3 && doc[0] == 0xEF && doc[1] == 0xBB && doc[2] == 0xBF;
}

#define RFQ_INIT_CLASS_FN()                                         \
  __shared__ uint32_t class_fn[8], class_ev[8];                     \
  __shared__ uint2 class_fe[8];   /* {next states, events} of a class in one 8-byte load */ \
  if (threadIdx.x < 8) {                                            \
    class_fn[threadIdx.x] = threadIdx.x < N_CLASSES ? csv_class_fn(threadIdx.x) : RFQ_FN_IDENTITY; \
    class_ev[threadIdx.x] = threadIdx.x < N_CLASSES ? csv_class_ev(threadIdx.x) : 0u;              \
    class_fe[threadIdx.x] = make_uint2(class_fn[threadIdx.x], class_ev[threadIdx.x]);              \
  }                                                                 \
  __syncthreads();

}  // namespace rfq
