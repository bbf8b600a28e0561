// Batched MD5 of the staged documents: what FileParser._get_file_hash_async
// computes per file (app/file_parser.py:122-130, hashlib.md5 over the file bytes -> 32 hex chars),
// for a whole batch in one launch.  MD5 (RFC 1321) is sequential inside a document, so the
// parallelism is across documents: one thread per document, 64-byte blocks read as four 16-byte
// loads (documents start 16-byte aligned in the arena).  Meant for batches of many small
// documents; the host mirror keeps hashlib for the rare multi-megabyte file.  Integer work, no
// tensor cores.
#include <cuda_runtime.h>
#include "launch.h"
#include "pipeline.h"

namespace rfq {

__device__ __forceinline__ uint32_t rotl32(uint32_t x, int s) { return __funnelshift_l(x, x, s); }

#define MD5_STEP(f, a, b, c, d, x, t, s) \
  a += f(b, c, d) + (x) + (t);           \
  a = rotl32(a, s) + b;
#define MD5_F(x, y, z) ((z) ^ ((x) & ((y) ^ (z))))
#define MD5_G(x, y, z) ((y) ^ ((z) & ((x) ^ (y))))
#define MD5_H(x, y, z) ((x) ^ (y) ^ (z))
#define MD5_I(x, y, z) ((y) ^ ((x) | ~(z)))

// one 64-byte block; w = the block as 16 little-endian words; constants T[i] = floor(2^32 |sin(i+1)|)
__device__ __forceinline__ void md5_block(uint32_t& A, uint32_t& B, uint32_t& C, uint32_t& D,
                                          const uint32_t* w) {
  uint32_t a = A, b = B, c = C, d = D;
  MD5_STEP(MD5_F, a, b, c, d, w[0], 0xd76aa478u, 7)   MD5_STEP(MD5_F, d, a, b, c, w[1], 0xe8c7b756u, 12)
  MD5_STEP(MD5_F, c, d, a, b, w[2], 0x242070dbu, 17)  MD5_STEP(MD5_F, b, c, d, a, w[3], 0xc1bdceeeu, 22)
  MD5_STEP(MD5_F, a, b, c, d, w[4], 0xf57c0fafu, 7)   MD5_STEP(MD5_F, d, a, b, c, w[5], 0x4787c62au, 12)
  MD5_STEP(MD5_F, c, d, a, b, w[6], 0xa8304613u, 17)  MD5_STEP(MD5_F, b, c, d, a, w[7], 0xfd469501u, 22)
  MD5_STEP(MD5_F, a, b, c, d, w[8], 0x698098d8u, 7)   MD5_STEP(MD5_F, d, a, b, c, w[9], 0x8b44f7afu, 12)
  MD5_STEP(MD5_F, c, d, a, b, w[10], 0xffff5bb1u, 17) MD5_STEP(MD5_F, b, c, d, a, w[11], 0x895cd7beu, 22)
  MD5_STEP(MD5_F, a, b, c, d, w[12], 0x6b901122u, 7)  MD5_STEP(MD5_F, d, a, b, c, w[13], 0xfd987193u, 12)
  MD5_STEP(MD5_F, c, d, a, b, w[14], 0xa679438eu, 17) MD5_STEP(MD5_F, b, c, d, a, w[15], 0x49b40821u, 22)

  MD5_STEP(MD5_G, a, b, c, d, w[1], 0xf61e2562u, 5)   MD5_STEP(MD5_G, d, a, b, c, w[6], 0xc040b340u, 9)
  MD5_STEP(MD5_G, c, d, a, b, w[11], 0x265e5a51u, 14) MD5_STEP(MD5_G, b, c, d, a, w[0], 0xe9b6c7aau, 20)
  MD5_STEP(MD5_G, a, b, c, d, w[5], 0xd62f105du, 5)   MD5_STEP(MD5_G, d, a, b, c, w[10], 0x02441453u, 9)
  MD5_STEP(MD5_G, c, d, a, b, w[15], 0xd8a1e681u, 14) MD5_STEP(MD5_G, b, c, d, a, w[4], 0xe7d3fbc8u, 20)
  MD5_STEP(MD5_G, a, b, c, d, w[9], 0x21e1cde6u, 5)   MD5_STEP(MD5_G, d, a, b, c, w[14], 0xc33707d6u, 9)
  MD5_STEP(MD5_G, c, d, a, b, w[3], 0xf4d50d87u, 14)  MD5_STEP(MD5_G, b, c, d, a, w[8], 0x455a14edu, 20)
  MD5_STEP(MD5_G, a, b, c, d, w[13], 0xa9e3e905u, 5)  MD5_STEP(MD5_G, d, a, b, c, w[2], 0xfcefa3f8u, 9)
  MD5_STEP(MD5_G, c, d, a, b, w[7], 0x676f02d9u, 14)  MD5_STEP(MD5_G, b, c, d, a, w[12], 0x8d2a4c8au, 20)

  MD5_STEP(MD5_H, a, b, c, d, w[5], 0xfffa3942u, 4)   MD5_STEP(MD5_H, d, a, b, c, w[8], 0x8771f681u, 11)
  MD5_STEP(MD5_H, c, d, a, b, w[11], 0x6d9d6122u, 16) MD5_STEP(MD5_H, b, c, d, a, w[14], 0xfde5380cu, 23)
  MD5_STEP(MD5_H, a, b, c, d, w[1], 0xa4beea44u, 4)   MD5_STEP(MD5_H, d, a, b, c, w[4], 0x4bdecfa9u, 11)
  MD5_STEP(MD5_H, c, d, a, b, w[7], 0xf6bb4b60u, 16)  MD5_STEP(MD5_H, b, c, d, a, w[10], 0xbebfbc70u, 23)
  MD5_STEP(MD5_H, a, b, c, d, w[13], 0x289b7ec6u, 4)  MD5_STEP(MD5_H, d, a, b, c, w[0], 0xeaa127fau, 11)
  MD5_STEP(MD5_H, c, d, a, b, w[3], 0xd4ef3085u, 16)  MD5_STEP(MD5_H, b, c, d, a, w[6], 0x04881d05u, 23)
  MD5_STEP(MD5_H, a, b, c, d, w[9], 0xd9d4d039u, 4)   MD5_STEP(MD5_H, d, a, b, c, w[12], 0xe6db99e5u, 11)
  MD5_STEP(MD5_H, c, d, a, b, w[15], 0x1fa27cf8u, 16) MD5_STEP(MD5_H, b, c, d, a, w[2], 0xc4ac5665u, 23)

  MD5_STEP(MD5_I, a, b, c, d, w[0], 0xf4292244u, 6)   MD5_STEP(MD5_I, d, a, b, c, w[7], 0x432aff97u, 10)
  MD5_STEP(MD5_I, c, d, a, b, w[14], 0xab9423a7u, 15) MD5_STEP(MD5_I, b, c, d, a, w[5], 0xfc93a039u, 21)
  MD5_STEP(MD5_I, a, b, c, d, w[12], 0x655b59c3u, 6)  MD5_STEP(MD5_I, d, a, b, c, w[3], 0x8f0ccc92u, 10)
  MD5_STEP(MD5_I, c, d, a, b, w[10], 0xffeff47du, 15) MD5_STEP(MD5_I, b, c, d, a, w[1], 0x85845dd1u, 21)
  MD5_STEP(MD5_I, a, b, c, d, w[8], 0x6fa87e4fu, 6)   MD5_STEP(MD5_I, d, a, b, c, w[15], 0xfe2ce6e0u, 10)
  MD5_STEP(MD5_I, c, d, a, b, w[6], 0xa3014314u, 15)  MD5_STEP(MD5_I, b, c, d, a, w[13], 0x4e0811a1u, 21)
  MD5_STEP(MD5_I, a, b, c, d, w[4], 0xf7537e82u, 6)   MD5_STEP(MD5_I, d, a, b, c, w[11], 0xbd3af235u, 10)
  MD5_STEP(MD5_I, c, d, a, b, w[2], 0x2ad7d2bbu, 15)  MD5_STEP(MD5_I, b, c, d, a, w[9], 0xeb86d391u, 21)
  A += a; B += b; C += c; D += d;
}

__global__ void __launch_bounds__(128) md5_kernel(const uint8_t* __restrict__ arena,
                                                  const uint32_t* __restrict__ doc_off,
                                                  const uint32_t* __restrict__ doc_len, uint32_t n_docs,
                                                  uint8_t* __restrict__ out,
                                                  unsigned long long* __restrict__ out_off,
                                                  unsigned long long* __restrict__ out_len) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= n_docs) return;
  const uint8_t* doc = arena + doc_off[d];
  const uint32_t len = doc_len[d];
  uint32_t A = 0x67452301u, B = 0xefcdab89u, C = 0x98badcfeu, D = 0x10325476u;
  uint32_t w[16];
  const uint32_t full = len >> 6;
  for (uint32_t b = 0; b < full; ++b) {
    const uint4* p = reinterpret_cast<const uint4*>(doc + ((size_t)b << 6));
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint4 v = __ldg(p + q);
      w[4 * q] = v.x; w[4 * q + 1] = v.y; w[4 * q + 2] = v.z; w[4 * q + 3] = v.w;
    }
    md5_block(A, B, C, D, w);
  }
  // tail: remaining bytes, 0x80, zero fill, 64-bit bit length
  const uint32_t rem = len & 63u;
#pragma unroll
  for (int q = 0; q < 16; ++q) w[q] = 0;
  const uint8_t* t = doc + ((size_t)full << 6);
  for (uint32_t i = 0; i < rem; ++i) w[i >> 2] |= (uint32_t)t[i] << ((i & 3) * 8);
  w[rem >> 2] |= 0x80u << ((rem & 3) * 8);
  if (rem >= 56) {
    md5_block(A, B, C, D, w);
#pragma unroll
    for (int q = 0; q < 16; ++q) w[q] = 0;
  }
  w[14] = len << 3;
  w[15] = len >> 29;
  md5_block(A, B, C, D, w);
  uint8_t* o = out + (size_t)d * 32;
  const uint32_t st[4] = {A, B, C, D};
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    uint32_t byte = (st[k >> 2] >> ((k & 3) * 8)) & 0xFFu;
    uint32_t hi = byte >> 4, lo = byte & 15u;
    o[2 * k] = (uint8_t)(hi < 10 ? '0' + hi : 'a' + hi - 10);
    o[2 * k + 1] = (uint8_t)(lo < 10 ? '0' + lo : 'a' + lo - 10);
  }
  out_off[d] = (unsigned long long)d * 32;
  out_len[d] = 32;
}

void md5_batch(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, uint32_t n_docs,
               uint8_t* out, unsigned long long* out_off, unsigned long long* out_len, cudaStream_t s) {
  if (n_docs == 0) return;
  prof_mark("md5_kernel", s);
  RFQ_LAUNCH(md5_kernel, (n_docs + 127) / 128, 128, 0, s, arena, doc_off, doc_len, n_docs, out, out_off,
             out_len);
}

}  // namespace rfq
