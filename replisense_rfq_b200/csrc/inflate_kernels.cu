// Device-side DEFLATE: the zip members of an .xlsx container (worksheet parts,
// sharedStrings) inflated on the GPU, so that the host neither spends its cores on zlib nor ships the
// inflated XML across PCIe — the reference does this step inside pd.read_excel (app/file_parser.py:225,
// zipfile + zlib).  RFC 1951: stored, fixed-Huffman and dynamic-Huffman blocks; RFC 1952 CRC-32 of the
// inflated bytes checked against the zip directory's.
//
// A DEFLATE stream is sequential (every code's position depends on all codes before it), so the
// parallelism is across members: one thread decodes one member with canonical-Huffman tables in its
// local memory (count-per-length + symbols in code order: 700 bytes), writes the literals, and copies
// matches out of its own earlier output.  A batch of workbooks has thousands of members in flight.
// Bit / byte work with data-dependent control flow: no tensor cores, by design.
#include <cuda_runtime.h>
#include "../../include/rfq_parse.h"
#include "launch.h"
#include "pipeline.h"

namespace rfq {

namespace {

struct Bits {
  const uint8_t* p;
  uint64_t pos, end;
  uint64_t buf;
  int cnt;
  bool bad;
  __device__ __forceinline__ void fill() {
    while (cnt <= 56 && pos < end) { buf |= (uint64_t)p[pos++] << cnt; cnt += 8; }
  }
  __device__ __forceinline__ uint32_t get(int n) {   // n <= 16
    if (cnt < n) { fill(); if (cnt < n) { bad = true; return 0; } }
    const uint32_t v = (uint32_t)buf & ((1u << n) - 1u);
    buf >>= n; cnt -= n;
    return v;
  }
};

struct Huff { uint16_t count[16]; uint16_t* symbol; };

// canonical code from code lengths; returns < 0 over-subscribed, > 0 incomplete, 0 complete
__device__ int huff_build(Huff& h, const uint16_t* length, int n) {
  for (int l = 0; l < 16; ++l) h.count[l] = 0;
  for (int s = 0; s < n; ++s) h.count[length[s]]++;
  if (h.count[0] == n) return 0;
  int left = 1;
  for (int l = 1; l < 16; ++l) { left <<= 1; left -= h.count[l]; if (left < 0) return left; }
  uint16_t offs[16];
  offs[1] = 0;
  for (int l = 1; l < 15; ++l) offs[l + 1] = offs[l] + h.count[l];
  for (int s = 0; s < n; ++s) if (length[s]) h.symbol[offs[length[s]]++] = (uint16_t)s;
  return left;
}

__device__ int huff_decode(Bits& b, const Huff& h) {
  int code = 0, first = 0, index = 0;
  for (int l = 1; l < 16; ++l) {
    code |= (int)b.get(1);
    if (b.bad) return -1;
    const int count = h.count[l];
    if (code - count < first) return h.symbol[index + (code - first)];
    index += count; first += count;
    first <<= 1; code <<= 1;
  }
  return -1;
}

__constant__ uint16_t c_lens[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
__constant__ uint8_t c_lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
__constant__ uint16_t c_dists[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
__constant__ uint8_t c_dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
__constant__ uint8_t c_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
__device__ uint32_t g_crc_table[256];

// literal / length and distance codes of one block: decode until end-of-block
__device__ int inflate_codes(Bits& b, const Huff& lc, const Huff& dc, uint8_t* out, uint64_t& n, uint64_t cap) {
  for (;;) {
    int sym = huff_decode(b, lc);
    if (sym < 0) return 1;
    if (sym < 256) { if (n >= cap) return 2; out[n++] = (uint8_t)sym; continue; }
    if (sym == 256) return 0;
    sym -= 257;
    if (sym >= 29) return 1;
    const uint32_t len = c_lens[sym] + b.get(c_lext[sym]);
    const int ds = huff_decode(b, dc);
    if (ds < 0 || ds >= 30) return 1;
    const uint32_t dist = c_dists[ds] + b.get(c_dext[ds]);
    if (b.bad || dist > n) return 1;
    if (n + len > cap) return 2;
    const uint8_t* src = out + n - dist;
    uint8_t* dst = out + n;
    for (uint32_t k = 0; k < len; ++k) dst[k] = src[k];   // overlapping copies repeat the pattern, byte by byte
    n += len;
  }
}

}  // namespace

__global__ void __launch_bounds__(64) crc_table_kernel() {
  for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
    g_crc_table[i] = c;
  }
}

// status: RFQ_DOC_OK, or RFQ_DOC_CORRUPT (bad stream, size or CRC mismatch, unknown method)
__global__ void __launch_bounds__(64) inflate_kernel(const uint8_t* __restrict__ arena, const uint32_t* __restrict__ doc_off,
                                                     const uint32_t* __restrict__ doc_len, const uint8_t* __restrict__ method,
                                                     const unsigned long long* __restrict__ out_off,
                                                     const unsigned long long* __restrict__ out_len,
                                                     const uint32_t* __restrict__ crc_want, uint32_t n_docs,
                                                     uint8_t* __restrict__ out, int32_t* __restrict__ status) {
  const uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= n_docs) return;
  const uint8_t* src = arena + doc_off[d];
  const uint64_t clen = doc_len[d], cap = out_len[d];
  uint8_t* dst = out + out_off[d];
  uint64_t n = 0;
  int err = 0;
  if (method[d] == 0) {                      // stored member
    if (clen != cap) err = 1;
    else for (; n < cap; ++n) dst[n] = src[n];
  } else if (method[d] == 8) {
    Bits b{src, 0, clen, 0, 0, false};
    uint16_t lsym[288], dsym[30], lengths[320];
    Huff lc, dc;
    lc.symbol = lsym; dc.symbol = dsym;
    int last = 0;
    while (!last && !err) {
      last = (int)b.get(1);
      const uint32_t type = b.get(2);
      if (b.bad) { err = 1; break; }
      if (type == 0) {                        // stored block: to the byte boundary, LEN, ~LEN
        b.buf >>= (b.cnt & 7); b.cnt -= (b.cnt & 7);
        const uint32_t len = b.get(16), nlen = b.get(16);
        if (b.bad || len != (~nlen & 0xFFFFu)) { err = 1; break; }
        if (n + len > cap) { err = 2; break; }
        for (uint32_t k = 0; k < len; ++k) { const uint32_t c = b.get(8); dst[n++] = (uint8_t)c; }
        if (b.bad) err = 1;
      } else if (type == 1) {                 // fixed codes
        int s = 0;
        for (; s < 144; ++s) lengths[s] = 8;
        for (; s < 256; ++s) lengths[s] = 9;
        for (; s < 280; ++s) lengths[s] = 7;
        for (; s < 288; ++s) lengths[s] = 8;
        huff_build(lc, lengths, 288);
        for (s = 0; s < 30; ++s) lengths[s] = 5;
        huff_build(dc, lengths, 30);
        err = inflate_codes(b, lc, dc, dst, n, cap);
      } else if (type == 2) {                 // dynamic codes
        const int nlen = (int)b.get(5) + 257, ndist = (int)b.get(5) + 1, ncode = (int)b.get(4) + 4;
        if (b.bad || nlen > 286 || ndist > 30) { err = 1; break; }
        int idx = 0;
        for (; idx < ncode; ++idx) lengths[c_order[idx]] = (uint16_t)b.get(3);
        for (; idx < 19; ++idx) lengths[c_order[idx]] = 0;
        if (b.bad || huff_build(lc, lengths, 19) != 0) { err = 1; break; }
        idx = 0;
        while (idx < nlen + ndist && !err) {
          int sym = huff_decode(b, lc);
          if (sym < 0) { err = 1; break; }
          if (sym < 16) lengths[idx++] = (uint16_t)sym;
          else {
            uint16_t prev = 0;
            int rep;
            if (sym == 16) { if (idx == 0) { err = 1; break; } prev = lengths[idx - 1]; rep = 3 + (int)b.get(2); }
            else if (sym == 17) rep = 3 + (int)b.get(3);
            else rep = 11 + (int)b.get(7);
            if (b.bad || idx + rep > nlen + ndist) { err = 1; break; }
            while (rep--) lengths[idx++] = prev;
          }
        }
        if (err) break;
        if (lengths[256] == 0) { err = 1; break; }
        int r = huff_build(lc, lengths, nlen);
        if (r < 0 || (r > 0 && nlen - lc.count[0] != 1)) { err = 1; break; }
        r = huff_build(dc, lengths + nlen, ndist);
        if (r < 0 || (r > 0 && ndist - dc.count[0] != 1)) { err = 1; break; }
        err = inflate_codes(b, lc, dc, dst, n, cap);
      } else err = 1;
    }
  } else err = 1;
  if (!err && n != cap) err = 1;
  if (!err) {                                 // CRC-32 of what was written
    uint32_t c = 0xFFFFFFFFu;
    for (uint64_t k = 0; k < n; ++k) c = g_crc_table[(c ^ dst[k]) & 0xFFu] ^ (c >> 8);
    if ((c ^ 0xFFFFFFFFu) != crc_want[d]) err = 1;
  }
  status[d] = err ? RFQ_DOC_CORRUPT : RFQ_DOC_OK;
}

void inflate_batch(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, const uint8_t* method,
                   const unsigned long long* out_off, const unsigned long long* out_len, const uint32_t* crc,
                   uint32_t n_docs, uint8_t* out, int32_t* status, cudaStream_t s) {
  if (n_docs == 0) return;
  prof_mark("inflate_kernel", s);
  RFQ_LAUNCH(crc_table_kernel, 1, 64, 0, s);
  RFQ_LAUNCH(inflate_kernel, (n_docs + 63) / 64, 64, 0, s, arena, doc_off, doc_len, method, out_off, out_len, crc, n_docs,
             out, status);
}

}  // namespace rfq
