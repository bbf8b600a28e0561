// Device-side DEFLATE: the zip members of an .xlsx container (worksheet parts,
// sharedStrings) inflated on the GPU, so that the host neither spends its cores on zlib nor ships the
// inflated XML across PCIe — the reference does this step inside pd.read_excel (app/file_parser.py:225,
// zipfile + zlib).  RFC 1951: stored, fixed-Huffman and dynamic-Huffman blocks; RFC 1952 CRC-32 of the
// inflated bytes checked against the zip directory's.
//
// A DEFLATE stream is sequential (every code's position depends on all codes before it), so the
// parallelism is across members: one WARP per member.  Lane 0 decodes — a 10-bit lookup table for the
// literal / length code and an 8-bit one for the distance code in shared memory (built by the whole warp
// from the canonical code lengths, longer codes fall back to the count-per-length walk), a 64-bit bit
// buffer refilled with aligned 32-bit loads — and broadcasts every symbol; literals are stored by lane 0,
// matches are copied by all 32 lanes (byte k of a match comes from out[n - dist + k mod dist], so
// overlapping matches need no special case).  The CRC-32 of the result is taken four bytes per step
// (slice-by-4 tables in global memory).  Members are handed out by an atomic counter, largest first is
// not needed: a batch of workbooks has thousands in flight.
// Bit / byte work with data-dependent control flow: no tensor cores, by design.
#include <cuda_runtime.h>
#include <algorithm>
#include "../../include/rfq_parse.h"
#include "launch.h"
#include "pipeline.h"

namespace rfq {

namespace {

#define INF_LBITS 10
#define INF_DBITS 8

struct WarpTabs {                 // per warp, shared memory
  uint16_t llut[1 << INF_LBITS];  // (symbol << 4) | code length, 0 = longer code / unused
  uint16_t dlut[1 << INF_DBITS];
  uint16_t lcount[16], dcount[16];
  uint16_t lsym[288], dsym[32];
  uint16_t code[320];             // canonical code of every symbol while the tables are built
  uint8_t len[320];               // code lengths
};

struct Bits {
  const uint32_t* w;              // aligned words of the member (it starts on a 16-byte boundary)
  uint32_t pos, nwords;           // next word, words that hold stream bytes
  uint64_t buf;
  int cnt;
  __device__ __forceinline__ void refill() {
    if (cnt <= 32) { buf |= (uint64_t)(pos < nwords ? __ldg(w + pos) : 0u) << cnt; ++pos; cnt += 32; }
  }
  __device__ __forceinline__ uint32_t peek(int n) const { return (uint32_t)buf & ((1u << n) - 1u); }
  __device__ __forceinline__ void drop(int n) { buf >>= n; cnt -= n; }
  __device__ __forceinline__ uint32_t get(int n) { refill(); const uint32_t v = peek(n); drop(n); return v; }
  __device__ __forceinline__ uint64_t consumed_bits() const { return (uint64_t)pos * 32ull - (uint64_t)cnt; }
};

__constant__ uint16_t c_lens[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
__constant__ uint8_t c_lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
__constant__ uint16_t c_dists[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
__constant__ uint8_t c_dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
__constant__ uint8_t c_order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
__device__ uint32_t g_crc_table[4][256];

// Canonical code from the lengths in T.len[base .. base + n): count / symbol arrays for the slow path and
// the lookup table for codes of at most `bits` bits.  Whole warp.  Returns < 0 over-subscribed,
// > 0 incomplete, 0 complete (all lanes the same value).
__device__ int build_tables(WarpTabs& T, int base, int n, uint16_t* count, uint16_t* symbol, uint16_t* lut, int bits, int lane) {
  for (int i = lane; i < (1 << bits); i += 32) lut[i] = 0;
  if (lane < 16) count[lane] = 0;
  __syncwarp();
  int left = 1;
  if (lane == 0) {
    for (int s = 0; s < n; ++s) count[T.len[base + s]]++;
    uint16_t offs[16], next[16];
    offs[1] = 0;
    int code = 0;
    next[0] = 0;
    for (int l = 1; l < 16; ++l) {
      left <<= 1; left -= count[l];
      if (l < 15) offs[l + 1] = offs[l] + count[l];
      code = (code + (l > 1 ? count[l - 1] : 0)) << 1;
      next[l] = (uint16_t)code;
      if (left < 0) break;
    }
    if (left >= 0)
      for (int s = 0; s < n; ++s) {
        const int l = T.len[base + s];
        if (l) { symbol[offs[l]++] = (uint16_t)s; T.code[s] = next[l]++; }
      }
    if (count[0] == n) left = 0;
  }
  left = __shfl_sync(0xffffffffu, left, 0);
  __syncwarp();
  if (left < 0) return left;
  for (int s = lane; s < n; s += 32) {
    const int l = T.len[base + s];
    if (l && l <= bits) {
      const uint32_t r = __brev((uint32_t)T.code[s]) >> (32 - l);       // the stream carries codes most significant bit first
      const uint16_t e = (uint16_t)((s << 4) | l);
      for (uint32_t k = r; k < (1u << bits); k += 1u << l) lut[k] = e;
    }
  }
  __syncwarp();
  return left;
}

// one symbol of a code: table hit, or the count-per-length walk for codes longer than the table
__device__ __forceinline__ int decode_sym(Bits& b, const uint16_t* lut, int bits, const uint16_t* count, const uint16_t* symbol) {
  b.refill();
  const uint32_t e = lut[b.peek(bits)];
  if (e & 15u) { b.drop((int)(e & 15u)); return (int)(e >> 4); }
  int code = 0, first = 0, index = 0;
  uint64_t v = b.buf;
  for (int l = 1; l < 16; ++l) {
    code |= (int)(v & 1u); v >>= 1;
    const int c = count[l];
    if (code - c < first) { b.drop(l); return symbol[index + (code - first)]; }
    index += c; first += c;
    first <<= 1; code <<= 1;
  }
  return -1;
}

}  // namespace

__global__ void __launch_bounds__(64) crc_table_kernel() {
  for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) {
    uint32_t c = i;
    for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1;
    g_crc_table[0][i] = c;
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < 256; i += blockDim.x) {
    uint32_t c = g_crc_table[0][i];
    for (int t = 1; t < 4; ++t) { c = g_crc_table[0][c & 0xFFu] ^ (c >> 8); g_crc_table[t][i] = c; }
  }
}

// status: RFQ_DOC_OK, or RFQ_DOC_CORRUPT (bad stream, size or CRC mismatch, unknown method)
__global__ void __launch_bounds__(128) inflate_kernel(const uint8_t* __restrict__ arena, const uint32_t* __restrict__ doc_off,
                                                      const uint32_t* __restrict__ doc_len, const uint8_t* __restrict__ method,
                                                      const unsigned long long* __restrict__ out_off,
                                                      const unsigned long long* __restrict__ out_len,
                                                      const uint32_t* __restrict__ crc_want, uint32_t n_docs,
                                                      uint8_t* __restrict__ out, int32_t* __restrict__ status,
                                                      uint32_t* __restrict__ ticket) {
  __shared__ WarpTabs tabs[4];
  const int lane = threadIdx.x & 31;
  WarpTabs& T = tabs[threadIdx.x >> 5];
  for (;;) {
    uint32_t d = 0;
    if (lane == 0) d = atomicAdd(ticket, 1u);
    d = __shfl_sync(0xffffffffu, d, 0);
    if (d >= n_docs) return;
    const uint8_t* src = arena + doc_off[d];
    const uint64_t clen = doc_len[d], cap = out_len[d];
    uint8_t* dst = out + out_off[d];
    uint64_t n = 0;
    int err = 0;
    if (method[d] == 0) {                      // stored member: a copy by the whole warp
      if (clen != cap) err = 1;
      else { for (uint64_t k = lane; k < cap; k += 32) dst[k] = src[k]; n = cap; }
    } else if (method[d] == 8) {
      Bits b{reinterpret_cast<const uint32_t*>(src), 0, (uint32_t)((clen + 3) >> 2), 0, 0};
      int last = 0;
      while (!last && !err) {
        uint32_t hdr = 0;
        if (lane == 0) hdr = b.get(3);
        hdr = __shfl_sync(0xffffffffu, hdr, 0);
        last = (int)(hdr & 1u);
        const uint32_t type = hdr >> 1;
        if (type == 0) {                        // stored block: to the byte boundary, LEN, ~LEN, the bytes
          uint32_t len = 0, bytepos = 0;
          if (lane == 0) {
            b.drop(b.cnt & 7);
            len = b.get(16);
            const uint32_t nlen = b.get(16);
            if (len != (~nlen & 0xFFFFu)) len = 0xFFFFFFFFu;
            bytepos = (uint32_t)(b.consumed_bits() >> 3);
          }
          len = __shfl_sync(0xffffffffu, len, 0);
          bytepos = __shfl_sync(0xffffffffu, bytepos, 0);
          if (len == 0xFFFFFFFFu || (uint64_t)bytepos + len > clen) { err = 1; break; }
          if (n + len > cap) { err = 1; break; }
          for (uint32_t k = lane; k < len; k += 32) dst[n + k] = src[bytepos + k];
          n += len;
          if (lane == 0) {                      // continue after the copied bytes
            const uint32_t np = bytepos + len;
            b.pos = np >> 2; b.buf = 0; b.cnt = 0;
            b.refill();
            b.drop((int)(np & 3u) * 8);
          }
          __syncwarp();
          continue;
        }
        int nlen = 288, ndist = 30;
        if (type == 1) {                        // fixed codes
          for (int s = lane; s < 288; s += 32) T.len[s] = s < 144 ? 8 : s < 256 ? 9 : s < 280 ? 7 : 8;
          for (int s = lane; s < 30; s += 32) T.len[288 + s] = 5;
          __syncwarp();
        } else if (type == 2) {                 // dynamic codes: the code-length code first
          int ok = 1;
          if (lane == 0) {
            nlen = (int)b.get(5) + 257; ndist = (int)b.get(5) + 1;
            const int ncode = (int)b.get(4) + 4;
            if (nlen > 286 || ndist > 30) ok = 0;
            int idx = 0;
            for (; idx < ncode; ++idx) T.len[c_order[idx]] = (uint8_t)b.get(3);
            for (; idx < 19; ++idx) T.len[c_order[idx]] = 0;
          }
          nlen = __shfl_sync(0xffffffffu, nlen, 0); ndist = __shfl_sync(0xffffffffu, ndist, 0);
          ok = __shfl_sync(0xffffffffu, ok, 0);
          __syncwarp();
          if (!ok || build_tables(T, 0, 19, T.lcount, T.lsym, T.llut, 7, lane) != 0) { err = 1; break; }
          if (lane == 0) {
            int idx = 0;
            while (idx < nlen + ndist && ok) {
              const int sym = decode_sym(b, T.llut, 7, T.lcount, T.lsym);
              if (sym < 0) { ok = 0; break; }
              if (sym < 16) T.code[idx++] = (uint16_t)sym;     // T.code as scratch: the table build below rewrites it
              else {
                uint16_t prev = 0;
                int rep;
                if (sym == 16) { if (idx == 0) { ok = 0; break; } prev = T.code[idx - 1]; rep = 3 + (int)b.get(2); }
                else if (sym == 17) rep = 3 + (int)b.get(3);
                else rep = 11 + (int)b.get(7);
                if (idx + rep > nlen + ndist) { ok = 0; break; }
                while (rep--) T.code[idx++] = prev;
              }
            }
            if (ok) {
              if (T.code[256] == 0) ok = 0;
              for (int s = 0; s < nlen; ++s) T.len[s] = (uint8_t)T.code[s];
              for (int s = 0; s < ndist; ++s) T.len[288 + s] = (uint8_t)T.code[nlen + s];
            }
          }
          ok = __shfl_sync(0xffffffffu, ok, 0);
          __syncwarp();
          if (!ok) { err = 1; break; }
        } else { err = 1; break; }
        {
          int r = build_tables(T, 0, nlen, T.lcount, T.lsym, T.llut, INF_LBITS, lane);
          if (r < 0 || (r > 0 && type == 2 && nlen - T.lcount[0] != 1)) { err = 1; break; }
          r = build_tables(T, 288, ndist, T.dcount, T.dsym, T.dlut, INF_DBITS, lane);
          if (r < 0 || (r > 0 && type == 2 && ndist - T.dcount[0] != 1)) { err = 1; break; }
        }
        // the block's symbols: lane 0 decodes, every lane copies
        for (;;) {
          uint32_t len = 0, dist = 0;          // len 0 = literal (lane 0 has stored it), 0xFFFFFFFF = end / error
          if (lane == 0) {
            for (;;) {                          // a run of literals stays on lane 0
              const int sym = decode_sym(b, T.llut, INF_LBITS, T.lcount, T.lsym);
              if (sym < 0) { len = 0xFFFFFFFFu; dist = 1; break; }
              if (sym < 256) {
                if (n >= cap) { len = 0xFFFFFFFFu; dist = 1; break; }
                dst[n++] = (uint8_t)sym;
                continue;
              }
              if (sym == 256) { len = 0xFFFFFFFFu; dist = 0; break; }
              const int ls = sym - 257;
              if (ls >= 29) { len = 0xFFFFFFFFu; dist = 1; break; }
              len = c_lens[ls] + b.get(c_lext[ls]);
              const int ds = decode_sym(b, T.dlut, INF_DBITS, T.dcount, T.dsym);
              if (ds < 0 || ds >= 30) { len = 0xFFFFFFFFu; dist = 1; break; }
              dist = c_dists[ds] + b.get(c_dext[ds]);
              if (dist > n || n + len > cap || b.consumed_bits() > clen * 8 + 64) { len = 0xFFFFFFFFu; dist = 1; }
              break;
            }
          }
          len = __shfl_sync(0xffffffffu, len, 0);
          dist = __shfl_sync(0xffffffffu, dist, 0);
          n = __shfl_sync(0xffffffffu, n, 0);
          if (len == 0xFFFFFFFFu) { if (dist) err = 1; break; }
          __syncwarp();                         // lane 0's literals are visible to the copying lanes
          const uint8_t* from = dst + n - dist;
          for (uint32_t k = lane; k < len; k += 32) dst[n + k] = from[dist >= len ? k : k % dist];
          n += len;
          __syncwarp();
        }
      }
      uint64_t used = 0;
      if (lane == 0) used = b.consumed_bits();
      used = __shfl_sync(0xffffffffu, used, 0);
      if (!err && used > clen * 8) err = 1;      // ran past the member's bytes
    } else err = 1;
    if (!err && n != cap) err = 1;
    __syncwarp();
    if (!err && lane == 0) {                     // CRC-32 of what was written, four bytes per step
      uint32_t c = 0xFFFFFFFFu;
      uint64_t k = 0;
      while (k < n && ((uintptr_t)(dst + k) & 3u)) { c = g_crc_table[0][(c ^ dst[k]) & 0xFFu] ^ (c >> 8); ++k; }
      for (; k + 4 <= n; k += 4) {
        c ^= *reinterpret_cast<const uint32_t*>(dst + k);
        c = g_crc_table[3][c & 0xFFu] ^ g_crc_table[2][(c >> 8) & 0xFFu] ^ g_crc_table[1][(c >> 16) & 0xFFu] ^ g_crc_table[0][c >> 24];
      }
      for (; k < n; ++k) c = g_crc_table[0][(c ^ dst[k]) & 0xFFu] ^ (c >> 8);
      if ((c ^ 0xFFFFFFFFu) != crc_want[d]) err = 1;
    }
    if (lane == 0) status[d] = err ? RFQ_DOC_CORRUPT : RFQ_DOC_OK;
    __syncwarp();
  }
}

void inflate_batch(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, const uint8_t* method,
                   const unsigned long long* out_off, const unsigned long long* out_len, const uint32_t* crc,
                   uint32_t n_docs, uint8_t* out, int32_t* status, uint32_t* ticket, cudaStream_t s) {
  if (n_docs == 0) return;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaMemsetAsync(ticket, 0, 4, s);
  prof_mark("inflate_kernel", s);
  RFQ_LAUNCH(crc_table_kernel, 1, 64, 0, s);
  const uint32_t grid = std::min<uint32_t>((n_docs + 3) / 4, (uint32_t)sms * 12u);
  RFQ_LAUNCH(inflate_kernel, grid, 128, 0, s, arena, doc_off, doc_len, method, out_off, out_len, crc, n_docs,
             out, status, ticket);
}

}  // namespace rfq
