// Text path (raw e-mail bodies / PDF-extracted text): UTF-8 decode with errors='ignore',
// universal newlines, str.strip() — reference app/file_parser.py:143-148.
//
// Layout: documents packed in one arena, each starting on a 16-byte boundary. A warp owns one
// 512-byte tile of one document (32 lanes x one 16-byte vector load); tiles never span
// documents. Three launches over the tiles:
//   txt_count_kernel   per-tile emitted-byte count + per-document first/last non-space char
//   (exclusive scan of tile counts)
//   txt_extent_kernel  per-document output length and rank of its first emitted byte
//   (exclusive scan of document lengths)
//   txt_emit_kernel    compaction through shared memory, coalesced stores
// HBM-bound byte work: no tensor cores.
#include <cuda_runtime.h>
#include <algorithm>
#include "rfq_common.cuh"
#include "utf8.cuh"
#include "scan.cuh"
#include "pipeline.h"

namespace rfq {

// 16 own bytes + 4 halo bytes on each side, in registers.
struct Win {
  uint32_t w[6];  // w[0] = bytes -4..-1, w[1..4] = own, w[5] = bytes 16..19
  __device__ __forceinline__ uint32_t operator()(int64_t i) const {  // i in [-4, 20)
    int j = (int)i + 4;
    return (w[j >> 2] >> ((j & 3) * 8)) & 0xFFu;
  }
};

// Eight bytes around a character, pulled out of the window ONCE (a dynamic index into the register
// window costs a select chain per byte): B(i) for i in [base, base + 8), offsets constant after inlining.
struct Near {
  uint64_t x;
  int base;
  __device__ __forceinline__ uint32_t operator()(int64_t i) const {
    return (uint32_t)(x >> (8 * ((int)i - base))) & 0xFFu;
  }
};
__device__ __forceinline__ Near near_at(const Win& w, int k) {   // bytes k .. k + 7, k in [-4, 15]; past byte 19: zero
  const int j = k + 4, q = j >> 2;
  const uint32_t sh = (uint32_t)(j & 3) * 8u;
  uint32_t a = w.w[0], b = w.w[1], c = w.w[2];
  if (q == 1) { a = w.w[1]; b = w.w[2]; c = w.w[3]; }
  else if (q == 2) { a = w.w[2]; b = w.w[3]; c = w.w[4]; }
  else if (q == 3) { a = w.w[3]; b = w.w[4]; c = w.w[5]; }
  else if (q >= 4) { a = w.w[4]; b = w.w[5]; c = 0u; }
  Near n;
  n.x = (unsigned long long)__funnelshift_r(a, b, sh) | ((unsigned long long)__funnelshift_r(b, c, sh) << 32);
  n.base = k;
  return n;
}

struct GlobalAcc {  // slow path: doc-relative accessor into global memory
  const uint8_t* p;
  int64_t len;
  __device__ __forceinline__ uint32_t operator()(int64_t i) const {
    return (i >= 0 && i < len) ? p[i] : 0u;
  }
};

__device__ __forceinline__ uint32_t mask_word(uint32_t v, int64_t pos, int64_t len) {
  // zero the bytes of a 4-byte word at doc-relative `pos` that lie outside [0, len)
  if (pos >= 0 && pos + 4 <= len) return v;
  uint32_t m = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k)
    if (pos + k >= 0 && pos + k < len) m |= 0xFFu << (8 * k);
  return v & m;
}

// Evaluate one lane's 16 bytes. `doc` points at the document start (16-byte aligned).
// emit: bit k = byte k is emitted; nsp: bit k = byte k starts an emitted non-space character (exact in its lowest
// and highest set bit, which is all anybody asks; non-ASCII characters between visible ASCII are not marked).
__device__ __forceinline__ void txt_lane_eval(const uint8_t* __restrict__ doc, uint32_t len,
                                              uint32_t tile, int lane, Win& win, uint32_t& emit,
                                              uint32_t& nsp) {
  int64_t base = (int64_t)tile * RFQ_TILE + lane * 16;
  uint4 v = make_uint4(0, 0, 0, 0);
  if (base < len) v = __ldg(reinterpret_cast<const uint4*>(doc + base));
  const int64_t left = (int64_t)len - base;
  const uint32_t valid = left >= 16 ? 0xFFFFu : left > 0 ? ((1u << left) - 1u) : 0u;
  if (valid != 0xFFFFu) {   // bytes beyond the document read as zero: four mask bits -> four mask bytes
#define RFQ_EXP4(m4) (((((m4) & 0xFu) * 0x00204081u) & 0x01010101u) * 0xFFu)
    v.x &= RFQ_EXP4(valid); v.y &= RFQ_EXP4(valid >> 4); v.z &= RFQ_EXP4(valid >> 8); v.w &= RFQ_EXP4(valid >> 12);
#undef RFQ_EXP4
  }
  uint32_t prev = __shfl_up_sync(0xffffffffu, v.w, 1);
  uint32_t next = __shfl_down_sync(0xffffffffu, v.x, 1);
  if (lane == 0)
    prev = base >= 4 ? __ldg(reinterpret_cast<const uint32_t*>(doc + base - 4)) : 0u;
  if (lane == 31) {
    next = 0u;
    if (base + 16 < len)
      next = mask_word(__ldg(reinterpret_cast<const uint32_t*>(doc + base + 16)), base + 16, len);
  }
  win.w[0] = prev; win.w[1] = v.x; win.w[2] = v.y; win.w[3] = v.z; win.w[4] = v.w;
  win.w[5] = next;
  // ASCII bytes are decided by 4-wide compares on the whole vector: a byte is emitted unless it is
  // the LF of a CR LF, and it starts visible text unless it is one of str.isspace()'s ASCII members
  // (\t-\r, \x1c-\x20).  Only non-ASCII CHARACTERS are walked, once each (their continuation bytes
  // are settled with the lead byte; a continuation byte whose lead sits in the previous lane looks
  // back through the halo).
  const uint32_t ww[4] = {v.x, v.y, v.z, v.w};
  uint32_t lf = 0, cr = 0, sp = 0, hi = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const uint32_t w = ww[q];
    uint32_t mlf = __vcmpeq4(w, 0x0A0A0A0Au), mcr = __vcmpeq4(w, 0x0D0D0D0Du);
    uint32_t msp = __vcmpleu4(__vsub4(w, 0x09090909u), 0x04040404u) |      // \t .. \r
                   __vcmpleu4(__vsub4(w, 0x1C1C1C1Cu), 0x04040404u);       // \x1c .. space
    uint32_t mhi = __vcmpgeu4(w, 0x80808080u);
    lf |= (((mlf & 0x08040201u) * 0x01010101u) >> 24) << (4 * q);
    cr |= (((mcr & 0x08040201u) * 0x01010101u) >> 24) << (4 * q);
    sp |= (((msp & 0x08040201u) * 0x01010101u) >> 24) << (4 * q);
    hi |= (((mhi & 0x08040201u) * 0x01010101u) >> 24) << (4 * q);
  }
  const uint32_t pb = prev >> 24;
  const uint32_t after_cr = ((cr << 1) | (pb == '\r' ? 1u : 0u)) & 0xFFFFu;
  const uint32_t after_hi = ((hi << 1) | (pb >= 0x80u ? 1u : 0u)) & 0xFFFFu;
  uint32_t swallowed = lf & ~after_hi & after_cr;
  for (uint32_t m = lf & after_hi & valid; m; m &= m - 1) {   // rare: dropped bytes may hide a CR
    const int k = __ffs(m) - 1;
    GlobalAcc g{doc, (int64_t)len};
    if (u8_lf_swallowed(g, base + k)) swallowed |= 1u << k;
  }
  emit = valid & ~hi & ~swallowed;
  nsp = emit & ~sp;
  for (uint32_t m = hi & valid; m;) {
    const int k = __ffs(m) - 1;
    m &= m - 1;
    const Near nb = near_at(win, k - 3);   // bytes k - 3 .. k + 4: all a lead or a continuation byte at k can ask for
    const uint32_t b = nb(k);
    if (u8_is_cont(b)) {              // only reached when no valid lead in this lane covers it
      if (k < 3 && u8_keep(nb, k)) emit |= 1u << k;
      continue;
    }
    const int n = u8_lead_len(b);
    if (n && u8_seq_valid(nb, k)) {
      const uint32_t span = (((1u << n) - 1u) << k) & 0xFFFFu;
      emit |= span;
      m &= ~span;
      // nsp is only ever asked for its lowest and highest bit: a character with visible ASCII on both sides in
      // this lane changes neither, whatever it is
      const uint32_t below = nsp & ((1u << k) - 1u), above = nsp >> k;
      if (!(below && above) && !u8_is_space_at(nb, k)) nsp |= 1u << k;
    }
  }
}

__device__ __forceinline__ uint32_t find_doc(const uint32_t* __restrict__ tile_base, uint32_t n,
                                             uint32_t t) {
  // largest d with tile_base[d] <= t  (tile_base has n+1 entries, ascending)
  uint32_t lo = 0, hi = n;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (__ldg(tile_base + mid) <= t) lo = mid; else hi = mid;
  }
  return lo;
}

__device__ __forceinline__ uint32_t below16(int64_t x) { return x <= 0 ? 0u : x >= 16 ? 0xFFFFu : ((1u << x) - 1u); }

__global__ void __launch_bounds__(256) txt_count_kernel(
    const uint8_t* __restrict__ arena, const uint32_t* __restrict__ doc_off,
    const uint32_t* __restrict__ doc_len, const uint32_t* __restrict__ tile_base, uint32_t n_docs,
    uint32_t n_tiles, uint32_t* __restrict__ tile_cnt, unsigned long long* __restrict__ first_key,
    unsigned long long* __restrict__ last_key, uint16_t* __restrict__ lane_emit) {
  // A warp owns a run of consecutive tiles (the grid is sized for the machine, not for the batch): the
  // document of a tile is found once per run and advanced from there, and no CTA lives for one tile only.
  const int lane = threadIdx.x & 31;
  const uint32_t nwarps = gridDim.x * 8u, gw = blockIdx.x * 8u + (threadIdx.x >> 5);
  const uint32_t per = (n_tiles + nwarps - 1u) / nwarps;
  uint32_t t = gw * per;
  const uint32_t t_end = min(t + per, n_tiles);
  if (t >= t_end) return;
  uint32_t d = warp_find_doc(tile_base, n_docs, t, lane);
  uint32_t next_tb = __ldg(tile_base + d + 1);
  for (; t < t_end; ++t) {
  while (t >= next_tb) { ++d; next_tb = __ldg(tile_base + d + 1); }
  uint32_t tile = t - __ldg(tile_base + d);
  uint32_t len = doc_len[d];
  const uint8_t* doc = arena + doc_off[d];
  Win win; uint32_t emit, nsp;
  txt_lane_eval(doc, len, tile, lane, win, emit, nsp);
  if (lane_emit) lane_emit[(size_t)t * 32u + lane] = (uint16_t)emit;
  uint32_t cnt = __popc(emit);
  uint32_t base = tile * RFQ_TILE + lane * 16;
  uint32_t fp = 0xFFFFFFFFu, le = 0u;
  if (nsp) {
    fp = base + (__ffs(nsp) - 1);
    int k = 31 - __clz(nsp);
    uint32_t b = near_at(win, k)(k);
    int cl = b < 0x80u ? 1 : u8_lead_len(b);
    le = base + k + cl;
  }
  cnt = __reduce_add_sync(0xffffffffu, cnt);     // one warp-wide reduction instruction each (redux.sync)
  fp = __reduce_min_sync(0xffffffffu, fp);
  le = __reduce_max_sync(0xffffffffu, le);
  if (fp == 0xFFFFFFFFu) { if (lane == 0) tile_cnt[t] = cnt; continue; }   // nothing visible in this tile
  // what the extent step needs from the tiles that win: the bytes this tile emits before its first visible
  // byte and before the end of its last visible character (which may reach into the next tile)
  const uint32_t tile_end = (tile + 1) * RFQ_TILE;
  uint32_t lead = __popc(emit & below16((int64_t)fp - base));
  uint32_t trail = __popc(emit & below16((int64_t)min(le, tile_end) - base));
  lead = __reduce_add_sync(0xffffffffu, lead);
  trail = __reduce_add_sync(0xffffffffu, trail);
  if (lane == 0) {
    tile_cnt[t] = cnt;
    const uint32_t spill = le > tile_end ? le - tile_end : 0u;           // at most 3
    atomicMin(first_key + d, ((unsigned long long)fp << 32) | lead);
    atomicMax(last_key + d, ((unsigned long long)le << 32) | (spill << 10) | trail);
  }
  }
}

// One thread per document: out_len and skip (global rank of the first emitted output byte), from the keys
// the count pass left (no second look at the bytes).
__global__ void __launch_bounds__(256) txt_extent_kernel(
    const uint32_t* __restrict__ tile_base, uint32_t n_docs, const uint32_t* __restrict__ tile_pre,
    const unsigned long long* __restrict__ first_key, const unsigned long long* __restrict__ last_key,
    unsigned long long* __restrict__ out_len, uint32_t* __restrict__ skip) {
  const uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= n_docs) return;
  const unsigned long long fk = first_key[d];
  if ((uint32_t)(fk >> 32) == 0xFFFFFFFFu) { out_len[d] = 0; skip[d] = 0; return; }
  const unsigned long long lk = last_key[d];
  const uint32_t tb = tile_base[d];
  const uint32_t fp = (uint32_t)(fk >> 32), le = (uint32_t)(lk >> 32);
  const uint32_t spill = ((uint32_t)lk >> 10) & 3u, trail = (uint32_t)lk & 1023u;
  const uint32_t r0 = tile_pre[tb + fp / RFQ_TILE] + (uint32_t)fk;
  const uint32_t r1 = tile_pre[tb + (le - spill - 1u) / RFQ_TILE] + trail + spill;
  out_len[d] = r1 - r0;
  skip[d] = r0;
}

__global__ void __launch_bounds__(256) txt_emit_kernel(
    const uint8_t* __restrict__ arena, const uint32_t* __restrict__ doc_off,
    const uint32_t* __restrict__ doc_len, const uint32_t* __restrict__ tile_base, uint32_t n_docs,
    uint32_t n_tiles, const uint32_t* __restrict__ tile_pre,
    const unsigned long long* __restrict__ first_key, const unsigned long long* __restrict__ last_key,
    const unsigned long long* __restrict__ out_off, const uint32_t* __restrict__ skip,
    uint8_t* __restrict__ out, const uint16_t* __restrict__ lane_emit) {
  __shared__ __align__(16) uint8_t stage[8][RFQ_TILE + 16];
  const int wib = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t nwarps = gridDim.x * 8u, gw = blockIdx.x * 8u + (uint32_t)wib;
  const uint32_t per = (n_tiles + nwarps - 1u) / nwarps;
  uint32_t t = gw * per;
  const uint32_t t_end = min(t + per, n_tiles);
  if (t >= t_end) return;
  uint32_t d = warp_find_doc(tile_base, n_docs, t, lane);
  uint32_t next_tb = __ldg(tile_base + d + 1);
  for (; t < t_end; ++t) {
  while (t >= next_tb) { ++d; next_tb = __ldg(tile_base + d + 1); }
  uint32_t fp = (uint32_t)(first_key[d] >> 32);
  if (fp == 0xFFFFFFFFu) continue;
  uint32_t le = (uint32_t)(last_key[d] >> 32);
  uint32_t tile = t - __ldg(tile_base + d);
  uint32_t tb0 = tile * RFQ_TILE;
  if (tb0 + RFQ_TILE <= fp || tb0 >= le) continue;  // tile entirely stripped
  uint32_t len = doc_len[d];
  const uint8_t* doc = arena + doc_off[d];
  Win win; uint32_t emit, nsp;
  uint32_t base = tb0 + lane * 16;
  if (lane_emit) {   // decided by the count pass: only the bytes themselves are needed
    emit = lane_emit[(size_t)t * 32u + lane];
    uint4 v = make_uint4(0, 0, 0, 0);
    if (base < len) v = __ldg(reinterpret_cast<const uint4*>(doc + base));
    win.w[0] = 0; win.w[1] = v.x; win.w[2] = v.y; win.w[3] = v.z; win.w[4] = v.w; win.w[5] = 0;
  } else {
    txt_lane_eval(doc, len, tile, lane, win, emit, nsp);
  }
  // restrict to [fp, le)
  uint32_t keep = emit;
  if (base < fp) keep &= (fp - base >= 16) ? 0u : ~((1u << (fp - base)) - 1u);
  if (base + 16 > le) keep &= (le <= base) ? 0u : ((1u << (le - base)) - 1u);
  uint32_t c = __popc(keep), incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t x = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += x;
  }
  uint32_t excl = incl - c;
  uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
  // rank of this tile's first kept byte within the document's output
  // emitted bytes in this tile that precede fp (non-zero only in the tile containing fp)
  uint32_t stripped_lead = 0u;
  if (base < fp)
    stripped_lead = (fp - base >= 16) ? __popc(emit) : __popc(emit & ((1u << (fp - base)) - 1u));
  stripped_lead = __reduce_add_sync(0xffffffffu, stripped_lead);
  // the tile's kept bytes are gathered in shared memory at the same 16-byte phase as their place in the
  // output, so that all but the ends leave as 16-byte vectors
  uint8_t* g = out + out_off[d] + (unsigned long long)(tile_pre[t] + stripped_lead - skip[d]);
  const uint32_t mis = (uint32_t)((uintptr_t)g & 15u);
  uint8_t* st = stage[wib] + mis;
  uint32_t pos = excl;
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    if (keep & (1u << k)) {
      uint32_t b = win(k);
      st[pos++] = (uint8_t)(b == '\r' ? '\n' : b);
    }
  }
  __syncwarp();
  {
    const uint32_t head = min(total, (16u - mis) & 15u);
    const uint32_t nvec = (total - head) >> 4, tail = (total - head) & 15u;
    if ((uint32_t)lane < head) g[lane] = st[lane];
    if ((uint32_t)lane < nvec)
      *reinterpret_cast<uint4*>(g + head + 16u * lane) = *reinterpret_cast<const uint4*>(st + head + 16u * lane);
    if ((uint32_t)lane >= 16u && (uint32_t)lane - 16u < tail)
      g[head + 16u * nvec + lane - 16u] = st[head + 16u * nvec + lane - 16u];
  }
  __syncwarp();   // the stage is the next tile's
  }
}

// ---------------------------------------------------------------------------------------------
static uint32_t txt_grid() {   // CTAs of 256 threads that fill the device once
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return (uint32_t)sms * 8u;
}

void txt_pipeline(const TxtArgs& a, cudaStream_t s) {
  if (a.n_docs == 0) return;
  cudaMemsetAsync(a.first_key, 0xFF, sizeof(unsigned long long) * a.n_docs, s);
  cudaMemsetAsync(a.last_key, 0, sizeof(unsigned long long) * a.n_docs, s);
  if (a.n_tiles) {
    const uint32_t nb = std::min<uint32_t>((a.n_tiles + 7) / 8, txt_grid());
    prof_mark("txt_count_kernel", s);
    RFQ_LAUNCH(txt_count_kernel, nb, 256, 0, s, a.arena, a.doc_off, a.doc_len, a.tile_base,
               a.n_docs, a.n_tiles, a.tile_cnt, a.first_key, a.last_key, a.lane_emit);
    prof_mark("scan(tile counts)", s);
    exclusive_scan<SumU32>(a.tile_cnt, a.tile_pre, a.n_tiles, a.scan_tmp32, nullptr, s);
  }
  prof_mark("txt_extent_kernel", s);
  RFQ_LAUNCH(txt_extent_kernel, (a.n_docs + 255) / 256, 256, 0, s, a.tile_base, a.n_docs, a.tile_pre, a.first_key,
             a.last_key, a.out_len, a.skip);
  prof_mark("scan(document sizes)", s);
  exclusive_scan<SumU64>(a.out_len, a.out_off, a.n_docs, a.scan_tmp64, a.out_total, s);
}

void txt_emit(const TxtArgs& a, uint8_t* out, cudaStream_t s) {
  if (a.n_tiles == 0) return;
  const uint32_t nb = std::min<uint32_t>((a.n_tiles + 7) / 8, txt_grid());
  RFQ_LAUNCH(txt_emit_kernel, nb, 256, 0, s, a.arena, a.doc_off, a.doc_len, a.tile_base, a.n_docs,
             a.n_tiles, a.tile_pre, a.first_key, a.last_key, a.out_off, a.skip, out, a.lane_emit);
}

}  // namespace rfq
