// Delimiter / quote-state scan and row / field splitter (SURVEY.md §7.1b K1 + K2).
//
// pandas' tokenizer state machine (csv_dfa.cuh) run in parallel: every lane turns its 16 bytes
// into a state->state transition function (8 states packed as nibbles, composed with PRMT),
// functions are composed across the warp (one 512-byte tile), tile functions are scanned across
// the arena, and a second sweep replays the DFA from the now-known start states to count and
// then index field ends, record ends and record starts.  Documents start on 16-byte boundaries
// and tiles never span documents; the end of each document is a virtual LF.
// HBM-bound byte work (the input is read three times, from L2 when it fits): no tensor cores.
#include <cuda_runtime.h>
#include <algorithm>
#include "csv_dfa.cuh"
#include "csv_scan_core.cuh"
#include "launch.h"
#include "pipeline.h"
#include "rfq_common.cuh"
#include "scan.cuh"

namespace rfq {

// ---- pass 1: transition function of every tile ------------------------------------------------
__global__ void __launch_bounds__(256) csv_tilefn_kernel(ScanArgs a) {
  RFQ_INIT_CLASS_FN();
  __shared__ uint32_t pair_fn[128];
  if (threadIdx.x < 128) pair_fn[threadIdx.x] = fn_compose(class_fn[threadIdx.x & 7u], class_fn[(threadIdx.x >> 4) & 7u]);
  __syncthreads();
  const int lane = threadIdx.x & 31;
  // a CTA keeps its tables and walks tiles grid-stride (8 tiles per round, one per warp)
  for (uint32_t t = blockIdx.x * 8 + (threadIdx.x >> 5); t < a.n_tiles; t += gridDim.x * 8) {
  uint32_t d = scan_find_doc(a.tile_base, a.n_docs, t);
  uint32_t tile = t - a.tile_base[d];
  uint32_t len = a.doc_len[d];
  const uint8_t* doc = a.arena + a.doc_off[d];
  bool bom = tile == 0 && doc_has_bom(doc, len);
  LaneBytes lb = load_lane_classes(doc, len, tile * RFQ_TILE + lane * 16, a.sep, bom);
  uint32_t f = lane_function(lb.cls, pair_fn);
  a.lane_fn[(size_t)t * 32 + lane] = f;   // pass 2 composes these instead of re-deriving them
  a.lane_cls[(size_t)t * 32 + lane] = lb.cls;   // passes 2 and 3 read 8 bytes of classes instead of 16 of text
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t g = __shfl_down_sync(0xffffffffu, f, o);
    if ((lane & (2 * o - 1)) == 0) f = fn_compose(f, g);
  }
  if (lane == 0) a.tile_fn[t] = tile == 0 ? fn_const(fn_apply(f, ST_R)) : f;
  }
}

// walk 16 bytes from a known state; calls ev(k, events, state_before) for bytes with events
// (table driven: next state and events are nibble look-ups in shared memory, no branches)
template <class F>
__device__ __forceinline__ int lane_walk(unsigned long long cls, int st,
                                         const uint2* __restrict__ class_fe, F&& ev) {
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    uint32_t c = (uint32_t)((cls >> (4 * k)) & 0xFu);
    const uint2 fe = class_fe[c];
    int e = (int)((fe.y >> (4 * st)) & 0xFu);
    if (e) ev(k, e, st);
    st = (int)((fe.x >> (4 * st)) & 0xFu);
  }
  return st;
}

// ---- pass 2: lane start states + per-tile counts ----------------------------------------------
__global__ void __launch_bounds__(256) csv_count_kernel(ScanArgs a) {
  RFQ_INIT_CLASS_FN();
  // two bytes per step: for classes (c0, c1) and every start state, one byte holding the state after
  // both bytes (bits 0-2), their field ends (bits 3-4) and their line ends (bits 5-6)
  __shared__ unsigned long long pair_walk[128];
  if (threadIdx.x < 128) {
    const uint32_t c0 = threadIdx.x & 7u, c1 = (threadIdx.x >> 4) & 7u;
    unsigned long long w = 0;
    for (int s0 = 0; s0 < 8; ++s0) {
      int e0 = (int)((class_ev[c0] >> (4 * s0)) & 0xFu), s1 = (int)((class_fn[c0] >> (4 * s0)) & 0xFu);
      int e1 = (int)((class_ev[c1] >> (4 * s1)) & 0xFu), s2 = (int)((class_fn[c1] >> (4 * s1)) & 0xFu);
      uint32_t f2 = ((e0 & EV_FIELD_END) ? 1u : 0u) + ((e1 & EV_FIELD_END) ? 1u : 0u);
      uint32_t l2 = ((e0 & EV_LINE_END) ? 1u : 0u) + ((e1 & EV_LINE_END) ? 1u : 0u);
      w |= (unsigned long long)((uint32_t)s2 | (f2 << 3) | (l2 << 5)) << (8 * s0);
    }
    pair_walk[threadIdx.x] = w;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  // a CTA keeps its tables and walks tiles grid-stride (8 tiles per round, one per warp)
  for (uint32_t t = blockIdx.x * 8 + (threadIdx.x >> 5); t < a.n_tiles; t += gridDim.x * 8) {
  uint32_t d = scan_find_doc(a.tile_base, a.n_docs, t);
  uint32_t tile = t - a.tile_base[d];
  uint32_t len = a.doc_len[d];
  const uint8_t* doc = a.arena + a.doc_off[d];
  bool bom = tile == 0 && doc_has_bom(doc, len);
  uint32_t base = tile * RFQ_TILE + lane * 16;
  LaneBytes lb{a.lane_cls[(size_t)t * 32 + lane]};
  uint32_t f = a.lane_fn[(size_t)t * 32 + lane];
  // inclusive scan of lane functions, then shift to exclusive
  uint32_t incl = f;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t g = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl = fn_compose(g, incl);
  }
  uint32_t excl = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) excl = RFQ_FN_IDENTITY;
  int tile_start = tile == 0 ? ST_R : fn_apply(a.tile_fn_pre[t], ST_R);
  int st = fn_apply(excl, tile_start);
  uint32_t b0 = __ballot_sync(0xffffffffu, st & 1);
  uint32_t b1 = __ballot_sync(0xffffffffu, st & 2);
  uint32_t b2 = __ballot_sync(0xffffffffu, st & 4);
  if (lane == 0) {
    a.lane_state[3 * t] = b0; a.lane_state[3 * t + 1] = b1; a.lane_state[3 * t + 2] = b2;
  }
  uint32_t nf = 0, nl = 0;
  uint32_t eof_k = (len >= base && len < base + 16) ? len - base : 0xFFu;
  int final_state = -1;
  {
    int s1 = st;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      uint32_t x = (uint32_t)(pair_walk[(lb.cls >> (8 * k)) & 0x7Fu] >> (8 * s1)) & 0xFFu;
      nf += (x >> 3) & 3u;
      nl += (x >> 5) & 3u;
      s1 = (int)(x & 7u);
    }
  }
  if (eof_k != 0xFFu) {  // state just before the virtual LF
    int s2 = st;
    for (uint32_t k = 0; k < eof_k; ++k) s2 = (int)((class_fn[(lb.cls >> (4 * k)) & 0xFu] >> (4 * s2)) & 0xFu);
    final_state = s2;
  }
  a.lane_cnt[(size_t)t * 32 + lane] = (uint16_t)(nf | (nl << 8));   // at most 16 of each per lane
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    nf += __shfl_xor_sync(0xffffffffu, nf, o);
    nl += __shfl_xor_sync(0xffffffffu, nl, o);
  }
  if (lane == 0) { a.tile_nf[t] = nf; a.tile_nl[t] = nl; }
  if (final_state >= 0) a.doc_final[d] = (uint32_t)final_state;
  }
}

// ---- pass 3: write field ends / record ends / record starts -----------------------------------
__global__ void __launch_bounds__(256) csv_index_kernel(ScanArgs a) {
  RFQ_INIT_CLASS_FN();
  const int lane = threadIdx.x & 31;
  // a CTA keeps its tables and walks tiles grid-stride (8 tiles per round, one per warp)
  for (uint32_t t = blockIdx.x * 8 + (threadIdx.x >> 5); t < a.n_tiles; t += gridDim.x * 8) {
  uint32_t d = scan_find_doc(a.tile_base, a.n_docs, t);
  uint32_t tile = t - a.tile_base[d];
  uint32_t len = a.doc_len[d];
  const uint8_t* doc = a.arena + a.doc_off[d];
  bool has_bom = doc_has_bom(doc, len);
  bool bom = tile == 0 && has_bom;
  uint32_t base = tile * RFQ_TILE + lane * 16;
  LaneBytes lb{a.lane_cls[(size_t)t * 32 + lane]};
  uint32_t b0 = a.lane_state[3 * t], b1 = a.lane_state[3 * t + 1], b2 = a.lane_state[3 * t + 2];
  int st = (int)(((b0 >> lane) & 1u) | (((b1 >> lane) & 1u) << 1) | (((b2 >> lane) & 1u) << 2));
  // this lane's counts (from pass 2), then exclusive prefix inside the warp
  const uint32_t cnt = a.lane_cnt[(size_t)t * 32 + lane];
  const uint32_t nf = cnt & 0xFFu, nl = cnt >> 8;
  uint32_t pf = nf, pl = nl;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t x = __shfl_up_sync(0xffffffffu, pf, o);
    uint32_t y = __shfl_up_sync(0xffffffffu, pl, o);
    if (lane >= o) { pf += x; pl += y; }
  }
  uint32_t g = a.tile_nf_pre[t] + pf - nf;          // global index of this lane's first field end
  uint32_t r = a.tile_nl_pre[t] + pl - nl + d;      // record slot currently open at this lane
  uint32_t doc_begin = has_bom ? 3u : 0u;
  auto on_event = [&](int k, int e, int sb) {
    uint32_t pos = base + k;
    if (e & EV_REC_START) a.rec_start[r] = pos;
    if (e & EV_REC_START_WS) {
      // the blanks that opened this line belong to its first field: walk back over them
      uint32_t p = pos;
      while (p > doc_begin) {
        uint32_t c = doc[p - 1];
        if ((c == ' ' || c == '\t') && c != a.sep) --p; else break;
      }
      a.rec_start[r] = p;
      // pandas finds the start of such a line by walking back to the previous '\n' (tokenizer.c,
      // WHITESPACE_LINE: `while (i + 1 > datapos && !IS_TERMINATOR(*buf))`, and only LF is a
      // terminator there).  After a bare CR, or after the delimiter a skipped blank line swallowed,
      // it runs into the earlier rows and what comes out depends on its read-buffer positions: not
      // reproduced, the document is reported unsupported instead.
      if (p > doc_begin && doc[p - 1] != '\n' &&
          r - (a.tile_nl_pre[a.tile_base[d]] + d) <= a.max_rows[d]) a.doc_unsup[d] = 1u;
    }
    if (e & EV_FIELD_END) a.fend[g++] = pos;
    if (e & EV_LINE_END) a.rec_fend[r++] = g;
  };
  {
    // An ordinary byte right after an ordinary byte changes nothing (the state is already U or Q and
    // neither raises an event on class O): only bytes of another class, and the first ordinary byte
    // after one (or at the start of the lane), are visited.
    unsigned long long t = lb.cls;
    t = (t | (t >> 1) | (t >> 2) | (t >> 3)) & 0x1111111111111111ull;   // bit 4k = nibble k is not CL_O
    t = (t | (t >> 3)) & 0x0303030303030303ull;
    t = (t | (t >> 6)) & 0x000F000F000F000Full;
    t = (t | (t >> 12)) & 0x000000FF000000FFull;
    const uint32_t nz = (uint32_t)(t | (t >> 24)) & 0xFFFFu;
    int s1 = st;
    for (uint32_t m = (nz | (nz << 1) | 1u) & 0xFFFFu; m; m &= m - 1) {
      const int kk = __ffs(m) - 1;
      const uint2 fe = class_fe[(uint32_t)(lb.cls >> (4 * kk)) & 0xFu];
      const int e = (int)((fe.y >> (4 * s1)) & 0xFu);
      if (e) on_event(kk, e, s1);
      s1 = (int)((fe.x >> (4 * s1)) & 0xFu);
    }
  }
  }
}

// ---- record shapes per document ---------------------------------------------------------------
__global__ void __launch_bounds__(256) csv_shape_init_kernel(ScanArgs a, DocShape* shape) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= a.n_docs) return;
  uint32_t t0 = a.tile_base[d], t1 = a.tile_base[d + 1];
  DocShape s;
  s.n_rec = a.tile_nl_pre[t1] - a.tile_nl_pre[t0];
  s.hdr_nf = 0; s.first_nf = 0; s.max_nf = 0;
  s.final_state = a.doc_final[d];
  s.r0 = a.tile_nl_pre[t0] + d;
  s.f0 = a.tile_nf_pre[t0];
  s.pad = a.doc_unsup[d];   // filled by the index pass, which runs first
  shape[d] = s;
}

__global__ void __launch_bounds__(256) csv_shape_kernel(ScanArgs a, const uint32_t* __restrict__ max_rows,
                                                        DocShape* shape, uint32_t n_slots) {
  uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
  if (slot >= n_slots) return;
  // slot -> document: largest d with r0[d] <= slot
  uint32_t lo = 0, hi = a.n_docs;
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (shape[mid].r0 <= slot) lo = mid; else hi = mid;
  }
  uint32_t d = lo;
  uint32_t L = slot - shape[d].r0;
  if (L >= shape[d].n_rec || L > max_rows[d]) return;  // spare slot or beyond nrows
  uint32_t fbeg = L == 0 ? shape[d].f0 : a.rec_fend[slot - 1];
  uint32_t nf = a.rec_fend[slot] - fbeg;
  if (L == 0) shape[d].hdr_nf = nf;
  else if (L == 1) shape[d].first_nf = nf;
  else if (nf > shape[d].max_nf) atomicMax(&shape[d].max_nf, nf);  // read first: rows rarely differ
}

// ---------------------------------------------------------------------------------------------
// resident CTAs of the scan kernels: 8 per SM
static uint32_t scan_grid() {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return (uint32_t)sms * 8u;
}

void csv_scan_count(const ScanArgs& a, cudaStream_t s) {
  if (a.n_tiles == 0) return;
  uint32_t nb = std::min<uint32_t>((a.n_tiles + 7) / 8, scan_grid());
  prof_mark("csv_tilefn_kernel", s);
  RFQ_LAUNCH(csv_tilefn_kernel, nb, 256, 0, s, a);
  prof_mark("scan(tile_fn)", s);
  exclusive_scan<FnCompose>(a.tile_fn, a.tile_fn_pre, a.n_tiles, a.scan_tmp, nullptr, s);
  cudaMemsetAsync(a.tile_nf + a.n_tiles, 0, 4, s);
  cudaMemsetAsync(a.tile_nl + a.n_tiles, 0, 4, s);
  prof_mark("csv_count_kernel", s);
  RFQ_LAUNCH(csv_count_kernel, nb, 256, 0, s, a);
  prof_mark("scan(counts)", s);
  exclusive_scan<SumU32>(a.tile_nf, a.tile_nf_pre, a.n_tiles + 1, a.scan_tmp, nullptr, s);
  exclusive_scan<SumU32>(a.tile_nl, a.tile_nl_pre, a.n_tiles + 1, a.scan_tmp, nullptr, s);
  prof_mark("host+sync", s);
}

void csv_scan_index(const ScanArgs& a, cudaStream_t s) {
  if (a.n_tiles == 0) return;
  prof_mark("csv_index_kernel", s);
  RFQ_LAUNCH(csv_index_kernel, (a.n_tiles + 7) / 8, 256, 0, s, a);   // no tables worth keeping: one round per CTA
}

void csv_doc_shape(const ScanArgs& a, const uint32_t* max_rows32, DocShape* shape, uint32_t n_slots,
                   cudaStream_t s) {
  if (a.n_docs == 0) return;
  prof_mark("csv_shape_kernels", s);
  RFQ_LAUNCH(csv_shape_init_kernel, (a.n_docs + 255) / 256, 256, 0, s, a, shape);
  if (n_slots)
    RFQ_LAUNCH(csv_shape_kernel, (n_slots + 255) / 256, 256, 0, s, a, max_rows32, shape, n_slots);
  prof_mark("host+sync", s);
}

}  // namespace rfq
