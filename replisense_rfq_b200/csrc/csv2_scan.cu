// One-pass row scan (SURVEY.md §7.1b K1 + K2): pandas' tokenizer state machine (csv_dfa.cuh; tokenizer.c
// `tokenize_bytes`, reached from app/file_parser.py:301) run over all documents of an attempt in ONE
// sweep of their bytes.  The only thing it leaves behind is where every record starts — fields are found
// again by the row kernels while they read the bytes anyway (csv2_rows.cuh).
//
// A CTA takes 16 KB chunks in ticket order.  Per chunk: every lane turns its 16 bytes into a
// state -> state transition function (8 states as nibbles, composed with PRMT), functions are
// composed per tile, per warp, per chunk; the chunk's function is published and the state the chunk
// starts in comes from a decoupled look-back over the chunks before it.  Then the lanes replay their
// bytes from the now-known states (only bytes that can change the state are visited), record starts
// and record ends are kept as bit masks in registers, the number of records since the start of the
// document is carried across chunks by a second (segmented) look-back, and every record start is
// written to its slot.  The first look-back is rarely walked: warp 0 also reads the tile BEFORE the chunk, and
// when that tile's function is constant (the automaton forgets its start within a record or two) the
// chunk's start state is known on the spot and published as an inclusive prefix at once.  Documents begin on 16-byte boundaries and tiles never span documents; the end
// of a document is a virtual LF.  The bytes are read from HBM exactly once: no tensor cores, by design.
#include <cuda_runtime.h>
#include <algorithm>
#include <stdlib.h>
#include "csv2.h"
#include "csv_scan_core.cuh"
#include "launch.h"

namespace rfq {

// look-back descriptor: bits 63..62 status, low bits payload
#define RFQ2_AGG 1ull
#define RFQ2_INC 2ull
#define RFQ2_SEG_FLAG (1ull << 40)            // segmented count: a document starts inside
#define RFQ2_SEG_MASK ((1ull << 40) - 1ull)

__device__ __forceinline__ unsigned long long ld_desc(const unsigned long long* p) {
  return *(const volatile unsigned long long*)p;
}
__device__ __forceinline__ void st_desc(unsigned long long* p, unsigned long long v) {
  *(volatile unsigned long long*)p = v;
}
// records since the last document start: a happens before b
__device__ __forceinline__ unsigned long long seg_op(unsigned long long a, unsigned long long b) {
  if (b & RFQ2_SEG_FLAG) return b;
  return (a & RFQ2_SEG_FLAG) | (((a & RFQ2_SEG_MASK) + (b & RFQ2_SEG_MASK)) & RFQ2_SEG_MASK);
}

// exclusive prefix of the transition functions of chunks [0, chunk); publishes this chunk's
// aggregate, then its inclusive prefix.  Called by all 32 lanes of one warp.
__device__ uint32_t lookback_fn(unsigned long long* desc, uint32_t chunk, uint32_t agg, int lane) {
  if (lane == 0) st_desc(desc + chunk, ((chunk == 0 ? RFQ2_INC : RFQ2_AGG) << 62) | agg);
  if (chunk == 0) return RFQ_FN_IDENTITY;
  uint32_t acc = RFQ_FN_IDENTITY;
  int base = (int)chunk - 1;
  for (;;) {
    const int idx = base - lane;
    unsigned long long v = (RFQ2_INC << 62) | RFQ_FN_IDENTITY;   // before chunk 0: nothing happened
    if (idx >= 0) { do { v = ld_desc(desc + idx); } while ((v >> 62) == 0ull); }
    const unsigned inc = __ballot_sync(0xffffffffu, (v >> 62) == RFQ2_INC);
    const int nl = inc ? __ffs(inc) : 32;                        // lanes [0, nl) take part
    uint32_t f = lane < nl ? (uint32_t)v : RFQ_FN_IDENTITY;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {                           // lane l + o is the earlier chunk
      const uint32_t g = __shfl_down_sync(0xffffffffu, f, o);
      if (lane + o < 32) f = fn_compose(g, f);
    }
    f = __shfl_sync(0xffffffffu, f, 0);
    acc = fn_compose(f, acc);
    if (inc) break;
    base -= 32;
  }
  if (lane == 0) st_desc(desc + chunk, (RFQ2_INC << 62) | fn_compose(acc, agg));
  return acc;
}

// the same for the segmented record counts (a chunk in which a document starts ends the walk)
__device__ unsigned long long lookback_seg(unsigned long long* desc, uint32_t chunk, unsigned long long agg, int lane) {
  if (lane == 0) st_desc(desc + chunk, ((chunk == 0 ? RFQ2_INC : RFQ2_AGG) << 62) | agg);
  if (chunk == 0) return 0ull;
  unsigned long long acc = 0ull;
  int base = (int)chunk - 1;
  for (;;) {
    const int idx = base - lane;
    unsigned long long v = (RFQ2_INC << 62);
    if (idx >= 0) { do { v = ld_desc(desc + idx); } while ((v >> 62) == 0ull); }
    const unsigned stop = __ballot_sync(0xffffffffu, (v >> 62) == RFQ2_INC || (v & RFQ2_SEG_FLAG));
    const int nl = stop ? __ffs(stop) : 32;
    unsigned long long f = lane < nl ? (v & (RFQ2_SEG_FLAG | RFQ2_SEG_MASK)) : 0ull;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long g = __shfl_down_sync(0xffffffffu, f, o);
      if (lane + o < 32) f = seg_op(g, f);
    }
    f = __shfl_sync(0xffffffffu, f, 0);
    acc = seg_op(f, acc);
    if (stop) break;
    base -= 32;
  }
  if (lane == 0) st_desc(desc + chunk, (RFQ2_INC << 62) | seg_op(acc, agg));
  return acc;
}

#define RFQ2_SCAN_WARPS (RFQ2_SCAN_TILES / 4u)
__global__ void __launch_bounds__(RFQ2_SCAN_WARPS * 32) csv2_scan_kernel(Csv2Args a) {
  RFQ_INIT_CLASS_FN();
  __shared__ uint32_t pair_fn[128];
  for (uint32_t i = threadIdx.x; i < 128; i += blockDim.x) pair_fn[i] = fn_compose(class_fn[i & 7u], class_fn[(i >> 4) & 7u]);
  __shared__ uint32_t s_chunk[2], s_wfn[RFQ2_SCAN_WARPS], s_start_state, s_behind;
  __shared__ unsigned long long s_wseg[RFQ2_SCAN_WARPS], s_seg_in;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  (void)class_ev;
  // four CTA barriers per chunk: ticket / functions known / records counted / records before the chunk known
  // (a fifth only when the state look-back has to be walked).  The ticket is drawn as late as possible:
  // a chunk whose CTA is still busy with the chunk before it keeps every look-back behind it polling
  // (drawn one phase early the sweep takes 0.36 ms instead of 0.26).
  for (uint32_t it = 0;; ++it) {
    if (threadIdx.x == 0) s_chunk[it & 1u] = atomicAdd(a.ticket, 1u);
    __syncthreads();
    const uint32_t chunk = s_chunk[it & 1u];
    if (chunk >= a.n_chunks) break;
    const uint32_t t0 = chunk * RFQ2_SCAN_TILES + (uint32_t)w * 4u;
    // ---- 1. byte classes, lane / tile / warp transition functions
    unsigned long long cls[4];
    uint32_t efn[4], tfn[4], dq[4];    // per tile: the lane's exclusive function, the tile's function, its document
    uint32_t wf = RFQ_FN_IDENTITY;
    uint32_t d_run = t0 < a.n_tiles ? warp_find_doc(a.tile_base, a.n_docs, t0, lane) : 0u;   // the warp's four tiles follow each other
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t t = t0 + q;
      cls[q] = 0ull; efn[q] = RFQ_FN_IDENTITY; tfn[q] = RFQ_FN_IDENTITY; dq[q] = 0;
      if (t < a.n_tiles) {
        while (d_run + 1 < a.n_docs && __ldg(a.tile_base + d_run + 1) <= t) ++d_run;
        const uint32_t d = d_run;
        const uint32_t tile = t - __ldg(a.tile_base + d);
        const uint32_t len = a.docs[d].len;
        const uint8_t* doc = a.arena + a.docs[d].off;
        const bool bom = tile == 0 && doc_has_bom(doc, len);
        const LaneBytes lb = load_lane_classes(doc, len, tile * RFQ_TILE + lane * 16, a.docs[d].sep, bom);
        uint32_t incl = lane_function(lb.cls, pair_fn);
        cls[q] = lb.cls; dq[q] = d;
        // one inclusive scan: lane 31 ends up with the tile's function, every lane with what happens before it
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t g = __shfl_up_sync(0xffffffffu, incl, o);
          const uint32_t c2 = fn_compose(g, incl);
          if (lane >= o) incl = c2;
        }
        uint32_t ex = __shfl_up_sync(0xffffffffu, incl, 1);
        if (lane == 0) ex = RFQ_FN_IDENTITY;
        efn[q] = ex;
        uint32_t f = __shfl_sync(0xffffffffu, incl, 31);
        if (tile == 0) f = fn_const(fn_apply(f, ST_R));   // a document starts in ST_R whatever came before
        tfn[q] = f;
        wf = fn_compose(wf, f);
      }
    }
    if (lane == 0) s_wfn[w] = wf;
    // pandas' automaton forgets where it started within a record or two (a quote opens a field only at the
    // field's first byte): when the function of the tile BEFORE the chunk is constant, the chunk's start
    // state is that constant and nobody has to be waited for.
    uint32_t behind = 0xFFFFFFFFu;
    if (w == 0) {
      const uint32_t d = dq[0];
      const uint32_t tile = t0 - __ldg(a.tile_base + d);
      if (tile == 0) behind = ST_R;               // a document starts here: what came before does not matter
      else {
        const uint32_t len = a.docs[d].len;
        const uint8_t* doc = a.arena + a.docs[d].off;
        const bool bom = tile == 1 && doc_has_bom(doc, len);
        const LaneBytes lb = load_lane_classes(doc, len, (tile - 1) * RFQ_TILE + lane * 16, a.docs[d].sep, bom);
        uint32_t f = lane_function(lb.cls, pair_fn);
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {          // lane l + o holds later bytes
          const uint32_t g = __shfl_down_sync(0xffffffffu, f, o);
          if (lane + o < 32) f = fn_compose(f, g);
        }
        f = __shfl_sync(0xffffffffu, f, 0);
        if (tile == 1) f = fn_const(fn_apply(f, ST_R));
        if (f == (f & 0xFu) * 0x11111111u) behind = f & 0xFu;
      }
      if (lane == 0) s_behind = behind;
    }
    __syncthreads();
    // ---- 2. the state this chunk starts in
    behind = s_behind;
    if (w == 0) {
      uint32_t cf = RFQ_FN_IDENTITY;
#pragma unroll
      for (int v = 0; v < (int)RFQ2_SCAN_WARPS; ++v) cf = fn_compose(cf, s_wfn[v]);
      if (behind != 0xFFFFFFFFu) {
        // known without a look-back: the chunk's inclusive prefix, as far as anybody can tell from applying it
        // to a document's start state, is the constant function "ends in cf(behind)"
        if (lane == 0) st_desc(a.fn_desc + chunk, (RFQ2_INC << 62) | fn_const(fn_apply(cf, (int)behind)));
      } else {
        const uint32_t pre = lookback_fn(a.fn_desc, chunk, cf, lane);
        if (lane == 0) s_start_state = (uint32_t)fn_apply(pre, ST_R);
      }
    }
    if (behind == 0xFFFFFFFFu) __syncthreads();   // (the same for every thread of the CTA)
    int st = behind != 0xFFFFFFFFu ? (int)behind : (int)s_start_state;
    for (int v = 0; v < w; ++v) st = fn_apply(s_wfn[v], st);
    // ---- 3. replay: record starts / ends of every lane as bit masks, counts
    uint32_t ev[4];      // bits 0-15 record starts, 16-31 starts after leading blanks
    uint32_t le[4];      // bits 0-15 record ends, 16-25 records of the lanes before this one in the tile
    uint32_t tot[4];     // records of the tile
    unsigned long long wseg = 0ull;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t t = t0 + q;
      ev[q] = 0; le[q] = 0; tot[q] = 0;
      if (t < a.n_tiles) {
        const uint32_t d = dq[q];
        const uint32_t tile = t - __ldg(a.tile_base + d);
        const uint32_t len = a.docs[d].len;
        const uint32_t base = tile * RFQ_TILE + lane * 16;
        int s1 = fn_apply(efn[q], tile == 0 ? ST_R : st);
        const uint32_t eof_k = (len >= base && len < base + 16) ? len - base : 0xFFu;
        const uint32_t vm = visit_mask16(cls[q]);
        uint32_t sm = 0, wm = 0, lm = 0;
        for (uint32_t m = vm; m; m &= m - 1) {
          const int kk = __ffs(m) - 1;
          const uint2 fe = class_fe[(uint32_t)(cls[q] >> (4 * kk)) & 0xFu];
          const uint32_t e = (fe.y >> (4 * s1)) & 0xFu;
          if ((uint32_t)kk == eof_k) a.run[d].final_state = (uint32_t)s1;   // state before the virtual LF
          if (e) {
            if (e & EV_REC_START) sm |= 1u << kk;
            if (e & EV_REC_START_WS) wm |= 1u << kk;
            if (e & EV_LINE_END) lm |= 1u << kk;
          }
          s1 = (int)((fe.x >> (4 * s1)) & 0xFu);
        }
        const uint32_t nl = (uint32_t)__popc(lm);
        uint32_t pl = nl;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t x = __shfl_up_sync(0xffffffffu, pl, o);
          if (lane >= o) pl += x;
        }
        ev[q] = sm | (wm << 16);
        le[q] = lm | ((pl - nl) << 16);
        tot[q] = __shfl_sync(0xffffffffu, pl, 31);
        wseg = seg_op(wseg, (unsigned long long)tot[q] | (tile == 0 ? RFQ2_SEG_FLAG : 0ull));
        st = fn_apply(tfn[q], st);
      }
    }
    if (lane == 0) s_wseg[w] = wseg;
    __syncthreads();
    // ---- 4. records since the start of the document, before this chunk
    if (w == 0) {
      unsigned long long cs = 0ull;
#pragma unroll
      for (int v = 0; v < (int)RFQ2_SCAN_WARPS; ++v) cs = seg_op(cs, s_wseg[v]);
      const unsigned long long pre = lookback_seg(a.cnt_desc, chunk, cs, lane);
      if (lane == 0) s_seg_in = pre;
    }
    __syncthreads();
    unsigned long long seg = s_seg_in;
    for (int v = 0; v < w; ++v) seg = seg_op(seg, s_wseg[v]);
    // ---- 5. write the record starts
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t t = t0 + q;
      if (t < a.n_tiles) {
        const uint32_t d = dq[q];
        const uint32_t tile = t - __ldg(a.tile_base + d);
        const Doc2& D = a.docs[d];
        const uint32_t len = D.len;
        const uint32_t base = tile * RFQ_TILE + lane * 16;
        const uint32_t rec0 = (tile == 0 ? 0u : (uint32_t)(seg & RFQ2_SEG_MASK)) + (le[q] >> 16);
        const uint32_t lm = le[q] & 0xFFFFu;
        uint32_t* out = a.rec_start + D.slot_base;
        for (uint32_t m = (ev[q] | (ev[q] >> 16)) & 0xFFFFu; m; m &= m - 1) {
          const uint32_t kk = (uint32_t)__ffs(m) - 1u;
          const uint32_t idx = rec0 + (uint32_t)__popc(lm & ((1u << kk) - 1u));
          if (idx > D.max_rows + 1u) continue;      // records beyond 1 + nrows are never read
          uint32_t pos = base + kk;
          if ((ev[q] >> 16) & (1u << kk)) {
            // the blanks that opened this line belong to its first field: walk back over them
            const uint8_t* doc = a.arena + D.off;
            const uint32_t doc_begin = doc_has_bom(doc, len) ? 3u : 0u;
            while (pos > doc_begin) {
              const uint32_t c = doc[pos - 1];
              if ((c == ' ' || c == '\t') && c != D.sep) --pos; else break;
            }
            // pandas finds the start of such a line by walking back to the previous '\n' (tokenizer.c,
            // WHITESPACE_LINE: only LF stops it).  After a bare CR, or after the delimiter a skipped
            // blank line swallowed, it runs into earlier rows and the outcome depends on its read-buffer
            // positions: not reproduced, the document is declined.
            if (pos > doc_begin && doc[pos - 1] != '\n') atomicOr(&a.run[d].flags, (uint32_t)RUN_UNSUP);
          }
          out[idx] = pos;
        }
        if (len >= base && len < base + 16) a.run[d].n_rec = rec0 + (uint32_t)__popc(lm);
        seg = seg_op(seg, (unsigned long long)tot[q] | (tile == 0 ? RFQ2_SEG_FLAG : 0ull));
      }
    }
  }
}

// ---- the same scan in three launches, without the count look-back -----------------------------
// csv2_scan_marks_kernel is the kernel above up to the replay: it leaves every lane's marks (8 bytes per 16
// input bytes), every tile's record count and every chunk's segmented count, and waits for nobody (the
// state look-back is only walked when the look-behind tile does not settle the state).  A one-CTA
// segmented scan turns chunk counts into "records since the document start, before the chunk";
// csv2_scan_write_kernel reads the marks back and writes rec_start.  ~90 MB of extra traffic for C2
// (30 us at HBM speed) against the barrier stalls of a look-back that is a true prefix sum.
__global__ void __launch_bounds__(RFQ2_SCAN_WARPS * 32) csv2_scan_marks_kernel(Csv2Args a) {
  RFQ_INIT_CLASS_FN();
  __shared__ uint32_t pair_fn[128];
  for (uint32_t i = threadIdx.x; i < 128; i += blockDim.x) pair_fn[i] = fn_compose(class_fn[i & 7u], class_fn[(i >> 4) & 7u]);
  __shared__ uint32_t s_chunk[2], s_wfn[RFQ2_SCAN_WARPS], s_start_state, s_behind;
  __shared__ unsigned long long s_wseg[RFQ2_SCAN_WARPS], s_seg_in;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  (void)class_ev;
  // four CTA barriers per chunk: ticket / functions known / records counted / records before the chunk known
  // (a fifth only when the state look-back has to be walked).  The ticket is drawn as late as possible:
  // a chunk whose CTA is still busy with the chunk before it keeps every look-back behind it polling
  // (drawn one phase early the sweep takes 0.36 ms instead of 0.26).
  for (uint32_t it = 0;; ++it) {
    if (threadIdx.x == 0) s_chunk[it & 1u] = atomicAdd(a.ticket, 1u);
    __syncthreads();
    const uint32_t chunk = s_chunk[it & 1u];
    if (chunk >= a.n_chunks) break;
    const uint32_t t0 = chunk * RFQ2_SCAN_TILES + (uint32_t)w * 4u;
    // ---- 1. byte classes, lane / tile / warp transition functions
    unsigned long long cls[4];
    uint32_t efn[4], tfn[4], dq[4];    // per tile: the lane's exclusive function, the tile's function, its document
    uint32_t wf = RFQ_FN_IDENTITY;
    uint32_t d_run = t0 < a.n_tiles ? warp_find_doc(a.tile_base, a.n_docs, t0, lane) : 0u;   // the warp's four tiles follow each other
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t t = t0 + q;
      cls[q] = 0ull; efn[q] = RFQ_FN_IDENTITY; tfn[q] = RFQ_FN_IDENTITY; dq[q] = 0;
      if (t < a.n_tiles) {
        while (d_run + 1 < a.n_docs && __ldg(a.tile_base + d_run + 1) <= t) ++d_run;
        const uint32_t d = d_run;
        const uint32_t tile = t - __ldg(a.tile_base + d);
        const uint32_t len = a.docs[d].len;
        const uint8_t* doc = a.arena + a.docs[d].off;
        const bool bom = tile == 0 && doc_has_bom(doc, len);
        const LaneBytes lb = load_lane_classes(doc, len, tile * RFQ_TILE + lane * 16, a.docs[d].sep, bom);
        uint32_t incl = lane_function(lb.cls, pair_fn);
        cls[q] = lb.cls; dq[q] = d;
        // one inclusive scan: lane 31 ends up with the tile's function, every lane with what happens before it
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t g = __shfl_up_sync(0xffffffffu, incl, o);
          const uint32_t c2 = fn_compose(g, incl);
          if (lane >= o) incl = c2;
        }
        uint32_t ex = __shfl_up_sync(0xffffffffu, incl, 1);
        if (lane == 0) ex = RFQ_FN_IDENTITY;
        efn[q] = ex;
        uint32_t f = __shfl_sync(0xffffffffu, incl, 31);
        if (tile == 0) f = fn_const(fn_apply(f, ST_R));   // a document starts in ST_R whatever came before
        tfn[q] = f;
        wf = fn_compose(wf, f);
      }
    }
    if (lane == 0) s_wfn[w] = wf;
    // pandas' automaton forgets where it started within a record or two (a quote opens a field only at the
    // field's first byte): when the function of the tile BEFORE the chunk is constant, the chunk's start
    // state is that constant and nobody has to be waited for.
    uint32_t behind = 0xFFFFFFFFu;
    if (w == 0) {
      const uint32_t d = dq[0];
      const uint32_t tile = t0 - __ldg(a.tile_base + d);
      if (tile == 0) behind = ST_R;               // a document starts here: what came before does not matter
      else {
        const uint32_t len = a.docs[d].len;
        const uint8_t* doc = a.arena + a.docs[d].off;
        const bool bom = tile == 1 && doc_has_bom(doc, len);
        const LaneBytes lb = load_lane_classes(doc, len, (tile - 1) * RFQ_TILE + lane * 16, a.docs[d].sep, bom);
        uint32_t f = lane_function(lb.cls, pair_fn);
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {          // lane l + o holds later bytes
          const uint32_t g = __shfl_down_sync(0xffffffffu, f, o);
          if (lane + o < 32) f = fn_compose(f, g);
        }
        f = __shfl_sync(0xffffffffu, f, 0);
        if (tile == 1) f = fn_const(fn_apply(f, ST_R));
        if (f == (f & 0xFu) * 0x11111111u) behind = f & 0xFu;
      }
      if (lane == 0) s_behind = behind;
    }
    __syncthreads();
    // ---- 2. the state this chunk starts in
    behind = s_behind;
    if (w == 0) {
      uint32_t cf = RFQ_FN_IDENTITY;
#pragma unroll
      for (int v = 0; v < (int)RFQ2_SCAN_WARPS; ++v) cf = fn_compose(cf, s_wfn[v]);
      if (behind != 0xFFFFFFFFu) {
        // known without a look-back: the chunk's inclusive prefix, as far as anybody can tell from applying it
        // to a document's start state, is the constant function "ends in cf(behind)"
        if (lane == 0) st_desc(a.fn_desc + chunk, (RFQ2_INC << 62) | fn_const(fn_apply(cf, (int)behind)));
      } else {
        const uint32_t pre = lookback_fn(a.fn_desc, chunk, cf, lane);
        if (lane == 0) s_start_state = (uint32_t)fn_apply(pre, ST_R);
      }
    }
    if (behind == 0xFFFFFFFFu) __syncthreads();   // (the same for every thread of the CTA)
    int st = behind != 0xFFFFFFFFu ? (int)behind : (int)s_start_state;
    for (int v = 0; v < w; ++v) st = fn_apply(s_wfn[v], st);
    // ---- 3. replay: record starts / ends of every lane as bit masks, counts
    unsigned long long wseg = 0ull;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t t = t0 + q;
      if (t < a.n_tiles) {
        const uint32_t d = dq[q];
        const uint32_t tile = t - __ldg(a.tile_base + d);
        const uint32_t len = a.docs[d].len;
        const uint32_t base = tile * RFQ_TILE + lane * 16;
        int s1 = fn_apply(efn[q], tile == 0 ? ST_R : st);
        const uint32_t eof_k = (len >= base && len < base + 16) ? len - base : 0xFFu;
        const uint32_t vm = visit_mask16(cls[q]);
        uint32_t sm = 0, wm = 0, lm = 0;
        for (uint32_t m = vm; m; m &= m - 1) {
          const int kk = __ffs(m) - 1;
          const uint2 fe = class_fe[(uint32_t)(cls[q] >> (4 * kk)) & 0xFu];
          const uint32_t e = (fe.y >> (4 * s1)) & 0xFu;
          if ((uint32_t)kk == eof_k) a.run[d].final_state = (uint32_t)s1;   // state before the virtual LF
          if (e) {
            if (e & EV_REC_START) sm |= 1u << kk;
            if (e & EV_REC_START_WS) wm |= 1u << kk;
            if (e & EV_LINE_END) lm |= 1u << kk;
          }
          s1 = (int)((fe.x >> (4 * s1)) & 0xFu);
        }
        // what the write sweep needs: the lane's marks, the tile's record count (bit 31: a document starts here), its document
        const uint32_t tot = __reduce_add_sync(0xffffffffu, (uint32_t)__popc(lm));
        a.lane_marks[(size_t)t * 32u + lane] = make_uint2(sm | (wm << 16), lm);
        if (lane == 0) { a.tile_tot[t] = tot | (tile == 0 ? 0x80000000u : 0u); a.tile_doc[t] = d; }
        wseg = seg_op(wseg, (unsigned long long)tot | (tile == 0 ? RFQ2_SEG_FLAG : 0ull));
        st = fn_apply(tfn[q], st);
      }
    }
    if (lane == 0) s_wseg[w] = wseg;
    __syncthreads();
    // ---- 4. the chunk's (segmented) record count, for the prefix kernel
    if (threadIdx.x == 0) {
      unsigned long long cs = 0ull;
#pragma unroll
      for (int v = 0; v < (int)RFQ2_SCAN_WARPS; ++v) cs = seg_op(cs, s_wseg[v]);
      a.chunk_seg[chunk] = cs;
    }
  }
}


__global__ void __launch_bounds__(1024) csv2_scan_prefix_kernel(Csv2Args a) {
  __shared__ unsigned long long s_warp[32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t per = (a.n_chunks + 1023u) / 1024u;
  const uint32_t lo = min(threadIdx.x * per, a.n_chunks), hi = min(lo + per, a.n_chunks);
  // up to eight items per thread (8 192 chunks = 128 MB of input) sit in registers: one round of independent loads
  unsigned long long own = 0ull, v[8];
  const bool in_regs = per <= 8u;
  if (in_regs) {
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = lo + k < hi ? a.chunk_seg[lo + k] : 0ull;
#pragma unroll
    for (int k = 0; k < 8; ++k) own = seg_op(own, v[k]);
  } else {
    for (uint32_t i = lo; i < hi; ++i) own = seg_op(own, a.chunk_seg[i]);
  }
  unsigned long long incl = own;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long g = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl = seg_op(g, incl);
  }
  if (lane == 31) s_warp[w] = incl;
  __syncthreads();
  unsigned long long pre = 0ull;
  for (int v = 0; v < w; ++v) pre = seg_op(pre, s_warp[v]);
  unsigned long long ex = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) ex = 0ull;
  unsigned long long run = seg_op(pre, ex);
  if (in_regs) {
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (lo + k < hi) { a.chunk_pre[lo + k] = run; run = seg_op(run, v[k]); }
  } else {
    for (uint32_t i = lo; i < hi; ++i) {
      a.chunk_pre[i] = run;
      run = seg_op(run, a.chunk_seg[i]);
    }
  }
}

__global__ void __launch_bounds__(RFQ2_SCAN_WARPS * 32) csv2_scan_write_kernel(Csv2Args a) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (uint32_t chunk = blockIdx.x; chunk < a.n_chunks; chunk += gridDim.x) {
    // records before every tile of the chunk: the chunk's prefix, then a warp-wide segmented scan of its 32 tile counts
    const uint32_t tl = chunk * RFQ2_SCAN_TILES + (uint32_t)lane;
    const uint32_t tv = tl < a.n_tiles ? __ldg(a.tile_tot + tl) : 0u;
    unsigned long long incl = (unsigned long long)(tv & 0x7FFFFFFFu) | ((tv >> 31) ? RFQ2_SEG_FLAG : 0ull);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long g = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl = seg_op(g, incl);
    }
    unsigned long long ex = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) ex = 0ull;
    ex = seg_op(__ldg(a.chunk_pre + chunk), ex);
    const uint32_t t0 = chunk * RFQ2_SCAN_TILES + (uint32_t)w * 4u;
    uint32_t dd[4];
    uint2 mks[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {       // all loads of the warp's four tiles first
      const uint32_t t = t0 + q;
      dd[q] = t < a.n_tiles ? __ldg(a.tile_doc + t) : 0u;
      mks[q] = t < a.n_tiles ? __ldg(reinterpret_cast<const uint2*>(a.lane_marks) + (size_t)t * 32u + lane) : make_uint2(0u, 0u);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t t = t0 + q;
      const unsigned long long seg = __shfl_sync(0xffffffffu, ex, w * 4 + q);
      const uint32_t tflag = __shfl_sync(0xffffffffu, tv, w * 4 + q) >> 31;
      if (t < a.n_tiles) {
        const uint32_t d = dd[q];
        const uint2 mk = mks[q];
        const uint32_t lm = mk.y & 0xFFFFu;
        const uint32_t nl = (uint32_t)__popc(lm);
        uint32_t pl = nl;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const uint32_t x = __shfl_up_sync(0xffffffffu, pl, o);
          if (lane >= o) pl += x;
        }
        const uint32_t tile = t - __ldg(a.tile_base + d);
        const Doc2& D = a.docs[d];
        const uint32_t len = D.len;
        const uint32_t base = tile * RFQ_TILE + lane * 16;
        const uint32_t rec0 = (tflag ? 0u : (uint32_t)(seg & RFQ2_SEG_MASK)) + (pl - nl);
        uint32_t* out = a.rec_start + D.slot_base;
        for (uint32_t m = (mk.x | (mk.x >> 16)) & 0xFFFFu; m; m &= m - 1) {
          const uint32_t kk = (uint32_t)__ffs(m) - 1u;
          const uint32_t idx = rec0 + (uint32_t)__popc(lm & ((1u << kk) - 1u));
          if (idx > D.max_rows + 1u) continue;      // records beyond 1 + nrows are never read
          uint32_t pos = base + kk;
          if ((mk.x >> 16) & (1u << kk)) {
            // the blanks that opened this line belong to its first field (see csv2_scan_kernel)
            const uint8_t* doc = a.arena + D.off;
            const uint32_t doc_begin = doc_has_bom(doc, len) ? 3u : 0u;
            while (pos > doc_begin) {
              const uint32_t c = doc[pos - 1];
              if ((c == ' ' || c == '\t') && c != D.sep) --pos; else break;
            }
            if (pos > doc_begin && doc[pos - 1] != '\n') atomicOr(&a.run[d].flags, (uint32_t)RUN_UNSUP);
          }
          out[idx] = pos;
        }
        if (len >= base && len < base + 16) a.run[d].n_rec = rec0 + (uint32_t)__popc(lm);
      }
    }
  }
}

void csv2_scan(const Csv2Args& a, cudaStream_t s) {
  if (a.n_chunks == 0) return;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  static const bool split_env = getenv("RFQ2_SCAN_SPLIT") ? atoi(getenv("RFQ2_SCAN_SPLIT")) != 0 : true;
  const bool split = a.lane_marks && split_env;
  if (split) {
    static const int per_sm_a = getenv("RFQ2_SCAN_CTAS") ? atoi(getenv("RFQ2_SCAN_CTAS")) : 6;   // 40 registers: 48 of 64 warps
    prof_mark("csv2_scan_kernel", s);
    RFQ_LAUNCH(csv2_scan_marks_kernel, std::min<uint32_t>(a.n_chunks, (uint32_t)sms * (uint32_t)per_sm_a), RFQ2_SCAN_WARPS * 32, 0, s, a);
    prof_mark("csv2_scan_prefix_kernel", s);
    RFQ_LAUNCH(csv2_scan_prefix_kernel, 1, 1024, 0, s, a);
    prof_mark("csv2_scan_write_kernel", s);
    RFQ_LAUNCH(csv2_scan_write_kernel, std::min<uint32_t>(a.n_chunks, (uint32_t)sms * 8u), RFQ2_SCAN_WARPS * 32, 0, s, a);
    return;
  }
  static const int per_sm = getenv("RFQ2_SCAN_CTAS") ? atoi(getenv("RFQ2_SCAN_CTAS")) : 5;   // 40 of 64 warps (48 registers): the look-back waits of one CTA are another's issue slots
  const uint32_t grid = std::min<uint32_t>(a.n_chunks, (uint32_t)sms * (uint32_t)per_sm);
  prof_mark("csv2_scan_kernel", s);
  RFQ_LAUNCH(csv2_scan_kernel, grid, RFQ2_SCAN_WARPS * 32, 0, s, a);
}

}  // namespace rfq
