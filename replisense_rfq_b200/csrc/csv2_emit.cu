// Emit sweep of the row pipeline: `df.to_string(index=False)` of
// app/file_parser.py:305 — every cell right-justified in its column, columns joined by one blank, rows
// by '\n'.
//
// Same work items and the same row walk as the typing sweep.  A work item's rows are one contiguous
// byte range of the output, so the CTA (1) brings the item's input span into shared memory with a bulk
// asynchronous copy, (2) builds the item's output range in a second shared-memory stage that was
// pre-filled with blanks (a cell only writes its own characters), (3) hands the range to the copy
// engine with one bulk store (cp.async.bulk shared -> global) and goes on with its next item while the
// store drains.  Cells are printed from their own bytes wherever the printed form is a copy of the
// token (text, canonical integers, decimal literals with the column's number of decimals); the few
// that need real formatting stay hot too when it is integer-only work.  Rows with anything else
// (exact "%.6f" of a parsed double, "%.6e", Python ints, text that needs the quote state machine) are
// appended to a list and written by csv2_emit_cold_kernel with the general formatters.
// HBM-bound byte work: no tensor cores, by design.
#include <algorithm>
#include <atomic>
#include "csv2.h"
#include "csv2_rows.cuh"
#include "csv2_stage.cuh"
#include "launch.h"
#include "scan.cuh"
#include "table_common.cuh"

namespace rfq {

// ---- any row, any cell: the general formatters, straight to global memory ----------------------
__device__ __noinline__ uint32_t emit_cell_general(CellSrc src, int ct, int kind, int dec, uint32_t width, uint8_t* field) {
  TextInfo t = cell_text(src);
  if (cell_is_nan(src, t, ct)) {
    for (uint32_t q = 0; q < width; ++q) field[q] = ' ';
    return 0;
  }
  uint32_t ncp, nb;
  if (kind == PR_STR) { ncp = t.ncp + t.nesc; nb = t.nbytes + t.nesc; }
  else printed_size(src, ct, kind, dec, &ncp, &nb);
  const uint32_t pad = width - ncp;
  for (uint32_t q = 0; q < pad; ++q) field[q] = ' ';
  print_cell(src, ct, kind, dec, field + pad);
  return nb - ncp;
}

__global__ void __launch_bounds__(128) csv2_emit_cold_kernel(Csv2Args a) {
  if (blockIdx.x == 0)   // everything in the block was final before this kernel started (it only reads it)
    for (uint32_t i = threadIdx.x; i < a.export_words; i += blockDim.x) a.export_dst[i] = a.export_src[i];
  const uint32_t n = min(a.counters[1], a.cold_rows_cap);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const ColdRow cr = a.cold_rows[i];
    const Doc2& D = a.docs[cr.doc];
    const DocRun& R = a.run[cr.doc];
    const uint32_t k_idx = cr.slot - D.slot_base;
    const uint32_t row = k_idx - 1u;
    const uint32_t rs = a.rec_start[cr.slot];
    const uint32_t limit = (k_idx + 1u < R.n_rec) ? a.rec_start[cr.slot + 1] : D.len;
    const uint8_t* ctrow = a.ctype + D.stat_base + (row / D.bl) * D.W;
    RowWalk w(a.arena + D.off, rs, limit, D.sep);
    for (uint32_t c = 0; c < D.lead; ++c) w.next();
    uint8_t* rowp = a.out + cr.out;
    rowp[0] = '\n';
    uint32_t xb = 0;
    for (uint32_t j = 0; j < D.h; ++j) {
      const ColOut2 co = a.col[D.col_base + j];
      const int ct = ctrow[D.lead + j];
      const CellSrc src = w.next();
      const int kind = print_kind(co.cls, ct, co.sci != 0);
      uint8_t* q = rowp + 1 + co.cp_off + xb;
      if (j) q[-1] = ' ';
      xb += emit_cell_general(src, ct, kind, co.dec, co.width, q);
    }
  }
}

// ---- the hot sweep ----------------------------------------------------------------------------------
__global__ void __launch_bounds__(RFQ2_ROWS) csv2_emit_kernel(Csv2Args a, uint32_t in_stage, uint32_t out_stage) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* s_in = smem;
  uint8_t* s_out = smem + in_stage + 32;
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ uint32_t s_doc, s_wtot[RFQ2_ROWS / 32];
  __shared__ ColOut2 s_col[32];
  __shared__ uint8_t s_ct[32];
  if (threadIdx.x == 0) mbar_init(&s_bar, 1);
  mbar_init_fence();
  __syncthreads();
  if (a.item_off[a.n_items] > a.out_cap) return;       // the host grows the buffer and runs the sweep again
  uint32_t phase = 0;
  for (uint32_t item = blockIdx.x; item < a.n_items; item += gridDim.x) {
    // the bulk store of the previous item (committed by thread 0) has read the output stage
    if (threadIdx.x == 0) bulk_store_wait_read();
    __syncthreads();
    if (threadIdx.x == 0) s_doc = ub_search(a.item_base, a.n_docs, item);
    __syncthreads();
    const uint32_t d = s_doc;
    const Doc2 D = a.docs[d];
    const DocRun R = a.run[d];
    if (R.flags & (RUN_FAIL | RUN_UNSUP | RUN_GENERAL)) continue;
    const uint32_t row0 = (item - D.item_base) * RFQ2_ROWS;
    const unsigned long long g0 = a.item_off[item], g1 = a.item_off[item + 1];
    unsigned long long gr = g0;                         // where this item's rows start
    if (item == D.item_base) {
      // banner + header line (or the "Empty DataFrame" text): warp 0, straight to global memory
      gr += R.hdr_bytes;
      if (threadIdx.x < 32) {
        const uint32_t lane = threadIdx.x;
        uint8_t* p = a.out + g0;
        for (uint32_t i = lane; i < D.banner_len; i += 32) p[i] = a.names[D.banner_off + i];
        p += D.banner_len;
        if (R.n_keep == 0) {
          for (uint32_t i = lane; i < D.empty_len; i += 32) p[i] = a.names[D.empty_off + i];
        } else {
          for (uint32_t j = lane; j < D.h; j += 32) {
            const ColOut2 co = a.col[D.col_base + j];
            uint32_t xb = 0;
            for (uint32_t j2 = 0; j2 < j; ++j2) xb += a.name_len[D.name_base + j2] - a.name_cp[D.name_base + j2];
            const bool numeric = co.cls == CC_INT || co.cls == CC_UINT || co.cls == CC_FLOAT || co.cls == CC_BOOL;
            const uint32_t cp = a.name_cp[D.name_base + j] + (numeric ? 1u : 0u);
            uint8_t* q = p + co.cp_off + xb;
            if (j) q[-1] = ' ';
            const uint32_t pad = co.width - cp;
            for (uint32_t i = 0; i < pad; ++i) q[i] = ' ';
            q += pad;
            if (numeric) *q++ = ' ';
            const uint8_t* nm = a.names + a.name_off[D.name_base + j];
            const uint32_t nl = a.name_len[D.name_base + j];
            for (uint32_t i = 0; i < nl; ++i) q[i] = nm[i];
          }
        }
      }
    }
    if (row0 >= R.n_data || R.n_keep == 0 || g1 == gr) continue;
    const uint32_t nrows = min(RFQ2_ROWS, R.n_data - row0);
    const uint32_t W = D.W;
    const uint32_t slot0 = D.slot_base + 1u + row0;
    const uint32_t e0 = D.stat_base + (row0 / D.bl) * W;
    const bool nf = a.chunk_nf[e0] != 0;
    // ---- input stage
    const uint32_t start = __ldg(a.rec_start + slot0);
    const uint32_t e_idx = 1u + row0 + nrows;
    const uint32_t end = e_idx < R.n_rec ? __ldg(a.rec_start + D.slot_base + e_idx) : D.len;
    const uint32_t g_lo = (D.off + start) & ~15u;
    const uint32_t g_hi = (D.off + end + 31u) & ~15u;
    const bool staged_in = g_hi - g_lo <= in_stage;
    if (staged_in && threadIdx.x == 0) bulk_load(s_in, a.arena + g_lo, g_hi - g_lo, &s_bar);
    const uint8_t* base = staged_in ? s_in - (size_t)(g_lo - D.off) : a.arena + D.off;
    // ---- output stage: rows kept, their offsets
    const unsigned long long nbytes = g1 - gr;
    const uint32_t mis = (uint32_t)((uintptr_t)(a.out + gr) & 15u);   // keep shared and global memory co-aligned
    const bool staged_out = nbytes + mis + 16 <= out_stage;
    const uint32_t slot = slot0 + threadIdx.x;
    bool kept = false;
    uint32_t rbytes = 0;
    if (threadIdx.x < nrows) {
      kept = !(__ldg(a.row_na + slot) == D.h && !nf);
      if (kept) rbytes = 1u + R.linew + __ldg(a.row_extra + slot);
    }
    // offsets of the rows inside the item: warp scans now, warp totals after the barrier below
    const uint32_t lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
    const uint32_t rincl = warp_incl_scan<SumU32>(rbytes, (int)lane);
    if (lane == 31) s_wtot[wid] = rincl;
    const bool small = D.h <= 32;
    if (small && threadIdx.x < D.h) {
      s_col[threadIdx.x] = a.col[D.col_base + threadIdx.x];
      s_ct[threadIdx.x] = a.ctype[e0 + D.lead + threadIdx.x];
    }
    if (staged_out) {
      const uint32_t nvec = (uint32_t)((mis + nbytes + 15) >> 4);
      const uint4 sp = make_uint4(0x20202020u, 0x20202020u, 0x20202020u, 0x20202020u);
      for (uint32_t i = threadIdx.x; i < nvec; i += blockDim.x) reinterpret_cast<uint4*>(s_out)[i] = sp;
    }
    if (staged_in) mbar_wait(&s_bar, phase), phase ^= 1u;
    __syncthreads();
    uint32_t roff = rincl - rbytes;
    for (uint32_t v = 0; v < wid; ++v) roff += s_wtot[v];
    if (kept) {
      bool cold = !staged_out;
      if (!cold) {
        uint8_t* rowp = s_out + mis + roff;
        const uint32_t rs = threadIdx.x == 0 ? start : __ldg(a.rec_start + slot);
        const uint32_t limit = (2u + row0 + threadIdx.x < R.n_rec) ? __ldg(a.rec_start + slot + 1) : D.len;
        const uint8_t* ctrow = a.ctype + e0 + D.lead;
        const uint4* mrow = reinterpret_cast<const uint4*>(a.cell_meta) + D.meta_base +
                            (size_t)(row0 + threadIdx.x) * RFQ2_META_VECS(W);
        rowp[0] = '\n';
        uint32_t xb = 0, pos = rs;
        uint4 mv = make_uint4(0u, 0u, 0u, 0u);
        for (uint32_t c = 0; c < W; ++c) {
          if ((c & 7u) == 0) mv = __ldg(mrow + (c >> 3));
          const uint32_t k = c & 7u;
          const uint32_t word = k < 2 ? mv.x : k < 4 ? mv.y : k < 6 ? mv.z : mv.w;
          const uint32_t me = (word >> (16u * (k & 1u))) & 0xFFFFu;
          const uint32_t mode = (me >> CMETA_MODE_SHIFT) & 3u;
          CellSrc src;
          src.mode = CM_RAW; src.missing = false; src.clean = mode == 0; src.bits = 0; src.p = base; src.n = 0;
          if (me & CMETA_MISSING) src.missing = true;
          else if (mode == 3u) { cold = true; break; }
          else if (mode == 2u) {                       // ""-pair text: its two lengths come from a fresh scan
            const FieldScan f = scan_field(base, pos, limit, D.sep);
            src = f.src; pos = f.next;
          } else {
            const uint32_t q = (me >> 11) & 1u, len = me & 0x7FFu;
            src.p = base + pos + q; src.n = len;
            pos += len + 1u + 2u * q;
          }
          if (c < D.lead) continue;
          const uint32_t j = c - D.lead;
          const ColOut2 co = small ? s_col[j] : a.col[D.col_base + j];
          const int ct = small ? s_ct[j] : ctrow[j];
          const int kind = print_kind(co.cls, ct, co.sci != 0);
          uint32_t ex;
          if (emit_hot(src, ct, kind, co.dec, co.width, rowp + 1 + co.cp_off + xb, &ex, (int)(me >> 15))) { cold = true; break; }
          xb += ex;
        }
      }
      if (cold) {
        const uint32_t k = atomicAdd(a.counters + 1, 1u);
        if (k < a.cold_rows_cap) a.cold_rows[k] = ColdRow{d, slot, gr + roff};
        else atomicOr(a.counters + 2, 1u);
      }
    }
    if (staged_out) {
      bulk_store_fence();                                 // generic-proxy writes -> visible to the bulk store
      __syncthreads();
      uint8_t* g = a.out + gr;
      const uint32_t head = mis ? (uint32_t)min((unsigned long long)(16 - mis), nbytes) : 0u;
      const unsigned long long body = (nbytes - head) >> 4;
      const uint32_t tail = (uint32_t)((nbytes - head) & 15u);
      if (threadIdx.x == 0 && body) bulk_store(g + head, s_out + mis + head, (uint32_t)(body << 4));
      if (threadIdx.x >= 32 && threadIdx.x < 32 + head) g[threadIdx.x - 32] = s_out[mis + threadIdx.x - 32];
      if (threadIdx.x >= 64 && threadIdx.x < 64 + tail) g[head + (body << 4) + threadIdx.x - 64] = s_out[mis + head + (body << 4) + threadIdx.x - 64];
    }
  }
  if (threadIdx.x == 0) bulk_store_wait_read();
}

static int sm_count2() {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms;
}

void csv2_emit(const Csv2Args& a, cudaStream_t s) {
  if (a.n_items == 0) return;
  const uint32_t in_stage = a.in_stage, out_stage = a.out_stage;
  const uint32_t smem = in_stage + 32u + out_stage + 32u;
  static std::atomic<unsigned long long> attr_done{0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (!(attr_done.load() >> (dev & 63) & 1ull)) {
    cudaFuncSetAttribute(csv2_emit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 << 10);
    attr_done.fetch_or(1ull << (dev & 63));
  }
  int per_sm = 1;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, csv2_emit_kernel, RFQ2_ROWS, smem);
  const uint32_t grid = std::min<uint32_t>(a.n_items, (uint32_t)(sm_count2() * std::max(per_sm, 1)));
  prof_mark("csv2_emit_kernel", s);
  RFQ_LAUNCH(csv2_emit_kernel, grid, RFQ2_ROWS, smem, s, a, in_stage, out_stage);
  prof_mark("csv2_emit_cold_kernel", s);
  RFQ_LAUNCH(csv2_emit_cold_kernel, sm_count2() * 4, 128, 0, s, a);
}

}  // namespace rfq
