// rfq_csv_run: host orchestration of the reference's (encoding x separator) trial loop
// (app/file_parser.py:298-308) over a staged batch.  Per attempt the host only does
// O(documents + columns) bookkeeping (header names, table shapes, buffer carving); every pass
// over document bytes or cells is a kernel.
#include <cuda_runtime.h>
#include <stdio.h>
#include <chrono>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/rfq_parse.h"
#include "col_decide.cuh"
#include "csv_dfa.cuh"
#include "csv2.h"
#include "ctx.h"
#include "header_names.h"

using namespace rfq;

namespace {

const uint32_t kSeps[3] = {',', ';', '\t'};
const char* kEncNames[3] = {"utf-8", "latin-1", "cp1252"};

// bump writer into a pinned upload buffer; returns offsets so device pointers can be derived
struct Uploader {
  uint8_t* h; size_t used = 0;
  explicit Uploader(uint8_t* p) : h(p) {}
  template <class T> size_t put(const T* src, size_t n) {
    used = (used + 255) & ~(size_t)255;
    size_t o = used;
    if (n) memcpy(h + o, src, n * sizeof(T));
    used += n * sizeof(T);
    return o;
  }
};

struct AttemptDoc {
  uint32_t doc;            // index in the staged batch
  HeaderInfo hdr;
  bool active = false;
  int fail = 0;            // 1 = attempt fails, 2 = unsupported
};

uint32_t pow2_floor(uint32_t x) { uint32_t p = 1; while ((uint64_t)p * 2 <= x) p *= 2; return p; }

}  // namespace

// One (encoding, separator) attempt over `sel` (indices into the staged batch).
// arena/off/len describe the documents for THIS encoding. Marks successes in ctx->r_*.
static int csv_attempt(rfq_ctx* c, const std::vector<uint32_t>& sel, const uint8_t* d_arena,
                       const std::vector<uint32_t>& off, const std::vector<uint32_t>& len,
                       int enc, int sep_i, const std::vector<int64_t>& max_rows,
                       std::vector<uint8_t>& resolved) {
  cudaStream_t s = c->stream;
  const uint32_t m = (uint32_t)sel.size();
  const uint32_t sep = kSeps[sep_i];
  // ---- 1. document tables for the tokenizer
  std::vector<uint32_t> a_off(m), a_len(m), a_tb(m + 1), a_mr(m);
  uint64_t tiles = 0;
  for (uint32_t i = 0; i < m; ++i) {
    a_off[i] = off[sel[i]]; a_len[i] = len[sel[i]];
    a_tb[i] = (uint32_t)tiles;
    tiles += ((uint64_t)a_len[i] + 1 + RFQ_TILE - 1) / RFQ_TILE;
    int64_t mr = max_rows[sel[i]];
    a_mr[i] = (uint32_t)std::min<int64_t>(std::max<int64_t>(mr, 0), 0x7FFFFFF0);
  }
  a_tb[m] = (uint32_t)tiles;
  const uint32_t nt = (uint32_t)tiles;
  RFQ_CK(c, c->h_up.reserve(16ull * (m + 2) + 4096));
  Uploader up((uint8_t*)c->h_up.p);
  size_t o_off = up.put(a_off.data(), m), o_len = up.put(a_len.data(), m);
  size_t o_tb = up.put(a_tb.data(), m + 1), o_mr = up.put(a_mr.data(), m);
  uint8_t* d_up = c->pool.alloc<uint8_t>(up.used + 256);
  ScanArgs sa{};
  sa.arena = d_arena;
  sa.n_docs = m; sa.n_tiles = nt; sa.sep = sep;
  sa.tile_fn = c->pool.alloc<uint32_t>(nt + 1);
  sa.tile_fn_pre = c->pool.alloc<uint32_t>(nt + 1);
  sa.lane_state = c->pool.alloc<uint32_t>(3ull * nt + 1);
  sa.lane_cls = c->pool.alloc<unsigned long long>(32ull * nt + 32);
  sa.lane_fn = c->pool.alloc<uint32_t>(32ull * nt + 32);
  sa.lane_cnt = c->pool.alloc<uint16_t>(32ull * nt + 32);
  sa.tile_nf = c->pool.alloc<uint32_t>(nt + 1);
  sa.tile_nl = c->pool.alloc<uint32_t>(nt + 1);
  sa.tile_nf_pre = c->pool.alloc<uint32_t>(nt + 1);
  sa.tile_nl_pre = c->pool.alloc<uint32_t>(nt + 1);
  sa.doc_final = c->pool.alloc<uint32_t>(m + 1);
  sa.doc_unsup = c->pool.alloc<uint32_t>(m + 1);
  sa.scan_tmp = c->pool.alloc<uint32_t>(scan_partials_needed(nt + 1));
  DocShape* d_shape = c->pool.alloc<DocShape>(m + 1);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  RFQ_CK(c, cudaMemcpyAsync(d_up, c->h_up.p, up.used, cudaMemcpyHostToDevice, s));
  sa.doc_off = (const uint32_t*)(d_up + o_off);
  sa.doc_len = (const uint32_t*)(d_up + o_len);
  sa.tile_base = (const uint32_t*)(d_up + o_tb);
  const uint32_t* d_mr = (const uint32_t*)(d_up + o_mr);
  sa.max_rows = d_mr;
  RFQ_CK(c, cudaMemsetAsync(sa.doc_final, 0, 4ull * (m + 1), s));
  RFQ_CK(c, cudaMemsetAsync(sa.doc_unsup, 0, 4ull * (m + 1), s));
  csv_scan_count(sa, s);
  RFQ_CK(c, c->h_down.reserve(sizeof(DocShape) * (m + 1) + 32ull * (m + 1) + 4096));
  uint32_t* h_tot = (uint32_t*)c->h_down.p;
  RFQ_CK(c, cudaMemcpyAsync(h_tot, sa.tile_nf_pre + nt, 4, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaMemcpyAsync(h_tot + 1, sa.tile_nl_pre + nt, 4, cudaMemcpyDeviceToHost, s));
  // ---- header names on the host while the scan runs (first record only)
  std::vector<AttemptDoc> ad(m);
  for (uint32_t i = 0; i < m; ++i) {
    ad[i].doc = sel[i];
    ad[i].hdr = parse_header(c->h_arena + c->h_doc_off[sel[i]], c->h_doc_len[sel[i]], sep, enc == 1);
  }
  RFQ_CK(c, cudaStreamSynchronize(s));
  const uint32_t NF = h_tot[0], NL = h_tot[1];
  const uint32_t n_slots = NL + m;
  // ---- 2. index pass + record shapes
  sa.fend = c->pool.alloc<uint32_t>((size_t)NF + 1);
  sa.rec_fend = c->pool.alloc<uint32_t>((size_t)n_slots + 1);
  sa.rec_start = c->pool.alloc<uint32_t>((size_t)n_slots + 1);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  csv_scan_index(sa, s);
  csv_doc_shape(sa, d_mr, d_shape, n_slots, s);
  DocShape* h_shape = (DocShape*)((uint8_t*)c->h_down.p + 256);
  RFQ_CK(c, cudaMemcpyAsync(h_shape, d_shape, sizeof(DocShape) * m, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaStreamSynchronize(s));
  // ---- 3. validate shapes, carve the table arrays
  std::vector<TDoc> td(m);
  std::vector<uint32_t> item_base(m + 1), slot_r0(m + 1);
  std::vector<uint32_t> name_off, name_len, name_cp;
  std::string names;
  uint64_t n_stat = 0, n_cols = 0, n_items = 0;
  for (uint32_t i = 0; i < m; ++i) {
    const DocShape& sh = h_shape[i];
    AttemptDoc& A = ad[i];
    TDoc& T = td[i];
    memset(&T, 0, sizeof(T));
    T.off = a_off[i]; T.len = a_len[i]; T.r0 = sh.r0; T.f0 = sh.f0;
    T.stat_base = (uint32_t)n_stat;   // non-decreasing over ALL documents: csv_decide_kernel searches it
    slot_r0[i] = sh.r0;
    item_base[i] = (uint32_t)n_items;
    uint32_t nrows = a_mr[i];
    uint32_t h = (uint32_t)A.hdr.names.size();
    if (sh.pad) A.fail = 2;                                                            // bare CR + blank-first record
    else if (sh.n_rec == 0 || !A.hdr.found || !A.hdr.utf8_ok) A.fail = 1;             // EmptyDataError / decode
    else if (sh.final_state == ST_Q && (uint64_t)sh.n_rec < (uint64_t)nrows + 1) A.fail = 1;  // EOF inside string
    else if (h != sh.hdr_nf) A.fail = 2;                                                // internal mismatch: never guess
    else {
      uint32_t W = h;
      if (sh.n_rec >= 2 && sh.first_nf > W) W = sh.first_nf;
      if (sh.max_nf > W) A.fail = 1;             // "Expected N fields in line K, saw M"
      else if (h <= 1) A.fail = 1;               // len(df.columns) > 1 is false
      else if (W > 1024) A.fail = 2;
      else {
        T.W = W; T.lead = W - h; T.h = h;
        T.n_data = std::min<uint32_t>(sh.n_rec - 1, nrows);
        T.bl = chunk_rows_for_width(W);
        T.n_chunks = (T.n_data + T.bl - 1) / T.bl;
        T.rb = std::min<uint32_t>(std::min<uint32_t>(256, pow2_floor(std::max<uint32_t>(1, 2048 / W))), T.bl);
        T.stat_base = (uint32_t)n_stat; n_stat += (uint64_t)T.n_chunks * W;
        T.col_base = (uint32_t)n_cols; n_cols += h;
        n_items += (T.n_data + T.rb - 1) / T.rb;
        std::string banner = std::string("=== CSV Data (using ") + kEncNames[enc] + ", separator '" +
                             (char)sep + "') ===\n";
        T.banner_off = (uint32_t)names.size(); T.banner_len = (uint32_t)banner.size();
        names += banner;
        std::string empty = empty_frame_text(A.hdr.names);
        T.empty_off = (uint32_t)names.size(); T.empty_len = (uint32_t)empty.size();
        names += empty;
        for (uint32_t j = 0; j < h; ++j) {
          name_off.push_back((uint32_t)names.size());
          name_len.push_back((uint32_t)A.hdr.disp[j].size());   // display label (escapes, trim_front)
          name_cp.push_back(A.hdr.ncp[j]);
          T.hdr_extra += (uint32_t)A.hdr.disp[j].size() - A.hdr.ncp[j];
          names += A.hdr.disp[j];
        }
        T.active = 1; A.active = true;
      }
    }
    T.item_base = item_base[i];
  }
  item_base[m] = (uint32_t)n_items;
  slot_r0[m] = n_slots;
  if (n_stat >= (1ull << 31) || n_items >= (1ull << 31)) { c->err = "batch too large"; return RFQ_ERR_ARG; }
  // ---- 4. upload the table description
  RFQ_CK(c, c->h_up.reserve(sizeof(TDoc) * (m + 1) + 8ull * (m + 2) + 12ull * (n_cols + 1) +
                            names.size() + 8192));
  Uploader up2((uint8_t*)c->h_up.p);
  size_t o_td = up2.put(td.data(), m), o_ib = up2.put(item_base.data(), m + 1);
  size_t o_sr = up2.put(slot_r0.data(), m + 1);
  size_t o_no = up2.put(name_off.data(), name_off.size());
  size_t o_nl = up2.put(name_len.data(), name_len.size());
  size_t o_nc = up2.put(name_cp.data(), name_cp.size());
  size_t o_nm = up2.put(names.data(), names.size());
  uint8_t* d_up2 = c->pool.alloc<uint8_t>(up2.used + 256);
  TableArgs ta{};
  ta.arena = d_arena;
  ta.n_docs = m; ta.n_items = (uint32_t)n_items;
  ta.fend = sa.fend; ta.rec_fend = sa.rec_fend; ta.rec_start = sa.rec_start;
  ta.n_slots = n_slots;
  ta.n_stat = (uint32_t)n_stat; ta.n_cols = (uint32_t)n_cols;
  ta.stat = c->pool.alloc<ChunkStat>(n_stat + 1);
  ta.meas = c->pool.alloc<ChunkMeasure>(n_stat + 1);
  ta.ctype = c->pool.alloc<uint8_t>(n_stat + 1);
  ta.chunk_nf = c->pool.alloc<uint8_t>(n_stat + 1);
  ta.row_na = c->pool.alloc<uint32_t>((size_t)n_slots + 1);
  ta.row_extra = c->pool.alloc<uint32_t>((size_t)n_slots + 1);
  ta.row_bytes = c->pool.alloc<unsigned long long>((size_t)n_slots + 2);
  ta.row_off = c->pool.alloc<unsigned long long>((size_t)n_slots + 2);
  ta.n_keep = c->pool.alloc<uint32_t>(m + 1);
  ta.doc_flags = c->pool.alloc<int32_t>(m + 1);
  ta.doc_linew = c->pool.alloc<uint32_t>(m + 1);
  ta.col = c->pool.alloc<ColOut>(n_cols + 1);
  ta.scan_tmp64 = c->pool.alloc<unsigned long long>(scan_partials_needed((uint64_t)n_slots + 1));
  ta.doc_out_off = c->pool.alloc<unsigned long long>(m + 1);
  ta.doc_out_len = c->pool.alloc<unsigned long long>(m + 1);
  // one sweep for typing AND measuring (cell_ops.cuh classify_measure_cell); RFQ_TWO_SWEEP=1 keeps
  // the separate exact measure sweep for every document (debugging / cross-checking)
  static const bool two_sweep = getenv("RFQ_TWO_SWEEP") && atoi(getenv("RFQ_TWO_SWEEP")) != 0;
  if (!two_sweep) {
    ta.cand = c->pool.alloc<ChunkCand>(n_stat + 1);
    ta.chunk_dropped = c->pool.alloc<uint32_t>(n_stat + 1);
    ta.doc_exact = c->pool.alloc<uint32_t>(m + 2);
    ta.cell_tag = c->pool.alloc<uint8_t>((size_t)NF + 16);
    uint32_t mw = 1;
    for (uint32_t i = 0; i < m; ++i) mw = std::max(mw, td[i].W);
    ta.max_w = mw;
  }
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  RFQ_CK(c, cudaMemcpyAsync(d_up2, c->h_up.p, up2.used, cudaMemcpyHostToDevice, s));
  ta.docs = (const TDoc*)(d_up2 + o_td);
  ta.item_doc_base = (const uint32_t*)(d_up2 + o_ib);
  ta.slot_doc_r0 = (const uint32_t*)(d_up2 + o_sr);
  ta.name_off = (const uint32_t*)(d_up2 + o_no);
  ta.name_len = (const uint32_t*)(d_up2 + o_nl);
  ta.name_cp = (const uint32_t*)(d_up2 + o_nc);
  ta.names = d_up2 + o_nm;
  RFQ_CK(c, cudaMemsetAsync(ta.row_na, 0, 4ull * (n_slots + 1), s));
  RFQ_CK(c, cudaMemsetAsync(ta.row_extra, 0, 4ull * (n_slots + 1), s));
  RFQ_CK(c, cudaMemsetAsync(ta.chunk_nf, 0, n_stat + 1, s));
  if (ta.cand) {
    RFQ_CK(c, cudaMemsetAsync(ta.chunk_dropped, 0, 4ull * (n_stat + 1), s));
    RFQ_CK(c, cudaMemsetAsync(ta.doc_exact, 0, 4ull * (m + 2), s));
  }
  RFQ_CK(c, cudaMemsetAsync(ta.n_keep, 0, 4ull * (m + 1), s));
  RFQ_CK(c, cudaMemsetAsync(ta.doc_flags, 0, 4ull * (m + 1), s));
  RFQ_CK(c, cudaMemsetAsync(ta.doc_linew, 0, 4ull * (m + 1), s));
  // ---- 5. classify, decide, measure, layout, sizes
  csv_table_classify(ta, s);
  csv_table_measure(ta, s);
  uint8_t* hd = (uint8_t*)c->h_down.p;
  unsigned long long* h_total = (unsigned long long*)hd;
  unsigned long long* h_ooff = (unsigned long long*)(hd + 256);
  unsigned long long* h_olen = h_ooff + (m + 1);
  uint32_t* h_keep = (uint32_t*)(h_olen + (m + 1));
  int32_t* h_flags = (int32_t*)(h_keep + (m + 1));
  RFQ_CK(c, cudaMemcpyAsync(h_total, ta.row_off + n_slots, 8, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaMemcpyAsync(h_ooff, ta.doc_out_off, 8ull * m, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaMemcpyAsync(h_olen, ta.doc_out_len, 8ull * m, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaMemcpyAsync(h_keep, ta.n_keep, 4ull * m, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaMemcpyAsync(h_flags, ta.doc_flags, 4ull * m, cudaMemcpyDeviceToHost, s));
  // slice boundaries for the overlapped device->host copy (e2e path): K slices of work items, the
  // output offset where each slice starts is row_off of its first row
  const int K = 4;
  const bool overlap = c->overlap_fetch && c->segs.empty() && c->copy_stream && n_items >= 64;
  uint32_t cut_item[K + 1];
  unsigned long long* h_cut = (unsigned long long*)(hd + 128);
  if (overlap) {
    for (int k = 0; k <= K; ++k) cut_item[k] = (uint32_t)(n_items * k / K);
    for (int k = 1; k < K; ++k) {
      uint32_t it = cut_item[k];
      uint32_t d = (uint32_t)(std::upper_bound(item_base.begin(), item_base.begin() + m, it) - item_base.begin()) - 1;
      uint32_t slot = td[d].r0 + 1 + (it - item_base[d]) * td[d].rb;
      RFQ_CK(c, cudaMemcpyAsync(h_cut + k, ta.row_off + slot, 8, cudaMemcpyDeviceToHost, s));
    }
  }
  RFQ_CK(c, cudaStreamSynchronize(s));
  const uint64_t total = *h_total;
  // ---- 6. emit
  uint8_t* d_text = c->pool.alloc<uint8_t>(total + 64);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  const uint64_t host_base = c->out_total;
  bool copied = false;
  if (total && overlap && total >= (4u << 20)) {
    RFQ_CK(c, c->h_out.reserve(total + 16));
    h_cut[0] = 0; h_cut[K] = total;
    csv_table_emit_header(ta, d_text, s);
    for (int k = 0; k < K; ++k) {
      csv_table_emit_items(ta, d_text, cut_item[k], cut_item[k + 1], s);
      RFQ_CK(c, cudaEventRecord(c->ev_slice[k], s));
      RFQ_CK(c, cudaStreamWaitEvent(c->copy_stream, c->ev_slice[k], 0));
      uint64_t b0 = h_cut[k], b1 = h_cut[k + 1];
      if (b1 > b0)
        RFQ_CK(c, cudaMemcpyAsync((uint8_t*)c->h_out.p + b0, d_text + b0, b1 - b0, cudaMemcpyDeviceToHost,
                                  c->copy_stream));
    }
    copied = true;
  } else if (total) {
    csv_table_emit(ta, d_text, s);
  }
  c->prof.mark("host+sync", s);
  RFQ_CK(c, cudaGetLastError());
  c->segs.push_back({d_text, total, host_base});
  c->segs.back().copied = copied;
  c->out_total += total;
  for (uint32_t i = 0; i < m; ++i) {
    uint32_t d = sel[i];
    int fail = ad[i].fail;
    if (!fail && (h_flags[i] & 1)) fail = 1;   // UnicodeDecodeError in a cell
    if (!fail && (h_flags[i] & 2)) fail = 2;
    if (fail == 1) continue;                    // next (encoding, separator)
    resolved[d] = 1;
    if (fail == 2) { c->r_status[d] = RFQ_DOC_UNSUPPORTED; continue; }
    c->r_status[d] = RFQ_DOC_OK;
    c->r_off[d] = host_base + h_ooff[i];
    c->r_len[d] = h_olen[i];
    c->r_rows[d] = h_keep[i];
    c->r_cols[d] = td[i].h;
    c->r_attempt[d] = (uint8_t)(enc * 3 + sep_i);
  }
  return RFQ_OK;
}

// One (encoding, separator) attempt through the row pipeline (csv2.h).  The host looks at the first
// two records of every document (header names, table width), carves every array from bounds, enqueues
// the whole attempt and synchronises ONCE, when it reads the results.  Documents whose header already
// decides the attempt (no header, one column) never reach the device.  Documents whose chunk types the
// one-sweep candidates do not cover come back in `general` and go through csv_attempt above.
namespace rfq {
__global__ void __launch_bounds__(128) csv2_header_scan_kernel(const uint8_t* __restrict__ arena, const uint32_t* __restrict__ doc_off,
                                                               const uint32_t* __restrict__ doc_len, const uint32_t* __restrict__ sel,
                                                               uint32_t n, uint32_t sep, bool latin1, HeaderScan* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t d = sel[i];
  out[i] = scan_header(arena + doc_off[d], (uint64_t)doc_len[d], DfaDirect{sep}, latin1);
}
void csv2_header_scan(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, const uint32_t* sel,
                      uint32_t n, uint32_t sep, bool latin1, void* out, cudaStream_t s) {
  if (n == 0) return;
  prof_mark("csv2_header_scan_kernel", s);
  RFQ_LAUNCH(csv2_header_scan_kernel, (n + 127) / 128, 128, 0, s, arena, doc_off, doc_len, sel, n, sep, latin1, (HeaderScan*)out);
}
}  // namespace rfq

static int csv2_attempt(rfq_ctx* c, const std::vector<uint32_t>& sel, const uint8_t* d_arena,
                        const std::vector<uint32_t>& off, const std::vector<uint32_t>& len,
                        int enc, int sep_i, const std::vector<int64_t>& max_rows,
                        std::vector<uint8_t>& resolved, std::vector<uint32_t>& general) {
  cudaStream_t s = c->stream;
  const uint32_t sep = kSeps[sep_i];
  static const bool host_prof = getenv("RFQ_HOST_PROFILE") != nullptr;   // host time of every stage on stderr
  auto t_prev = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!host_prof) return;
    const auto t = std::chrono::steady_clock::now();
    fprintf(stderr, "csv2_attempt host %-22s %8.3f ms (%u documents)\n", what,
            std::chrono::duration<double, std::milli>(t - t_prev).count(), (unsigned)sel.size());
    t_prev = t;
  };
  // ---- 1. headers on the host; documents the header already decides.  The scan of the first two
  // records allocates nothing; names are built once per DISTINCT header record (batches of files made
  // from one template share it), and their texts are stored once in the names arena.
  struct HeaderTexts {
    HeaderInfo info; uint32_t banner_off, banner_len, empty_off, empty_len, hdr_extra, name_base;
    const uint8_t* key_p; size_t key_n;             // the header record it was built from (in the caller's arena)
  };
  std::unordered_map<uint64_t, uint32_t> seen_hash; // hash of the header record (computed on the threads) -> hts
  std::unordered_map<std::string, uint32_t> seen;   // header record bytes -> hts, for records that share a hash
  std::vector<HeaderTexts> hts;
  std::vector<uint32_t> act, act_ht, act_first_nf;  // indices into sel / hts, second-record fields
  std::vector<uint32_t> name_off, name_len, name_cp;
  std::string names, key;
  std::vector<HeaderScan> scans_host;
  const HeaderScan* scans = nullptr;
  if (sel.size() >= 4096) {
    // thousands of documents: their first bytes are already on the device, one thread each reads them there
    // (on the host this walk misses the cache once per document: 5-7 ms for 40 000 files)
    const uint32_t ns = (uint32_t)sel.size();
    RFQ_CK(c, c->h_up.reserve(4ull * ns + 64));
    memcpy(c->h_up.p, sel.data(), 4ull * ns);
    uint32_t* d_sel = c->pool.alloc<uint32_t>(ns + 1);
    HeaderScan* d_scans = c->pool.alloc<HeaderScan>(ns + 1);
    if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
    RFQ_CK(c, c->h_down.reserve(sizeof(HeaderScan) * (ns + 1) + 64));
    RFQ_CK(c, cudaMemcpyAsync(d_sel, c->h_up.p, 4ull * ns, cudaMemcpyHostToDevice, s));
    csv2_header_scan((const uint8_t*)c->d_arena.p, c->d_doc_off, c->d_doc_len, d_sel, ns, sep, enc == 1, d_scans, s);
    RFQ_CK(c, cudaMemcpyAsync(c->h_down.p, d_scans, sizeof(HeaderScan) * ns, cudaMemcpyDeviceToHost, s));
    RFQ_CK(c, cudaStreamSynchronize(s));
    scans_host.assign((const HeaderScan*)c->h_down.p, (const HeaderScan*)c->h_down.p + ns);   // h_down is reused below
    scans = scans_host.data();
  } else {
    scans_host.resize(sel.size());
    const DfaTables dfa(sep);
    parallel_for((uint32_t)sel.size(), [&](uint32_t i) {
      const uint32_t d = sel[i];
      scans_host[i] = scan_header(c->h_arena + c->h_doc_off[d], c->h_doc_len[d], dfa, enc == 1);
    });
    scans = scans_host.data();
  }
  for (uint32_t i = 0; i < sel.size(); ++i) {
    const uint32_t d = sel[i];
    const uint8_t* hdoc = c->h_arena + c->h_doc_off[d];
    const HeaderScan& hs = scans[i];
    if (!hs.found || hs.h <= 1) continue;                       // EmptyDataError / len(df.columns) <= 1
    const size_t key_n = (size_t)(hs.end - hs.begin);
    uint32_t hi = 0xFFFFFFFFu;
    bool hash_taken = false;
    {
      auto f = seen_hash.find(hs.hash);
      if (f != seen_hash.end()) {
        hash_taken = true;
        const HeaderTexts& T0 = hts[f->second];
        if (T0.key_n == key_n && memcmp(T0.key_p, hdoc + hs.begin, key_n) == 0) hi = f->second;
      }
    }
    if (hi == 0xFFFFFFFFu && hash_taken) {
      key.assign((const char*)hdoc + hs.begin, key_n);
      auto it = seen.find(key);
      if (it != seen.end()) hi = it->second;
    }
    if (hi == 0xFFFFFFFFu) {
      HeaderTexts T;
      T.info = parse_header(hdoc, c->h_doc_len[d], sep, enc == 1);
      T.banner_off = T.banner_len = T.empty_off = T.empty_len = T.hdr_extra = T.name_base = 0;
      if (T.info.found && T.info.utf8_ok) {
        const HeaderInfo& H = T.info;
        std::string banner = std::string("=== CSV Data (using ") + kEncNames[enc] + ", separator '" + (char)sep + "') ===\n";
        T.banner_off = (uint32_t)names.size(); T.banner_len = (uint32_t)banner.size();
        names += banner;
        std::string empty = empty_frame_text(H.names);
        T.empty_off = (uint32_t)names.size(); T.empty_len = (uint32_t)empty.size();
        names += empty;
        T.name_base = (uint32_t)name_off.size();
        for (size_t j = 0; j < H.names.size(); ++j) {
          name_off.push_back((uint32_t)names.size());
          name_len.push_back((uint32_t)H.disp[j].size());
          name_cp.push_back(H.ncp[j]);
          T.hdr_extra += (uint32_t)H.disp[j].size() - H.ncp[j];
          names += H.disp[j];
        }
      }
      T.key_p = hdoc + hs.begin; T.key_n = key_n;
      hi = (uint32_t)hts.size();
      hts.push_back(std::move(T));
      if (!hash_taken) seen_hash.emplace(hs.hash, hi);
      else { key.assign((const char*)hdoc + hs.begin, key_n); seen.emplace(key, hi); }
    }
    const HeaderInfo& H = hts[hi].info;
    if (!H.found || !H.utf8_ok || H.names.size() <= 1) continue;   // decode error in a label
    const uint32_t W = std::max<uint32_t>((uint32_t)H.names.size(), hs.first_nf);
    if (W > 1024) { resolved[d] = 1; c->r_status[d] = RFQ_DOC_UNSUPPORTED; continue; }
    act.push_back(i); act_ht.push_back(hi); act_first_nf.push_back(hs.first_nf);
  }
  const uint32_t m = (uint32_t)act.size();
  if (m == 0) return RFQ_OK;
  lap("1 headers");
  // ---- 2. carve
  std::vector<Doc2> docs(m);
  std::vector<uint32_t> tile_base(m + 1), item_base(m + 1);
  uint64_t tiles = 0, n_slots = 0, n_items = 0, n_stat = 0, n_cols = 0, in_bytes = 0, cells = 0, meta_vecs = 0;
  uint32_t max_w = 1;
  for (uint32_t k = 0; k < m; ++k) {
    const uint32_t d = sel[act[k]];
    const HeaderTexts& T = hts[act_ht[k]];
    Doc2& D = docs[k];
    memset(&D, 0, sizeof(D));
    D.off = off[d]; D.len = len[d]; D.sep = sep;
    D.max_rows = (uint32_t)std::min<int64_t>(std::max<int64_t>(max_rows[d], 0), 0x7FFFFFF0);
    D.h = (uint32_t)T.info.names.size();
    D.W = std::max<uint32_t>(D.h, act_first_nf[k]);
    D.lead = D.W - D.h;
    D.bl = chunk_rows_for_width(D.W);
    D.latin1 = enc == 1;
    const uint64_t rows_bound = std::min<uint64_t>(D.max_rows, (uint64_t)D.len / 2 + 2);   // a record takes two bytes at least
    D.slot_base = (uint32_t)n_slots; n_slots += rows_bound + 2;
    D.item_base = (uint32_t)n_items; D.n_items = (uint32_t)std::max<uint64_t>(1, (rows_bound + RFQ2_ROWS - 1) / RFQ2_ROWS);
    n_items += D.n_items;
    D.stat_base = (uint32_t)n_stat; D.n_chunks = (uint32_t)std::max<uint64_t>(1, (rows_bound + D.bl - 1) / D.bl);
    n_stat += (uint64_t)D.n_chunks * D.W;
    D.col_base = (uint32_t)n_cols; n_cols += D.h;      // layout of this document's columns
    D.name_base = T.name_base;                          // labels: shared by every document with this header record
    D.meta_base = (uint32_t)meta_vecs; meta_vecs += rows_bound * RFQ2_META_VECS(D.W);
    cells += rows_bound * D.W;
    max_w = std::max(max_w, D.W);
    tile_base[k] = (uint32_t)tiles; item_base[k] = D.item_base;
    tiles += ((uint64_t)D.len + 1 + RFQ_TILE - 1) / RFQ_TILE;
    in_bytes += D.len;
    D.banner_off = T.banner_off; D.banner_len = T.banner_len;
    D.empty_off = T.empty_off; D.empty_len = T.empty_len; D.hdr_extra = T.hdr_extra;
  }
  tile_base[m] = (uint32_t)tiles; item_base[m] = (uint32_t)n_items;
  if (n_slots >= (1ull << 31) || n_items >= (1ull << 31) || n_stat >= (1ull << 31) || tiles >= (1ull << 31) ||
      meta_vecs >= (1ull << 31)) {
    c->err = "batch too large"; return RFQ_ERR_ARG;
  }
  Csv2Args a{};
  a.arena = d_arena; a.n_docs = m;
  a.n_tiles = (uint32_t)tiles; a.n_chunks = (uint32_t)((tiles + RFQ2_SCAN_TILES - 1) / RFQ2_SCAN_TILES);
  a.n_items = (uint32_t)n_items; a.n_slots = (uint32_t)n_slots; a.n_stat = (uint32_t)n_stat; a.n_cols = (uint32_t)n_cols;
  a.max_w = max_w;
  lap("2 carve");
  // ---- 3. upload the description
  const size_t z_fn = 0, z_cnt = z_fn + 8ull * a.n_chunks, z_tick = z_cnt + 8ull * a.n_chunks, z_cntr = z_tick + 16;
  // counters | run | doc_out_off | doc_out_len are what the host reads back: contiguous, one copy
  const size_t z_run = z_cntr + 32, z_ooff = z_run + sizeof(DocRun) * m, z_olen = z_ooff + 8ull * (m + 1), z_down_end = z_olen + 8ull * (m + 1);
  const size_t z_stat = (z_down_end + 255) & ~(size_t)255;
  const size_t z_ctype = (z_stat + sizeof(StatV2) * n_stat + 255) & ~(size_t)255, z_allna = (z_ctype + n_stat + 255) & ~(size_t)255;
  const size_t z_nf = (z_allna + 4ull * n_stat + 255) & ~(size_t)255, z_end = (z_nf + n_stat + 255) & ~(size_t)255;
  RFQ_CK(c, c->h_up.reserve(sizeof(Doc2) * (m + 1) + 8ull * (m + 2) + 12ull * (name_off.size() + 1) + names.size() + 8192));
  Uploader up((uint8_t*)c->h_up.p);
  const size_t o_docs = up.put(docs.data(), m), o_tb = up.put(tile_base.data(), m + 1), o_ib = up.put(item_base.data(), m + 1);
  const size_t o_no = up.put(name_off.data(), name_off.size()), o_nl = up.put(name_len.data(), name_len.size());
  const size_t o_nc = up.put(name_cp.data(), name_cp.size()), o_nm = up.put(names.data(), names.size());
  // ---- 4. scratch: one zeroed region, then arrays that are written before they are read.  (Sending the zeros
  // with the description in one copy instead of a copy and a memset was measured: 0.07 instead of 0.02 ms
  // before the scan can start.)
  const size_t up_al = (up.used + 255) & ~(size_t)255;
  uint8_t* d_up = c->pool.alloc<uint8_t>(up_al + z_end + 256);
  uint8_t* z = d_up ? d_up + up_al : nullptr;
  a.rec_start = c->pool.alloc<uint32_t>(n_slots + 2);
  {   // scratch of the three-launch scan (half the input size for the lanes' marks)
    static const bool split = getenv("RFQ2_SCAN_SPLIT") ? atoi(getenv("RFQ2_SCAN_SPLIT")) != 0 : true;   // RFQ2_SCAN_SPLIT=0: the one-launch scan with its count look-back
    if (split) {
      a.lane_marks = c->pool.alloc<uint2>(32ull * a.n_tiles + 32);
      a.tile_tot = c->pool.alloc<uint32_t>((size_t)a.n_tiles + 32);
      a.tile_doc = c->pool.alloc<uint32_t>((size_t)a.n_tiles + 32);
      a.chunk_seg = c->pool.alloc<unsigned long long>((size_t)a.n_chunks + 1);
      a.chunk_pre = c->pool.alloc<unsigned long long>((size_t)a.n_chunks + 1);
    }
  }
  a.row_na = c->pool.alloc<uint32_t>(n_slots + 2);
  a.row_extra = c->pool.alloc<uint32_t>(n_slots + 2);
  a.cell_meta = (uint16_t*)c->pool.alloc<uint4>(meta_vecs + 1);
  a.meas = c->pool.alloc<ChunkMeasure>(n_stat + 1);
  a.item_allna = c->pool.alloc<uint32_t>(n_items + 1);
  a.item_extra = c->pool.alloc<uint32_t>(n_items + 1);
  a.item_bytes = c->pool.alloc<unsigned long long>(n_items + 2);
  a.item_off = c->pool.alloc<unsigned long long>(n_items + 2);
  a.scan_tmp64 = c->pool.alloc<unsigned long long>(scan_partials_needed(n_items + 1));
  a.col = c->pool.alloc<ColOut2>(n_cols + 1);
  a.cold_cells_cap = (uint32_t)std::min<uint64_t>(cells + 1, 4u << 20);
  a.cold_cells = c->pool.alloc<ColdCell>(a.cold_cells_cap);
  a.cold_rows_cap = (uint32_t)n_slots;
  a.cold_rows = c->pool.alloc<ColdRow>(a.cold_rows_cap + 1);
  // output: sized from what the last run produced, else twice the input; a sweep that does not fit
  // reports it and is run again in a larger buffer
  unsigned long long cap = std::max<unsigned long long>(2ull * in_bytes, c->csv2_out_hint + c->csv2_out_hint / 16) + (64u << 10);
  uint8_t* d_text = c->pool.alloc<uint8_t>(cap + 64);
  if (c->pool.failed || !z || !d_up) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  a.fn_desc = (unsigned long long*)(z + z_fn); a.cnt_desc = (unsigned long long*)(z + z_cnt);
  a.ticket = (uint32_t*)(z + z_tick); a.counters = (uint32_t*)(z + z_cntr);
  a.run = (DocRun*)(z + z_run); a.stat = (StatV2*)(z + z_stat); a.ctype = z + z_ctype;
  a.doc_out_off = (unsigned long long*)(z + z_ooff); a.doc_out_len = (unsigned long long*)(z + z_olen);
  a.chunk_allna = (uint32_t*)(z + z_allna); a.chunk_nf = z + z_nf;
  RFQ_CK(c, cudaMemcpyAsync(d_up, c->h_up.p, up.used, cudaMemcpyHostToDevice, s));
  RFQ_CK(c, cudaMemsetAsync(z, 0, z_end, s));
  a.docs = (const Doc2*)(d_up + o_docs); a.tile_base = (const uint32_t*)(d_up + o_tb);
  a.item_base = (const uint32_t*)(d_up + o_ib);
  a.name_off = (const uint32_t*)(d_up + o_no); a.name_len = (const uint32_t*)(d_up + o_nl);
  a.name_cp = (const uint32_t*)(d_up + o_nc); a.names = d_up + o_nm;
  a.out = d_text; a.out_cap = cap;
  // shared-memory stages of the emit sweep: input span of an item from the average row, output from the
  // last run's largest item (rows that do not fit are written by the fix-up kernel)
  {
    const uint64_t avg_item = n_items ? (in_bytes + n_items - 1) / n_items : 0;
    uint64_t in_st = std::max<uint64_t>(avg_item + avg_item / 4 + 512, c->csv2_in_hint + c->csv2_in_hint / 8 + 64);
    in_st = std::min<uint64_t>(std::max<uint64_t>(in_st, 4096), 48u << 10);
    uint64_t out_st = c->csv2_item_out_hint ? c->csv2_item_out_hint + c->csv2_item_out_hint / 8 + 64 : 2 * in_st;
    out_st = std::min<uint64_t>(std::max<uint64_t>(out_st, 8192), 96u << 10);
    a.in_stage = (uint32_t)((in_st + 15) & ~15ull); a.out_stage = (uint32_t)((out_st + 15) & ~15ull);
  }
  lap("3-4 upload, scratch");
  // ---- 5. the whole attempt, one synchronisation
  RFQ_CK(c, c->h_down.reserve(z_down_end - z_cntr + 64));
  if (z_down_end - z_cntr <= (64u << 10)) {   // a few documents: the last kernel writes the results to pinned memory itself
    a.export_src = (const uint32_t*)(z + z_cntr);
    a.export_dst = (uint32_t*)c->h_down.p;
    a.export_words = (uint32_t)((z_down_end - z_cntr) / 4);
  }
  csv2_scan(a, s);
  csv2_type(a, s);
  csv2_plan(a, s);
  csv2_emit(a, s);
  uint8_t* hd = (uint8_t*)c->h_down.p;
  const bool by_kernel = a.export_words != 0;
  uint32_t* h_cntr = (uint32_t*)hd;                                    // [6..7] = total bytes (64 bits)
  const unsigned long long* h_total = (const unsigned long long*)(hd + 24);
  const DocRun* h_run = (const DocRun*)(hd + (z_run - z_cntr));
  const unsigned long long* h_ooff = (const unsigned long long*)(hd + (z_ooff - z_cntr));
  const unsigned long long* h_olen = (const unsigned long long*)(hd + (z_olen - z_cntr));
  if (!by_kernel) RFQ_CK(c, cudaMemcpyAsync(hd, z + z_cntr, z_down_end - z_cntr, cudaMemcpyDeviceToHost, s));
  c->prof.mark("host+sync", s);
  lap("5 enqueue");
  RFQ_CK(c, cudaStreamSynchronize(s));
  lap("5 wait for the GPU");
  unsigned long long total = *h_total;
  if (total > cap) {   // the output did not fit: larger buffer, emit again
    cap = total + total / 16 + (64u << 10);
    d_text = c->pool.alloc<uint8_t>(cap + 64);
    if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
    a.out = d_text; a.out_cap = cap;
    RFQ_CK(c, cudaMemsetAsync(a.counters + 1, 0, 8, s));
    csv2_emit(a, s);
    RFQ_CK(c, cudaMemcpyAsync(h_cntr, a.counters, 32, cudaMemcpyDeviceToHost, s));
    RFQ_CK(c, cudaStreamSynchronize(s));
  }
  RFQ_CK(c, cudaGetLastError());
  if (h_cntr[2]) { c->err = "csv2: fix-up list overflow"; return RFQ_ERR_STATE; }
  c->csv2_out_hint = total;
  c->csv2_in_hint = h_cntr[3];
  c->csv2_item_out_hint = h_cntr[4];
  const uint64_t host_base = c->out_total;
  c->segs.push_back({d_text, total, host_base});
  c->out_total += total;
  for (uint32_t k = 0; k < m; ++k) {
    const uint32_t d = sel[act[k]];
    const uint32_t fl = h_run[k].flags;
    if (fl & RUN_UNSUP) { resolved[d] = 1; c->r_status[d] = RFQ_DOC_UNSUPPORTED; continue; }
    if (fl & RUN_FAIL) continue;                  // next (encoding, separator)
    if (fl & RUN_GENERAL) { general.push_back(d); continue; }
    resolved[d] = 1;
    c->r_status[d] = RFQ_DOC_OK;
    c->r_off[d] = host_base + h_ooff[k];
    c->r_len[d] = h_olen[k];
    c->r_rows[d] = h_run[k].n_keep;
    c->r_cols[d] = docs[k].h;
    {   // code points of the text: bytes minus the continuation bytes of the header line and of the rows
      uint64_t ex = h_run[k].extra;
      if (h_run[k].n_keep) ex += docs[k].hdr_extra;
      else { uint32_t cp = 0; std::string e = empty_frame_text(hts[act_ht[k]].info.names); utf8_valid_count(e, &cp); ex += e.size() - cp; }
      c->r_chars[d] = h_olen[k] - ex;
    }
    c->r_attempt[d] = (uint8_t)(enc * 3 + sep_i);
  }
  return RFQ_OK;
}

extern "C" int rfq_csv_run(rfq_ctx* c, const int64_t* max_rows_in) {
  if (!c) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_csv_run: nothing staged"; return RFQ_ERR_STATE; }
  if (!c->h_arena) { c->err = "rfq_csv_run: the staged batch exists on the device only (rfq_restage_output); the CSV path reads header records on the host"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  const uint32_t n = c->n_docs;
  if (!c->nested) c->prof.begin();
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  if (!c->nested) { RFQ_CK(c, cudaEventRecord(c->ev0, s)); c->pool.reset(); }
  c->prof.mark("host prep + upload", s);
  c->status_override.clear(); c->out_n = 0;
  c->segs.clear();
  c->out_total = 0;
  c->host_tables = true;
  c->r_off.assign(n, 0); c->r_len.assign(n, 0);
  c->r_status.assign(n, RFQ_DOC_CSV_UNPARSEABLE);
  c->r_rows.assign(n, 0); c->r_cols.assign(n, 0); c->r_attempt.assign(n, 0xFF);
  c->r_chars.assign(n, ~0ull);   // unknown unless the row pipeline produced the document
  std::vector<int64_t> max_rows(n, 1000);
  if (max_rows_in) for (uint32_t d = 0; d < n; ++d) max_rows[d] = max_rows_in[d];
  std::vector<uint8_t> resolved(n, 0);
  std::vector<uint32_t> pending(n);
  for (uint32_t d = 0; d < n; ++d) pending[d] = d;
  // latin-1 view of the pending documents (built only if the utf-8 attempts leave any)
  std::vector<uint32_t> l_off, l_len;
  const uint8_t* d_latin = nullptr;
  for (int enc = 0; enc < 2 && !pending.empty(); ++enc) {
    const uint8_t* arena = (const uint8_t*)c->d_arena.p;
    const std::vector<uint32_t>* off = &c->h_doc_off;
    const std::vector<uint32_t>* len = &c->h_doc_len;
    if (enc == 1) {
      const uint32_t m = (uint32_t)pending.size();
      std::vector<uint32_t> p_off(m), p_len(m);
      for (uint32_t i = 0; i < m; ++i) { p_off[i] = c->h_doc_off[pending[i]]; p_len[i] = c->h_doc_len[pending[i]]; }
      RFQ_CK(c, c->h_up.reserve(16ull * (m + 2) + 4096));
      Uploader up((uint8_t*)c->h_up.p);
      size_t o_off = up.put(p_off.data(), m), o_len = up.put(p_len.data(), m);
      uint8_t* d_up = c->pool.alloc<uint8_t>(up.used + 256);
      uint32_t* d_hi = c->pool.alloc<uint32_t>(m + 1);
      if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
      RFQ_CK(c, cudaMemcpyAsync(d_up, c->h_up.p, up.used, cudaMemcpyHostToDevice, s));
      c->prof.mark("latin1_transcode", s);
      latin1_count(arena, (const uint32_t*)(d_up + o_off), (const uint32_t*)(d_up + o_len), m, d_hi, s);
      RFQ_CK(c, c->h_down.reserve(4ull * (m + 1) + 4096));
      uint32_t* h_hi = (uint32_t*)c->h_down.p;
      RFQ_CK(c, cudaMemcpyAsync(h_hi, d_hi, 4ull * m, cudaMemcpyDeviceToHost, s));
      RFQ_CK(c, cudaStreamSynchronize(s));
      l_off.assign(n, 0); l_len.assign(n, 0);
      std::vector<uint32_t> n_off(m);
      uint64_t total = 0;
      for (uint32_t i = 0; i < m; ++i) {
        uint64_t nl = (uint64_t)p_len[i] + h_hi[i];
        n_off[i] = (uint32_t)total;
        l_off[pending[i]] = (uint32_t)total; l_len[pending[i]] = (uint32_t)nl;
        total += (nl + RFQ_ALIGN - 1) / RFQ_ALIGN * RFQ_ALIGN;
      }
      if (total + 64 >= (1ull << 32)) { c->err = "latin-1 arena must be < 4 GiB"; return RFQ_ERR_ARG; }
      RFQ_CK(c, c->d_arena2.reserve(total + 64));
      RFQ_CK(c, cudaMemsetAsync(c->d_arena2.p, 0, total + 64, s));
      Uploader up2((uint8_t*)c->h_up.p);
      size_t o_no = up2.put(n_off.data(), m);
      uint32_t* d_no = c->pool.alloc<uint32_t>(m + 64);
      if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
      RFQ_CK(c, cudaMemcpyAsync(d_no, (uint8_t*)c->h_up.p + o_no, 4ull * m, cudaMemcpyHostToDevice, s));
      latin1_expand(arena, (const uint32_t*)(d_up + o_off), (const uint32_t*)(d_up + o_len), d_no, m,
                    (uint8_t*)c->d_arena2.p, s);
      RFQ_CK(c, cudaStreamSynchronize(s));  // h_up is reused by the attempts below
      d_latin = (const uint8_t*)c->d_arena2.p;
      arena = d_latin; off = &l_off; len = &l_len;
    }
    // RFQ_CSV_V1=1 keeps every document on the three-sweep path (cross-checking, sheets use it anyway)
    static const bool v1 = getenv("RFQ_CSV_V1") && atoi(getenv("RFQ_CSV_V1")) != 0;
    for (int sep_i = 0; sep_i < 3 && !pending.empty(); ++sep_i) {
      int rc;
      if (v1) rc = csv_attempt(c, pending, arena, *off, *len, enc, sep_i, max_rows, resolved);
      else {
        std::vector<uint32_t> general;
        rc = csv2_attempt(c, pending, arena, *off, *len, enc, sep_i, max_rows, resolved, general);
        if (!rc && !general.empty()) rc = csv_attempt(c, general, arena, *off, *len, enc, sep_i, max_rows, resolved);
      }
      if (rc) return rc;
      std::vector<uint32_t> next;
      for (uint32_t d : pending) if (!resolved[d]) next.push_back(d);
      pending.swap(next);
    }
  }
  c->prof.mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  c->d_out_off = nullptr; c->d_out_len = nullptr;
  c->have_output = true;
  c->timed = true;
  return RFQ_OK;
}


// ---- CSV parse fused with prompt assembly (the /upload/ flow: app/main.py:243-273 never returns raw_text,
// only the prompt built from it and its length) ----------------------------------------------------------
extern "C" int rfq_csv_prompt_run(rfq_ctx* c, const int64_t* max_rows, const uint8_t* head, uint32_t head_len,
                                  const uint8_t* tail, uint32_t tail_len, const uint8_t* mark, uint32_t mark_len,
                                  uint32_t max_chars) {
  if (!c || (head_len && !head) || (tail_len && !tail) || (mark_len && !mark)) return RFQ_ERR_ARG;
  int rc = rfq_csv_run(c, max_rows);
  if (rc) return rc;
  cudaStream_t s = c->stream;
  const uint32_t n = c->n_docs;
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  // what the parse stage produced; where every parsed text lies on the device
  c->raw_bytes.assign(c->r_len.begin(), c->r_len.end());
  c->raw_chars.assign(c->r_chars.begin(), c->r_chars.end());
  c->status_override.assign(n, -1);
  const uint64_t cap_bytes = 4ull * max_chars + 4;     // max_chars + 1 code points never take more
  std::vector<unsigned long long> src(n, 0);
  std::vector<uint32_t> take(n, 0), dst_off(n, 0);
  uint64_t total_g = 0;
  for (uint32_t d = 0; d < n; ++d) {
    if (c->r_status[d] != RFQ_DOC_OK) { c->status_override[d] = c->r_status[d]; c->raw_bytes[d] = 0; c->raw_chars[d] = 0; }
    else {
      const uint64_t off = c->r_off[d];
      for (const auto& sg : c->segs)
        if (off >= sg.host_base && off < sg.host_base + std::max<uint64_t>(sg.bytes, 1)) { src[d] = (unsigned long long)(uintptr_t)(sg.d + (off - sg.host_base)); break; }
      take[d] = (uint32_t)std::min<uint64_t>(c->r_len[d], cap_bytes);
    }
    dst_off[d] = (uint32_t)total_g;
    total_g += ((uint64_t)take[d] + 15) & ~15ull;
  }
  if (total_g + 64 >= (1ull << 32)) { c->err = "prompt arena must be < 4 GiB"; return RFQ_ERR_ARG; }
  const size_t o_tail = ((size_t)head_len + 15) & ~(size_t)15, o_mark = (o_tail + tail_len + 15) & ~(size_t)15;
  const size_t pieces = o_mark + mark_len + 16;
  RFQ_CK(c, c->h_up.reserve(pieces + 16ull * (n + 2) + 4096));
  Uploader up((uint8_t*)c->h_up.p);
  const size_t o_src = up.put(src.data(), n), o_take = up.put(take.data(), n), o_dst = up.put(dst_off.data(), n);
  up.used = (up.used + 255) & ~(size_t)255;
  const size_t o_pieces = up.used;
  uint8_t* hp = (uint8_t*)c->h_up.p + o_pieces;
  if (head_len) memcpy(hp, head, head_len);
  if (tail_len) memcpy(hp + o_tail, tail, tail_len);
  if (mark_len) memcpy(hp + o_mark, mark, mark_len);
  up.used += pieces;
  uint8_t* d_up = c->pool.alloc<uint8_t>(up.used + 256);
  uint8_t* d_gather = c->pool.alloc<uint8_t>(total_g + 64);
  PromptArgs a{};
  a.n_docs = n;
  a.max_chars = max_chars;
  a.cut = c->pool.alloc<uint32_t>(n + 1);
  a.trunc = c->pool.alloc<uint32_t>(n + 1);
  a.status = c->pool.alloc<int32_t>(n + 1);
  a.out_len = c->pool.alloc<unsigned long long>(n + 1);
  a.out_off = c->pool.alloc<unsigned long long>(n + 1);
  a.out_total = c->pool.alloc<unsigned long long>(1);
  a.scan_tmp64 = c->pool.alloc<unsigned long long>(scan_partials_needed(n));
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  RFQ_CK(c, cudaMemcpyAsync(d_up, c->h_up.p, up.used, cudaMemcpyHostToDevice, s));
  RFQ_CK(c, cudaMemsetAsync(d_gather + total_g, 0, 64, s));
  prompt_gather((const unsigned long long*)(d_up + o_src), (const uint32_t*)(d_up + o_take),
                (const uint32_t*)(d_up + o_dst), d_gather, n, s);
  a.arena = d_gather;
  a.doc_off = (const uint32_t*)(d_up + o_dst); a.doc_len = (const uint32_t*)(d_up + o_take);
  a.head = d_up + o_pieces; a.head_len = head_len;
  a.tail = d_up + o_pieces + o_tail; a.tail_len = tail_len;
  a.mark = d_up + o_pieces + o_mark; a.mark_len = mark_len;
  prompt_measure(a, s);
  unsigned long long total = 0;
  if (n) {
    RFQ_CK(c, cudaMemcpyAsync(c->h_scalar(), a.out_total, 8, cudaMemcpyDeviceToHost, s));
    RFQ_CK(c, cudaStreamSynchronize(s));
    total = *c->h_scalar();
  }
  uint8_t* d_text = c->pool.alloc<uint8_t>(total + 16);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  prompt_emit(a, d_text, s);
  // the prompts are this run's output; the parsed texts stay where they are until the next run
  c->segs.clear();
  c->segs.push_back({d_text, total, 0});
  c->host_tables = false;
  prof_mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  RFQ_CK(c, cudaGetLastError());
  c->out_total = total;
  c->d_out_off = a.out_off; c->d_out_len = a.out_len;
  c->d_status = a.status; c->d_rows = nullptr; c->d_cols = nullptr; c->d_attempt = nullptr;
  c->have_output = true;
  c->timed = true;
  return RFQ_OK;
}

extern "C" int rfq_csv_prompt_parse(rfq_ctx* c, const uint8_t* arena, const uint64_t* offsets, const uint64_t* lengths,
                                    uint32_t n, const int64_t* max_rows, const uint8_t* head, uint32_t head_len,
                                    const uint8_t* tail, uint32_t tail_len, const uint8_t* mark, uint32_t mark_len,
                                    uint32_t max_chars) {
  int r = rfq_stage(c, arena, offsets, lengths, n);
  if (r) return r;
  r = rfq_csv_prompt_run(c, max_rows, head, head_len, tail, tail_len, mark, mark_len, max_chars);
  if (r) return r;
  return rfq_fetch(c);
}

extern "C" const uint64_t* rfq_out_raw_bytes(const rfq_ctx* c) { return c && !c->raw_bytes.empty() ? c->raw_bytes.data() : nullptr; }
extern "C" const uint64_t* rfq_out_raw_chars(const rfq_ctx* c) { return c && !c->raw_chars.empty() ? c->raw_chars.data() : nullptr; }
extern "C" const uint64_t* rfq_out_chars(const rfq_ctx* c) { return c && c->host_tables && !c->r_chars.empty() ? c->r_chars.data() : nullptr; }
