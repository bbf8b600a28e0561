// rfq_sheets_run: host orchestration of the sheet path (the GPU side of pd.read_excel for
// .xlsx, app/file_parser.py:225-243).  A batch mixes worksheet parts and sharedStrings parts
// (the host inflated the zip and read workbook.xml / rels / styles.xml);
// every pass over XML bytes or cells is a kernel, the host does O(documents + columns) work.
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
#include "cell_ops.cuh"
#include "col_decide.cuh"
#include "ctx.h"
#include "header_names.h"
#include "xlsx_cells.cuh"

using namespace rfq;

namespace {
struct Up {
  uint8_t* h; size_t used = 0;
  explicit Up(uint8_t* p) : h(p) {}
  template <class T> size_t put(const T* src, size_t n) {
    used = (used + 255) & ~(size_t)255;
    size_t o = used;
    if (n) memcpy(h + o, src, n * sizeof(T));
    used += n * sizeof(T);
    return o;
  }
};
uint32_t p2floor(uint32_t x) { uint32_t p = 1; while ((uint64_t)p * 2 <= x) p *= 2; return p; }
}  // namespace

extern "C" int rfq_sheets_run(rfq_ctx* c, const uint8_t* kind, const int32_t* sst_of,
                              const int64_t* max_rows, const uint32_t* date_off,
                              const uint32_t* date_ids) {
  if (!c || (c->n_docs && (!kind || !sst_of))) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_sheets_run: nothing staged"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  const uint32_t n = c->n_docs;
  static const bool why = getenv("RFQ_SHEET_WHY") != nullptr;   // say on stderr why a sheet is declined
  static const bool host_prof = getenv("RFQ_HOST_PROFILE") != nullptr;   // host time of every stage on stderr
  auto t_prev = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!host_prof) return;
    const auto t = std::chrono::steady_clock::now();
    fprintf(stderr, "rfq_sheets_run host %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t - t_prev).count());
    t_prev = t;
  };
  if (!c->nested) c->prof.begin();
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  if (!c->nested) { RFQ_CK(c, cudaEventRecord(c->ev0, s)); c->pool.reset(); }
  c->status_override.clear(); c->out_n = 0;
  c->segs.clear();
  c->out_total = 0;
  c->host_tables = true;
  c->r_off.assign(n, 0); c->r_len.assign(n, 0);
  c->r_status.assign(n, RFQ_DOC_OK);
  c->r_rows.assign(n, 0); c->r_cols.assign(n, 0); c->r_attempt.assign(n, 0xFF);
  // ---- 1. per-document parameters
  std::vector<uint32_t> cap(n), doff(n + 1, 0);
  for (uint32_t d = 0; d < n; ++d) {
    int64_t mr = max_rows ? max_rows[d] : 1000;
    cap[d] = (uint32_t)std::min<int64_t>(std::max<int64_t>(mr, 0), 0x7FFFFFF0) + 1;  // file_rows_needed
    if (date_off) doff[d + 1] = date_off[d + 1];
    if (!(kind[d] & 1) && sst_of[d] >= 0 && ((uint32_t)sst_of[d] >= n || !(kind[sst_of[d]] & 1))) {
      c->err = "rfq_sheets_run: sst_of must point at a sharedStrings document";
      return RFQ_ERR_ARG;
    }
  }
  const uint32_t n_date = date_off ? doff[n] : 0;
  RFQ_CK(c, c->h_up.reserve(16ull * (n + 2) + 4ull * n_date + 8192));
  Up up((uint8_t*)c->h_up.p);
  size_t o_kind = up.put(kind, n), o_sst = up.put(sst_of, n), o_cap = up.put(cap.data(), n);
  size_t o_doff = up.put(doff.data(), n + 1), o_dids = up.put(date_ids, n_date);
  uint8_t* d_up = c->pool.alloc<uint8_t>(up.used + 256);
  XmlArgs xa{};
  xa.arena = (const uint8_t*)c->d_arena.p;
  xa.doc_off = c->d_doc_off; xa.doc_len = c->d_doc_len; xa.tile_base = c->d_tile_base;
  xa.n_docs = n; xa.n_tiles = c->n_tiles;
  const uint32_t nt = c->n_tiles;
  for (int q = 0; q < 3; ++q) {
    if (q == 0) xa.lane_kinds = c->pool.alloc<uint32_t>(32ull * nt + 32);
    xa.tile_cnt[q] = c->pool.alloc<uint32_t>(nt + 1);
    xa.tile_pre[q] = c->pool.alloc<uint32_t>(nt + 1);
  }
  xa.shape = c->pool.alloc<XDocShape>(n + 1);
  xa.row_base = c->pool.alloc<uint32_t>(n + 2);
  xa.cell_base = c->pool.alloc<uint32_t>(n + 2);
  xa.si_base = c->pool.alloc<uint32_t>(n + 2);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  RFQ_CK(c, cudaMemcpyAsync(d_up, c->h_up.p, up.used, cudaMemcpyHostToDevice, s));
  xa.kind = d_up + o_kind;
  xa.sst_of = (const int32_t*)(d_up + o_sst);
  xa.cap_rows = (const uint32_t*)(d_up + o_cap);
  xa.date_off = (const uint32_t*)(d_up + o_doff);
  xa.date_ids = (const uint32_t*)(d_up + o_dids);
  RFQ_CK(c, cudaMemsetAsync(xa.shape, 0, sizeof(XDocShape) * (n + 1), s));
  // scan partials must cover the largest scan of this run; element counts are <= arena bytes / 2
  xa.scan_tmp = c->pool.alloc<uint32_t>(scan_partials_needed(c->arena_bytes + nt + 16));
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  xml_count(xa, s);
  RFQ_CK(c, c->h_down.reserve(sizeof(XDocShape) * (n + 1) + 4096));
  uint32_t* h_tot = (uint32_t*)c->h_down.p;
  RFQ_CK(c, cudaMemcpyAsync(h_tot, xa.row_base + n, 4, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaMemcpyAsync(h_tot + 1, xa.cell_base + n, 4, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaMemcpyAsync(h_tot + 2, xa.si_base + n, 4, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaStreamSynchronize(s));
  lap("sync");
  xa.n_rows = h_tot[0]; xa.n_cells = h_tot[1]; xa.n_si = h_tot[2];
  lap("before 2");
  // ---- 2. positions, shared strings, rows, cells
  uint64_t sst_bytes = 0;
  for (uint32_t d = 0; d < n; ++d) if (kind[d] & 1) sst_bytes += c->h_doc_len[d];
  xa.rowpos = c->pool.alloc<uint32_t>((size_t)xa.n_rows + 1);
  xa.row_c0 = c->pool.alloc<uint32_t>((size_t)xa.n_rows + 1);
  xa.rownum = c->pool.alloc<uint32_t>((size_t)xa.n_rows + 1);
  xa.cpos = c->pool.alloc<uint32_t>((size_t)xa.n_cells + 1);
  xa.cell_doc = c->pool.alloc<uint32_t>((size_t)xa.n_cells + 1);
  xa.cell_row = c->pool.alloc<uint32_t>((size_t)xa.n_cells + 1);
  xa.xcells = c->pool.alloc<XCell>((size_t)xa.n_cells + 1);
  xa.num_list = c->pool.alloc<uint2>((size_t)xa.n_cells + 1);
  xa.num_count = c->pool.alloc<uint32_t>(4);
  xa.sipos = c->pool.alloc<uint32_t>((size_t)xa.n_si + 1);
  xa.si_len = c->pool.alloc<uint32_t>((size_t)xa.n_si + 1);
  xa.si_off = c->pool.alloc<uint32_t>((size_t)xa.n_si + 1);
  xa.str_arena = c->pool.alloc<uint8_t>(sst_bytes + 64);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  xml_decode(xa, s);
  XDocShape* h_shape = (XDocShape*)((uint8_t*)c->h_down.p + 256);
  RFQ_CK(c, cudaMemcpyAsync(h_shape, xa.shape, sizeof(XDocShape) * n, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaStreamSynchronize(s));
  lap("sync");
  lap("before 3");
  // ---- 3. table shapes
  std::vector<TDoc> td(n);
  std::vector<uint32_t> item_base(n + 1), slot_r0(n + 1), r2_base(n + 1);
  uint64_t n_cols = 0, n_items = 0, n_slots = 0, n_r2 = 0;
  for (uint32_t d = 0; d < n; ++d) {
    TDoc& T = td[d];
    memset(&T, 0, sizeof(T));
    const XDocShape& sh = h_shape[d];
    T.off = c->h_doc_off[d]; T.len = c->h_doc_len[d];
    T.r0 = (uint32_t)n_slots; T.f0 = (uint32_t)n_r2;
    T.col_base = (uint32_t)n_cols; T.stat_base = (uint32_t)n_cols;
    slot_r0[d] = (uint32_t)n_slots; r2_base[d] = (uint32_t)n_r2; item_base[d] = (uint32_t)n_items;
    T.item_base = (uint32_t)n_items;
    if (kind[d] & 1) { n_slots += 1; continue; }           // sharedStrings part: no output of its own
    n_r2 += (uint64_t)sh.lastrow + 1;                       // every sheet owns its row map (the scatter
                                                            // kernel writes it for all unflagged sheets)
    uint32_t fl = sh.flags;
    if (sst_of[d] >= 0) fl |= h_shape[sst_of[d]].flags;
    if (xu_problem(fl) || sh.W > 1024) {
      if (why) fprintf(stderr, "rfq sheet %u declined: part flags 0x%x width %u\n", d, fl, sh.W);
      c->r_status[d] = RFQ_DOC_UNSUPPORTED; n_slots += 1; continue;
    }
    if (sh.lastrow <= 1) { c->r_status[d] = RFQ_DOC_EMPTY_SHEET; n_slots += 1; continue; }
    T.W = sh.W; T.h = sh.W; T.lead = 0;
    T.n_data = sh.lastrow - 1;
    T.bl = std::max<uint32_t>(T.n_data, 1); T.n_chunks = 1;   // the python engine infers over the whole column
    T.rb = std::min<uint32_t>(256, p2floor(std::max<uint32_t>(1, 2048 / T.W)));
    T.active = 1;
    n_cols += T.W;
    n_items += (T.n_data + T.rb - 1) / T.rb;
    n_slots += (uint64_t)T.n_data + 2;
  }
  item_base[n] = (uint32_t)n_items; slot_r0[n] = (uint32_t)n_slots; r2_base[n] = (uint32_t)n_r2;
  if (n_slots >= (1ull << 31) || n_items >= (1ull << 31)) { c->err = "batch too large"; return RFQ_ERR_ARG; }
  lap("before 4");
  // ---- 4. sheet-row -> <row> element map, header cells
  RFQ_CK(c, c->h_up.reserve(sizeof(TDoc) * (n + 1) + 16ull * (n + 2) + 8192));
  Up up2((uint8_t*)c->h_up.p);
  size_t o_td = up2.put(td.data(), n), o_ib = up2.put(item_base.data(), n + 1);
  size_t o_sr = up2.put(slot_r0.data(), n + 1), o_r2 = up2.put(r2_base.data(), n + 1);
  uint8_t* d_up2 = c->pool.alloc<uint8_t>(up2.used + 256);
  uint32_t* d_r2elem = c->pool.alloc<uint32_t>(n_r2 + 1);
  uint8_t* d_hdr = c->pool.alloc<uint8_t>(n_cols * 256 + 256);
  uint32_t* d_hmeta = c->pool.alloc<uint32_t>(2 * n_cols + 2);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  RFQ_CK(c, cudaMemcpyAsync(d_up2, c->h_up.p, up2.used, cudaMemcpyHostToDevice, s));
  RFQ_CK(c, cudaMemsetAsync(d_r2elem, 0xFF, 4ull * (n_r2 + 1), s));
  xlsx_scatter_rows(xa, (const uint32_t*)(d_up2 + o_r2), d_r2elem, s);
  TableArgs ta{};
  ta.arena = xa.arena;
  ta.docs = (const TDoc*)(d_up2 + o_td); ta.n_docs = n;
  ta.item_doc_base = (const uint32_t*)(d_up2 + o_ib);
  ta.slot_doc_r0 = (const uint32_t*)(d_up2 + o_sr);
  ta.n_items = (uint32_t)n_items; ta.n_slots = (uint32_t)n_slots;
  ta.n_cols = (uint32_t)n_cols; ta.n_stat = (uint32_t)n_cols;
  ta.xcells = xa.xcells; ta.x_row_c0 = xa.row_c0; ta.x_r2elem = d_r2elem; ta.x_str_arena = xa.str_arena;
  prof_mark("xheader_kernel", s);
  xlsx_header_cells(ta, d_hdr, d_hmeta, s);
  prof_mark("host+sync", s);
  RFQ_CK(c, c->h_down.reserve(n_cols * 264 + 4096));
  uint8_t* h_hdr = (uint8_t*)c->h_down.p;
  uint32_t* h_hmeta = (uint32_t*)(h_hdr + n_cols * 256 + 256);
  if (n_cols) {
    RFQ_CK(c, cudaMemcpyAsync(h_hdr, d_hdr, n_cols * 256, cudaMemcpyDeviceToHost, s));
    RFQ_CK(c, cudaMemcpyAsync(h_hmeta, d_hmeta, 8ull * n_cols, cudaMemcpyDeviceToHost, s));
  }
  RFQ_CK(c, cudaStreamSynchronize(s));
  lap("sync");
  lap("before 5");
  // ---- 5. column names (python engine: blank -> "Unnamed: k", duplicates mangled)
  std::vector<uint32_t> name_off(n_cols), name_len(n_cols), name_cp(n_cols);
  std::string names;
  // sheets cut from one template share their header row: labels are built once per distinct row and stored
  // once; a batch of many small sheets spends its host time here, so the per-sheet parts run on host threads
  struct HdrTexts {
    bool ok = true; std::vector<std::string> disp; std::vector<uint32_t> ncp; std::string empty;
    uint32_t empty_off = 0, hdr_extra = 0; std::vector<uint32_t> off;   // where the texts sit in `names`
  };
  // looked up by a 64-bit hash of the header row computed on the threads (the row itself is compared on a hit;
  // two rows with one hash go through the string-keyed map)
  std::unordered_map<uint64_t, std::pair<std::string, HdrTexts>> hdr_fast;
  std::unordered_map<std::string, HdrTexts> hdr_cache;
  // the header row as bytes (type + text of every label), walked without being copied
  auto key_walk = [&](uint32_t d, auto&& sink) {
    const TDoc& T = td[d];
    for (uint32_t j = 0; j < T.W; ++j) {
      const uint32_t g = T.col_base + j, len = h_hmeta[2 * g + 1];
      sink((const char*)&h_hmeta[2 * g], (size_t)8);
      if (len != 0xFFFFFFFFu) sink((const char*)h_hdr + (size_t)g * 256, (size_t)std::min<uint32_t>(len, 256));
      if (h_hmeta[2 * g] == CM_FLOAT) sink((const char*)h_hdr + (size_t)g * 256 + 240, (size_t)8);   // the value itself
    }
  };
  std::vector<uint64_t> hhash(n, 0);
  parallel_for(n, [&](uint32_t d) {
    if (!td[d].active) return;
    uint64_t h = 1469598103934665603ull;                 // FNV-1a
    key_walk(d, [&](const char* p, size_t k) { for (size_t q = 0; q < k; ++q) { h ^= (unsigned char)p[q]; h *= 1099511628211ull; } });
    hhash[d] = h;
  });
  lap("5 header keys (threads)");
  std::vector<const HdrTexts*> hdr_of(n, nullptr);
  for (uint32_t d = 0; d < n; ++d) {
    TDoc& T = td[d];
    if (!T.active) continue;
    {
      auto f = hdr_fast.find(hhash[d]);
      if (f != hdr_fast.end()) {
        const std::string& have = f->second.first;
        size_t pos = 0; bool same = true;
        key_walk(d, [&](const char* p, size_t k) {
          if (same && (pos + k > have.size() || memcmp(have.data() + pos, p, k) != 0)) same = false;
          pos += k;
        });
        if (same && pos == have.size()) { hdr_of[d] = &f->second.second; continue; }
      }
    }
    std::string hkey;
    key_walk(d, [&](const char* p, size_t k) { hkey.append(p, k); });
    auto hit = hdr_cache.find(hkey);
    if (hit == hdr_cache.end()) {
      HdrTexts H;
      std::vector<std::string> raw(T.W);
      std::vector<size_t> unnamed;
      bool all_int = true;
      std::vector<int> lkind(T.W, 0);
      std::vector<std::string> ftext(T.W);          // float labels as an object Index prints them
      std::vector<uint8_t> no_blank(T.W, 0);
      std::vector<double> lval(T.W, 0.0);
      for (uint32_t j = 0; j < T.W; ++j) {
        uint32_t g = T.col_base + j, mode = h_hmeta[2 * g], len = h_hmeta[2 * g + 1];
        if (mode != CM_INT) all_int = false;
        lkind[j] = mode == CM_INT ? 1 : mode == CM_BOOL ? 2 : mode == CM_DATE ? 3 : mode == CM_TIME ? 4
                   : mode == CM_FLOAT ? 5 : mode == CM_TDELTA ? 6 : (mode == CM_RAW || mode == CM_XML || mode == CM_NULL) ? 0 : 9;
        if (mode == CM_NULL) { raw[j] = "Unnamed: " + std::to_string(j); unnamed.push_back(j); }
        else if (lkind[j] == 5 && len != 0xFFFFFFFFu && len <= 200) {
          double v;
          memcpy(&v, h_hdr + (size_t)g * 256 + 240, 8);
          ftext[j].assign((const char*)h_hdr + (size_t)g * 256, len);
          raw[j] = py_float_repr(v);                // the label as a name (de-duplication, "Empty DataFrame")
          lval[j] = v;
        }
        else if (lkind[j] != 9 && lkind[j] != 5 && len != 0xFFFFFFFFu) {
          raw[j].assign((const char*)h_hdr + (size_t)g * 256, len);
          if (lkind[j] == 1) lval[j] = strtod(raw[j].c_str(), nullptr);
          if (lkind[j] == 2) lval[j] = raw[j] == "True" ? 1.0 : 0.0;
        }
        else H.ok = false;  // NaN / oversized labels
      }
      if (H.ok && !sheet_labels_supported(lkind, raw, &lval)) H.ok = false;
      if (H.ok) {
        std::vector<std::string> before = raw;
        dedupe_names(raw, unnamed);
        if (all_int && before == raw) {
          // an all-integer header is an Int64 Index: pandas left-justifies its labels to one width
          size_t mx = 0;
          for (auto& x : raw) mx = std::max(mx, x.size());
          for (auto& x : raw) x.append(mx - x.size(), ' ');
        }
        H.empty = empty_frame_text(raw);
        H.ncp.assign(T.W, 0);
        for (uint32_t j = 0; j < T.W; ++j)
          if (!utf8_valid_count(raw[j], &H.ncp[j])) H.ok = false;
        // labels as to_string prints them: escapes + trim_front (header_names.h display_labels)
        H.disp = raw;
        if (H.ok && !(all_int && before == raw)) {
          // a float label that kept its name prints through the float formatter (a renamed one is a str now)
          std::vector<std::string> shown = raw;
          for (uint32_t j = 0; j < T.W; ++j)
            if (lkind[j] == 5 && raw[j] == before[j]) { shown[j] = ftext[j]; no_blank[j] = ftext[j][0] == '-'; }
          display_labels(shown, H.disp, H.ncp, &no_blank);
        }
      }
      if (H.ok) {
        H.empty_off = (uint32_t)names.size();
        names += H.empty;
        H.off.resize(T.W);
        for (uint32_t j = 0; j < T.W; ++j) {
          H.off[j] = (uint32_t)names.size();
          H.hdr_extra += (uint32_t)H.disp[j].size() - H.ncp[j];
          names += H.disp[j];
        }
      }
      if (!hdr_fast.count(hhash[d])) {       // first row with this hash: the fast map owns it
        auto ins = hdr_fast.emplace(hhash[d], std::make_pair(hkey, std::move(H))).first;
        hdr_of[d] = &ins->second.second;
        continue;
      }
      hit = hdr_cache.emplace(hkey, std::move(H)).first;
    }
    hdr_of[d] = &hit->second;       // (node addresses of an unordered_map are stable)
  }
  lap("5 distinct headers");
  parallel_for(n, [&](uint32_t d) {
    TDoc& T = td[d];
    if (!T.active) return;
    const HdrTexts& H = *hdr_of[d];
    if (!H.ok) {
      if (why) fprintf(stderr, "rfq sheet %u declined: header labels\n", d);
      T.active = 0; c->r_status[d] = RFQ_DOC_UNSUPPORTED; return;
    }
    T.empty_off = H.empty_off; T.empty_len = (uint32_t)H.empty.size();
    T.banner_off = 0; T.banner_len = 0;
    T.hdr_extra += H.hdr_extra;
    for (uint32_t j = 0; j < T.W; ++j) {
      const uint32_t g = T.col_base + j;
      name_off[g] = H.off[j]; name_len[g] = (uint32_t)H.disp[j].size(); name_cp[g] = H.ncp[j];
    }
  });
  // a document that turned unsupported keeps its slots and columns but gets no work items
  {
    uint64_t it = 0;
    for (uint32_t d = 0; d < n; ++d) {
      item_base[d] = (uint32_t)it; td[d].item_base = (uint32_t)it;
      if (td[d].active) it += (td[d].n_data + td[d].rb - 1) / td[d].rb;
    }
    item_base[n] = (uint32_t)it; n_items = it;
  }
  RFQ_CK(c, c->h_up.reserve(sizeof(TDoc) * (n + 1) + 8ull * (n + 2) + 12ull * (n_cols + 1) + names.size() + 8192));
  Up up3((uint8_t*)c->h_up.p);
  o_td = up3.put(td.data(), n); o_ib = up3.put(item_base.data(), n + 1);
  size_t o_no = up3.put(name_off.data(), n_cols), o_nl = up3.put(name_len.data(), n_cols);
  size_t o_nc = up3.put(name_cp.data(), n_cols), o_nm = up3.put(names.data(), names.size());
  uint8_t* d_up3 = c->pool.alloc<uint8_t>(up3.used + 256);
  ta.n_items = (uint32_t)n_items;
  ta.stat = c->pool.alloc<ChunkStat>(n_cols + 1);
  ta.meas = c->pool.alloc<ChunkMeasure>(n_cols + 1);
  ta.ctype = c->pool.alloc<uint8_t>(n_cols + 1);
  ta.chunk_nf = c->pool.alloc<uint8_t>(n_cols + 1);
  ta.row_na = c->pool.alloc<uint32_t>(n_slots + 1);
  ta.row_extra = c->pool.alloc<uint32_t>(n_slots + 1);
  ta.row_bytes = c->pool.alloc<unsigned long long>(n_slots + 2);
  ta.row_off = c->pool.alloc<unsigned long long>(n_slots + 2);
  ta.n_keep = c->pool.alloc<uint32_t>(n + 1);
  ta.doc_flags = c->pool.alloc<int32_t>(n + 1);
  ta.doc_linew = c->pool.alloc<uint32_t>(n + 1);
  ta.col = c->pool.alloc<ColOut>(n_cols + 1);
  ta.x_colflags = c->pool.alloc<uint32_t>(n_cols + 1);
  ta.scan_tmp64 = c->pool.alloc<unsigned long long>(scan_partials_needed(n_slots + 1));
  ta.doc_out_off = c->pool.alloc<unsigned long long>(n + 1);
  ta.doc_out_len = c->pool.alloc<unsigned long long>(n + 1);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  RFQ_CK(c, cudaMemcpyAsync(d_up3, c->h_up.p, up3.used, cudaMemcpyHostToDevice, s));
  ta.docs = (const TDoc*)(d_up3 + o_td);
  ta.item_doc_base = (const uint32_t*)(d_up3 + o_ib);
  ta.name_off = (const uint32_t*)(d_up3 + o_no); ta.name_len = (const uint32_t*)(d_up3 + o_nl);
  ta.name_cp = (const uint32_t*)(d_up3 + o_nc); ta.names = d_up3 + o_nm;
  RFQ_CK(c, cudaMemsetAsync(ta.row_na, 0, 4ull * (n_slots + 1), s));
  RFQ_CK(c, cudaMemsetAsync(ta.row_extra, 0, 4ull * (n_slots + 1), s));
  RFQ_CK(c, cudaMemsetAsync(ta.chunk_nf, 0, n_cols + 1, s));
  RFQ_CK(c, cudaMemsetAsync(ta.ctype, 0, n_cols + 1, s));
  RFQ_CK(c, cudaMemsetAsync(ta.n_keep, 0, 4ull * (n + 1), s));
  RFQ_CK(c, cudaMemsetAsync(ta.doc_flags, 0, 4ull * (n + 1), s));
  RFQ_CK(c, cudaMemsetAsync(ta.doc_linew, 0, 4ull * (n + 1), s));
  RFQ_CK(c, cudaMemsetAsync(ta.x_colflags, 0, 4ull * (n_cols + 1), s));
  lap("before 6");
  // ---- 6. infer, measure, lay out, emit (shared with the CSV path)
  table_stat_init(ta, s);
  xlsx_table_classify(ta, s);
  csv_table_measure(ta, s);
  uint8_t* hd = (uint8_t*)c->h_down.p;
  RFQ_CK(c, c->h_down.reserve(32ull * (n + 2) + 4096));
  hd = (uint8_t*)c->h_down.p;
  unsigned long long* h_total = (unsigned long long*)hd;
  unsigned long long* h_ooff = (unsigned long long*)(hd + 256);
  unsigned long long* h_olen = h_ooff + (n + 1);
  uint32_t* h_keep = (uint32_t*)(h_olen + (n + 1));
  int32_t* h_flags = (int32_t*)(h_keep + (n + 1));
  RFQ_CK(c, cudaMemcpyAsync(h_total, ta.row_off + n_slots, 8, cudaMemcpyDeviceToHost, s));
  if (n) {
    RFQ_CK(c, cudaMemcpyAsync(h_ooff, ta.doc_out_off, 8ull * n, cudaMemcpyDeviceToHost, s));
    RFQ_CK(c, cudaMemcpyAsync(h_olen, ta.doc_out_len, 8ull * n, cudaMemcpyDeviceToHost, s));
    RFQ_CK(c, cudaMemcpyAsync(h_keep, ta.n_keep, 4ull * n, cudaMemcpyDeviceToHost, s));
    RFQ_CK(c, cudaMemcpyAsync(h_flags, ta.doc_flags, 4ull * n, cudaMemcpyDeviceToHost, s));
  }
  RFQ_CK(c, cudaStreamSynchronize(s));
  lap("sync");
  const uint64_t total = *h_total;
  uint8_t* d_text = c->pool.alloc<uint8_t>(total + 64);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  if (total) csv_table_emit(ta, d_text, s);
  prof_mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  RFQ_CK(c, cudaGetLastError());
  c->segs.push_back({d_text, total, 0});
  c->out_total = total;
  for (uint32_t d = 0; d < n; ++d) {
    if (!td[d].active) continue;
    if (h_flags[d]) {
      if (why) fprintf(stderr, "rfq sheet %u declined: column flags 0x%x\n", d, h_flags[d]);
      c->r_status[d] = RFQ_DOC_UNSUPPORTED; continue;
    }
    c->r_off[d] = h_ooff[d]; c->r_len[d] = h_olen[d];
    c->r_rows[d] = h_keep[d]; c->r_cols[d] = td[d].W;
  }
  c->d_out_off = nullptr; c->d_out_len = nullptr;
  c->have_output = true;
  c->timed = true;
  return RFQ_OK;
}

extern "C" int rfq_sheets_parse(rfq_ctx* c, const uint8_t* arena, const uint64_t* offsets,
                                const uint64_t* lengths, uint32_t n, const uint8_t* kind,
                                const int32_t* sst_of, const int64_t* max_rows,
                                const uint32_t* date_off, const uint32_t* date_ids) {
  int r = rfq_stage(c, arena, offsets, lengths, n);
  if (r) return r;
  r = rfq_sheets_run(c, kind, sst_of, max_rows, date_off, date_ids);
  if (r) return r;
  return rfq_fetch(c);
}
