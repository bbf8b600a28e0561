// rfq_mixed_run: one staged batch that mixes raw text, CSV and sheet-XML documents (the
// reference dispatches on the file suffix, app/file_parser.py:62-77).  The arena is uploaded once; the
// text, CSV and sheet pipelines then run over their own subsets of it (views of the document tables, no
// second copy of any byte), and their results are merged into one per-document table in input order.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <chrono>
#include <algorithm>
#include <thread>
#include <vector>

#include "../../include/rfq_parse.h"
#include "ctx.h"

using namespace rfq;

namespace {
struct Saved {
  uint32_t n_docs, n_tiles;
  std::vector<uint32_t> off, len, tb;
  uint32_t *d_off, *d_len, *d_tb;
};
}

extern "C" int rfq_mixed_run(rfq_ctx* c, const uint8_t* kind, const int64_t* max_rows, const int32_t* sst_of,
                             const uint32_t* date_off, const uint32_t* date_ids) {
  if (!c || (c->n_docs && !kind)) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_mixed_run: nothing staged"; return RFQ_ERR_STATE; }
  if (c->nested) { c->err = "rfq_mixed_run: already inside a mixed run"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  const uint32_t n = c->n_docs;
  c->prof.begin();
  RFQ_CK(c, cudaEventRecord(c->ev0, s));
  c->pool.reset();
  c->status_override.clear(); c->out_n = 0;
  static const bool host_prof = getenv("RFQ_HOST_PROFILE") != nullptr;   // host time of every stage on stderr
  auto t_prev = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!host_prof) return;
    const auto t = std::chrono::steady_clock::now();
    fprintf(stderr, "rfq_mixed_run host %-28s %8.3f ms\n", what, std::chrono::duration<double, std::milli>(t - t_prev).count());
    t_prev = t;
  };
  // ---- the three subsets, in input order
  std::vector<uint32_t> idx[3];          // 0 text, 1 csv, 2 sheets + sharedStrings
  std::vector<int32_t> pos_in_sheets(n, -1);
  for (uint32_t d = 0; d < n; ++d) {
    switch (kind[d]) {
      case RFQ_KIND_TEXT: idx[0].push_back(d); break;
      case RFQ_KIND_CSV: idx[1].push_back(d); break;
      case RFQ_KIND_SHEET: case RFQ_KIND_SHEET_1904: case RFQ_KIND_SST:
        pos_in_sheets[d] = (int32_t)idx[2].size(); idx[2].push_back(d); break;
      default: c->err = "rfq_mixed_run: unknown document kind"; return RFQ_ERR_ARG;
    }
  }
  Saved full{c->n_docs, c->n_tiles, c->h_doc_off, c->h_doc_len, c->h_tile_base, c->d_doc_off, c->d_doc_len, c->d_tile_base};
  // device copies of the three views (one upload)
  size_t words = 0;
  for (int k = 0; k < 3; ++k) words += 3 * idx[k].size() + 1 + 64;
  RFQ_CK(c, c->d_viewtab.reserve(4 * words + 1024));
  RFQ_CK(c, c->h_viewtab.reserve(4 * words + 1024));
  uint32_t* hv = (uint32_t*)c->h_viewtab.p;
  uint32_t* dv = (uint32_t*)c->d_viewtab.p;
  size_t base[3];
  uint32_t view_tiles[3];
  {
    size_t w = 0;
    for (int k = 0; k < 3; ++k) {
      const uint32_t m = (uint32_t)idx[k].size();
      base[k] = w;
      uint64_t tiles = 0;
      for (uint32_t i = 0; i < m; ++i) {
        hv[w + i] = full.off[idx[k][i]];
        hv[w + m + i] = full.len[idx[k][i]];
        hv[w + 2 * m + i] = (uint32_t)tiles;
        tiles += ((uint64_t)full.len[idx[k][i]] + RFQ_TILE - 1) / RFQ_TILE;
      }
      hv[w + 3 * m] = (uint32_t)tiles;
      view_tiles[k] = (uint32_t)tiles;
      w += 3 * m + 1;
      w = (w + 63) & ~(size_t)63;
    }
    RFQ_CK(c, cudaMemcpyAsync(dv, hv, 4 * w, cudaMemcpyHostToDevice, s));
  }
  // merged results
  std::vector<uint64_t> m_off(n, 0), m_len(n, 0), m_chars(n, ~0ull);
  std::vector<int32_t> m_status(n, RFQ_DOC_OK);
  std::vector<uint32_t> m_rows(n, 0), m_cols(n, 0);
  std::vector<uint8_t> m_attempt(n, 0xFF);
  std::vector<rfq_ctx::OutSeg> m_segs;
  uint64_t m_total = 0;
  int rc = RFQ_OK;
  c->nested = true;
  lap("views");
  // ---- the sheet documents on a second context, beside the CSV run: the sheet path is a chain of small
  // kernels and host stages between synchronisations, the CSV path one long run of kernels, and neither
  // needs anything of the other but the staged arena
  static const bool serial = getenv("RFQ_MIXED_SERIAL") != nullptr;
  const bool side = !serial && !idx[1].empty() && idx[2].size() >= 256;
  std::vector<uint8_t> sk; std::vector<int32_t> so; std::vector<int64_t> smr; std::vector<uint32_t> sdoff, sdids;
  int rc_side = RFQ_OK;
  std::thread side_thread;
  rfq_ctx* h = nullptr;
  struct Joiner {            // an early return (a CUDA error on this thread) still waits for the other one
    std::thread& t; rfq_ctx*& hh;
    ~Joiner() { if (t.joinable()) t.join(); if (hh) { hh->d_arena.p = nullptr; hh->d_arena.cap = 0; } }
  } joiner{side_thread, h};
  auto sheet_args = [&]() -> int {
    const uint32_t m = (uint32_t)idx[2].size();
    sk.assign(m, 0); so.assign(m, -1); smr.assign(m, 1000); sdoff.assign(m + 1, 0); sdids.clear();
    for (uint32_t i = 0; i < m; ++i) {
      const uint32_t d = idx[2][i];
      sk[i] = kind[d] == RFQ_KIND_SST ? 1 : kind[d] == RFQ_KIND_SHEET_1904 ? 2 : 0;
      if (sk[i] != 1 && sst_of && sst_of[d] >= 0) {
        if ((uint32_t)sst_of[d] >= n || pos_in_sheets[sst_of[d]] < 0) { c->err = "rfq_mixed_run: sst_of must point at a sharedStrings document"; return RFQ_ERR_ARG; }
        so[i] = pos_in_sheets[sst_of[d]];
      }
      if (max_rows) smr[i] = max_rows[d];
      if (date_off && date_ids) sdids.insert(sdids.end(), date_ids + date_off[d], date_ids + date_off[d + 1]);
      sdoff[i + 1] = (uint32_t)sdids.size();
    }
    if (sdids.empty()) sdids.push_back(0);
    return RFQ_OK;
  };
  if (side) {
    if (!c->helper) {
      if (rfq_create(c->device, &c->helper) != RFQ_OK) { c->nested = false; c->err = "rfq_mixed_run: second context"; return RFQ_ERR_CUDA; }
      cudaEventCreateWithFlags(&c->ev_staged, cudaEventDisableTiming);
      cudaEventCreateWithFlags(&c->ev_helper, cudaEventDisableTiming);
    }
    h = c->helper;
    rc = sheet_args();
    if (rc != RFQ_OK) { c->nested = false; return rc; }
    const uint32_t m = (uint32_t)idx[2].size();
    cudaEventRecord(c->ev_staged, s);                       // the arena and the views are on the device after this
    cudaStreamWaitEvent(h->stream, c->ev_staged, 0);
    h->staged = true; h->nested = false;
    h->n_docs = m; h->n_tiles = view_tiles[2]; h->arena_bytes = c->arena_bytes; h->h_arena = c->h_arena;
    h->h_doc_off.assign(hv + base[2], hv + base[2] + m);
    h->h_doc_len.assign(hv + base[2] + m, hv + base[2] + 2 * m);
    h->h_tile_base.assign(hv + base[2] + 2 * m, hv + base[2] + 3 * m + 1);
    h->d_doc_off = dv + base[2]; h->d_doc_len = dv + base[2] + m; h->d_tile_base = dv + base[2] + 2 * m;
    h->d_arena.p = c->d_arena.p; h->d_arena.cap = 0;       // borrowed
    side_thread = std::thread([&, h]() {
      rc_side = rfq_sheets_run(h, sk.data(), so.data(), smr.data(), sdoff.data(), sdids.data());
      if (rc_side == RFQ_OK) cudaEventRecord(c->ev_helper, h->stream);
    });
  }
  for (int k = 0; k < 3 && rc == RFQ_OK; ++k) {
    const uint32_t m = (uint32_t)idx[k].size();
    if (m == 0 || (k == 2 && side)) continue;
    // the view: same arena, this kind's documents only
    c->n_docs = m; c->n_tiles = view_tiles[k];
    c->h_doc_off.assign(hv + base[k], hv + base[k] + m);
    c->h_doc_len.assign(hv + base[k] + m, hv + base[k] + 2 * m);
    c->h_tile_base.assign(hv + base[k] + 2 * m, hv + base[k] + 3 * m + 1);
    c->d_doc_off = dv + base[k]; c->d_doc_len = dv + base[k] + m; c->d_tile_base = dv + base[k] + 2 * m;
    c->segs.clear(); c->out_total = 0;
    if (k == 0) {
      rc = rfq_text_run(c);
      if (rc == RFQ_OK) {   // the text run keeps its per-document table on the device
        RFQ_CK(c, c->h_down.reserve(16ull * (m + 1) + 4096));
        uint64_t* ho = (uint64_t*)c->h_down.p;
        RFQ_CK(c, cudaMemcpyAsync(ho, c->d_out_off, 8ull * m, cudaMemcpyDeviceToHost, s));
        RFQ_CK(c, cudaMemcpyAsync(ho + m, c->d_out_len, 8ull * m, cudaMemcpyDeviceToHost, s));
        RFQ_CK(c, cudaStreamSynchronize(s));
        for (uint32_t i = 0; i < m; ++i) { m_off[idx[0][i]] = m_total + ho[i]; m_len[idx[0][i]] = ho[m + i]; }
      }
    } else if (k == 1) {
      std::vector<int64_t> mr(m, 1000);
      if (max_rows) for (uint32_t i = 0; i < m; ++i) mr[i] = max_rows[idx[1][i]];
      rc = rfq_csv_run(c, mr.data());
      if (rc == RFQ_OK)
        for (uint32_t i = 0; i < m; ++i) {
          const uint32_t d = idx[1][i];
          m_off[d] = m_total + c->r_off[i]; m_len[d] = c->r_len[i]; m_status[d] = c->r_status[i];
          m_rows[d] = c->r_rows[i]; m_cols[d] = c->r_cols[i]; m_attempt[d] = c->r_attempt[i]; m_chars[d] = c->r_chars[i];
        }
    } else {
      rc = sheet_args();
      if (rc == RFQ_OK) rc = rfq_sheets_run(c, sk.data(), so.data(), smr.data(), sdoff.data(), sdids.data());
      if (rc == RFQ_OK)
        for (uint32_t i = 0; i < m; ++i) {
          const uint32_t d = idx[2][i];
          m_off[d] = m_total + c->r_off[i]; m_len[d] = c->r_len[i]; m_status[d] = c->r_status[i];
          m_rows[d] = c->r_rows[i]; m_cols[d] = c->r_cols[i];
        }
    }
    if (rc == RFQ_OK) {
      for (auto sg : c->segs) { sg.host_base += m_total; sg.copied = false; m_segs.push_back(sg); }
      m_total += c->out_total;
    }
    lap(k == 0 ? "text run + merge" : k == 1 ? "csv run + merge" : "sheets run + merge");
  }
  if (side) {
    side_thread.join();
    h->d_arena.p = nullptr; h->d_arena.cap = 0;
    if (rc == RFQ_OK && rc_side != RFQ_OK) { rc = rc_side; c->err = h->err; }
    if (rc == RFQ_OK) {
      cudaStreamWaitEvent(s, c->ev_helper, 0);             // this stream's later copies read what the helper wrote
      const uint32_t m = (uint32_t)idx[2].size();
      for (uint32_t i = 0; i < m; ++i) {
        const uint32_t d = idx[2][i];
        m_off[d] = m_total + h->r_off[i]; m_len[d] = h->r_len[i]; m_status[d] = h->r_status[i];
        m_rows[d] = h->r_rows[i]; m_cols[d] = h->r_cols[i];
      }
      for (auto sg : h->segs) { sg.host_base += m_total; sg.copied = false; m_segs.push_back(sg); }
      m_total += h->out_total;
    }
    lap("sheets on the second context: join + merge");
  }
  c->nested = false;
  // ---- back to the full batch
  c->n_docs = full.n_docs; c->n_tiles = full.n_tiles;
  c->h_doc_off = full.off; c->h_doc_len = full.len; c->h_tile_base = full.tb;
  c->d_doc_off = full.d_off; c->d_doc_len = full.d_len; c->d_tile_base = full.d_tb;
  if (rc != RFQ_OK) { c->have_output = false; return rc; }
  c->segs = m_segs;
  c->out_total = m_total;
  c->host_tables = true;
  c->r_off = m_off; c->r_len = m_len; c->r_status = m_status; c->r_rows = m_rows; c->r_cols = m_cols;
  c->r_attempt = m_attempt; c->r_chars = m_chars;
  c->d_out_off = nullptr; c->d_out_len = nullptr; c->d_status = nullptr; c->d_rows = nullptr; c->d_cols = nullptr; c->d_attempt = nullptr;
  c->prof.mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  c->have_output = true;
  c->timed = true;
  lap("restore + tables");
  return RFQ_OK;
}

extern "C" int rfq_mixed_parse(rfq_ctx* c, const uint8_t* arena, const uint64_t* offsets, const uint64_t* lengths,
                               uint32_t n, const uint8_t* kind, const int64_t* max_rows, const int32_t* sst_of,
                               const uint32_t* date_off, const uint32_t* date_ids) {
  int r = rfq_stage(c, arena, offsets, lengths, n);
  if (r) return r;
  r = rfq_mixed_run(c, kind, max_rows, sst_of, date_off, date_ids);
  if (r) return r;
  return rfq_fetch(c);
}
