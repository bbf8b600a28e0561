// C-ABI layer (include/rfq_parse.h): context, device/pinned memory pools, staging, fetch.
// Host-side orchestration only; every byte of parsing happens in the kernels.
#include <cuda_runtime.h>
#include <sys/mman.h>
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <atomic>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rfq_parse.h"
#include "ctx.h"

namespace rfq {
std::atomic<uint64_t> g_launch_count{0};
thread_local Prof* g_prof = nullptr;
}

using namespace rfq;

extern "C" {

const char* rfq_version(void) { return "rfq-parse-b200 0.1 (sm_100a)"; }

int rfq_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int rfq_create(int device, rfq_ctx** out) {
  if (!out) return RFQ_ERR_ARG;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0 || device < 0 || device >= n) {
    cudaGetLastError();
    fprintf(stderr, "rfq_create: no usable CUDA device %d (%s); there is no CPU fallback\n",
            device, e != cudaSuccess ? cudaGetErrorString(e) : "bad index");
    return RFQ_ERR_CUDA;
  }
  rfq_ctx* c = new rfq_ctx();
  c->device = device;
  if (cudaSetDevice(device) != cudaSuccess ||
      cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&c->ev0) != cudaSuccess || cudaEventCreate(&c->ev1) != cudaSuccess) {
    cudaGetLastError();
    delete c;
    return RFQ_ERR_CUDA;
  }
  c->stream = c->own_stream;
  cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
  for (auto& e : c->ev_slice) cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
  *out = c;
  return RFQ_OK;
}

void rfq_destroy(rfq_ctx* c) {
  if (!c) return;
  if (c->helper) { c->helper->d_arena.p = nullptr; c->helper->d_arena.cap = 0; rfq_destroy(c->helper); c->helper = nullptr; }
  cudaSetDevice(c->device);
  if (c->ev_staged) cudaEventDestroy(c->ev_staged);
  if (c->ev_helper) cudaEventDestroy(c->ev_helper);
  cudaStreamSynchronize(c->stream);
  c->release_all();
  if (c->ev0) cudaEventDestroy(c->ev0);
  if (c->ev1) cudaEventDestroy(c->ev1);
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
  for (auto e : c->ev_slice) if (e) cudaEventDestroy(e);
  delete c;
}

const char* rfq_last_error(const rfq_ctx* c) { return c ? c->err.c_str() : "null ctx"; }

int rfq_set_stream(rfq_ctx* c, void* s) {
  if (!c) return RFQ_ERR_ARG;
  c->stream = s ? (cudaStream_t)s : c->own_stream;
  return RFQ_OK;
}
void* rfq_get_stream(rfq_ctx* c) { return c ? (void*)c->stream : nullptr; }

void* rfq_host_alloc(rfq_ctx* c, uint64_t bytes) {
  if (!c) return nullptr;
  cudaSetDevice(c->device);
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
    c->err = "cudaHostAlloc failed";
    cudaGetLastError();
    return nullptr;
  }
  return p;
}
void rfq_host_free(rfq_ctx* c, void* p) {
  if (c && p) { cudaSetDevice(c->device); cudaFreeHost(p); }
}

int rfq_stage(rfq_ctx* c, const uint8_t* arena, const uint64_t* offsets, const uint64_t* lengths,
              uint32_t n) {
  if (!c || (n && (!arena || !offsets || !lengths))) return RFQ_ERR_ARG;
  RFQ_CK(c, cudaSetDevice(c->device));
  RFQ_CK(c, cudaStreamSynchronize(c->stream));  // previous run must be done before buffers move
  c->n_docs = n;
  c->staged = false;
  c->have_output = false;
  c->status_override.clear(); c->out_n = 0; c->raw_bytes.clear(); c->raw_chars.clear();
  bool aligned = true;
  uint64_t span = 0;
  for (uint32_t d = 0; d < n; ++d) {
    if (offsets[d] % RFQ_ALIGN) aligned = false;
    if (d && offsets[d] < offsets[d - 1] + lengths[d - 1]) aligned = false;  // overlapping/unordered
    span = std::max<uint64_t>(span, offsets[d] + lengths[d]);
  }
  c->h_doc_off.resize(n); c->h_doc_len.resize(n); c->h_tile_base.resize(n + 1);
  uint64_t total = 0, tiles = 0;
  if (aligned) {
    total = (span + RFQ_ALIGN - 1) / RFQ_ALIGN * RFQ_ALIGN;
    for (uint32_t d = 0; d < n; ++d) {
      c->h_doc_off[d] = (uint32_t)offsets[d];
      c->h_doc_len[d] = (uint32_t)lengths[d];
    }
  } else {
    for (uint32_t d = 0; d < n; ++d) {
      c->h_doc_off[d] = (uint32_t)total;
      c->h_doc_len[d] = (uint32_t)lengths[d];
      total += (lengths[d] + RFQ_ALIGN - 1) / RFQ_ALIGN * RFQ_ALIGN;
    }
  }
  if (total + 64 >= (1ull << 32)) {
    c->err = "batch arena must be < 4 GiB (split the batch)";
    return RFQ_ERR_ARG;
  }
  for (uint32_t d = 0; d < n; ++d) {
    c->h_tile_base[d] = (uint32_t)tiles;
    tiles += (lengths[d] + RFQ_TILE - 1) / RFQ_TILE;
  }
  c->h_tile_base[n] = (uint32_t)tiles;
  c->n_tiles = (uint32_t)tiles;
  c->arena_bytes = total;
  // device arena with 64 bytes of zeroed slack so vector loads past a document's end are safe
  RFQ_CK(c, c->d_arena.reserve(total + 64));
  RFQ_CK(c, cudaMemsetAsync((uint8_t*)c->d_arena.p + total, 0, 64, c->stream));
  if (aligned) {
    c->h_arena = arena;
    if (total) {
      uint64_t ncopy = std::min<uint64_t>(span, total);
      RFQ_CK(c, cudaMemcpyAsync(c->d_arena.p, arena, ncopy, cudaMemcpyHostToDevice, c->stream));
      if (ncopy < total)
        RFQ_CK(c, cudaMemsetAsync((uint8_t*)c->d_arena.p + ncopy, 0, total - ncopy, c->stream));
    }
  } else {
    RFQ_CK(c, c->h_stage.reserve(total + 64));
    uint8_t* st = (uint8_t*)c->h_stage.p;
    for (uint32_t d = 0; d < n; ++d) {
      memcpy(st + c->h_doc_off[d], arena + offsets[d], lengths[d]);
      uint64_t pad = (RFQ_ALIGN - lengths[d] % RFQ_ALIGN) % RFQ_ALIGN;
      memset(st + c->h_doc_off[d] + lengths[d], 0, pad);
    }
    c->h_arena = st;
    if (total)
      RFQ_CK(c, cudaMemcpyAsync(c->d_arena.p, st, total, cudaMemcpyHostToDevice, c->stream));
  }
  c->h2d_bytes = total;
  // document tables
  size_t tab = sizeof(uint32_t) * (3ull * n + 1);
  RFQ_CK(c, c->d_doctab.reserve(tab));
  RFQ_CK(c, c->h_doctab.reserve(tab));
  uint32_t* ht = (uint32_t*)c->h_doctab.p;
  if (n) {
    memcpy(ht, c->h_doc_off.data(), 4ull * n);
    memcpy(ht + n, c->h_doc_len.data(), 4ull * n);
  }
  memcpy(ht + 2ull * n, c->h_tile_base.data(), 4ull * (n + 1));
  RFQ_CK(c, cudaMemcpyAsync(c->d_doctab.p, ht, tab, cudaMemcpyHostToDevice, c->stream));
  c->d_doc_off = (uint32_t*)c->d_doctab.p;
  c->d_doc_len = c->d_doc_off + n;
  c->d_tile_base = c->d_doc_len + n;
  c->h2d_bytes += tab;
  c->staged = true;
  return RFQ_OK;
}

int rfq_text_run(rfq_ctx* c) {
  if (!c) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_text_run: nothing staged"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  uint32_t n = c->n_docs, nt = c->n_tiles;
  if (!c->nested) c->prof.begin();
  if (!c->nested) { RFQ_CK(c, cudaEventRecord(c->ev0, s)); c->pool.reset(); }
  c->status_override.clear(); c->out_n = 0;
  TxtArgs a{};
  a.arena = (const uint8_t*)c->d_arena.p;
  a.doc_off = c->d_doc_off; a.doc_len = c->d_doc_len; a.tile_base = c->d_tile_base;
  a.n_docs = n; a.n_tiles = nt;
  a.tile_cnt = c->pool.alloc<uint32_t>(nt + 1);
  a.tile_pre = c->pool.alloc<uint32_t>(nt + 1);
  a.lane_emit = c->pool.alloc<uint16_t>(32ull * nt + 32);
  a.first_key = c->pool.alloc<unsigned long long>(n + 1);
  a.last_key = c->pool.alloc<unsigned long long>(n + 1);
  a.out_len = c->pool.alloc<unsigned long long>(n + 1);
  a.out_off = c->pool.alloc<unsigned long long>(n + 1);
  a.out_total = c->pool.alloc<unsigned long long>(1);
  a.skip = c->pool.alloc<uint32_t>(n + 1);
  a.scan_tmp32 = c->pool.alloc<uint32_t>(scan_partials_needed(nt));
  a.scan_tmp64 = c->pool.alloc<unsigned long long>(scan_partials_needed(n));
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  txt_pipeline(a, s);
  unsigned long long total = 0;
  if (n) {
    RFQ_CK(c, cudaMemcpyAsync(c->h_scalar(), a.out_total, 8, cudaMemcpyDeviceToHost, s));
    RFQ_CK(c, cudaStreamSynchronize(s));
    total = *c->h_scalar();
  }
  uint8_t* d_text = c->pool.alloc<uint8_t>(total + 16);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  c->prof.mark("txt_emit", s);
  txt_emit(a, d_text, s);
  c->segs.clear();
  c->segs.push_back({d_text, total, 0});
  c->host_tables = false;
  c->prof.mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  RFQ_CK(c, cudaGetLastError());
  c->out_total = total;
  c->d_out_off = a.out_off; c->d_out_len = a.out_len;
  c->d_status = nullptr; c->d_rows = nullptr; c->d_cols = nullptr; c->d_attempt = nullptr;
  c->have_output = true;
  c->timed = true;
  return RFQ_OK;
}

int rfq_md5_run(rfq_ctx* c) {
  if (!c) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_md5_run: nothing staged"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  uint32_t n = c->n_docs;
  if (!c->nested) c->prof.begin();
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  if (!c->nested) { RFQ_CK(c, cudaEventRecord(c->ev0, s)); c->pool.reset(); }
  c->status_override.clear(); c->out_n = 0;
  unsigned long long* out_off = c->pool.alloc<unsigned long long>(n + 1);
  unsigned long long* out_len = c->pool.alloc<unsigned long long>(n + 1);
  uint8_t* d_text = c->pool.alloc<uint8_t>(32ull * n + 16);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  md5_batch((const uint8_t*)c->d_arena.p, c->d_doc_off, c->d_doc_len, n, d_text, out_off, out_len, s);
  c->segs.clear();
  c->segs.push_back({d_text, 32ull * n, 0});
  c->host_tables = false;
  prof_mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  RFQ_CK(c, cudaGetLastError());
  c->out_total = 32ull * n;
  c->d_out_off = out_off; c->d_out_len = out_len;
  c->d_status = nullptr; c->d_rows = nullptr; c->d_cols = nullptr; c->d_attempt = nullptr;
  c->have_output = true;
  c->timed = true;
  return RFQ_OK;
}

int rfq_inflate_run(rfq_ctx* c, const uint8_t* method, const uint64_t* out_len, const uint32_t* crc32) {
  if (!c || (c->n_docs && (!method || !out_len || !crc32))) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_inflate_run: nothing staged"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  const uint32_t n = c->n_docs;
  if (!c->nested) c->prof.begin();
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  if (!c->nested) { RFQ_CK(c, cudaEventRecord(c->ev0, s)); c->pool.reset(); }
  c->status_override.clear(); c->out_n = 0;
  c->inflated_valid = false;
  c->inf_off.assign(n, 0); c->inf_len.assign(out_len, out_len + n);
  uint64_t total = 0;
  for (uint32_t d = 0; d < n; ++d) { c->inf_off[d] = total; total += (out_len[d] + RFQ_ALIGN - 1) / RFQ_ALIGN * RFQ_ALIGN; }
  if (total + 64 >= (1ull << 32)) { c->err = "inflated batch must be < 4 GiB (split the batch)"; return RFQ_ERR_ARG; }
  c->inf_total = total;
  RFQ_CK(c, c->d_inflated.reserve(total + 64));
  RFQ_CK(c, cudaMemsetAsync(c->d_inflated.p, 0, total + 64, s));
  RFQ_CK(c, c->h_up.reserve(21ull * (n + 2) + 4096));
  uint8_t* hu = (uint8_t*)c->h_up.p;
  const size_t o_off = 0, o_len = (8ull * n + 255) & ~255ull, o_crc = (o_len + 8ull * n + 255) & ~255ull,
               o_m = (o_crc + 4ull * n + 255) & ~255ull, used = o_m + n;
  RFQ_CK(c, c->h_up.reserve(used + 256));
  hu = (uint8_t*)c->h_up.p;
  if (n) {
    memcpy(hu + o_off, c->inf_off.data(), 8ull * n);
    memcpy(hu + o_len, out_len, 8ull * n);
    memcpy(hu + o_crc, crc32, 4ull * n);
    memcpy(hu + o_m, method, n);
  }
  uint8_t* d_up = c->pool.alloc<uint8_t>(used + 256);
  int32_t* d_status = c->pool.alloc<int32_t>(n + 1);
  uint32_t* d_ticket = c->pool.alloc<uint32_t>(4);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  RFQ_CK(c, cudaMemcpyAsync(d_up, hu, used, cudaMemcpyHostToDevice, s));
  inflate_batch((const uint8_t*)c->d_arena.p, c->d_doc_off, c->d_doc_len, d_up + o_m,
                (const unsigned long long*)(d_up + o_off), (const unsigned long long*)(d_up + o_len),
                (const uint32_t*)(d_up + o_crc), n, (uint8_t*)c->d_inflated.p, d_status, d_ticket, s);
  c->segs.clear();
  c->segs.push_back({(const uint8_t*)c->d_inflated.p, total, 0});
  c->host_tables = false;
  prof_mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  RFQ_CK(c, cudaGetLastError());
  c->out_total = total;
  c->d_out_off = (unsigned long long*)(d_up + o_off); c->d_out_len = (unsigned long long*)(d_up + o_len);
  c->d_status = d_status; c->d_rows = nullptr; c->d_cols = nullptr; c->d_attempt = nullptr;
  c->have_output = true;
  c->timed = true;
  c->inflated_valid = true;
  return RFQ_OK;
}

int rfq_restage_output(rfq_ctx* c) {
  if (!c) return RFQ_ERR_ARG;
  if (!c->inflated_valid || !c->have_output) { c->err = "rfq_restage_output: the last run was not rfq_inflate_run"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  const uint32_t n = c->n_docs;
  // statuses decide which members carry text; the inflate itself must be finished before the buffers swap roles
  RFQ_CK(c, c->h_down.reserve(4ull * (n + 1) + 256));
  int32_t* hs = (int32_t*)c->h_down.p;
  if (n) RFQ_CK(c, cudaMemcpyAsync(hs, c->d_status, 4ull * n, cudaMemcpyDeviceToHost, s));
  RFQ_CK(c, cudaStreamSynchronize(s));
  std::swap(c->d_arena, c->d_inflated);
  uint64_t tiles = 0;
  c->h_doc_off.resize(n); c->h_doc_len.resize(n); c->h_tile_base.resize(n + 1);
  for (uint32_t d = 0; d < n; ++d) {
    c->h_doc_off[d] = (uint32_t)c->inf_off[d];
    c->h_doc_len[d] = hs[d] == RFQ_DOC_OK ? (uint32_t)c->inf_len[d] : 0u;
    c->h_tile_base[d] = (uint32_t)tiles;
    tiles += ((uint64_t)c->h_doc_len[d] + RFQ_TILE - 1) / RFQ_TILE;
  }
  c->h_tile_base[n] = (uint32_t)tiles;
  c->n_tiles = (uint32_t)tiles;
  c->arena_bytes = c->inf_total;
  c->h_arena = nullptr;            // the bytes exist on the device only
  size_t tab = sizeof(uint32_t) * (3ull * n + 1);
  RFQ_CK(c, c->d_doctab.reserve(tab));
  RFQ_CK(c, c->h_doctab.reserve(tab));
  uint32_t* ht = (uint32_t*)c->h_doctab.p;
  if (n) { memcpy(ht, c->h_doc_off.data(), 4ull * n); memcpy(ht + n, c->h_doc_len.data(), 4ull * n); }
  memcpy(ht + 2ull * n, c->h_tile_base.data(), 4ull * (n + 1));
  RFQ_CK(c, cudaMemcpyAsync(c->d_doctab.p, ht, tab, cudaMemcpyHostToDevice, s));
  c->d_doc_off = (uint32_t*)c->d_doctab.p;
  c->d_doc_len = c->d_doc_off + n;
  c->d_tile_base = c->d_doc_len + n;
  c->inflated_valid = false;
  c->have_output = false;
  c->staged = true;
  return RFQ_OK;
}

int rfq_frame_run(rfq_ctx* c, uint32_t n_out, const uint32_t* doc_first_part, uint32_t n_parts, const uint8_t* part_kind,
                  const uint32_t* part_first_piece, const uint8_t* piece_strip) {
  if (!c || !doc_first_part || (n_parts && (!part_kind || !part_first_piece)) || (c->n_docs && !piece_strip)) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_frame_run: nothing staged"; return RFQ_ERR_STATE; }
  const uint32_t n = c->n_docs;
  if (doc_first_part[n_out] != n_parts || (n_parts && part_first_piece[n_parts] != n)) { c->err = "rfq_frame_run: tables do not cover the staged pieces"; return RFQ_ERR_ARG; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  if (!c->nested) c->prof.begin();
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  if (!c->nested) { RFQ_CK(c, cudaEventRecord(c->ev0, s)); c->pool.reset(); }
  c->status_override.clear(); c->out_n = 0;
  const size_t o_dfp = 0, o_pk = (4ull * (n_out + 1) + 255) & ~255ull, o_pfp = (o_pk + n_parts + 255) & ~255ull,
               o_ps = (o_pfp + 4ull * (n_parts + 1) + 255) & ~255ull, used = o_ps + n;
  RFQ_CK(c, c->h_up.reserve(used + 256));
  uint8_t* hu = (uint8_t*)c->h_up.p;
  memcpy(hu + o_dfp, doc_first_part, 4ull * (n_out + 1));
  if (n_parts) { memcpy(hu + o_pk, part_kind, n_parts); memcpy(hu + o_pfp, part_first_piece, 4ull * (n_parts + 1)); }
  if (n) memcpy(hu + o_ps, piece_strip, n);
  uint8_t* d_up = c->pool.alloc<uint8_t>(used + 256);
  FrameArgs a{};
  a.arena = (const uint8_t*)c->d_arena.p; a.doc_off = c->d_doc_off; a.doc_len = c->d_doc_len;
  a.n_pieces = n; a.n_parts = n_parts; a.n_out = n_out;
  a.piece_a = c->pool.alloc<uint32_t>(n + 1); a.piece_n = c->pool.alloc<uint32_t>(n + 1);
  a.piece_dst = c->pool.alloc<unsigned long long>(n + 1);
  a.out_len = c->pool.alloc<unsigned long long>(n_out + 1); a.out_off = c->pool.alloc<unsigned long long>(n_out + 1);
  a.out_total = c->pool.alloc<unsigned long long>(1);
  a.scan_tmp64 = c->pool.alloc<unsigned long long>(scan_partials_needed(n_out));
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  RFQ_CK(c, cudaMemcpyAsync(d_up, hu, used, cudaMemcpyHostToDevice, s));
  a.doc_first_part = (const uint32_t*)(d_up + o_dfp); a.part_kind = d_up + o_pk;
  a.part_first_piece = (const uint32_t*)(d_up + o_pfp); a.piece_strip = d_up + o_ps;
  frame_measure(a, s);
  unsigned long long total = 0;
  if (n_out) {
    RFQ_CK(c, cudaMemcpyAsync(c->h_scalar(), a.out_total, 8, cudaMemcpyDeviceToHost, s));
    RFQ_CK(c, cudaStreamSynchronize(s));
    total = *c->h_scalar();
  }
  uint8_t* d_text = c->pool.alloc<uint8_t>(total + 16);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  frame_emit(a, d_text, s);
  c->segs.clear();
  c->segs.push_back({d_text, total, 0});
  c->host_tables = false;
  prof_mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  RFQ_CK(c, cudaGetLastError());
  c->out_total = total;
  c->out_n = n_out;
  c->d_out_off = a.out_off; c->d_out_len = a.out_len;
  c->d_status = nullptr; c->d_rows = nullptr; c->d_cols = nullptr; c->d_attempt = nullptr;
  c->have_output = true;
  c->timed = true;
  return RFQ_OK;
}

int rfq_json_spans_run(rfq_ctx* c) {
  if (!c) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_json_spans_run: nothing staged"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  uint32_t n = c->n_docs;
  if (!c->nested) c->prof.begin();
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  if (!c->nested) { RFQ_CK(c, cudaEventRecord(c->ev0, s)); c->pool.reset(); }
  c->status_override.clear(); c->out_n = 0;
  unsigned long long* out_off = c->pool.alloc<unsigned long long>(n + 1);
  unsigned long long* out_len = c->pool.alloc<unsigned long long>(n + 1);
  uint32_t* d_val = c->pool.alloc<uint32_t>(4ull * n + 4);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  json_spans_batch((const uint8_t*)c->d_arena.p, c->d_doc_off, c->d_doc_len, n, d_val, out_off, out_len, s);
  c->segs.clear();
  c->segs.push_back({(uint8_t*)d_val, 16ull * n, 0});
  c->host_tables = false;
  prof_mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  RFQ_CK(c, cudaGetLastError());
  c->out_total = 16ull * n;
  c->d_out_off = out_off; c->d_out_len = out_len;
  c->d_status = nullptr; c->d_rows = nullptr; c->d_cols = nullptr; c->d_attempt = nullptr;
  c->have_output = true;
  c->timed = true;
  return RFQ_OK;
}

// The .json path (app/file_parser.py:315-331): strict UTF-8, json.loads, json.dumps(indent=2, ensure_ascii=False)
// behind the "=== JSON Data ===" banner.  Two launches around one host synchronisation (output sizes).
int rfq_json_run(rfq_ctx* c) {
  if (!c) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_json_run: nothing staged"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  const uint32_t n = c->n_docs;
  if (!c->nested) c->prof.begin();
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  if (!c->nested) { RFQ_CK(c, cudaEventRecord(c->ev0, s)); c->pool.reset(); }
  c->status_override.clear(); c->out_n = 0;
  c->segs.clear();
  c->out_total = 0;
  c->host_tables = true;
  c->r_off.assign(n, 0); c->r_len.assign(n, 0); c->r_chars.assign(n, ~0ull);
  c->r_status.assign(n, RFQ_DOC_OK);
  c->r_rows.assign(n, 0); c->r_cols.assign(n, 0); c->r_attempt.assign(n, 0);
  uint32_t* d_keys = c->pool.alloc<uint32_t>(c->arena_bytes / 2 + 4ull * n + 64);
  JsonDocRes* d_res = c->pool.alloc<JsonDocRes>(n + 1);
  unsigned long long* d_off = c->pool.alloc<unsigned long long>(n + 1);
  uint32_t* d_ticket = c->pool.alloc<uint32_t>(4);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  RFQ_CK(c, cudaMemsetAsync(d_ticket, 0, 16, s));
  json_pretty_measure((const uint8_t*)c->d_arena.p, c->d_doc_off, c->d_doc_len, n, d_keys, d_res, d_ticket, s);
  RFQ_CK(c, c->h_down.reserve(sizeof(JsonDocRes) * (n + 1) + 64));
  JsonDocRes* h_res = (JsonDocRes*)c->h_down.p;
  if (n) RFQ_CK(c, cudaMemcpyAsync(h_res, d_res, sizeof(JsonDocRes) * n, cudaMemcpyDeviceToHost, s));
  prof_mark("host+sync", s);
  RFQ_CK(c, cudaStreamSynchronize(s));
  RFQ_CK(c, c->h_up.reserve(8ull * (n + 1) + 64));
  unsigned long long* h_off = (unsigned long long*)c->h_up.p;
  unsigned long long total = 0;
  for (uint32_t d = 0; d < n; ++d) {
    const JsonDocRes& r = h_res[d];
    h_off[d] = total;
    if (r.status == 0) {
      c->r_off[d] = total; c->r_len[d] = 18 + r.out_bytes;      // "=== JSON Data ===\n" + the text
      total += (c->r_len[d] + 15) & ~15ull;
    } else if (r.status == 1) {
      // json.JSONDecodeError: which message, and the position it reports (include/rfq_parse.h)
      c->r_status[d] = RFQ_DOC_JSON_INVALID;
      c->r_attempt[d] = (uint8_t)r.err; c->r_rows[d] = r.line; c->r_cols[d] = r.col; c->r_chars[d] = r.pos;
    } else c->r_status[d] = RFQ_DOC_UNSUPPORTED;
  }
  uint8_t* d_text = c->pool.alloc<uint8_t>(total + 64);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  if (total) {
    RFQ_CK(c, cudaMemcpyAsync(d_off, h_off, 8ull * n, cudaMemcpyHostToDevice, s));
    RFQ_CK(c, cudaMemsetAsync(d_ticket + 1, 0, 4, s));
    json_pretty_emit((const uint8_t*)c->d_arena.p, c->d_doc_off, c->d_doc_len, n, d_keys, d_res, d_off, d_text, d_ticket + 1, s);
  }
  c->segs.push_back({d_text, total, 0});
  c->out_total = total;
  prof_mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  RFQ_CK(c, cudaGetLastError());
  c->d_out_off = nullptr; c->d_out_len = nullptr;
  c->have_output = true;
  c->timed = true;
  return RFQ_OK;
}

int rfq_numeric_run(rfq_ctx* c, const uint8_t* kind) {
  if (!c || (c->n_docs && !kind)) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_numeric_run: nothing staged"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  uint32_t n = c->n_docs;
  if (!c->nested) c->prof.begin();
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  if (!c->nested) { RFQ_CK(c, cudaEventRecord(c->ev0, s)); c->pool.reset(); }
  c->status_override.clear(); c->out_n = 0;
  RFQ_CK(c, c->h_up.reserve(n + 16));
  if (n) memcpy(c->h_up.p, kind, n);
  uint8_t* d_kind = c->pool.alloc<uint8_t>(n + 16);
  unsigned long long* out_off = c->pool.alloc<unsigned long long>(n + 1);
  unsigned long long* out_len = c->pool.alloc<unsigned long long>(n + 1);
  unsigned long long* d_val = c->pool.alloc<unsigned long long>(n + 2);
  int32_t* d_status = c->pool.alloc<int32_t>(n + 1);
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  if (n) RFQ_CK(c, cudaMemcpyAsync(d_kind, c->h_up.p, n, cudaMemcpyHostToDevice, s));
  numeric_batch((const uint8_t*)c->d_arena.p, c->d_doc_off, c->d_doc_len, d_kind, n, d_val, d_status,
                out_off, out_len, s);
  c->segs.clear();
  c->segs.push_back({(uint8_t*)d_val, 8ull * n, 0});
  c->host_tables = false;
  prof_mark("end", s);
  RFQ_CK(c, cudaEventRecord(c->ev1, s));
  RFQ_CK(c, cudaGetLastError());
  c->out_total = 8ull * n;
  c->d_out_off = out_off; c->d_out_len = out_len;
  c->d_status = d_status; c->d_rows = nullptr; c->d_cols = nullptr; c->d_attempt = nullptr;
  c->have_output = true;
  c->timed = true;
  return RFQ_OK;
}

int rfq_prompt_run(rfq_ctx* c, const uint8_t* head, uint32_t head_len, const uint8_t* tail,
                   uint32_t tail_len, const uint8_t* mark, uint32_t mark_len, uint32_t max_chars) {
  if (!c || (head_len && !head) || (tail_len && !tail) || (mark_len && !mark)) return RFQ_ERR_ARG;
  if (!c->staged) { c->err = "rfq_prompt_run: nothing staged"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  uint32_t n = c->n_docs;
  if (!c->nested) c->prof.begin();
  g_prof = &c->prof;
  struct ProfGuard { ~ProfGuard() { g_prof = nullptr; } } prof_guard;
  if (!c->nested) { RFQ_CK(c, cudaEventRecord(c->ev0, s)); c->pool.reset(); }
  c->status_override.clear(); c->out_n = 0;
  // template pieces: one small pinned upload
  const size_t o_tail = ((size_t)head_len + 15) & ~(size_t)15, o_mark = (o_tail + tail_len + 15) & ~(size_t)15;
  const size_t pieces = o_mark + mark_len + 16;
  RFQ_CK(c, c->h_up.reserve(pieces));
  uint8_t* hp = (uint8_t*)c->h_up.p;
  if (head_len) memcpy(hp, head, head_len);
  if (tail_len) memcpy(hp + o_tail, tail, tail_len);
  if (mark_len) memcpy(hp + o_mark, mark, mark_len);
  uint8_t* d_pieces = c->pool.alloc<uint8_t>(pieces);
  PromptArgs a{};
  a.arena = (const uint8_t*)c->d_arena.p;
  a.doc_off = c->d_doc_off; a.doc_len = c->d_doc_len; a.n_docs = n;
  a.head = d_pieces; a.head_len = head_len;
  a.tail = d_pieces + o_tail; a.tail_len = tail_len;
  a.mark = d_pieces + o_mark; a.mark_len = mark_len;
  a.max_chars = max_chars;
  a.cut = c->pool.alloc<uint32_t>(n + 1);
  a.trunc = c->pool.alloc<uint32_t>(n + 1);
  a.status = c->pool.alloc<int32_t>(n + 1);
  a.out_len = c->pool.alloc<unsigned long long>(n + 1);
  a.out_off = c->pool.alloc<unsigned long long>(n + 1);
  a.out_total = c->pool.alloc<unsigned long long>(1);
  a.scan_tmp64 = c->pool.alloc<unsigned long long>(scan_partials_needed(n));
  if (c->pool.failed) { c->err = "device pool allocation failed"; return RFQ_ERR_CUDA; }
  RFQ_CK(c, cudaMemcpyAsync(d_pieces, hp, pieces, cudaMemcpyHostToDevice, s));
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

static int fetch_impl(rfq_ctx* c, bool with_text);
int rfq_fetch(rfq_ctx* c) { return fetch_impl(c, true); }
int rfq_fetch_tables(rfq_ctx* c) { return fetch_impl(c, false); }
static int fetch_impl(rfq_ctx* c, bool with_text) {
  if (!c) return RFQ_ERR_ARG;
  if (!c->have_output) { c->err = "rfq_fetch: no run to fetch"; return RFQ_ERR_STATE; }
  RFQ_CK(c, cudaSetDevice(c->device));
  cudaStream_t s = c->stream;
  uint32_t n = c->out_n ? c->out_n : c->n_docs;
  bool pre = false;
  for (auto& sg : c->segs) pre |= sg.copied;
  if (pre && c->h_out.cap < c->out_total + 16) {   // must grow: the early copies are void
    RFQ_CK(c, cudaStreamSynchronize(c->copy_stream));
    for (auto& sg : c->segs) sg.copied = false;
    pre = false;
  }
  if (with_text) RFQ_CK(c, c->h_out.reserve(c->out_total + 16));
  RFQ_CK(c, c->h_tabs.reserve((8ull + 8 + 4 + 4 + 4 + 1) * (n + 1) + 64));
  uint8_t* base = (uint8_t*)c->h_tabs.p;
  c->h_out_off = (uint64_t*)base;
  c->h_out_len = (uint64_t*)(base + 8ull * (n + 1));
  c->h_status = (int32_t*)(base + 16ull * (n + 1));
  c->h_rows = (uint32_t*)(base + 20ull * (n + 1));
  c->h_cols = (uint32_t*)(base + 24ull * (n + 1));
  c->h_attempt = (uint8_t*)(base + 28ull * (n + 1));
  memset(c->h_status, 0, 4ull * (n + 1) * 3 + (n + 1));
  for (auto& sg : c->segs)
    if (with_text && sg.bytes && !sg.copied)
      RFQ_CK(c, cudaMemcpyAsync((uint8_t*)c->h_out.p + sg.host_base, sg.d, sg.bytes,
                                cudaMemcpyDeviceToHost, s));
  c->d2h_bytes = with_text ? c->out_total : 0;
  if (c->host_tables) {
    if (n) {
      memcpy(c->h_out_off, c->r_off.data(), 8ull * n);
      memcpy(c->h_out_len, c->r_len.data(), 8ull * n);
      memcpy(c->h_status, c->r_status.data(), 4ull * n);
      memcpy(c->h_rows, c->r_rows.data(), 4ull * n);
      memcpy(c->h_cols, c->r_cols.data(), 4ull * n);
      memcpy(c->h_attempt, c->r_attempt.data(), n);
    }
  } else if (n) {
    RFQ_CK(c, cudaMemcpyAsync(c->h_out_off, c->d_out_off, 8ull * n, cudaMemcpyDeviceToHost, s));
    RFQ_CK(c, cudaMemcpyAsync(c->h_out_len, c->d_out_len, 8ull * n, cudaMemcpyDeviceToHost, s));
    c->d2h_bytes += 16ull * n;
    if (c->d_status) {
      RFQ_CK(c, cudaMemcpyAsync(c->h_status, c->d_status, 4ull * n, cudaMemcpyDeviceToHost, s));
      c->d2h_bytes += 4ull * n;
    }
    if (c->d_rows) {
      RFQ_CK(c, cudaMemcpyAsync(c->h_rows, c->d_rows, 4ull * n, cudaMemcpyDeviceToHost, s));
      RFQ_CK(c, cudaMemcpyAsync(c->h_cols, c->d_cols, 4ull * n, cudaMemcpyDeviceToHost, s));
      c->d2h_bytes += 8ull * n;
    }
    if (c->d_attempt) {
      RFQ_CK(c, cudaMemcpyAsync(c->h_attempt, c->d_attempt, n, cudaMemcpyDeviceToHost, s));
      c->d2h_bytes += n;
    }
  }
  RFQ_CK(c, cudaStreamSynchronize(s));
  if (pre) RFQ_CK(c, cudaStreamSynchronize(c->copy_stream));
  if (!c->host_tables && c->status_override.size() == n)   // fused parse -> prompt run: parse failures win
    for (uint32_t d = 0; d < n; ++d)
      if (c->status_override[d] >= 0) c->h_status[d] = c->status_override[d];
  if (c->timed) {
    cudaEventElapsedTime(&c->last_ms, c->ev0, c->ev1);
    c->prof.finish();
    c->timed = false;
  }
  c->fetched = true;
  return RFQ_OK;
}

int rfq_text_parse(rfq_ctx* c, const uint8_t* arena, const uint64_t* offsets,
                   const uint64_t* lengths, uint32_t n) {
  int r = rfq_stage(c, arena, offsets, lengths, n);
  if (r) return r;
  r = rfq_text_run(c);
  if (r) return r;
  return rfq_fetch(c);
}

int rfq_prompt_build(rfq_ctx* c, const uint8_t* arena, const uint64_t* offsets,
                     const uint64_t* lengths, uint32_t n, const uint8_t* head, uint32_t head_len,
                     const uint8_t* tail, uint32_t tail_len, const uint8_t* mark, uint32_t mark_len,
                     uint32_t max_chars) {
  int r = rfq_stage(c, arena, offsets, lengths, n);
  if (r) return r;
  r = rfq_prompt_run(c, head, head_len, tail, tail_len, mark, mark_len, max_chars);
  if (r) return r;
  return rfq_fetch(c);
}

int rfq_csv_parse(rfq_ctx* c, const uint8_t* arena, const uint64_t* offsets,
                  const uint64_t* lengths, uint32_t n, const int64_t* max_rows) {
  int r = rfq_stage(c, arena, offsets, lengths, n);
  if (r) return r;
  c->overlap_fetch = true;   // stream finished output slices to the host while the rest is emitted
  r = rfq_csv_run(c, max_rows);
  c->overlap_fetch = false;
  if (r) return r;
  return rfq_fetch(c);
}

uint32_t rfq_out_count(const rfq_ctx* c) { return c ? (c->out_n ? c->out_n : c->n_docs) : 0; }
const uint8_t* rfq_out_text(const rfq_ctx* c) { return c ? (const uint8_t*)c->h_out.p : nullptr; }
const uint64_t* rfq_out_offsets(const rfq_ctx* c) { return c ? c->h_out_off : nullptr; }
const uint64_t* rfq_out_lengths(const rfq_ctx* c) { return c ? c->h_out_len : nullptr; }
const int32_t* rfq_out_status(const rfq_ctx* c) { return c ? c->h_status : nullptr; }
uint64_t rfq_out_total_bytes(const rfq_ctx* c) { return c ? c->out_total : 0; }
const uint32_t* rfq_out_rows(const rfq_ctx* c) { return c ? c->h_rows : nullptr; }
const uint32_t* rfq_out_cols(const rfq_ctx* c) { return c ? c->h_cols : nullptr; }
const uint8_t* rfq_out_attempt(const rfq_ctx* c) { return c ? c->h_attempt : nullptr; }

// ---- fixed-width code units of a fetched text, for bindings whose string type stores them that way
// (CPython: 1, 2 or 4 bytes per code point).  Host-side, a few threads: a language-boundary conversion,
// not part of the parse.
}  // extern "C"
namespace {
template <class F> void host_parallel(uint64_t n, uint64_t min_chunk, F&& fn) {
  unsigned hw = std::thread::hardware_concurrency();
  uint64_t nt = std::min<uint64_t>(std::min<uint64_t>(hw ? hw : 4u, 16u), std::max<uint64_t>(1, n / min_chunk));
  if (nt <= 1) { fn(0, n, 0u); return; }
  std::vector<std::thread> th;
  const uint64_t per = (n + nt - 1) / nt;
  for (uint64_t t = 0; t < nt; ++t) th.emplace_back([=, &fn] { fn(std::min(n, t * per), std::min(n, (t + 1) * per), (unsigned)t); });
  for (auto& x : th) x.join();
}
}  // namespace
extern "C" {

int rfq_out_maxchar(const rfq_ctx* c, uint32_t d, uint32_t* maxchar) {
  const uint32_t n = c ? (c->out_n ? c->out_n : c->n_docs) : 0;
  if (!c || !c->fetched || d >= n || !maxchar) return RFQ_ERR_ARG;
  const uint8_t* p = (const uint8_t*)c->h_out.p + c->h_out_off[d];
  const uint64_t len = c->h_out_len[d];
  uint8_t part[16] = {0};
  host_parallel(len, 4u << 20, [&](uint64_t a, uint64_t b, unsigned t) {
    uint8_t m = 0;
    for (uint64_t i = a; i < b; ++i) m = std::max(m, p[i]);
    part[t] = m;
  });
  uint8_t m = 0;
  for (uint8_t x : part) m = std::max(m, x);
  // the largest byte of valid UTF-8 is a lead byte (or ASCII): it bounds the largest code point
  *maxchar = m < 0x80 ? 0x7Fu : m < 0xC4 ? 0xFFu : m < 0xF0 ? 0xFFFFu : 0x10FFFFu;
  return RFQ_OK;
}

int rfq_out_widen(const rfq_ctx* c, uint32_t d, int unit, void* dst, uint64_t n_chars) {
  const uint32_t n = c ? (c->out_n ? c->out_n : c->n_docs) : 0;
  if (!c || !c->fetched || d >= n || !dst || (unit != 1 && unit != 2 && unit != 4)) return RFQ_ERR_ARG;
  const uint8_t* p = (const uint8_t*)c->h_out.p + c->h_out_off[d];
  const uint64_t len = c->h_out_len[d];
  // chunk boundaries on character starts; characters before every chunk
  uint64_t cut[17], before[17];
  unsigned nt = 0;
  {
    unsigned hw = std::thread::hardware_concurrency();
    nt = (unsigned)std::min<uint64_t>(std::min<uint64_t>(hw ? hw : 4u, 16u), std::max<uint64_t>(1, len / (4u << 20)));
    for (unsigned t = 0; t <= nt; ++t) {
      uint64_t x = len * t / nt;
      while (x < len && (p[x] & 0xC0) == 0x80) ++x;
      cut[t] = x;
    }
  }
  std::vector<std::thread> th;
  // the pages of a fresh str have never been touched: ask for huge ones where the kernel hands them out on request
  {
    const uintptr_t a0 = ((uintptr_t)dst + (2u << 20) - 1) & ~(uintptr_t)((2u << 20) - 1);
    const uintptr_t a1 = ((uintptr_t)dst + n_chars * (uint64_t)unit) & ~(uintptr_t)((2u << 20) - 1);
    if (a1 > a0) madvise((void*)a0, a1 - a0, MADV_HUGEPAGE);
  }
  if (unit == 1 && n_chars == len) {
    // every byte starts a character: the text is ASCII and the str is a copy of it
    auto copy = [&](unsigned t) { memcpy((uint8_t*)dst + cut[t], p + cut[t], cut[t + 1] - cut[t]); };
    if (nt == 1) copy(0); else { for (unsigned t = 0; t < nt; ++t) th.emplace_back(copy, t); for (auto& x : th) x.join(); }
    return RFQ_OK;
  }
  uint64_t cnt[16] = {0};
  auto count = [&](unsigned t) {
    uint64_t k = 0, i = cut[t];
    const uint64_t e = cut[t + 1];
    for (; i + 8 <= e; i += 8) {     // eight bytes at a time: continuation bytes are 10xxxxxx
      uint64_t x; memcpy(&x, p + i, 8);
      k += 8 - (uint64_t)__builtin_popcountll((x >> 7) & ~(x >> 6) & 0x0101010101010101ull);
    }
    for (; i < e; ++i) k += (p[i] & 0xC0) != 0x80;
    cnt[t] = k;
  };
  if (nt == 1) count(0); else { for (unsigned t = 0; t < nt; ++t) th.emplace_back(count, t); for (auto& x : th) x.join(); th.clear(); }
  before[0] = 0;
  for (unsigned t = 0; t < nt; ++t) before[t + 1] = before[t] + cnt[t];
  if (before[nt] != n_chars) return RFQ_ERR_ARG;
  auto widen = [&](unsigned t) {
    uint64_t o = before[t];
    uint8_t* d1 = (uint8_t*)dst; uint16_t* d2 = (uint16_t*)dst; uint32_t* d4 = (uint32_t*)dst;
    for (uint64_t i = cut[t]; i < cut[t + 1];) {
      if (i + 8 <= cut[t + 1]) {       // a run of ASCII is widened eight characters at a time
        uint64_t x; memcpy(&x, p + i, 8);
        if (!(x & 0x8080808080808080ull)) {
          if (unit == 1) memcpy(d1 + o, &x, 8);
          else if (unit == 2) for (int q = 0; q < 8; ++q) d2[o + q] = (uint16_t)((x >> (8 * q)) & 0xFF);
          else for (int q = 0; q < 8; ++q) d4[o + q] = (uint32_t)((x >> (8 * q)) & 0xFF);
          o += 8; i += 8;
          continue;
        }
      }
      uint32_t b = p[i], cp; int l;
      if (b < 0x80) { cp = b; l = 1; }
      else if (b < 0xE0) { cp = ((b & 0x1Fu) << 6) | (p[i + 1] & 0x3Fu); l = 2; }
      else if (b < 0xF0) { cp = ((b & 0x0Fu) << 12) | ((p[i + 1] & 0x3Fu) << 6) | (p[i + 2] & 0x3Fu); l = 3; }
      else { cp = ((b & 0x07u) << 18) | ((p[i + 1] & 0x3Fu) << 12) | ((p[i + 2] & 0x3Fu) << 6) | (p[i + 3] & 0x3Fu); l = 4; }
      if (unit == 1) d1[o] = (uint8_t)cp; else if (unit == 2) d2[o] = (uint16_t)cp; else d4[o] = cp;
      ++o; i += l;
    }
  };
  if (nt == 1) widen(0); else { for (unsigned t = 0; t < nt; ++t) th.emplace_back(widen, t); for (auto& x : th) x.join(); }
  return RFQ_OK;
}

// Many documents at once (a batch of sheets, bodies, small files): the two steps above for every listed
// document, spread over host threads by bytes.  plan: code points and widest code point of each; fill: each
// text widened into the buffer the caller allocated for it (units[i] bytes per character, n_chars[i] of them).
int rfq_out_plan_many(const rfq_ctx* c, uint32_t n_list, const uint32_t* docs, uint64_t* n_chars, uint32_t* maxchar) {
  const uint32_t n = c ? (c->out_n ? c->out_n : c->n_docs) : 0;
  if (!c || !c->fetched || (n_list && (!docs || !n_chars || !maxchar))) return RFQ_ERR_ARG;
  for (uint32_t i = 0; i < n_list; ++i) if (docs[i] >= n) return RFQ_ERR_ARG;
  std::atomic<uint32_t> next{0};
  auto work = [&]() {
    for (;;) {
      const uint32_t i = next.fetch_add(1);
      if (i >= n_list) return;
      const uint8_t* p = (const uint8_t*)c->h_out.p + c->h_out_off[docs[i]];
      const uint64_t len = c->h_out_len[docs[i]];
      uint64_t k = 0, q = 0; uint8_t m = 0;
      for (; q + 8 <= len; q += 8) {
        uint64_t x; memcpy(&x, p + q, 8);
        k += 8 - (uint64_t)__builtin_popcountll((x >> 7) & ~(x >> 6) & 0x0101010101010101ull);
        if (x & 0x8080808080808080ull) for (int b = 0; b < 8; ++b) m = std::max<uint8_t>(m, (uint8_t)(x >> (8 * b)));
      }
      for (; q < len; ++q) { k += (p[q] & 0xC0) != 0x80; m = std::max(m, p[q]); }
      n_chars[i] = k;
      maxchar[i] = m < 0x80 ? 0x7Fu : m < 0xC4 ? 0xFFu : m < 0xF0 ? 0xFFFFu : 0x10FFFFu;
    }
  };
  unsigned hw = std::thread::hardware_concurrency();
  const unsigned nt = n_list < 64 ? 1u : std::min<unsigned>(hw ? hw : 4u, 16u);
  if (nt <= 1) { work(); return RFQ_OK; }
  std::vector<std::thread> th;
  for (unsigned t = 0; t < nt; ++t) th.emplace_back(work);
  for (auto& x : th) x.join();
  return RFQ_OK;
}

int rfq_out_widen_many(const rfq_ctx* c, uint32_t n_list, const uint32_t* docs, const uint8_t* units, void* const* dsts,
                       const uint64_t* n_chars) {
  const uint32_t n = c ? (c->out_n ? c->out_n : c->n_docs) : 0;
  if (!c || !c->fetched || (n_list && (!docs || !units || !dsts || !n_chars))) return RFQ_ERR_ARG;
  for (uint32_t i = 0; i < n_list; ++i)
    if (docs[i] >= n || !dsts[i] || (units[i] != 1 && units[i] != 2 && units[i] != 4)) return RFQ_ERR_ARG;
  std::atomic<uint32_t> next{0};
  std::atomic<int> bad{0};
  auto work = [&]() {
    for (;;) {
      const uint32_t i = next.fetch_add(1);
      if (i >= n_list) return;
      const uint8_t* p = (const uint8_t*)c->h_out.p + c->h_out_off[docs[i]];
      const uint64_t len = c->h_out_len[docs[i]];
      const int unit = units[i];
      if (unit == 1 && n_chars[i] == len) { memcpy(dsts[i], p, len); continue; }   // ASCII: the str is a copy
      uint8_t* d1 = (uint8_t*)dsts[i]; uint16_t* d2 = (uint16_t*)dsts[i]; uint32_t* d4 = (uint32_t*)dsts[i];
      uint64_t o = 0;
      for (uint64_t q = 0; q < len;) {
        uint32_t b = p[q], cp; int l;
        if (b < 0x80) { cp = b; l = 1; }
        else if (b < 0xE0) { cp = ((b & 0x1Fu) << 6) | (p[q + 1] & 0x3Fu); l = 2; }
        else if (b < 0xF0) { cp = ((b & 0x0Fu) << 12) | ((p[q + 1] & 0x3Fu) << 6) | (p[q + 2] & 0x3Fu); l = 3; }
        else { cp = ((b & 0x07u) << 18) | ((p[q + 1] & 0x3Fu) << 12) | ((p[q + 2] & 0x3Fu) << 6) | (p[q + 3] & 0x3Fu); l = 4; }
        if (o >= n_chars[i]) { bad = 1; break; }
        if (unit == 1) d1[o] = (uint8_t)cp; else if (unit == 2) d2[o] = (uint16_t)cp; else d4[o] = cp;
        ++o; q += l;
      }
      if (o != n_chars[i]) bad = 1;
    }
  };
  unsigned hw = std::thread::hardware_concurrency();
  const unsigned nt = n_list < 64 ? 1u : std::min<unsigned>(hw ? hw : 4u, 16u);
  if (nt <= 1) work();
  else {
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; ++t) th.emplace_back(work);
    for (auto& x : th) x.join();
  }
  return bad ? RFQ_ERR_ARG : RFQ_OK;
}

float rfq_last_run_ms(const rfq_ctx* c) { return c ? c->last_ms : 0.f; }
uint64_t rfq_launch_count(const rfq_ctx*) { return g_launch_count.load(); }
int rfq_set_profile(rfq_ctx* c, int on) { if (!c) return RFQ_ERR_ARG; c->prof.on = on != 0; return RFQ_OK; }
int rfq_profile_count(const rfq_ctx* c) { return c ? (int)c->prof.ms.size() : 0; }
const char* rfq_profile_name(const rfq_ctx* c, int i) {
  return (c && i >= 0 && i < (int)c->prof.names.size()) ? c->prof.names[i].c_str() : "";
}
float rfq_profile_ms(const rfq_ctx* c, int i) {
  return (c && i >= 0 && i < (int)c->prof.ms.size()) ? c->prof.ms[i] : 0.f;
}

}  // extern "C"
