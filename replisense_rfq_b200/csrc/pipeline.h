// Internal launch interfaces between the C-ABI layer (rfq_api.cu) and the kernel files.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rfq {

// scratch elements an exclusive_scan over n items needs (scan.cuh: 4096 items per CTA)
inline uint32_t scan_partials_needed(uint64_t n) { return (uint32_t)((n + 4095) / 4096) + 1; }

struct TxtArgs {
  const uint8_t* arena;       // device, packed documents (16-byte aligned starts)
  const uint32_t* doc_off;    // [n_docs] byte offset of each document in arena
  const uint32_t* doc_len;    // [n_docs]
  const uint32_t* tile_base;  // [n_docs+1] exclusive prefix of ceil(len/512)
  uint32_t n_docs, n_tiles;
  uint32_t* tile_cnt;         // [n_tiles]
  uint32_t* tile_pre;         // [n_tiles]
  uint16_t* lane_emit;        // [32 * n_tiles] or null: which of a lane's 16 bytes the count pass found emitted, so that
                              // the emit pass does not decide them again (2 B per 16 input bytes)
  // strip extents, reduced over the tiles of a document with 64-bit atomics so that the winning tile also
  // hands over what the extent step needs from it (no second look at the bytes):
  unsigned long long* first_key;  // [n_docs] min of (first visible byte << 32 | bytes the tile emits before it)
  unsigned long long* last_key;   // [n_docs] max of (end of the last visible character << 32 | bytes of it in the
                                  //          next tile << 10 | bytes the tile emits before that end)
  unsigned long long* out_len;  // [n_docs]
  unsigned long long* out_off;  // [n_docs]
  unsigned long long* out_total;  // [1]
  uint32_t* skip;             // [n_docs]
  uint32_t* scan_tmp32;       // scan partials
  unsigned long long* scan_tmp64;
};

void txt_pipeline(const TxtArgs& a, cudaStream_t s);   // sizes: fills out_len/out_off/out_total
void txt_emit(const TxtArgs& a, uint8_t* out, cudaStream_t s);

// ------------------------------------------------------------------------------------------
// batched MD5 (md5_kernels.cu): 32 lowercase hex characters per document at out + 32 * d
void md5_batch(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, uint32_t n_docs,
               uint8_t* out, unsigned long long* out_off, unsigned long long* out_len, cudaStream_t s);

// ------------------------------------------------------------------------------------------
// quantity / price validators (numeric_kernels.cu): value[d] = int64 or double bits
void numeric_batch(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len,
                   const uint8_t* kind, uint32_t n_docs, unsigned long long* value, int32_t* status,
                   unsigned long long* out_off, unsigned long long* out_len, cudaStream_t s);

// ------------------------------------------------------------------------------------------
// text framing (frame_kernels.cu): documents assembled from pieces (page texts, table rows, literal headers)
struct FrameArgs {
  const uint8_t* arena; const uint32_t* doc_off; const uint32_t* doc_len;   // the staged pieces
  uint32_t n_pieces, n_parts, n_out;
  const uint32_t* doc_first_part;     // [n_out+1]
  const uint8_t* part_kind;           // [n_parts] RFQ_PART_*
  const uint32_t* part_first_piece;   // [n_parts+1]
  const uint8_t* piece_strip;         // [n_pieces] 1 = str.strip() the piece
  uint32_t* piece_a; uint32_t* piece_n;          // [n_pieces] kept span of every piece
  unsigned long long* piece_dst;      // [n_pieces] where it goes in the output, ~0 = nowhere
  unsigned long long* out_len; unsigned long long* out_off; unsigned long long* out_total;   // [n_out]
  unsigned long long* scan_tmp64;
};
void frame_measure(const FrameArgs& a, cudaStream_t s);
void frame_emit(const FrameArgs& a, uint8_t* out, cudaStream_t s);

// ------------------------------------------------------------------------------------------
// device-side DEFLATE (inflate_kernels.cu): member d of the arena -> out + out_off[d] (out_len[d] bytes expected)
void inflate_batch(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, const uint8_t* method,
                   const unsigned long long* out_off, const unsigned long long* out_len, const uint32_t* crc,
                   uint32_t n_docs, uint8_t* out, int32_t* status, uint32_t* ticket, cudaStream_t s);

// ------------------------------------------------------------------------------------------
// where the JSON sits in an LLM reply (json_kernels.cu): out[4*d..] = first '{', last '}', cleaned [begin, end)
void json_spans_batch(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, uint32_t n_docs,
                      uint32_t* out, unsigned long long* out_off, unsigned long long* out_len, cudaStream_t s);

// ------------------------------------------------------------------------------------------
// the .json path (json_pretty.cu): measure = validate + error position + output size, emit = write
struct JsonDocRes { uint32_t status, err, pos, line, col; unsigned long long out_bytes; };   // = JsonResult (json_pretty.cuh)
void json_pretty_measure(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, uint32_t n_docs, uint32_t* keys,
                         JsonDocRes* res, uint32_t* ticket, cudaStream_t s);
void json_pretty_emit(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, uint32_t n_docs, uint32_t* keys,
                      const JsonDocRes* res, const unsigned long long* out_off, uint8_t* out, uint32_t* ticket, cudaStream_t s);

// ------------------------------------------------------------------------------------------
// prompt assembly (prompt_kernels.cu)
struct PromptArgs {
  const uint8_t* arena;       // device, packed raw_text documents (UTF-8)
  const uint32_t* doc_off;
  const uint32_t* doc_len;
  uint32_t n_docs;
  const uint8_t* head; uint32_t head_len;   // device copies of the host's template pieces
  const uint8_t* tail; uint32_t tail_len;
  const uint8_t* mark; uint32_t mark_len;   // appended when the text was cut
  uint32_t max_chars;                       // code points kept
  uint32_t* cut;              // [n_docs] bytes of the document kept
  uint32_t* trunc;            // [n_docs]
  int32_t* status;            // [n_docs]
  unsigned long long* out_len;  // [n_docs]
  unsigned long long* out_off;  // [n_docs]
  unsigned long long* out_total;  // [1]
  unsigned long long* scan_tmp64;
};
// the first take[d] bytes of every document, wherever it lies on the device, into one 16-byte aligned arena
void prompt_gather(const unsigned long long* src_ptr, const uint32_t* take, const uint32_t* dst_off, uint8_t* dst,
                   uint32_t n_docs, cudaStream_t s);
void prompt_measure(const PromptArgs& a, cudaStream_t s);
void prompt_emit(const PromptArgs& a, uint8_t* out, cudaStream_t s);

// ------------------------------------------------------------------------------------------
// CSV / table pipeline

// tokeniser (csv_scan.cu): one attempt = one separator over a set of documents
struct ScanArgs {
  const uint8_t* arena;
  const uint32_t* doc_off;    // [n_docs]
  const uint32_t* doc_len;    // [n_docs]
  const uint32_t* tile_base;  // [n_docs+1] prefix of ceil((len+1)/512): the virtual LF at `len` has a tile
  uint32_t n_docs, n_tiles, sep;
  uint32_t* tile_fn;          // [n_tiles+1] transition function of each tile (nibble form)
  uint32_t* tile_fn_pre;      // [n_tiles+1] exclusive composition
  uint32_t* lane_state;       // [n_tiles*3] bit planes of the 32 lane start states
  unsigned long long* lane_cls;  // [n_tiles*32] byte classes of each lane's 16 bytes, a nibble each (pass 1 -> 2, 3)
  uint32_t* lane_fn;          // [n_tiles*32] transition function of each lane's 16 bytes (pass 1 -> pass 2)
  uint16_t* lane_cnt;         // [n_tiles*32] field ends | line ends << 8 of each lane (pass 2 -> pass 3)
  uint32_t* tile_nf;          // [n_tiles+1] field ends per tile      (last = 0)
  uint32_t* tile_nl;          // [n_tiles+1] line ends per tile       (last = 0)
  uint32_t* tile_nf_pre;      // [n_tiles+1] exclusive prefix, last = total
  uint32_t* tile_nl_pre;      // [n_tiles+1]
  uint32_t* doc_final;        // [n_docs] DFA state before the virtual LF (ST_Q = EOF in string)
  const uint32_t* max_rows;   // [n_docs] nrows of each document (records beyond 1 + nrows are never used)
  uint32_t* doc_unsup;        // [n_docs] set when the document holds a construct pandas tokenises by accident (see csv_index_kernel)
  uint32_t* scan_tmp;         // scan partials
  // outputs of the index pass (allocated after the totals are known)
  uint32_t* fend;             // [n_fields] doc-relative position of each field-end byte
  uint32_t* rec_fend;         // [n_lines + n_docs] one-past-last global field index of each record
  uint32_t* rec_start;        // [n_lines + n_docs] doc-relative position where each record starts
};
void csv_scan_count(const ScanArgs& a, cudaStream_t s);  // tile fns, states, counts, prefixes
void csv_scan_index(const ScanArgs& a, cudaStream_t s);  // fend / rec_fend / rec_start

// per-document record shape (device -> host after the index pass)
struct DocShape {
  uint32_t n_rec;      // complete records
  uint32_t hdr_nf;     // fields of record 0
  uint32_t first_nf;   // fields of record 1 (0 if none)
  uint32_t max_nf;     // max fields over records 2..nrows
  uint32_t final_state;
  uint32_t r0;         // global record slot of record 0
  uint32_t f0;         // global field index of the document's first field
  uint32_t pad;        // != 0: a record pandas tokenises by accident (bare CR + blank-first line): unsupported
};
void csv_doc_shape(const ScanArgs& a, const uint32_t* max_rows32, DocShape* shape, uint32_t n_slots,
                   cudaStream_t s);

// table description handed back to the device by the host after it validated the shapes
struct TDoc {
  uint32_t off, len;          // document bytes in the arena
  uint32_t r0, f0;            // global record slot of the header, first field index
  uint32_t W, lead, h;        // table width, implicit-index columns, output columns
  uint32_t n_data;            // data rows used (<= nrows)
  uint32_t bl, rb, n_chunks;  // low_memory chunk rows, rows per work item, chunks
  uint32_t stat_base;         // (chunk, col) arrays: stat_base + chunk * W + col
  uint32_t col_base;          // per output column arrays: col_base + j
  uint32_t item_base;         // work items of this document start here
  uint32_t banner_len;        // bytes of "=== CSV Data (...) ===\n"
  uint32_t hdr_extra;         // header-name bytes minus code points
  uint32_t empty_off, empty_len;  // "Empty DataFrame..." text in the names arena
  uint32_t banner_off;        // banner text in the names arena
  uint32_t active;            // 0 = attempt already failed on the host side (skip)
};

struct TableArgs {
  const uint8_t* arena;
  const TDoc* docs; uint32_t n_docs;
  const uint32_t* item_doc_base;  // [n_docs+1] prefix of items
  uint32_t n_items;
  const uint32_t* fend; const uint32_t* rec_fend; const uint32_t* rec_start;
  uint32_t n_slots;               // record slots (lines + docs)
  const uint32_t* slot_doc_r0;    // [n_docs+1] r0 of each document (for slot -> doc search)
  // (chunk, col)
  void* stat;        // ChunkStat[n_stat]
  void* meas;        // ChunkMeasure[n_stat]
  uint8_t* ctype;    // [n_stat]
  uint8_t* chunk_nf; // [n_stat] (slot of column 0 of each chunk row): chunk has a CT_STRNF output column
  uint32_t n_stat;
  // rows (indexed by record slot)
  uint32_t* row_na; uint32_t* row_extra;
  unsigned long long* row_bytes; unsigned long long* row_off;  // [n_slots+1]
  // per document
  uint32_t* n_keep; int32_t* doc_flags;  // bit0 decode error, bit1 unsupported
  uint32_t* doc_linew;            // code points of one output line
  // per output column
  void* col;         // ColOut[n_cols]
  const uint32_t* name_off; const uint32_t* name_len; const uint32_t* name_cp;  // [n_cols]
  const uint8_t* names;  // names / banner / empty-frame text arena
  uint32_t n_cols;
  unsigned long long* scan_tmp64; unsigned long long* out_total;
  unsigned long long* doc_out_off; unsigned long long* doc_out_len;  // [n_docs]
  // sheet mode (xcells != nullptr): cells are typed records instead of CSV fields
  const void* xcells;             // XCell[]
  const uint32_t* x_row_c0;       // [n_row_elems+1] first cell of each <row> element
  const uint32_t* x_r2elem;       // per document (TDoc.f0 + sheet row r) -> <row> element, ~0u = none
  const uint8_t* x_str_arena;     // decoded shared strings
  uint32_t* x_colflags;           // [n_cols] XF_* per output column
  uint32_t item_lo, item_hi;      // emit: work-item range of this launch (0, n_items = all)
  // fused classify + measure (CSV path): candidates per (chunk, col), see cell_ops.cuh ChunkCand
  void* cand;                     // ChunkCand[n_stat], nullptr = two-sweep mode
  uint32_t* chunk_dropped;        // [n_stat] (entry of column 0 of each chunk): rows dropna removes
  uint32_t* doc_exact;            // [n_docs] 1 = a chunk type the candidates do not cover: exact sweep
  uint32_t max_w;                 // widest table of the batch (shared-memory size of the fused sweep)
  uint8_t* cell_tag;              // [n_fields] how the fused sweep found each field (table_common.cuh), nullptr = none
};

struct ColOut { int32_t cls, dec, sci; uint32_t width, cp_off; };

void table_stat_init(const TableArgs& a, cudaStream_t s);
void csv_table_classify(const TableArgs& a, cudaStream_t s);
void csv_table_measure(const TableArgs& a, cudaStream_t s);   // decide + measure + layout + sizes
void csv_table_emit(const TableArgs& a, uint8_t* out, cudaStream_t s);
void csv_table_emit_header(const TableArgs& a, uint8_t* out, cudaStream_t s);
void csv_table_emit_items(const TableArgs& a, uint8_t* out, uint32_t lo, uint32_t hi, cudaStream_t s);

// ------------------------------------------------------------------------------------------
// sheet XML / sharedStrings (xlsx_kernels.cu)
struct XDocShape {
  uint32_t n_rows, n_cells, n_si;        // <row>, <c>, <si> elements of the document
  uint32_t row_base, cell_base, si_base; // global indices of its first element of each kind
  uint32_t flags;                        // XU_* problems
  uint32_t W, lastrow;                   // table width / last sheet row with data (within the cap)
  uint32_t pad[3];
};
struct XmlArgs {
  const uint8_t* arena;
  const uint32_t* doc_off; const uint32_t* doc_len; const uint32_t* tile_base;  // ceil(len/512) tiles
  uint32_t n_docs, n_tiles;
  const uint8_t* kind;          // [n_docs] 0 = sheet, 1 = sharedStrings
  const int32_t* sst_of;        // [n_docs] sharedStrings document of a sheet, -1 = none
  const uint32_t* cap_rows;     // [n_docs] 1 + nrows (file_rows_needed)
  const uint32_t* date_off;     // [n_docs+1] into date_ids
  const uint32_t* date_ids;
  uint32_t* tile_cnt[3];        // [n_tiles+1] per tile: rows, cells, si
  uint32_t* lane_kinds;         // [n_tiles * 32] what the count pass found in a lane's 16 bytes (2 bits per byte:
                                // 1 <row, 2 <c, 3 <si): the index pass reads this instead of the document
  uint32_t* tile_pre[3];        // [n_tiles+1]
  uint32_t* scan_tmp;
  XDocShape* shape;             // [n_docs]
  uint32_t* row_base; uint32_t* cell_base; uint32_t* si_base;  // [n_docs+1] (for element -> doc search)
  // after the totals are known
  uint32_t n_rows, n_cells, n_si;
  uint32_t* rowpos; uint32_t* row_c0; uint32_t* rownum;  // [n_rows+1]
  uint32_t* cpos;                                        // [n_cells]
  uint32_t* cell_doc; uint32_t* cell_row;                // [n_cells] document / <row> element of each <c> (~0u = none)
  void* xcells;                                          // XCell[n_cells]
  uint2* num_list; uint32_t* num_count;                  // cells whose number needs the general parsers: {cell, date style}
  uint32_t* sipos; uint32_t* si_len; uint32_t* si_off;   // [n_si+1]
  uint8_t* str_arena;                                    // decoded shared strings
};
void xml_count(const XmlArgs& a, cudaStream_t s);    // tile counts, prefixes, per-document bases
void xml_decode(const XmlArgs& a, cudaStream_t s);   // positions, strings, rows, cells, W / lastrow
void xlsx_table_classify(const TableArgs& a, cudaStream_t s);
void xlsx_header_cells(const TableArgs& a, uint8_t* hdr_buf /* n_cols * 256 */, uint32_t* hdr_meta /* n_cols * 2 */,
                       cudaStream_t s);
void xlsx_scatter_rows(const XmlArgs& x, const uint32_t* r2_base /* [n_docs] */, uint32_t* r2elem, cudaStream_t s);

// latin-1 -> UTF-8 stream transcode (second encoding of the trial loop)
void latin1_count(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len,
                  uint32_t n_docs, uint32_t* hi_count, cudaStream_t s);
void latin1_expand(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len,
                   const uint32_t* new_off, uint32_t n_docs, uint8_t* out, cudaStream_t s);

}  // namespace rfq
