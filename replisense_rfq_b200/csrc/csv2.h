// The row pipeline of the CSV path (csv2_*.cu): launch interfaces between csv2_api.cu and the kernels.
//
//   csv2_scan   one pass over the bytes: tokenizer DFA by decoupled look-back, output = where every
//               record starts (nothing per field)
//   csv2_type   a thread walks one row out of a shared-memory stage of its work item's bytes: field
//               ends, NA / int / decimal / text typing and the printed widths of every candidate type;
//               cells it does not decide itself go to a fix-up list (csv2_type_cold)
//   csv2_plan   per document: chunk types, rows dropna removes, column classes / widths / offsets,
//               bytes of every work item -> output offsets
//   csv2_emit   the same walk, printing right-justified cells into a shared-memory stage of the item's
//               output range, stored with one bulk copy; rows with a cell it does not print itself
//               go to a fix-up list (csv2_emit_cold)
//
// The host carves every array from bounds it knows before the first kernel runs (header width from
// the first two records, rows <= min(nrows, bytes / 2)), so one attempt needs no host
// synchronisation until its results are read.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rfq {

#define RFQ2_ROWS 128u            // data rows per work item = threads per CTA of the row kernels
#define RFQ2_SCAN_TILES 32u
// 16-byte units of cell_meta per row: 8 cells each
#define RFQ2_META_VECS(W) (((W) + 7u) / 8u)       // 512-byte tiles per scan chunk (16 KB per CTA round)

struct Doc2 {
  uint32_t off, len;              // document bytes in the arena
  uint32_t meta_base;             // cell_meta of record slot k at meta_base + (k - 1) * meta_stride(W), in 16-byte units
  uint32_t sep, max_rows;         // separator of this attempt, nrows
  uint32_t W, lead, h, bl;        // table width, implicit-index columns, output columns, low_memory chunk rows
  uint32_t slot_base;             // rec_start[slot_base + k] = start of record k (k <= max_rows + 1)
  uint32_t item_base, n_items;    // work items (RFQ2_ROWS data rows each), at least one
  uint32_t stat_base, n_chunks;   // (chunk, col) entries at stat_base + chunk * W + col; chunks carved
  uint32_t col_base;              // layout of output column j at col_base + j
  uint32_t name_base;             // its label at name_base + j (shared by documents with the same header record)
  uint32_t banner_off, banner_len, empty_off, empty_len, hdr_extra;   // texts in the names arena
  uint32_t latin1;                // the bytes are latin-1 (every high byte is one character, two bytes of output)
};

// what the kernels find out about a document
struct DocRun {
  uint32_t n_rec;                 // records (uncapped)
  uint32_t final_state;           // DFA state before the end of the document (ST_Q = EOF inside string)
  uint32_t flags;                 // RUN_* below
  uint32_t n_data, n_keep;        // data rows used, rows left after dropna
  uint32_t linew;                 // code points of one output line
  uint32_t hdr_bytes;             // banner + header line (or the empty-frame text)
  uint32_t extra;                 // printed bytes - printed code points of the data rows (0 = they are ASCII)
};
enum { RUN_FAIL = 1,              // this (encoding, separator) attempt fails the way pandas fails it
       RUN_UNSUP = 2,             // a construct that is declined (RFQ_DOC_UNSUPPORTED)
       RUN_GENERAL = 4 };         // chunk types the candidates do not cover: re-run through the three-sweep path

struct StatV2 {                   // per (chunk, col), all-zero = empty (one memset clears everything)
  uint32_t flags;                 // SF_*
  uint32_t inv_bad, inv_ovf, inv_uinv, inv_uovf;   // ~(first row ...), reduced with max (0 = none)
  uint32_t str_len, int_len, uint_len, bool_len, ip_len, inv_tz /* 6 - min_tz */, inf_len, obj_len, sci_len, fl;
  uint32_t na_count, need_exact, pad;
};

struct ColOut2 { int32_t cls, dec, sci; uint32_t width, cp_off; };

// a cell / a row handed to the fix-up kernels
struct ColdCell { uint32_t doc, slot, col; };
struct ColdRow { uint32_t doc, slot; unsigned long long out; };   // out = where the row's '\n' goes

struct Csv2Args {
  const uint8_t* arena;
  const Doc2* docs; uint32_t n_docs;
  const uint32_t* tile_base;      // [n_docs+1] prefix of ceil((len + 1) / 512)
  uint32_t n_tiles, n_chunks;     // scan chunks of RFQ2_SCAN_TILES tiles
  const uint32_t* item_base;      // [n_docs+1]
  uint32_t n_items;
  uint32_t n_slots, n_stat, n_cols;
  uint32_t max_w;                 // widest table of the attempt (shared-memory size of the typing sweep)
  // scan
  unsigned long long* fn_desc;    // [n_chunks] look-back descriptors (transition functions)
  unsigned long long* cnt_desc;   // [n_chunks] (records since the document start)
  uint32_t* ticket;               // [2]
  // three-launch scan (csv2_scan.cu): what the marks kernel leaves for the write kernel; null = one-launch scan
  uint2* lane_marks;              // [32 * n_tiles] {record starts | starts after blanks << 16, record ends} of a lane's 16 bytes
  uint32_t* tile_tot;             // [n_tiles] records that end in the tile, bit 31: a document starts here
  uint32_t* tile_doc;             // [n_tiles]
  unsigned long long* chunk_seg;  // [n_chunks] segmented record count of a chunk
  unsigned long long* chunk_pre;  // [n_chunks] records since the document start, before the chunk
  uint32_t* rec_start;            // [n_slots]
  DocRun* run;                    // [n_docs]
  // rows
  uint32_t* row_na;               // [n_slots] NA cells among the output columns
  uint32_t* row_extra;            // [n_slots] printed bytes - printed code points
  uint16_t* cell_meta;            // [n_slots * meta_stride(W)] what the typing sweep found per cell (csv2_rows.cuh)
  StatV2* stat;                   // [n_stat]
  uint8_t* ctype;                 // [n_stat]
  uint32_t* chunk_allna;          // [n_stat] (entry of column 0 of each chunk) rows whose output cells are all NA
  uint8_t* chunk_nf;              // [n_stat] (same entry) the chunk has a CT_STRNF output column
  void* meas;                     // ChunkMeasure[n_stat]
  uint32_t* item_allna;           // [n_items]
  uint32_t* item_extra;           // [n_items]
  unsigned long long* item_bytes; // [n_items+1]
  unsigned long long* item_off;   // [n_items+1] exclusive scan; last = total
  unsigned long long* scan_tmp64;
  ColOut2* col;                   // [n_cols]
  const uint32_t* name_off; const uint32_t* name_len; const uint32_t* name_cp;   // [n_cols]
  const uint8_t* names;
  unsigned long long* doc_out_off; unsigned long long* doc_out_len;   // [n_docs]
  // fix-up lists
  ColdCell* cold_cells; uint32_t cold_cells_cap;
  ColdRow* cold_rows; uint32_t cold_rows_cap;
  uint32_t* counters;             // [0] cold cells, [1] cold rows, [2] list overflow, [3] largest input span of an item, [4] largest output range
  // emit
  uint8_t* out; unsigned long long out_cap;
  uint32_t in_stage, out_stage;   // shared-memory bytes of the emit sweep's input / output stage (multiples of 16)
  // a small result block (counters, per-document results) is written to pinned host memory by the last kernel
  // of the attempt instead of being fetched by a copy of its own (export_words = 0: the host copies it)
  uint32_t* export_dst; const uint32_t* export_src; uint32_t export_words;
};

// The first two records of many documents, one thread each (header_names.h scan_header on the device): for
// batches of thousands of small files, where walking every file's first bytes in pinned host memory is the
// largest host cost of an attempt.  out[i] (a HeaderScan) describes document sel[i].
void csv2_header_scan(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len, const uint32_t* sel,
                      uint32_t n, uint32_t sep, bool latin1, void* out, cudaStream_t s);
void csv2_scan(const Csv2Args& a, cudaStream_t s);
void csv2_type(const Csv2Args& a, cudaStream_t s);     // hot sweep + fix-up of the cold cells
void csv2_plan(const Csv2Args& a, cudaStream_t s);     // decide / layout / item bytes / offsets
void csv2_emit(const Csv2Args& a, cudaStream_t s);     // hot emit + fix-up of the cold rows

}  // namespace rfq
