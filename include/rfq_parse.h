/* rfq_parse.h — C ABI of the B200 (sm_100a) document→text engine.
 *
 * Drop-in boundary for the ONE hot path of Replisense-rfq: `FileParser` in
 * app/file_parser.py (bytes of CSV / XLSX sheet XML / raw text → one normalised `raw_text`
 * per document).  The reference has no FFI of its own; these entry points are
 * what a ctypes binding placed at the reference's three `run_in_executor(self.thread_pool,
 * self._parse_*_sync, …)` seams (app/file_parser.py:163,216,290) and at `_parse_txt_async`
 * (app/file_parser.py:143) binds.  INTEGRATION.md shows that binding.
 *
 * Conventions
 *   - plain pointers and sizes, no C++/torch types; every function returns an int error code
 *     (RFQ_OK = 0) except accessors; nothing throws across the ABI; rfq_last_error() gives text.
 *   - a batch is a packed byte arena + per-document offsets/lengths.  Offsets that are multiples
 *     of 16 let the library DMA straight out of the caller's buffer (use rfq_host_alloc() for
 *     pinned memory); other offsets are accepted and repacked through a pinned staging buffer.
 *   - outputs are library-owned pinned host buffers, valid until the next run on the same ctx.
 *   - one ctx = one device + one stream; a ctx is not re-entrant (use one per thread), distinct
 *     ctxs are independent.  There is NO CPU fallback: without a CUDA device rfq_create fails.
 */
#ifndef RFQ_PARSE_H_
#define RFQ_PARSE_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RFQ_OK 0
#define RFQ_ERR_CUDA (-1)
#define RFQ_ERR_ARG (-2)
#define RFQ_ERR_STATE (-3)
#define RFQ_ERR_TOO_LARGE (-4)

/* per-document status codes (rfq_out_status) */
#define RFQ_DOC_OK 0
/* every (encoding, separator) attempt failed → the reference raises FileParsingError("CSV
 * parsing failed: Unable to parse CSV with any supported encoding/separator combination"),
 * app/file_parser.py:310,313 */
#define RFQ_DOC_CSV_UNPARSEABLE 1
/* input uses a construct this build does not decode on the GPU (reported, never guessed) */
#define RFQ_DOC_UNSUPPORTED 2
/* sheet produced an empty frame (reference skips it: app/file_parser.py:228-229) */
#define RFQ_DOC_EMPTY_SHEET 3
/* prompt path: raw_text is empty or whitespace only → the reference answers "Empty or invalid
 * input text" without building a prompt (app/rfq_agent.py:138-140) */
#define RFQ_DOC_EMPTY_TEXT 4
/* numeric path: Python's int() / float() raises ValueError for this field */
#define RFQ_DOC_VALUE_ERROR 5
/* inflate path: bad DEFLATE stream, size or CRC-32 differs from the zip directory's, unknown method
 * (zipfile raises BadZipFile / zlib.error there) */
#define RFQ_DOC_CORRUPT 6
/* .json path: json.loads raises json.JSONDecodeError.  Which message and where, as Python reports them:
 * rfq_out_attempt[d] = RFQ_JSON_E_* (the message), rfq_out_rows[d] = lineno, rfq_out_cols[d] = colno,
 * rfq_out_chars[d] = the character index ("<message>: line L column C (char P)") */
#define RFQ_DOC_JSON_INVALID 7
#define RFQ_JSON_E_VALUE 1          /* Expecting value */
#define RFQ_JSON_E_PROPNAME 2       /* Expecting property name enclosed in double quotes */
#define RFQ_JSON_E_COLON 3          /* Expecting ':' delimiter */
#define RFQ_JSON_E_COMMA 4          /* Expecting ',' delimiter */
#define RFQ_JSON_E_EXTRA 5          /* Extra data */
#define RFQ_JSON_E_UNTERMINATED 6   /* Unterminated string starting at */
#define RFQ_JSON_E_CTRL 7           /* Invalid control character at */
#define RFQ_JSON_E_ESCAPE 8         /* Invalid \escape */
#define RFQ_JSON_E_UESCAPE 9        /* Invalid \uXXXX escape */
#define RFQ_JSON_E_BOM 10           /* Unexpected UTF-8 BOM (decode using utf-8-sig) */

typedef struct rfq_ctx rfq_ctx;

/* lifetime — replaces FileParser.__init__'s thread pool (app/file_parser.py:30-42) */
int rfq_create(int device, rfq_ctx** out);
void rfq_destroy(rfq_ctx* ctx);
const char* rfq_last_error(const rfq_ctx* ctx);
const char* rfq_version(void);
int rfq_device_count(void);
/* run on a caller-owned CUDA stream (e.g. torch's current stream) instead of the ctx's own
 * (plumbing of the boundary: the reference has no counterpart) */
int rfq_set_stream(rfq_ctx* ctx, void* cuda_stream);
void* rfq_get_stream(rfq_ctx* ctx);
/* pinned host memory for input arenas (plumbing of the boundary: the reference reads files into Python objects,
 * app/file_parser.py:146,225,301) */
void* rfq_host_alloc(rfq_ctx* ctx, uint64_t bytes);
void rfq_host_free(rfq_ctx* ctx, void* p);

/* ---- staging: host → device copy of one batch (kept resident until the next rfq_stage) ----
 * When every offset is a multiple of 16 the library copies straight out of `arena` (asynchronously if
 * it is pinned) and keeps the pointer: rfq_csv_run / rfq_mixed_run read the first records of every
 * document from it on the host.  The caller must keep `arena` alive and unchanged until the runs over
 * this batch have returned; other offsets are repacked into a library-owned buffer at once. */
int rfq_stage(rfq_ctx* ctx, const uint8_t* arena, const uint64_t* offsets,
              const uint64_t* lengths, uint32_t n_docs);

/* ---- text path: replaces FileParser._parse_txt_async (app/file_parser.py:143-159):
 * UTF-8 decode errors='ignore', universal newlines, str.strip().  Kernels only; the staged
 * batch is the input and the result stays on the device until rfq_fetch(). */
int rfq_text_run(rfq_ctx* ctx);

/* ---- CSV path: replaces FileParser._parse_csv_sync (app/file_parser.py:288-313): the
 * (utf-8, latin-1, cp1252) x (',', ';', TAB) trial loop around pd.read_csv(nrows=max_rows),
 * dropna(how='all'), fillna(''), banner + to_string(index=False).  max_rows[d] is the
 * reference's `max_excel_rows` (NULL → 1000 for every document). */
int rfq_csv_run(rfq_ctx* ctx, const int64_t* max_rows);

/* ---- sheet path: the GPU side of FileParser._parse_excel_sync (app/file_parser.py:214-252 →
 * pd.read_excel → openpyxl read-only reader → pandas TextParser → dropna/fillna/to_string).
 * The staged batch mixes worksheet parts (kind 0; kind 2 = a worksheet of a workbook with
 * <workbookPr date1904="1">) and sharedStrings parts (kind 1) that the host inflated from the
 * .xlsx zip; sst_of[d] is the batch index of the sharedStrings part of sheet d
 * (-1 = none); max_rows[d] = the reference's max_excel_rows; date_ids[date_off[d]..date_off[d+1])
 * are the cellXfs indices of sheet d's workbook whose number format is a date (styles.xml, read
 * on the host; bit 31 set = a timedelta format such as [h]:mm:ss).  Date-styled numbers become
 * datetimes / times exactly as openpyxl's from_excel makes them.  Output per sheet = df.to_string(index=False) (the host adds the
 * "=== Sheet: name ===" banner and joins sheets); status RFQ_DOC_EMPTY_SHEET where the reference
 * skips the sheet (df.empty), RFQ_DOC_UNSUPPORTED for constructs not decoded on the GPU. */
int rfq_sheets_run(rfq_ctx* ctx, const uint8_t* kind, const int32_t* sst_of, const int64_t* max_rows,
                   const uint32_t* date_off, const uint32_t* date_ids);
int rfq_sheets_parse(rfq_ctx* ctx, const uint8_t* arena, const uint64_t* offsets,
                     const uint64_t* lengths, uint32_t n_docs, const uint8_t* kind,
                     const int32_t* sst_of, const int64_t* max_rows, const uint32_t* date_off,
                     const uint32_t* date_ids);

/* ---- batched file hash: FileParser._get_file_hash_async (app/file_parser.py:122-130) for every
 * staged document at once — MD5 of the document bytes as 32 lowercase hex characters (the
 * envelope's `file_hash`).  One thread per document (MD5 is sequential inside a document): meant
 * for batches of many small files. */
int rfq_md5_run(rfq_ctx* ctx);

/* ---- quantity / price validators: LineItem.parse_quantity / parse_target_price
 * (app/rfq_agent.py:27-39) over a batch of field strings: kind[d] 0 → int(v.replace(",", "")),
 * 1 → float(v.replace(",", "")).  Output: 8 bytes per field (int64, or the bits of the double;
 * rfq_out_text at 8*d) and rfq_out_status: RFQ_DOC_OK, RFQ_DOC_VALUE_ERROR (Python raises
 * ValueError), RFQ_DOC_UNSUPPORTED (non-ASCII field, integer beyond int64, > 160 characters). */
int rfq_numeric_run(rfq_ctx* ctx, const uint8_t* kind);

/* ---- where the JSON sits in an LLM reply: the byte work of
 * RFQFieldGenerator._extract_json_from_string (app/rfq_agent.py:208-236) over a batch of staged replies.
 * Output: 16 bytes per reply (rfq_out_text at 16*d), four little-endian u32: position of the first '{',
 * of the last '}' (strategy 2: re.search(r'\{.*\}', s, re.DOTALL); 0xFFFFFFFF = none), and [begin, end) of
 * strategy 3's cleaned text (strip, drop a leading ```json and a trailing ```, strip).  json.loads of
 * those spans stays with the caller, as it is CPython's parser in the reference. */
int rfq_json_spans_run(rfq_ctx* ctx);

/* The .json file type of FileParser (replaces _parse_json_async, app/file_parser.py:315-331): the staged
 * documents are read as strict UTF-8 with universal newlines, parsed like json.loads and printed like
 * json.dumps(data, indent=2, ensure_ascii=False) behind "=== JSON Data ===\n".  Status per document:
 * RFQ_DOC_OK, RFQ_DOC_JSON_INVALID (with the reference's message and position, see above), or
 * RFQ_DOC_UNSUPPORTED for what is declined (bytes that are not UTF-8, an object with two equal keys, a
 * lone surrogate escape, an integer of more than 4300 digits, nesting beyond 96). */
int rfq_json_run(rfq_ctx* ctx);

/* ---- prompt assembly (the step right after the parser): the text handling of
 * RFQFieldGenerator._generate_sync (app/rfq_agent.py:138-151).  The staged batch holds raw_text
 * documents (UTF-8).  Per document: whitespace-only (str.strip() empty) → RFQ_DOC_EMPTY_TEXT and
 * no output; more than max_chars code points (the reference: 8000) → the first max_chars code
 * points + mark ("... [truncated]"); output = head + text + tail, where the host passes
 * head = EXTRACTION_PROMPT_TEMPLATE + '\n"""\n' and tail = '\n"""'. */
int rfq_prompt_run(rfq_ctx* ctx, const uint8_t* head, uint32_t head_len, const uint8_t* tail,
                   uint32_t tail_len, const uint8_t* mark, uint32_t mark_len, uint32_t max_chars);

/* ---- CSV parse fused with prompt assembly: the /upload/ flow of the reference (app/main.py:243-273) parses
 * the file, builds the prompt from raw_text (app/rfq_agent.py:138-151) and returns the LLM's fields plus
 * `text_length`; raw_text itself never leaves the service.  This entry runs rfq_csv_run and
 * rfq_prompt_run back to back on the device: the parsed text stays in HBM, the output of the run (what
 * rfq_fetch downloads) is the prompts.  Per-document status: RFQ_DOC_OK, RFQ_DOC_CSV_UNPARSEABLE /
 * RFQ_DOC_UNSUPPORTED from the parse stage, RFQ_DOC_EMPTY_TEXT from the prompt stage.
 * rfq_out_raw_bytes / rfq_out_raw_chars give the size of each raw_text in bytes and in code points
 * (the reference's `text_length`; ~0 where the parse stage took the three-sweep path and did not count). */
int rfq_csv_prompt_run(rfq_ctx* ctx, const int64_t* max_rows, const uint8_t* head, uint32_t head_len,
                       const uint8_t* tail, uint32_t tail_len, const uint8_t* mark, uint32_t mark_len,
                       uint32_t max_chars);
int rfq_csv_prompt_parse(rfq_ctx* ctx, const uint8_t* arena, const uint64_t* offsets, const uint64_t* lengths,
                         uint32_t n_docs, const int64_t* max_rows, const uint8_t* head, uint32_t head_len,
                         const uint8_t* tail, uint32_t tail_len, const uint8_t* mark, uint32_t mark_len,
                         uint32_t max_chars);
const uint64_t* rfq_out_raw_bytes(const rfq_ctx* ctx);
const uint64_t* rfq_out_raw_chars(const rfq_ctx* ctx);
/* code points of every text of the last CSV run (NULL after other runs; ~0 = not counted): lets a
 * binding build its string object without scanning the bytes when chars == bytes (ASCII) */
const uint64_t* rfq_out_chars(const rfq_ctx* ctx);

/* ---- mixed batch: text, CSV and sheet documents in ONE arena, one upload.  The
 * reference dispatches on the file suffix (app/file_parser.py:62-77); kind[d] carries that decision:
 * RFQ_KIND_TEXT → the text path, RFQ_KIND_CSV → the CSV path (max_rows[d] = max_excel_rows),
 * RFQ_KIND_SHEET / RFQ_KIND_SHEET_1904 / RFQ_KIND_SST → the sheet path (sst_of, date_off, date_ids as in
 * rfq_sheets_run, indices into THIS batch; date_off has n_docs + 1 entries).  Results come back in input
 * order with the statuses of the respective paths; a sharedStrings part has no output of its own. */
#define RFQ_KIND_TEXT 0
#define RFQ_KIND_CSV 1
#define RFQ_KIND_SHEET 2
#define RFQ_KIND_SHEET_1904 3
#define RFQ_KIND_SST 4
int rfq_mixed_run(rfq_ctx* ctx, const uint8_t* kind, const int64_t* max_rows, const int32_t* sst_of,
                  const uint32_t* date_off, const uint32_t* date_ids);
int rfq_mixed_parse(rfq_ctx* ctx, const uint8_t* arena, const uint64_t* offsets, const uint64_t* lengths,
                    uint32_t n_docs, const uint8_t* kind, const int64_t* max_rows, const int32_t* sst_of,
                    const uint32_t* date_off, const uint32_t* date_ids);

/* ---- text framing: documents assembled on the device from pieces, the way
 * FileParser._parse_pdf_sync frames what the PDF engine extracted (app/file_parser.py:166-212; the
 * extraction itself — MuPDF — is out of scope): per page "=== Page N ===\n" + page_text.strip() when that
 * is not empty, table blocks, image notes, parts joined by "\n\n", a fixed text when nothing is left.
 * The staged batch holds the PIECES (page texts, table blocks, literal headers) in output order;
 * piece_strip[i] = 1 applies str.strip() (Unicode whitespace) to piece i.  Pieces form parts
 * (part_first_piece, n_parts + 1 entries), parts form output documents (doc_first_part, n_out + 1 entries):
 *   RFQ_PART_ALWAYS    printed as it is
 *   RFQ_PART_IF_TEXT   printed only if its LAST piece is not empty after stripping (a page header + its text)
 *   RFQ_PART_FALLBACK  printed only if no other part of the document is
 * Printed parts of a document are joined by "\n\n".  Output: n_out texts. */
#define RFQ_PART_ALWAYS 0
#define RFQ_PART_IF_TEXT 1
#define RFQ_PART_FALLBACK 2
int rfq_frame_run(rfq_ctx* ctx, uint32_t n_out, const uint32_t* doc_first_part, uint32_t n_parts,
                  const uint8_t* part_kind, const uint32_t* part_first_piece, const uint8_t* piece_strip);

/* ---- device-side DEFLATE: the zip members of .xlsx containers inflated on the GPU —
 * the step pd.read_excel does with zipfile + zlib inside app/file_parser.py:225.  The staged batch holds
 * the members' RAW data as it lies in the container (the host walks the zip directory: O(members));
 * method[d] is the zip compression method (0 stored, 8 deflate), out_len[d] the uncompressed size and
 * crc32[d] the CRC-32 from the directory.  Output = the inflated members (rfq_out_text etc.), status
 * RFQ_DOC_OK or RFQ_DOC_CORRUPT.
 * rfq_restage_output makes the inflated members the staged batch without leaving the device (each starts
 * on a 16-byte boundary; corrupt members become empty documents), so rfq_sheets_run / rfq_text_run /
 * rfq_mixed_run (text and sheet kinds) follow directly; rfq_fetch_tables downloads only the per-document
 * tables of the last run (statuses, lengths), not its text. */
int rfq_inflate_run(rfq_ctx* ctx, const uint8_t* method, const uint64_t* out_len, const uint32_t* crc32);
int rfq_restage_output(rfq_ctx* ctx);
int rfq_fetch_tables(rfq_ctx* ctx);

/* ---- device → host copy of the last run's output (plumbing of the boundary: no reference counterpart) ---- */
int rfq_fetch(rfq_ctx* ctx);

/* convenience: stage + run + fetch.  rfq_text_parse = _parse_txt_async (app/file_parser.py:143-148),
 * rfq_csv_parse = _parse_csv_sync (app/file_parser.py:288-313), rfq_prompt_build = the prompt assembly of
 * app/rfq_agent.py:138-151 */
int rfq_prompt_build(rfq_ctx* ctx, const uint8_t* arena, const uint64_t* offsets,
                     const uint64_t* lengths, uint32_t n_docs, const uint8_t* head,
                     uint32_t head_len, const uint8_t* tail, uint32_t tail_len,
                     const uint8_t* mark, uint32_t mark_len, uint32_t max_chars);
int rfq_text_parse(rfq_ctx* ctx, const uint8_t* arena, const uint64_t* offsets,
                   const uint64_t* lengths, uint32_t n_docs);
int rfq_csv_parse(rfq_ctx* ctx, const uint8_t* arena, const uint64_t* offsets,
                  const uint64_t* lengths, uint32_t n_docs, const int64_t* max_rows);

/* ---- results of the last run (host pointers, valid after rfq_fetch): what the reference returns as
 * result["raw_text"] (app/file_parser.py:135-141), per document, plus the status a binding maps to the
 * reference's exceptions ---- */
uint32_t rfq_out_count(const rfq_ctx* ctx);
const uint8_t* rfq_out_text(const rfq_ctx* ctx);     /* UTF-8, all documents */
const uint64_t* rfq_out_offsets(const rfq_ctx* ctx); /* [n] start of document d in out_text */
const uint64_t* rfq_out_lengths(const rfq_ctx* ctx); /* [n] bytes of document d */
const int32_t* rfq_out_status(const rfq_ctx* ctx);   /* [n] RFQ_DOC_* */
uint64_t rfq_out_total_bytes(const rfq_ctx* ctx);
/* per-document table shape of the last CSV / sheet run: rows (after dropna) and columns */
const uint32_t* rfq_out_rows(const rfq_ctx* ctx);
const uint32_t* rfq_out_cols(const rfq_ctx* ctx);
/* CSV: winning attempt, encoding index (0 utf-8, 1 latin-1, 2 cp1252) * 3 + separator index */
const uint8_t* rfq_out_attempt(const rfq_ctx* ctx);

/* ---- a fetched text as fixed-width code units, for bindings whose string type stores text that way
 * (CPython: 1, 2 or 4 bytes per code point).  rfq_out_maxchar: an upper bound of the largest code point
 * of document d (0x7F, 0xFF, 0xFFFF or 0x10FFFF).  rfq_out_widen: its n_chars code points written to dst
 * as 1-, 2- or 4-byte units (n_chars must be the document's code-point count: rfq_out_chars).  Host code,
 * a few threads: the conversion at the language boundary, not part of the parse. */
int rfq_out_maxchar(const rfq_ctx* ctx, uint32_t d, uint32_t* maxchar);
int rfq_out_widen(const rfq_ctx* ctx, uint32_t d, int unit, void* dst, uint64_t n_chars);
/* The same two steps for many documents of the fetched result at once (the texts of a batch of sheets, bodies or
 * small files), spread over host threads: n_chars[i] / maxchar[i] of document docs[i]; then every text widened
 * into dsts[i] (units[i] = 1, 2 or 4 bytes per character, room for n_chars[i] characters). */
int rfq_out_plan_many(const rfq_ctx* ctx, uint32_t n_list, const uint32_t* docs, uint64_t* n_chars, uint32_t* maxchar);
int rfq_out_widen_many(const rfq_ctx* ctx, uint32_t n_list, const uint32_t* docs, const uint8_t* units, void* const* dsts,
                       const uint64_t* n_chars);

/* ---- measurement helpers (bench.py): device time of the last *_run, and launch counters (no reference
 * counterpart) ---- */
float rfq_last_run_ms(const rfq_ctx* ctx);
uint64_t rfq_launch_count(const rfq_ctx* ctx);
/* name/ms of the dominant kernel family of the last CSV run, timed with CUDA events on the
 * ctx stream when profiling is enabled with rfq_set_profile(ctx, 1) */
int rfq_set_profile(rfq_ctx* ctx, int on);
int rfq_profile_count(const rfq_ctx* ctx);
const char* rfq_profile_name(const rfq_ctx* ctx, int i);
float rfq_profile_ms(const rfq_ctx* ctx, int i);

/* ---- text path without any GPU memory allocation -------------------------------------------
 * No counterpart in the reference: added for embedders that must not allocate on the GPU — a
 * plugin living inside another process's memory budget (a serving stack whose allocator owns the
 * whole device), a handler on a latency-critical path where cudaMalloc's implicit device-wide
 * synchronisation is unacceptable, or a host that froze its memory map after start-up.
 *
 * Same result as the text path above (UTF-8 decode errors='ignore', universal newlines,
 * str.strip(); app/file_parser.py:143-148) for ONE document of at most RFQ_NOALLOC_MAX_BYTES
 * bytes; larger input is refused with RFQ_ERR_TOO_LARGE (and a message on stderr).  All scratch
 * lives in a fixed RFQ_NOALLOC_SCRATCH_BYTES array inside the library image, so every GPU that
 * loads the library carries that much until the process exits, and the call performs no
 * cudaMalloc / cudaFree / allocating library call of any kind.
 *
 *   device      CUDA device ordinal the pointers belong to
 *   d_in        DEVICE pointer, 16-byte aligned, readable up to len rounded up to 16 bytes
 *   len         document bytes
 *   d_out       DEVICE pointer with room for len bytes (the result is never longer than the input)
 *   d_out_len   DEVICE pointer, 8-byte aligned: receives the number of bytes written to d_out
 *   cuda_stream stream to run on (NULL = the default stream); the call only enqueues work —
 *               synchronise the stream before reading d_out / d_out_len
 *
 * The first call on each device also resolves the device address of the scratch array (a
 * one-off, host-side set-up step), so make that first call before starting any stream capture.
 * Every call on a device shares the one array: calls whose GPU work can overlap (different
 * streams, or different threads without ordering between them) race on it — serialise them. */
#define RFQ_NOALLOC_MAX_BYTES (4u << 20)
#define RFQ_NOALLOC_SCRATCH_BYTES (96u << 10)
int rfq_text_normalise_noalloc(int device, const uint8_t* d_in, uint64_t len, uint8_t* d_out,
                               uint64_t* d_out_len, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* RFQ_PARSE_H_ */
