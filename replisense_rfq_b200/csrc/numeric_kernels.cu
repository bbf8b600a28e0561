// Quantity / price field parser: the validators the reference applies to the
// line items that come back from the LLM (app/rfq_agent.py:27-39):
//   quantity      int(v.replace(",", ""))
//   target_price  float(v.replace(",", ""))
// One thread per field: drop the commas, strip white space, check Python's literal grammar
// (one optional sign, digits with single '_' between digits, for floats a fraction, an exponent and
// the words inf / infinity / nan), then an exact integer or a correctly rounded double.
// Results: 8 bytes per field (int64, or the bits of the double) + a status per field.
// Declined (never guessed): non-ASCII input (Python also accepts other scripts' digits and
// blanks), integers beyond int64, literals longer than the scratch buffer.
#include <cuda_runtime.h>
#include "../../include/rfq_parse.h"
#include "cell_ops.cuh"
#include "launch.h"
#include "pipeline.h"

namespace rfq {

#define NUM_MAX_LITERAL 160

// What int() / float() strip from an ASCII str: Py_ISSPACE (\t \n \v \f \r and the blank).  The
// wider str.isspace() set (\x1c-\x1f, U+0085, ...) only applies to non-ASCII strings, which go
// through _PyUnicode_TransformDecimalAndSpaceToASCII first — and are declined here.
__device__ __forceinline__ bool py_space(uint32_t c) { return (c >= 0x09 && c <= 0x0D) || c == 0x20; }
__device__ __forceinline__ bool dig(uint32_t c) { return c - '0' < 10u; }
__device__ __forceinline__ uint32_t lower(uint32_t c) { return (c >= 'A' && c <= 'Z') ? c + 32 : c; }

// copies s[0..n) without commas, stripped, into buf; returns length, or -1 (non-ASCII), -2 (too long)
__device__ int clean_literal(const uint8_t* s, uint32_t n, uint8_t* buf) {
  uint32_t a = 0, b = n;
  // the commas go first (replace), then strip: a comma between blanks does not stop the strip
  int k = 0;
  bool lead = true;
  for (uint32_t i = a; i < b; ++i) {
    uint32_t c = s[i];
    if (c >= 0x80) return -1;
    if (c == ',') continue;
    if (lead && py_space(c)) continue;
    lead = false;
    if (k >= NUM_MAX_LITERAL) return -2;
    buf[k++] = (uint8_t)c;
  }
  while (k > 0 && py_space(buf[k - 1])) --k;
  return k;
}

// removes '_' that sit between two digits; any other '_' is a syntax error (returns -1)
__device__ int drop_underscores(uint8_t* buf, int k) {
  int o = 0;
  for (int i = 0; i < k; ++i) {
    if (buf[i] == '_') {
      if (i == 0 || i + 1 >= k || !dig(buf[i - 1]) || !dig(buf[i + 1])) return -1;
      continue;
    }
    buf[o++] = buf[i];
  }
  return o;
}

__global__ void __launch_bounds__(128) numeric_kernel(const uint8_t* __restrict__ arena,
                                                      const uint32_t* __restrict__ doc_off,
                                                      const uint32_t* __restrict__ doc_len,
                                                      const uint8_t* __restrict__ kind, uint32_t n_docs,
                                                      unsigned long long* __restrict__ value,
                                                      int32_t* __restrict__ status,
                                                      unsigned long long* __restrict__ out_off,
                                                      unsigned long long* __restrict__ out_len) {
  uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= n_docs) return;
  out_off[d] = 8ull * d;
  out_len[d] = 8;
  value[d] = 0;
  uint8_t buf[NUM_MAX_LITERAL];
  int k = clean_literal(arena + doc_off[d], doc_len[d], buf);
  if (k < 0) { status[d] = RFQ_DOC_UNSUPPORTED; return; }
  const bool is_float = kind[d] != 0;
  if (is_float) {
    // words: [+-]? (inf | infinity | nan), any case
    int i = 0;
    bool neg = false;
    if (i < k && (buf[i] == '+' || buf[i] == '-')) { neg = buf[i] == '-'; ++i; }
    int wl = k - i;
    if (wl == 3 || wl == 8) {
      const char* inf = "infinity";
      bool is_inf = true, is_nan = wl == 3;
      for (int q = 0; q < wl; ++q) if (lower(buf[i + q]) != (uint32_t)inf[q]) is_inf = false;
      if (is_nan) { const char* nn = "nan"; for (int q = 0; q < 3; ++q) if (lower(buf[i + q]) != (uint32_t)nn[q]) is_nan = false; }
      if (is_inf) { value[d] = neg ? 0xFFF0000000000000ull : 0x7FF0000000000000ull; status[d] = RFQ_DOC_OK; return; }
      if (is_nan) { value[d] = neg ? 0xFFF8000000000000ull : 0x7FF8000000000000ull; status[d] = RFQ_DOC_OK; return; }
    }
  }
  k = drop_underscores(buf, k);
  if (k <= 0) { status[d] = RFQ_DOC_VALUE_ERROR; return; }
  if (!is_float) {
    int i = 0;
    bool neg = false;
    if (buf[i] == '+' || buf[i] == '-') { neg = buf[i] == '-'; ++i; }
    if (i >= k) { status[d] = RFQ_DOC_VALUE_ERROR; return; }
    unsigned long long mag = 0;
    const unsigned long long lim = neg ? 9223372036854775808ull : 9223372036854775807ull;
    bool big = false;
    for (; i < k; ++i) {
      if (!dig(buf[i])) { status[d] = RFQ_DOC_VALUE_ERROR; return; }
      uint32_t dg = buf[i] - '0';
      if (!big && mag > (lim - dg) / 10) big = true;
      if (!big) mag = mag * 10 + dg;
    }
    if (big) { status[d] = RFQ_DOC_UNSUPPORTED; return; }   // a Python int beyond int64
    value[d] = neg ? 0ull - mag : mag;
    status[d] = RFQ_DOC_OK;
    return;
  }
  // float literal: Python also takes "1." and ".5" (strtod_exact does); a lone "." has no digit
  double v;
  if (!strtod_exact(buf, (uint32_t)k, &v)) { status[d] = RFQ_DOC_VALUE_ERROR; return; }
  value[d] = dbl_bits(v);
  status[d] = RFQ_DOC_OK;
}

void numeric_batch(const uint8_t* arena, const uint32_t* doc_off, const uint32_t* doc_len,
                   const uint8_t* kind, uint32_t n_docs, unsigned long long* value, int32_t* status,
                   unsigned long long* out_off, unsigned long long* out_len, cudaStream_t s) {
  if (n_docs == 0) return;
  prof_mark("numeric_kernel", s);
  RFQ_LAUNCH(numeric_kernel, (n_docs + 127) / 128, 128, 0, s, arena, doc_off, doc_len, kind, n_docs,
             value, status, out_off, out_len);
}

}  // namespace rfq
