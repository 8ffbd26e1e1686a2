// b200ps N6 — native libsvm text parser.  The reference parses every line in CPython
// (Sample.parse_line_libsvm, lr2.py:57-67) and then rebuilds python lists per batch
// (format_samples_sparse, lr2.py:69-84); that loop is its real CPU hot spot.  This parser
// turns a whole text buffer into CSR arrays in one pass, multi-buffer safe (no globals).
//
// Grammar per line:  <int label> (SEP <int index> ':' <float value>)*   with SEP = ' ' | '\t'
// Lines shorter than 2 characters are skipped, as in lr2.py:105.
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>

#include <climits>

#include "fast_parse.h"

namespace {
using b200ps_parse::fast_float;
using b200ps_parse::fast_int;
using b200ps_parse::is_sep;

// The common line — plain decimal label, ids and values — straight from the buffer, no copy, no libc call
// (fast_parse.h: bit-identical to strtol / strtoll / strtof or it declines).  Returns the number of non-zeros written
// at fid / fval, -1 when anything about the line is unusual (the caller then runs the general path on the whole
// line), -2 when the line does not fit the arrays.
inline int64_t parse_line_fast(const char* p, const char* le, float* label, int64_t* fid, float* fval, uint64_t room) {
  const char* q = p;
  int64_t lab;
  if (!fast_int(q, le, &lab) || !b200ps_parse::is_end(q, le)) return -1;
  uint64_t n = 0;
  while (q < le) {
    while (q < le && (is_sep(*q) || *q == '\r')) ++q;
    if (q >= le) break;
    int64_t idx;
    if (!fast_int(q, le, &idx) || q >= le || *q != ':') return -1;
    ++q;
    float val;
    if (!fast_float(q, le, &val)) return -1;
    if (n >= room) return -2;
    fid[n] = idx;
    fval[n] = val;
    ++n;
  }
  *label = static_cast<float>(lab);
  return static_cast<int64_t>(n);
}

// strtol / strtoll / strtof need a terminator and skip leading white space INCLUDING '\n'; a line is only a
// (pointer, length) range of a buffer that may be an mmap or a C caller's array.  So every line is parsed from a
// NUL-terminated copy: short lines on the stack, long ones in a reusable heap block.
struct LineCopy {
  char stack[4096];
  char* heap = nullptr;
  size_t heap_cap = 0;
  ~LineCopy() { free(heap); }
  const char* copy(const char* p, size_t n) {
    char* dst = stack;
    if (n + 1 > sizeof(stack)) {
      if (n + 1 > heap_cap) {
        free(heap);
        heap_cap = (n + 1) * 2;
        heap = static_cast<char*>(malloc(heap_cap));
        if (!heap) { heap_cap = 0; return nullptr; }
      }
      dst = heap;
    }
    memcpy(dst, p, n);
    dst[n] = 0;
    return dst;
  }
};
}

// Pass 1: count rows and non-zeros so the caller can size the arrays.
extern "C" int b200ps_libsvm_count(const char* buf, uint64_t len, uint64_t* rows, uint64_t* nnz) {
  uint64_t r = 0, z = 0;
  const char* p = buf;
  const char* end = buf + len;
  while (p < end) {
    const char* nl = static_cast<const char*>(memchr(p, '\n', end - p));
    const char* le = nl ? nl : end;
    if (le - p >= 2) {
      ++r;
      uint32_t c = 0;                       // a byte-compare-and-add loop the compiler vectorises
      for (const char* q = p; q < le; ++q) c += (*q == ':');
      z += c;
    }
    p = nl ? nl + 1 : end;
  }
  *rows = r;
  *nnz = z;
  return 0;
}

// Pass 2: fill labels[rows], row_ptr[rows+1], fid[nnz], fval[nnz].
// Returns 0, or -(line number) of the first malformed line.
extern "C" int64_t b200ps_libsvm_parse(const char* buf, uint64_t len, float* labels, int32_t* row_ptr,
                                       int64_t* fid, float* fval, uint64_t max_rows, uint64_t max_nnz) {
  uint64_t r = 0, z = 0;
  int64_t line_no = 0;
  const char* p = buf;
  const char* end = buf + len;
  LineCopy lc;
  row_ptr[0] = 0;
  while (p < end) {
    const char* nl = static_cast<const char*>(memchr(p, '\n', end - p));
    const char* line_end = nl ? nl : end;
    ++line_no;
    if (line_end - p >= 2) {
      if (r >= max_rows) return -line_no;
      {
        const uint64_t cap = max_nnz < static_cast<uint64_t>(INT32_MAX) ? max_nnz : static_cast<uint64_t>(INT32_MAX);
        const int64_t n = parse_line_fast(p, line_end, labels + r, fid + z, fval + z, cap - z);
        if (n == -2) return -line_no;
        if (n >= 0) {
          z += static_cast<uint64_t>(n);
          ++r;
          row_ptr[r] = static_cast<int32_t>(z);
          p = nl ? nl + 1 : end;
          continue;
        }
      }
      // anything unusual (exponents, inf / nan, long digit strings, malformed text): the general path
      const char* q = lc.copy(p, static_cast<size_t>(line_end - p));   // NUL-terminated: no strto* call can leave it
      if (!q) return -line_no;
      const char* le = q + (line_end - p);
      char* e = nullptr;
      if (is_sep(*q) || *q == '\r') return -line_no;      // the label starts the line
      const long lab = strtol(q, &e, 10);
      if (e == q) return -line_no;
      labels[r] = static_cast<float>(lab);
      q = e;
      while (q < le) {
        while (q < le && (is_sep(*q) || *q == '\r')) ++q;
        if (q >= le) break;
        const long long idx = strtoll(q, &e, 10);
        if (e == q || e >= le || *e != ':') return -line_no;
        q = e + 1;
        if (q >= le || is_sep(*q) || *q == '\r') return -line_no;   // "idx:" with no value
        const float val = strtof(q, &e);
        if (e == q || e > le) return -line_no;
        q = e;
        if (z >= max_nnz || z >= static_cast<uint64_t>(INT32_MAX)) return -line_no;   // row_ptr is int32
        fid[z] = idx;
        fval[z] = val;
        ++z;
      }
      ++r;
      row_ptr[r] = static_cast<int32_t>(z);
    }
    p = nl ? nl + 1 : end;
  }
  return 0;
}
