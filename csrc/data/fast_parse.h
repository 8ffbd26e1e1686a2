// b200ps — number parsing for the libsvm readers (libsvm_parse.cpp, loader.cpp): the int(...) / float(...) calls of
// Sample.parse_line_libsvm (lr2.py:57-67), which the reference runs once per feature in CPython.
//
// strtol / strtoll / strtof cost 50-150 ns per token, want a NUL terminator and skip leading white space including
// new lines, so a bounded (pointer, end) reader has to copy every line first.  Almost every token of a libsvm file is
// a plain decimal — "1739201", "0.731058", "-1" — and those convert exactly with a handful of integer operations:
//
//   value = m / 10^k   with m < 2^53 (at most 15 significant digits) and k <= 22: both operands are exact doubles,
//                      so the quotient is the correctly rounded double of the decimal
//   float(value)       equals the correctly rounded float of the decimal unless the double sits exactly on a float
//                      rounding boundary (then the decimal may lie on either side): that case, and everything else
//                      that is not a plain short decimal (exponents, inf / nan, hex, > 15 digits, float subnormals),
//                      returns false and the caller takes its general strtof path.
//
// So the fast path returns bit for bit what strtof returns, or declines.
#ifndef B200PS_FAST_PARSE_H_
#define B200PS_FAST_PARSE_H_

#include <cstdint>
#include <cstring>

namespace b200ps_parse {

inline bool is_sep(char c) { return c == ' ' || c == '\t'; }
inline bool is_end(const char* p, const char* le) { return p >= le || *p == ' ' || *p == '\t' || *p == '\r'; }

// [+-]?digits, at most 18 of them, ending anywhere.  On success q is advanced past the digits.
inline bool fast_int(const char*& q, const char* le, int64_t* out) {
  const char* p = q;
  bool neg = false;
  if (p < le && (*p == '-' || *p == '+')) { neg = *p == '-'; ++p; }
  const char* d0 = p;
  uint64_t v = 0;
  while (p < le && (unsigned)(*p - '0') <= 9u) { v = v * 10 + (uint64_t)(*p - '0'); ++p; }
  const int nd = (int)(p - d0);
  if (nd == 0 || nd > 18) return false;
  *out = neg ? -(int64_t)v : (int64_t)v;
  q = p;
  return true;
}

// [+-]?digits[.digits] ending at a separator, '\r' or le.  On success q is advanced past the number.
inline bool fast_float(const char*& q, const char* le, float* out) {
  static const double p10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
  const char* p = q;
  bool neg = false;
  if (p < le && (*p == '-' || *p == '+')) { neg = *p == '-'; ++p; }
  uint64_t m = 0;
  int sig = 0, nd = 0, nf = 0;              // significant digits (leading zeros dropped), integer / fraction digits
  while (p < le && (unsigned)(*p - '0') <= 9u) {
    m = m * 10 + (uint64_t)(*p - '0');
    sig += (m != 0);
    ++p; ++nd;
    if (sig > 15) return false;
  }
  if (p < le && *p == '.') {
    ++p;
    while (p < le && (unsigned)(*p - '0') <= 9u) {
      m = m * 10 + (uint64_t)(*p - '0');
      sig += (m != 0);
      ++p; ++nf;
      if (sig > 15 || nf > 22) return false;
    }
  }
  if (nd + nf == 0 || !is_end(p, le)) return false;
  const double d = (double)m / p10[nf];
  if (m != 0) {
    if (d < 1.1754943508222875e-38 || d > 3.4028234663852886e+38) return false;   // float subnormal / overflow
    uint64_t bits;
    std::memcpy(&bits, &d, sizeof(bits));
    if ((bits & 0x1FFFFFFFull) == 0x10000000ull) return false;                    // exactly on a float boundary
  }
  *out = (float)(neg ? -d : d);
  q = p;
  return true;
}

}  // namespace b200ps_parse

#endif  // B200PS_FAST_PARSE_H_
