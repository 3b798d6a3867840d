// text_table.cpp -- multithreaded parser for the numeric body of phin / phout text files.
//
// Host-side replacement for the reference's  pd.read_csv(path, skiprows=skip, header=None, sep="[ ]+",
// engine="python")  (SimuBox/toolkit/reader.py:170-172; 0.40 s per 37 k rows).
// The file body is one row per grid point, blank-separated decimal numbers, every row the same width.
// Tokens are converted with the arithmetic of the reference's own conversion (parse_token below).  The buffer is cut into per-thread pieces at line ends; pass 1 counts the rows of each
// piece (for the output offsets), pass 2 converts.
#include <algorithm>
#include <atomic>
#include <charconv>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "common.h"

namespace sbx {
namespace {

inline bool is_blank(char c) { return c == ' ' || c == '\t' || c == '\r' || c == '\v' || c == '\f'; }

// first byte after `n` line ends (or `end`)
const char* skip_lines(const char* p, const char* end, int n) {
  for (; n > 0 && p < end; --n) {
    const void* nl = memchr(p, '\n', (size_t)(end - p));
    p = nl ? static_cast<const char*>(nl) + 1 : end;
  }
  return p;
}

// [p, line_end) of the line starting at p
inline const char* line_end(const char* p, const char* end) {
  const void* nl = memchr(p, '\n', (size_t)(end - p));
  return nl ? static_cast<const char*>(nl) : end;
}

inline bool blank_line(const char* p, const char* e) {
  for (; p < e; ++p)
    if (!is_blank(*p)) return false;
  return true;
}

int count_tokens(const char* p, const char* e) {
  int n = 0;
  while (p < e) {
    while (p < e && is_blank(*p)) ++p;
    if (p == e) break;
    ++n;
    while (p < e && !is_blank(*p)) ++p;
  }
  return n;
}

int64_t count_rows(const char* p, const char* end) {
  int64_t rows = 0;
  while (p < end) {
    const char* e = line_end(p, end);
    if (!blank_line(p, e)) ++rows;
    p = e + 1;
  }
  return rows;
}

// 1e0 .. 1e308, correctly rounded literals
struct Pow10 {
  double e[309];
  Pow10() {
    for (int i = 0; i <= 308; ++i) {
      char lit[8];
      const int n = snprintf(lit, sizeof lit, "1e%d", i);
      std::from_chars(lit, lit + n, e[i]);
    }
  }
};
const Pow10 kPow10;

inline bool digit(char c) { return c >= '0' && c <= '9'; }
inline char lower(char c) { return (c >= 'A' && c <= 'Z') ? (char)(c + 32) : c; }

bool word_is(const char* p, const char* e, const char* w) {
  for (; *w; ++w, ++p)
    if (p == e || lower(*p) != *w) return false;
  return p == e;
}

// One token -> double with the arithmetic of the reference's parse.  The pandas python engine converts a column
// with lib.maybe_convert_numeric -> floatify -> precise_xstrtod (pandas/_libs/src/parser/tokenizer.c; pandas is a
// pinned dependency of the reference, pyproject / poetry.lock, not vendored): at most 17 digits -- leading zeros
// included -- are accumulated in a double, later digits only move the exponent (integer part) or are dropped
// (fraction), then ONE multiplication or division by a tabulated power of ten.  Correctly rounded for <= 15
// significant digits and |exponent| <= 22 (every "%lf" / "%.10e" file the SCFT codes write); for longer mantissas
// it is the reference's value, which may be one ulp from the correctly rounded one.  tests/test_text_table.py
// checks this restatement against the live pandas call.
inline bool parse_token(const char* p, const char* e, double* out) {
  bool neg = false;
  if (p < e && (*p == '+' || *p == '-')) {
    neg = *p == '-';
    ++p;
  }
  if (p == e) return false;
  if (!digit(*p) && *p != '.') {
    double v;
    if (word_is(p, e, "inf") || word_is(p, e, "infinity")) v = __builtin_inf();
    else if (word_is(p, e, "nan")) v = __builtin_nan("");
    else return false;
    *out = neg ? -v : v;
    return true;
  }
  constexpr int kMaxDigits = 17;
  double number = 0.0;
  int exponent = 0, num_digits = 0, num_decimals = 0;
  for (; p < e && digit(*p); ++p) {
    if (num_digits < kMaxDigits) {
      number = number * 10.0 + (*p - '0');
      ++num_digits;
    } else {
      ++exponent;
    }
  }
  if (p < e && *p == '.') {
    ++p;
    for (; p < e && digit(*p) && num_digits < kMaxDigits; ++p) {
      number = number * 10.0 + (*p - '0');
      ++num_digits;
      ++num_decimals;
    }
    while (p < e && digit(*p)) ++p;
    exponent -= num_decimals;
  }
  if (num_digits == 0) return false;
  if (neg) number = -number;
  if (p < e && (*p == 'e' || *p == 'E')) {
    ++p;
    bool eneg = false;
    if (p < e && (*p == '+' || *p == '-')) {
      eneg = *p == '-';
      ++p;
    }
    int n = 0, nd = 0;
    for (; p < e && digit(*p) && nd < kMaxDigits; ++p, ++nd) n = n < 100000 ? n * 10 + (*p - '0') : n;
    if (nd == 0) return false;
    exponent += eneg ? -n : n;
  }
  if (p != e) return false;
  if (exponent > 308) number = number == 0.0 ? number : (number < 0 ? -__builtin_inf() : __builtin_inf());
  else if (exponent > 0) number *= kPow10.e[exponent];
  else if (exponent < -308) {
    if (exponent < -616) number = 0.0;
    else {
      number /= kPow10.e[-308 - exponent];
      number /= kPow10.e[308];
    }
  } else number /= kPow10.e[-exponent];
  *out = number;
  return true;
}

// cut [begin, end) into `parts` pieces that start at line starts
std::vector<const char*> cut_at_lines(const char* begin, const char* end, int parts) {
  std::vector<const char*> cuts(parts + 1, end);
  cuts[0] = begin;
  const int64_t len = end - begin;
  for (int t = 1; t < parts; ++t) {
    const char* p = begin + len * t / parts;
    if (p < cuts[t - 1]) p = cuts[t - 1];
    const char* e = line_end(p, end);
    cuts[t] = e < end ? e + 1 : end;
  }
  return cuts;
}

int pick_threads(int threads, int64_t len) {
  int hw = (int)std::thread::hardware_concurrency();
  if (hw < 1) hw = 1;
  int t = threads > 0 ? threads : hw;
  const int64_t by_size = len / (256 * 1024) + 1;  // below 256 KB per thread the spawn costs more than it saves
  if (t > by_size) t = (int)by_size;
  return std::max(1, std::min(t, 256));
}

template <class F>
void run_parts(int parts, F&& fn) {
  if (parts == 1) {
    fn(0);
    return;
  }
  std::vector<std::thread> pool;
  pool.reserve(parts);
  for (int t = 0; t < parts; ++t) pool.emplace_back([&fn, t] { fn(t); });
  for (auto& th : pool) th.join();
}

}  // namespace
}  // namespace sbx

using namespace sbx;

extern "C" int sbx_text_is_ascii(const char* buf, int64_t len) {
  if (!buf && len > 0) return fail(SBX_EINVAL, "NULL buffer");
  const unsigned char* p = reinterpret_cast<const unsigned char*>(buf);
  int64_t i = 0;
  uint64_t acc = 0;
  for (; i + 8 <= len; i += 8) {
    uint64_t w;
    memcpy(&w, p + i, 8);
    acc |= w;
  }
  for (; i < len; ++i) acc |= (uint64_t)p[i] << 0;
  return (acc & 0x8080808080808080ull) ? 0 : 1;
}

extern "C" int sbx_text_table_shape(const char* buf, int64_t len, int skip, int threads, int64_t* rows, int32_t* cols) {
  if (!buf || len < 0 || !rows || !cols) return fail(SBX_EINVAL, "NULL argument");
  const char* end = buf + len;
  const char* body = skip_lines(buf, end, skip);
  *rows = 0;
  *cols = 0;
  for (const char* p = body; p < end;) {  // width = tokens of the first non-blank row
    const char* e = line_end(p, end);
    if (!blank_line(p, e)) {
      *cols = count_tokens(p, e);
      break;
    }
    p = e + 1;
  }
  const int parts = pick_threads(threads, end - body);
  const auto cuts = cut_at_lines(body, end, parts);
  std::vector<int64_t> n(parts, 0);
  run_parts(parts, [&](int t) { n[t] = count_rows(cuts[t], cuts[t + 1]); });
  for (int t = 0; t < parts; ++t) *rows += n[t];
  return SBX_OK;
}

extern "C" int sbx_text_table_parse(const char* buf, int64_t len, int skip, int64_t rows, int32_t cols, double* out,
                                    int threads) {
  if (!buf || len < 0 || !out) return fail(SBX_EINVAL, "NULL argument");
  if (rows < 0 || cols < 1) return fail(SBX_EINVAL, "bad table shape %lld x %d", (long long)rows, cols);
  const char* end = buf + len;
  const char* body = skip_lines(buf, end, skip);
  const int parts = pick_threads(threads, end - body);
  const auto cuts = cut_at_lines(body, end, parts);
  std::vector<int64_t> first(parts + 1, 0);
  {
    std::vector<int64_t> n(parts, 0);
    run_parts(parts, [&](int t) { n[t] = count_rows(cuts[t], cuts[t + 1]); });
    for (int t = 0; t < parts; ++t) first[t + 1] = first[t] + n[t];
  }
  if (first[parts] != rows) return fail(SBX_EINVAL, "table has %lld rows, caller expected %lld", (long long)first[parts], (long long)rows);
  std::atomic<int64_t> bad_row{-1};  // smallest offending row (0-based among the data rows)
  std::atomic<int> bad_kind{0};      // 1 = wrong number of columns, 2 = not a number
  run_parts(parts, [&](int t) {
    int64_t row = first[t];
    for (const char* p = cuts[t]; p < cuts[t + 1];) {
      const char* e = line_end(p, cuts[t + 1]);
      const char* q = p;
      p = e + 1;
      if (blank_line(q, e)) continue;
      double* dst = out + row * cols;
      int c = 0, kind = 0;
      while (q < e) {
        while (q < e && is_blank(*q)) ++q;
        if (q == e) break;
        const char* tok = q;
        while (q < e && !is_blank(*q)) ++q;
        if (c >= cols) { kind = 1; break; }
        if (!parse_token(tok, q, dst + c)) { kind = 2; break; }
        ++c;
      }
      if (!kind && c != cols) kind = 1;
      if (kind) {
        int64_t cur = bad_row.load();
        while ((cur < 0 || row < cur) && !bad_row.compare_exchange_weak(cur, row)) {}
        if (bad_row.load() == row) bad_kind.store(kind);
        return;
      }
      ++row;
    }
  });
  if (bad_row.load() >= 0)
    return fail(SBX_EINVAL, bad_kind.load() == 1 ? "data row %lld does not have %d columns" : "data row %lld holds a token that is not a number (%d columns)",
                (long long)bad_row.load(), cols);
  return SBX_OK;
}
