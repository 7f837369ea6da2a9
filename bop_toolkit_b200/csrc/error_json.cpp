// Error-file writer of the drop-in driver (host code, no CUDA).
//
// Writes exactly the text `inout.save_json` of the reference produces for the record lists of
// scripts/eval_calc_errors.py:380-418 (bop_toolkit_lib/inout.py:85-113: one element per line,
// `json.dumps(elem, sort_keys=True)`), straight from the driver's flat arrays.  A YCB-V-size VSD
// run writes 1.6 M records into 120 files; json.dumps (and any Python formatting) holds the GIL
// for seconds, this call does not.  Floats are formatted like Python's float.__repr__ (shortest
// round-trip digits, exponent notation iff decpt <= -4 or decpt > 16; json's Infinity / NaN).
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

#include "../../include/bop_b200.h"

int bop_set_error_msg(int code, const char* msg);

namespace {

// float.__repr__(v) as json.dumps writes it.  Returns the number of chars written (<= 32).
inline int py_repr(double v, char* out) {
  if (std::isnan(v)) { memcpy(out, "NaN", 3); return 3; }
  if (std::isinf(v)) {
    if (v < 0) { memcpy(out, "-Infinity", 9); return 9; }
    memcpy(out, "Infinity", 8);
    return 8;
  }
  char* p = out;
  if (std::signbit(v)) { *p++ = '-'; v = -v; }
  if (v == 0.0) { memcpy(p, "0.0", 3); return (int)(p - out) + 3; }
  // shortest round-trip digits in scientific form: d[.ddd]e[+-]XX
  char sci[40];
  const auto res = std::to_chars(sci, sci + sizeof(sci), v, std::chars_format::scientific);
  const char* e = sci;
  while (e < res.ptr && *e != 'e') ++e;
  char digits[24];
  int nd = 0;
  for (const char* c = sci; c < e; ++c)
    if (*c != '.') digits[nd++] = *c;
  int exp10 = 0;
  {
    const char* c = e + 1;
    const bool neg = *c == '-';
    if (*c == '-' || *c == '+') ++c;
    for (; c < res.ptr; ++c) exp10 = exp10 * 10 + (*c - '0');
    if (neg) exp10 = -exp10;
  }
  const int decpt = exp10 + 1;          // value = 0.d1d2... x 10^decpt
  if (decpt <= -4 || decpt > 16) {      // Python's repr switches to exponent notation here
    *p++ = digits[0];
    if (nd > 1) {
      *p++ = '.';
      memcpy(p, digits + 1, (size_t)(nd - 1));
      p += nd - 1;
    }
    *p++ = 'e';
    int x = decpt - 1;
    *p++ = x < 0 ? '-' : '+';
    if (x < 0) x = -x;
    char tmp[8];
    int nt = 0;
    do { tmp[nt++] = (char)('0' + x % 10); x /= 10; } while (x);
    if (nt < 2) tmp[nt++] = '0';
    while (nt) *p++ = tmp[--nt];
  } else if (decpt <= 0) {
    *p++ = '0'; *p++ = '.';
    for (int k = 0; k < -decpt; ++k) *p++ = '0';
    memcpy(p, digits, (size_t)nd);
    p += nd;
  } else if (decpt >= nd) {
    memcpy(p, digits, (size_t)nd);
    p += nd;
    for (int k = 0; k < decpt - nd; ++k) *p++ = '0';
    *p++ = '.'; *p++ = '0';
  } else {
    memcpy(p, digits, (size_t)decpt);
    p += decpt;
    *p++ = '.';
    memcpy(p, digits + decpt, (size_t)(nd - decpt));
    p += nd - decpt;
  }
  return (int)(p - out);
}

inline void put_int(std::string& s, long long v) {
  char b[24];
  const auto r = std::to_chars(b, b + sizeof(b), v);
  s.append(b, (size_t)(r.ptr - b));
}
inline void put_float(std::string& s, double v) {
  char b[40];
  s.append(b, (size_t)py_repr(v, b));
}

}  // namespace

extern "C" int bop_format_double_host(double v, char* out32) {
  if (!out32) return bop_set_error_msg(-1, "bop_format_double_host: null pointer");
  const int n = py_repr(v, out32);
  out32[n] = 0;
  return 0;
}

extern "C" int bop_write_error_file_host(const char* path, int64_t n_rec, const int64_t* im_id,
                                         const int64_t* obj_id, const int64_t* est_id,
                                         const double* score, int64_t n_pairs,
                                         const int64_t* pair_rec, const int64_t* pair_gt,
                                         const double* errs, int n_cols, int col) {
  if (!path || n_rec < 0 || n_pairs < 0 || n_cols < 1 || col >= n_cols ||
      (n_rec > 0 && !(im_id && obj_id && est_id && score)) ||
      (n_pairs > 0 && !(pair_rec && pair_gt && errs)))
    return bop_set_error_msg(-1, "bop_write_error_file_host: bad arguments");
  std::string s;
  s.reserve((size_t)n_rec * 110 + (size_t)n_pairs * (size_t)(col < 0 ? n_cols : 1) * 22 + 16);
  s += "[\n";
  int64_t k = 0;
  for (int64_t r = 0; r < n_rec; ++r) {
    s += "  {\"errors\": {";
    bool first = true;
    while (k < n_pairs && pair_rec[k] == r) {
      if (k > 0 && pair_rec[k - 1] == r && pair_gt[k - 1] >= pair_gt[k])
        return bop_set_error_msg(-1, "bop_write_error_file_host: GT ids of a record must ascend");
      if (!first) s += ", ";
      first = false;
      s += '"';
      put_int(s, pair_gt[k]);
      s += "\": [";
      const double* e = errs + (size_t)k * (size_t)n_cols;
      if (col >= 0) {
        put_float(s, e[col]);
      } else {
        for (int c = 0; c < n_cols; ++c) {
          if (c) s += ", ";
          put_float(s, e[c]);
        }
      }
      s += ']';
      ++k;
    }
    if (k < n_pairs && pair_rec[k] < r)
      return bop_set_error_msg(-1, "bop_write_error_file_host: pairs must be sorted by record");
    s += "}, \"est_id\": ";
    put_int(s, est_id[r]);
    s += ", \"im_id\": ";
    put_int(s, im_id[r]);
    s += ", \"obj_id\": ";
    put_int(s, obj_id[r]);
    s += ", \"score\": ";
    put_float(s, score[r]);
    s += r + 1 < n_rec ? "},\n" : "}\n";
  }
  if (k != n_pairs) return bop_set_error_msg(-1, "bop_write_error_file_host: pair_rec out of range");
  s += ']';
  FILE* f = fopen(path, "wb");
  if (!f) return bop_set_error_msg(-2, "bop_write_error_file_host: cannot open the output file");
  const size_t w = fwrite(s.data(), 1, s.size(), f);
  if (fclose(f) != 0 || w != s.size())
    return bop_set_error_msg(-2, "bop_write_error_file_host: short write");
  return 0;
}
