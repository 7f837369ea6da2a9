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

// ---------------------------------------------------------------------------------------------
// Result-file reader: the numbers of a BOP19 results CSV (inout.load_bop_results,
// bop_toolkit_lib/inout.py:221-261: `scene_id,im_id,obj_id,score,R(9 space-sep),t(3 space-sep),
// time`, header line skipped) into flat arrays.  Only the plain forms are taken (digits, sign,
// decimal point, exponent, inf / nan as std::from_chars reads them); any line that Python's
// int() / float() might read differently -- or reject -- makes the call return 1 so that the
// caller parses the file line by line the reference's way and raises the reference's errors.
// ---------------------------------------------------------------------------------------------
namespace {
inline const char* skip_ws(const char* p, const char* e) {
  while (p < e && (*p == ' ' || *p == '\t' || *p == '\r')) ++p;
  return p;
}
inline bool read_i64(const char*& p, const char* e, int64_t& v) {
  p = skip_ws(p, e);
  const auto r = std::from_chars(p, e, v);
  if (r.ec != std::errc() || r.ptr == p) return false;
  p = skip_ws(r.ptr, e);
  return true;
}
inline bool read_f64(const char*& p, const char* e, double& v) {
  p = skip_ws(p, e);
  if (p < e && *p == '+') return false;
  const auto r = std::from_chars(p, e, v, std::chars_format::general);
  if (r.ec != std::errc() || r.ptr == p) return false;
  // "1.5x" / "1_0" / hex and the like: not ours to judge
  if (r.ptr < e && *r.ptr != ' ' && *r.ptr != '\t' && *r.ptr != '\r' && *r.ptr != ',') return false;
  p = skip_ws(r.ptr, e);
  return true;
}
inline bool eat(const char*& p, const char* e, char c) {
  if (p < e && *p == c) { ++p; return true; }
  return false;
}
}  // namespace

// Returns 0 and *n_out rows on success, 1 if the text holds anything irregular (nothing is
// guaranteed about the outputs then), < 0 on bad arguments.  Arrays hold `cap` rows.
extern "C" int bop_parse_results_csv_host(const char* text, size_t nbytes, int64_t cap,
                                          int64_t* ids, double* score, double* R, double* t,
                                          double* time, int64_t* n_out) {
  if (!text || !n_out || cap < 0 || (cap > 0 && !(ids && score && R && t && time)))
    return bop_set_error_msg(-1, "bop_parse_results_csv_host: bad arguments");
  static const char header[] = "scene_id,im_id,obj_id,score,R,t,time";
  const char* p = text;
  const char* const end = text + nbytes;
  int64_t n = 0;
  bool first = true;
  while (p < end) {
    const char* eol = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
    if (!eol) eol = end;
    if (first) {
      first = false;
      const std::string line(p, (size_t)(eol - p));
      if (line.find(header) != std::string::npos) { p = eol < end ? eol + 1 : end; continue; }
    }
    if (n >= cap) return 1;
    const char* q = p;
    int64_t* id = ids + 3 * n;
    if (!read_i64(q, eol, id[0]) || !eat(q, eol, ',') || !read_i64(q, eol, id[1]) || !eat(q, eol, ',') ||
        !read_i64(q, eol, id[2]) || !eat(q, eol, ',') || !read_f64(q, eol, score[n]) || !eat(q, eol, ','))
      return 1;
    for (int k = 0; k < 9; ++k)
      if (!read_f64(q, eol, R[9 * n + k])) return 1;
    if (!eat(q, eol, ',')) return 1;
    for (int k = 0; k < 3; ++k)
      if (!read_f64(q, eol, t[3 * n + k])) return 1;
    if (!eat(q, eol, ',') || !read_f64(q, eol, time[n]) || q != eol) return 1;
    ++n;
    p = eol < end ? eol + 1 : end;
  }
  *n_out = n;
  return 0;
}
