// mm_reader.cpp — Matrix Market coordinate-real reader, dialect-compatible with the reference's
// read_file (reference test/mm_reader.cpp:154-171 and :45-107) but built differently: the file is
// read in one block, numbers are parsed with from_chars, and the CSR is assembled with a counting
// sort over rows (O(nnz)) plus a per-row column sort, instead of a comparison sort over tuples.
//
// Dialect kept on purpose:
//   * the banner line is consumed and ignored (:160, parse_header result unused);
//   * the line after the banner is discarded unconditionally (:163-164);
//   * following lines starting with '%' are skipped (:146-152);
//   * the size line is consumed, its numbers are not used (:47-52);
//   * triples "i j v" are read until one fails to parse (:62-66), 1-based;
//   * EVERY off-diagonal entry is mirrored, whatever the banner says (:68-74);
//   * entries are ordered by (row, column) (:76-86);
//   * a row offset is emitted per DISTINCT row id, so empty rows vanish (:92-104).
#include "../../include/cgb200.h"

#include <algorithm>
#include <charconv>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <vector>

namespace cgb {
int fail(int code, const char *fmt, ...);
}

struct cgb_mm {
  std::vector<double> vals;
  std::vector<int32_t> cols;
  std::vector<int32_t> rowptr;
};

namespace {

inline const char *skip_line(const char *p, const char *end) {
  while (p < end && *p != '\n') p++;
  return p < end ? p + 1 : end;
}
inline const char *skip_ws(const char *p, const char *end) {
  while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r' || *p == '\f' || *p == '\v')) p++;
  return p;
}
inline bool parse_int(const char *&p, const char *end, long &out) {
  p = skip_ws(p, end);
  if (p < end && *p == '+') p++;
  auto r = std::from_chars(p, end, out);
  if (r.ec != std::errc()) return false;
  p = r.ptr;
  return true;
}
inline bool parse_double(const char *&p, const char *end, double &out) {
  p = skip_ws(p, end);
  if (p < end && *p == '+') p++;
  auto r = std::from_chars(p, end, out);
  if (r.ec != std::errc()) return false;
  p = r.ptr;
  return true;
}

} // namespace

extern "C" {

int cgb_mm_read(const char *path, cgb_mm **out) {
  if (!path || !out) return cgb::fail(CGB_ERR_INVALID, "cgb_mm_read: NULL argument");
  *out = nullptr;
  FILE *f = std::fopen(path, "rb");
  if (!f) return cgb::fail(CGB_ERR_IO, "cgb_mm_read: cannot open '%s'", path);
  std::fseek(f, 0, SEEK_END);
  const long sz = std::ftell(f);
  std::fseek(f, 0, SEEK_SET);
  std::vector<char> buf((size_t)(sz > 0 ? sz : 0));
  if (sz > 0 && std::fread(buf.data(), 1, (size_t)sz, f) != (size_t)sz) {
    std::fclose(f);
    return cgb::fail(CGB_ERR_IO, "cgb_mm_read: short read on '%s'", path);
  }
  std::fclose(f);
  const char *p = buf.data(), *end = p + buf.size();
  p = skip_line(p, end);                          // banner
  p = skip_line(p, end);                          // unconditional discard
  while (p < end && *p == '%') p = skip_line(p, end); // comments
  p = skip_line(p, end);                          // size line (unused)

  std::vector<int32_t> ri, ci;
  std::vector<double> vv;
  {
    size_t guess = (size_t)(end - p) / 12 + 16;
    ri.reserve(guess);
    ci.reserve(guess);
    vv.reserve(guess);
  }
  long a, b;
  double v;
  while (true) {
    const char *q = p;
    if (!parse_int(q, end, a) || !parse_int(q, end, b) || !parse_double(q, end, v)) break;
    p = q;
    ri.push_back((int32_t)(a - 1));
    ci.push_back((int32_t)(b - 1));
    vv.push_back(v);
  }
  const size_t h = ri.size();
  if (h == 0) return cgb::fail(CGB_ERR_IO, "cgb_mm_read: no entries in '%s'", path);
  int32_t lo = ri[0], hi = ri[0];
  size_t off_diag = 0;
  for (size_t i = 0; i < h; i++) {
    lo = std::min(lo, std::min(ri[i], ci[i]));
    hi = std::max(hi, std::max(ri[i], ci[i]));
    off_diag += ri[i] != ci[i];
  }
  if (lo < 0) return cgb::fail(CGB_ERR_IO, "cgb_mm_read: index below 1 in '%s'", path);
  const size_t total = h + off_diag;
  if (total >= ((size_t)1 << 31)) return cgb::fail(CGB_ERR_IO, "cgb_mm_read: too many entries for 32-bit offsets");
  // counting sort by row over the entries and their mirrors
  const size_t span = (size_t)hi + 1;
  std::vector<int64_t> count(span + 1, 0);
  for (size_t i = 0; i < h; i++) {
    count[(size_t)ri[i] + 1]++;
    if (ri[i] != ci[i]) count[(size_t)ci[i] + 1]++;
  }
  for (size_t r = 0; r < span; r++) count[r + 1] += count[r];
  cgb_mm *m = new cgb_mm();
  m->vals.resize(total);
  m->cols.resize(total);
  {
    std::vector<int64_t> cursor(count.begin(), count.end() - 1);
    for (size_t i = 0; i < h; i++) {
      int64_t o = cursor[(size_t)ri[i]]++;
      m->cols[(size_t)o] = ci[i];
      m->vals[(size_t)o] = vv[i];
      if (ri[i] != ci[i]) {
        o = cursor[(size_t)ci[i]]++;
        m->cols[(size_t)o] = ri[i];
        m->vals[(size_t)o] = vv[i];
      }
    }
  }
  // ascending columns inside each row; offsets only for rows that have entries
  std::vector<std::pair<int32_t, double>> tmp;
  m->rowptr.reserve(span + 1);
  for (size_t r = 0; r < span; r++) {
    const int64_t s = count[r], e = count[r + 1];
    if (s == e) continue;
    m->rowptr.push_back((int32_t)s);
    bool sorted = true;
    for (int64_t j = s + 1; j < e; j++)
      if (m->cols[(size_t)j] < m->cols[(size_t)j - 1]) {
        sorted = false;
        break;
      }
    if (sorted) continue;
    tmp.resize((size_t)(e - s));
    for (int64_t j = s; j < e; j++) tmp[(size_t)(j - s)] = {m->cols[(size_t)j], m->vals[(size_t)j]};
    std::stable_sort(tmp.begin(), tmp.end(), [](const auto &x, const auto &y) { return x.first < y.first; });
    for (int64_t j = s; j < e; j++) {
      m->cols[(size_t)j] = tmp[(size_t)(j - s)].first;
      m->vals[(size_t)j] = tmp[(size_t)(j - s)].second;
    }
  }
  m->rowptr.push_back((int32_t)total);
  *out = m;
  return CGB_OK;
}

int64_t cgb_mm_n(const cgb_mm *m) { return m ? (int64_t)m->rowptr.size() - 1 : 0; }
int64_t cgb_mm_nnz(const cgb_mm *m) { return m ? (int64_t)m->vals.size() : 0; }
const double *cgb_mm_vals(const cgb_mm *m) { return m ? m->vals.data() : nullptr; }
const int32_t *cgb_mm_cols(const cgb_mm *m) { return m ? m->cols.data() : nullptr; }
const int32_t *cgb_mm_rowptr(const cgb_mm *m) { return m ? m->rowptr.data() : nullptr; }
int cgb_mm_free(cgb_mm *m) {
  delete m;
  return CGB_OK;
}

} // extern "C"
