// mm_reader.cpp — Matrix Market coordinate-real reader, dialect-compatible with the reference's
// read_file (reference test/mm_reader.cpp:154-171 and :45-107) but built differently: the file is
// memory-mapped and its body parsed with from_chars on all host cores (one line-aligned chunk per
// thread), and the host CSR is assembled with a counting sort over rows (O(nnz)) plus a per-row
// column sort, instead of a comparison sort over tuples.  cgb_csr_from_mm_file (coo.cu) feeds the
// same triplets to the GPU assembly instead.
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
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>
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

// One chunk of the body, line-aligned: every non-blank line must be exactly "i j v".  Returns false when a line
// does not fit that shape (the caller then re-parses the whole body as a token stream, which is the reference's
// own reading loop, test/mm_reader.cpp:62-66).
bool parse_lines(const char *p, const char *end, std::vector<int32_t> &ri_out, std::vector<int32_t> &ci_out,
                 std::vector<double> &vv_out) {
  // thread-local vectors, handed over at the end: the callers' vector headers sit next to each other in memory,
  // and push_back rewrites the header on every line (false sharing made 2 threads slower than 1)
  std::vector<int32_t> ri, ci;
  std::vector<double> vv;
  const size_t guess = (size_t)(end - p) / 12 + 16;
  ri.reserve(guess);
  ci.reserve(guess);
  vv.reserve(guess);
  long a, b;
  double v;
  while (p < end) {
    while (p < end && (*p == ' ' || *p == '\t' || *p == '\r')) p++;
    if (p < end && *p == '\n') { // blank line
      p++;
      continue;
    }
    if (p >= end) break;
    if (!parse_int(p, end, a) || !parse_int(p, end, b) || !parse_double(p, end, v)) return false;
    while (p < end && (*p == ' ' || *p == '\t' || *p == '\r')) p++;
    if (p < end && *p != '\n') return false; // a fourth token on the line
    ri.push_back((int32_t)a);
    ci.push_back((int32_t)b);
    vv.push_back(v);
  }
  ri_out = std::move(ri);
  ci_out = std::move(ci);
  vv_out = std::move(vv);
  return true;
}

} // namespace

namespace cgb {

// Body of a Matrix Market file in the reference dialect -> 1-based triplets, in file order, one chunk per thread.  The file is mapped,
// cut at line ends into one chunk per host thread and parsed with from_chars; files whose lines are not plain
// "i j v" fall back to the serial token stream.
int mm_parse_chunks(const char *path, std::vector<std::vector<int32_t>> &pr, std::vector<std::vector<int32_t>> &pc,
                    std::vector<std::vector<double>> &pv) {
  pr.clear();
  pc.clear();
  pv.clear();
  const int fd = ::open(path, O_RDONLY);
  if (fd < 0) return fail(CGB_ERR_IO, "cgb_mm_read: cannot open '%s'", path);
  struct stat st;
  if (::fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) {
    ::close(fd);
    return fail(CGB_ERR_IO, "cgb_mm_read: cannot stat '%s'", path);
  }
  const size_t sz = (size_t)st.st_size;
  if (sz == 0) {
    ::close(fd);
    return CGB_OK;
  }
  void *map = ::mmap(nullptr, sz, PROT_READ, MAP_PRIVATE, fd, 0);
  ::close(fd);
  if (map == MAP_FAILED) return fail(CGB_ERR_IO, "cgb_mm_read: cannot map '%s'", path);
  ::madvise(map, sz, MADV_SEQUENTIAL);
  const char *p = (const char *)map, *end = p + sz;
  p = skip_line(p, end);                              // banner (test/mm_reader.cpp:160; result unused there too)
  p = skip_line(p, end);                              // unconditional discard (:163-164)
  while (p < end && *p == '%') p = skip_line(p, end); // comments (:146-152)
  p = skip_line(p, end);                              // size line (:47-52, numbers unused)

  unsigned threads = std::thread::hardware_concurrency();
  if (const char *e = std::getenv("CGB_MM_THREADS")) threads = (unsigned)std::max(1, std::atoi(e));
  const size_t len = (size_t)(end - p);
  threads = (unsigned)std::max<size_t>(1, std::min<size_t>(threads ? threads : 1, len / (256 * 1024) + 1));
  std::vector<const char *> cut(threads + 1, end);
  cut[0] = p;
  for (unsigned k = 1; k < threads; k++) {
    const char *q = p + len * k / threads;
    if (q < cut[k - 1]) q = cut[k - 1];
    cut[k] = skip_line(q, end); // the chunk starts after the next line end
  }
  pr.resize(threads);
  pc.resize(threads);
  pv.resize(threads);
  std::vector<int> ok(threads, 1);
  {
    std::vector<std::thread> pool;
    for (unsigned k = 1; k < threads; k++)
      pool.emplace_back([&, k] { ok[k] = parse_lines(cut[k], cut[k + 1], pr[k], pc[k], pv[k]) ? 1 : 0; });
    ok[0] = parse_lines(cut[0], cut[1], pr[0], pc[0], pv[0]) ? 1 : 0;
    for (auto &t : pool) t.join();
  }
  bool all_ok = true;
  for (unsigned k = 0; k < threads; k++) all_ok = all_ok && ok[k];
  if (!all_ok) { // the reference's loop: whitespace-separated tokens, stop at the first triple that does not parse
    pr.assign(1, {});
    pc.assign(1, {});
    pv.assign(1, {});
    long a, b;
    double v;
    while (true) {
      const char *q = p;
      if (!parse_int(q, end, a) || !parse_int(q, end, b) || !parse_double(q, end, v)) break;
      p = q;
      pr[0].push_back((int32_t)a);
      pc[0].push_back((int32_t)b);
      pv[0].push_back(v);
    }
  }
  ::munmap(map, sz);
  return CGB_OK;
}


// the same, concatenated in file order
int mm_parse_triplets(const char *path, std::vector<int32_t> &ri, std::vector<int32_t> &ci, std::vector<double> &vv) {
  std::vector<std::vector<int32_t>> pr, pc;
  std::vector<std::vector<double>> pv;
  const int rc = mm_parse_chunks(path, pr, pc, pv);
  if (rc != CGB_OK) return rc;
  size_t total = 0;
  for (auto &x : pr) total += x.size();
  ri.resize(total);
  ci.resize(total);
  vv.resize(total);
  size_t o = 0;
  for (size_t k = 0; k < pr.size(); k++) {
    if (pr[k].empty()) continue;
    std::memcpy(ri.data() + o, pr[k].data(), pr[k].size() * sizeof(int32_t));
    std::memcpy(ci.data() + o, pc[k].data(), pc[k].size() * sizeof(int32_t));
    std::memcpy(vv.data() + o, pv[k].data(), pv[k].size() * sizeof(double));
    o += pr[k].size();
  }
  return CGB_OK;
}

} // namespace cgb

extern "C" {

int cgb_mm_read(const char *path, cgb_mm **out) {
  if (!path || !out) return cgb::fail(CGB_ERR_INVALID, "cgb_mm_read: NULL argument");
  *out = nullptr;
  std::vector<int32_t> ri, ci;
  std::vector<double> vv;
  {
    const int rc = cgb::mm_parse_triplets(path, ri, ci, vv);
    if (rc != CGB_OK) return rc;
  }
  for (auto &x : ri) x -= 1; // 1-based in the file (test/mm_reader.cpp:62-66)
  for (auto &x : ci) x -= 1;
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
