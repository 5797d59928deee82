// kvm_csv.cu - host-side columnar ingest of requests.csv and the classified-CSV writer
// (SURVEY.md section 8f, F1).  No device code: this file only lives in the library so that the
// drop-in scripts have one binding.
//
// Replaces the per-row work of analyze.py:34-56 (`csv.DictReader` + `float()` on eight named
// cells) and analyze.py:597-604 (`csv.DictWriter` of every row + is_cold_start).  The contract
// with the Python side (kserve-vllm-mini_b200/columns.py) keeps the reference's coercion rules
// exact without re-implementing Python's float() here:
//   * a cell that is a PLAIN DECIMAL literal  [+-]digits[.digits][(e|E)[+-]digits]  is parsed
//     with std::from_chars, which is correctly rounded like CPython's float() -> same double;
//   * "" and "NaN" are reported as such (the reference leaves them as strings);
//   * anything else (whitespace, inf/nan spellings, underscores, overflow, garbage) is reported
//     as IRREGULAR and the Python side runs the reference's own `float()` on the raw text;
//   * files the fast path does not model (a lone '\r' row terminator, rows with more or fewer
//     fields than the header) are flagged and the Python side falls back to its csv-module path.

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <charconv>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/kvmini_b200.h"
#include "kvm_common.cuh"

struct kvm_csv {
  const char* data = nullptr;
  size_t size = 0;
  int fd = -1;
  bool mapped = false;
  std::vector<std::string> header;
  std::vector<int64_t> row_start;  // byte offset of every data row (empty lines skipped)
  std::vector<int64_t> row_end;    // one past the last byte of the row (terminator excluded)
  std::vector<uint8_t> row_quoted; // the row contains a '"'
  int32_t flags = 0;               // KVM_CSV_* irregularities
};

namespace {

struct Field {
  const char* p;
  size_t n;
  bool quoted;  // raw text still carries its quotes
};

// Splits one row (no terminator) into raw fields.  `quoted_row` rows go through the csv state
// machine (quotes may hide delimiters); others are split on ','.
void split_row(const char* b, const char* e, bool quoted_row, std::vector<Field>& out) {
  out.clear();
  if (!quoted_row) {
    const char* s = b;
    for (;;) {
      const char* c = static_cast<const char*>(memchr(s, ',', e - s));
      if (!c) {
        out.push_back({s, (size_t)(e - s), false});
        return;
      }
      out.push_back({s, (size_t)(c - s), false});
      s = c + 1;
    }
  }
  const char* s = b;
  while (true) {
    const char* f = s;
    bool q = false, in = false;
    while (s < e) {
      const char ch = *s;
      if (in) {
        if (ch == '"') {
          if (s + 1 < e && s[1] == '"') ++s;  // escaped quote
          else in = false;
        }
      } else if (ch == '"') {
        q = true;
        in = true;
      } else if (ch == ',') {
        break;
      }
      ++s;
    }
    out.push_back({f, (size_t)(s - f), q});
    if (s >= e) return;
    ++s;  // the delimiter
    if (s == e) {  // trailing delimiter: one more empty field
      out.push_back({s, 0, false});
      return;
    }
  }
}

// csv.reader semantics for a raw field (non-strict dialect): a quote opens a quoted section,
// "" inside it is a literal quote, text after the closing quote is appended as is.
void unquote(const Field& f, std::string& out) {
  out.clear();
  if (!f.quoted) {
    out.assign(f.p, f.n);
    return;
  }
  bool in = false;
  for (size_t i = 0; i < f.n; ++i) {
    const char ch = f.p[i];
    if (in) {
      if (ch == '"') {
        if (i + 1 < f.n && f.p[i + 1] == '"') {
          out.push_back('"');
          ++i;
        } else {
          in = false;
        }
      } else {
        out.push_back(ch);
      }
    } else if (ch == '"' && (i == 0 || out.empty())) {
      in = true;  // csv only opens a quoted section at the start of a field
    } else {
      out.push_back(ch);
    }
  }
}

// Quoted rows the fast path models: every field is either unquoted without any '"' or one
// quoted section spanning the whole field (what csv.writer produces).
bool well_formed_quotes(const char* b, const char* e) {
  const char* s = b;
  while (true) {
    if (s < e && *s == '"') {
      ++s;
      for (;;) {
        const char* q = static_cast<const char*>(memchr(s, '"', e - s));
        if (!q) return false;  // unterminated
        if (q + 1 < e && q[1] == '"') {
          s = q + 2;
          continue;
        }
        s = q + 1;
        break;
      }
      if (s < e && *s != ',') return false;  // text after the closing quote
    } else {
      const char* c = static_cast<const char*>(memchr(s, ',', e - s));
      const char* fe = c ? c : e;
      if (memchr(s, '"', fe - s)) return false;  // bare quote inside an unquoted field
      s = fe;
    }
    if (s >= e) return true;
    ++s;  // delimiter
    if (s == e) return true;
  }
}

bool plain_decimal(const char* p, size_t n) {
  size_t i = 0;
  if (i < n && (p[i] == '+' || p[i] == '-')) ++i;
  size_t digits = 0;
  while (i < n && p[i] >= '0' && p[i] <= '9') ++i, ++digits;
  if (i < n && p[i] == '.') {
    ++i;
    while (i < n && p[i] >= '0' && p[i] <= '9') ++i, ++digits;
  }
  if (!digits) return false;
  if (i < n && (p[i] == 'e' || p[i] == 'E')) {
    ++i;
    if (i < n && (p[i] == '+' || p[i] == '-')) ++i;
    size_t ed = 0;
    while (i < n && p[i] >= '0' && p[i] <= '9') ++i, ++ed;
    if (!ed) return false;
  }
  return i == n;
}

// kind of one cell + its value when plain
uint8_t classify_cell(const char* p, size_t n, double* v) {
  if (n == 0) return KVM_CSV_CELL_EMPTY;
  if (n == 3 && p[0] == 'N' && p[1] == 'a' && p[2] == 'N') return KVM_CSV_CELL_NAN_TEXT;
  if (!plain_decimal(p, n)) return KVM_CSV_CELL_IRREGULAR;
  const char* s = p;
  if (*s == '+') ++s;  // from_chars does not take a leading plus, float() does
  double x = 0.0;
  const auto r = std::from_chars(s, p + n, x, std::chars_format::general);
  if (r.ec != std::errc() || r.ptr != p + n) return KVM_CSV_CELL_IRREGULAR;  // overflow / underflow: float() decides
  *v = x;
  return KVM_CSV_CELL_NUMBER;
}

int worker_count(int32_t threads, int64_t rows) {
  int t = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
  if (t < 1) t = 1;
  if (t > 64) t = 64;
  const int64_t by_rows = rows / 20000 + 1;
  return (int)std::min<int64_t>(t, by_rows);
}

template <class F>
void parallel_rows(int64_t rows, int32_t threads, F&& fn) {
  const int t = worker_count(threads, rows);
  if (t == 1) {
    fn(0, (int64_t)0, rows);
    return;
  }
  std::vector<std::thread> pool;
  for (int w = 0; w < t; ++w) {
    const int64_t lo = rows * w / t, hi = rows * (w + 1) / t;
    pool.emplace_back([&fn, w, lo, hi] { fn(w, lo, hi); });
  }
  for (auto& th : pool) th.join();
}

// repr(float) of CPython ('r' format, Py_DTSF_ADD_DOT_0) from the shortest round-trip digits
void py_float_repr(double v, std::string& out) {
  if (v != v) {
    out += "nan";
    return;
  }
  if (v == HUGE_VAL || v == -HUGE_VAL) {
    out += v < 0 ? "-inf" : "inf";
    return;
  }
  char buf[48];
  const auto r = std::to_chars(buf, buf + sizeof buf, v, std::chars_format::scientific);
  // buf = [-]d[.ddd]e(+|-)XX
  const char* p = buf;
  if (*p == '-') {
    out.push_back('-');
    ++p;
  }
  const char* epos = static_cast<const char*>(memchr(p, 'e', r.ptr - p));
  char digits[24];
  int nd = 0;
  for (const char* c = p; c < epos; ++c)
    if (*c != '.') digits[nd++] = *c;
  int ex = 0;
  {
    const char* c = epos + 1;
    const bool neg = *c == '-';
    ++c;
    for (; c < r.ptr; ++c) ex = ex * 10 + (*c - '0');
    if (neg) ex = -ex;
  }
  if (nd == 1 && digits[0] == '0') {  // zero
    out += "0.0";
    return;
  }
  const int decpt = ex + 1;
  if (decpt <= -4 || decpt > 16) {
    out.push_back(digits[0]);
    if (nd > 1) {
      out.push_back('.');
      out.append(digits + 1, nd - 1);
    }
    out.push_back('e');
    int e10 = decpt - 1;
    out.push_back(e10 < 0 ? '-' : '+');
    if (e10 < 0) e10 = -e10;
    char eb[8];
    int k = 0;
    do {
      eb[k++] = (char)('0' + e10 % 10);
      e10 /= 10;
    } while (e10);
    if (k < 2) eb[k++] = '0';
    while (k) out.push_back(eb[--k]);
  } else if (decpt <= 0) {
    out += "0.";
    out.append((size_t)(-decpt), '0');
    out.append(digits, nd);
  } else if (decpt >= nd) {
    out.append(digits, nd);
    out.append((size_t)(decpt - nd), '0');
    out += ".0";
  } else {
    out.append(digits, decpt);
    out.push_back('.');
    out.append(digits + decpt, nd - decpt);
  }
}

// csv.writer QUOTE_MINIMAL with the default dialect
void write_field(const std::string& v, std::string& out) {
  bool need = false;
  for (const char ch : v)
    if (ch == ',' || ch == '"' || ch == '\r' || ch == '\n') {
      need = true;
      break;
    }
  if (!need) {
    out += v;
    return;
  }
  out.push_back('"');
  for (const char ch : v) {
    if (ch == '"') out.push_back('"');
    out.push_back(ch);
  }
  out.push_back('"');
}

}  // namespace

extern "C" int kvm_csv_open(const char* path, kvm_csv** out) {
  KVM_REQUIRE(path && out, "kvm_csv_open: null argument");
  *out = nullptr;
  const int fd = open(path, O_RDONLY);
  if (fd < 0) {
    kvm_set_error("kvm_csv_open: cannot open %s", path);
    return KVM_ERR_INVALID;
  }
  struct stat st;
  if (fstat(fd, &st) != 0) {
    close(fd);
    kvm_set_error("kvm_csv_open: cannot stat %s", path);
    return KVM_ERR_INVALID;
  }
  kvm_csv* h = new kvm_csv;
  h->fd = fd;
  h->size = (size_t)st.st_size;
  if (h->size) {
    void* m = mmap(nullptr, h->size, PROT_READ, MAP_PRIVATE, fd, 0);
    if (m == MAP_FAILED) {
      close(fd);
      delete h;
      kvm_set_error("kvm_csv_open: mmap of %s failed", path);
      return KVM_ERR_INVALID;
    }
    madvise(m, h->size, MADV_WILLNEED);
    h->data = static_cast<const char*>(m);
    h->mapped = true;
  }
  // ---- row index: physical lines, merged while a quoted section is open
  const char* b = h->data;
  const char* const end = b + h->size;
  bool first = true;
  std::vector<Field> fields;
  std::string tmp;
  while (b < end) {
    const char* row_b = b;
    bool quoted = false, open_q = false;
    const char* e;
    for (;;) {
      const char* nl = static_cast<const char*>(memchr(b, '\n', end - b));
      e = nl ? nl : end;
      // quote parity of this physical line
      for (const char* q = static_cast<const char*>(memchr(b, '"', e - b)); q;
           q = static_cast<const char*>(memchr(q + 1, '"', e - q - 1))) {
        quoted = true;
        open_q = !open_q;
      }
      b = nl ? nl + 1 : end;
      if (!open_q || !nl) break;  // row complete (or file ends inside a quote)
    }
    const char* row_e = e;
    if (row_e > row_b && row_e[-1] == '\r') --row_e;
    if (memchr(row_b, '\r', row_e - row_b) && !quoted) h->flags |= KVM_CSV_LONE_CR;
    if (row_e == row_b) {  // csv skips empty lines - but DictReader takes a blank FIRST line as "no fieldnames"
      if (first) {
        first = false;
        h->flags |= KVM_CSV_RAGGED;
      }
      continue;
    }
    if (first) {
      split_row(row_b, row_e, quoted, fields);
      for (const Field& f : fields) {
        unquote(f, tmp);
        h->header.push_back(tmp);
      }
      first = false;
      continue;
    }
    h->row_start.push_back(row_b - h->data);
    h->row_end.push_back(row_e - h->data);
    h->row_quoted.push_back(quoted ? 1 : 0);
  }
  // ---- field counts: every row must have exactly the header's number of fields
  const int64_t rows = (int64_t)h->row_start.size();
  const size_t ncols = h->header.size();
  std::vector<int> ragged(65, 0);
  parallel_rows(rows, 0, [&](int w, int64_t lo, int64_t hi) {
    std::vector<Field> fl;
    for (int64_t i = lo; i < hi; ++i) {
      const char* rb = h->data + h->row_start[i];
      const char* re = h->data + h->row_end[i];
      size_t cnt;
      if (!h->row_quoted[i]) {
        cnt = 1;
        for (const char* c = static_cast<const char*>(memchr(rb, ',', re - rb)); c;
             c = static_cast<const char*>(memchr(c + 1, ',', re - c - 1)))
          ++cnt;
      } else {
        if (!well_formed_quotes(rb, re)) ragged[w] |= 2;
        split_row(rb, re, true, fl);
        cnt = fl.size();
      }
      if (cnt != ncols) ragged[w] |= 1;
    }
  });
  for (int r : ragged) {
    if (r & 1) h->flags |= KVM_CSV_RAGGED;
    if (r & 2) h->flags |= KVM_CSV_BARE_QUOTE;
  }
  *out = h;
  return KVM_OK;
}

extern "C" int kvm_csv_close(kvm_csv* h) {
  if (!h) return KVM_OK;
  if (h->mapped) munmap(const_cast<char*>(h->data), h->size);
  if (h->fd >= 0) close(h->fd);
  delete h;
  return KVM_OK;
}

extern "C" int64_t kvm_csv_rows(const kvm_csv* h) { return h ? (int64_t)h->row_start.size() : 0; }
extern "C" int32_t kvm_csv_ncols(const kvm_csv* h) { return h ? (int32_t)h->header.size() : 0; }
extern "C" int32_t kvm_csv_flags(const kvm_csv* h) { return h ? h->flags : 0; }
extern "C" const char* kvm_csv_col_name(const kvm_csv* h, int32_t i) {
  return (h && i >= 0 && i < (int32_t)h->header.size()) ? h->header[i].c_str() : nullptr;
}
extern "C" int32_t kvm_csv_find_col(const kvm_csv* h, const char* name) {
  if (!h || !name) return -1;
  // DictReader: a repeated header name keeps the LAST column's value
  for (int32_t i = (int32_t)h->header.size() - 1; i >= 0; --i)
    if (h->header[i] == name) return i;
  return -1;
}

extern "C" int kvm_csv_parse_columns(const kvm_csv* h, const int32_t* cols, int32_t ncols, double* const* values,
                                     uint8_t* const* kinds, int32_t threads) {
  KVM_REQUIRE(h && cols && values && kinds && ncols >= 1 && ncols <= 16, "kvm_csv_parse_columns: bad argument");
  KVM_REQUIRE(!(h->flags & KVM_CSV_NEEDS_CSV_MODULE), "kvm_csv_parse_columns: file needs the csv-module path");
  const int32_t width = (int32_t)h->header.size();
  // requested columns in file order, so that one walk over the delimiters of a row serves all
  int32_t order[16];
  for (int32_t k = 0; k < ncols; ++k) {
    KVM_REQUIRE(cols[k] >= 0 && cols[k] < width && values[k] && kinds[k],
                "kvm_csv_parse_columns: column %d out of range or null output", cols[k]);
    order[k] = k;
  }
  std::sort(order, order + ncols, [&](int32_t a, int32_t b) { return cols[a] < cols[b]; });
  const int64_t rows = (int64_t)h->row_start.size();
  parallel_rows(rows, threads, [&](int, int64_t lo, int64_t hi) {
    std::vector<Field> fl;
    std::string tmp;
    for (int64_t i = lo; i < hi; ++i) {
      const char* rb = h->data + h->row_start[i];
      const char* re = h->data + h->row_end[i];
      if (!h->row_quoted[i]) {
        const char* p = rb;
        int32_t at = 0;  // column p points at
        for (int32_t k = 0; k < ncols; ++k) {
          const int32_t o = order[k], c = cols[o];
          for (; at < c; ++at) p = static_cast<const char*>(memchr(p, ',', re - p)) + 1;
          const char* q = static_cast<const char*>(memchr(p, ',', re - p));
          double v = __builtin_nan("");
          kinds[o][i] = classify_cell(p, (size_t)((q ? q : re) - p), &v);
          values[o][i] = v;
        }
      } else {
        split_row(rb, re, true, fl);
        for (int32_t k = 0; k < ncols; ++k) {
          const Field& f = fl[cols[k]];
          const char* p = f.p;
          size_t n = f.n;
          if (f.quoted) {
            unquote(f, tmp);
            p = tmp.data();
            n = tmp.size();
          }
          double v = __builtin_nan("");
          kinds[k][i] = classify_cell(p, n, &v);
          values[k][i] = v;
        }
      }
    }
  });
  return KVM_OK;
}

extern "C" int kvm_csv_parse_f64(const kvm_csv* h, int32_t col, double* out, uint8_t* kind, int32_t threads) {
  return kvm_csv_parse_columns(h, &col, 1, &out, &kind, threads);
}

extern "C" int64_t kvm_csv_cell(const kvm_csv* h, int64_t row, int32_t col, char* dst, int64_t cap) {
  if (!h || row < 0 || row >= (int64_t)h->row_start.size() || col < 0 || col >= (int32_t)h->header.size()) return -1;
  std::vector<Field> fl;
  split_row(h->data + h->row_start[row], h->data + h->row_end[row], h->row_quoted[row] != 0, fl);
  if (col >= (int32_t)fl.size()) return -1;
  std::string tmp;
  unquote(fl[col], tmp);
  if (dst && cap > 0) memcpy(dst, tmp.data(), (size_t)std::min<int64_t>(cap, (int64_t)tmp.size()));
  return (int64_t)tmp.size();
}

extern "C" int kvm_csv_match_any(const kvm_csv* h, int32_t col, const char* const* options, int32_t noptions,
                                 uint8_t* out, int32_t threads) {
  KVM_REQUIRE(h && options && out && noptions >= 0, "kvm_csv_match_any: null argument");
  KVM_REQUIRE(col >= 0 && col < (int32_t)h->header.size(), "kvm_csv_match_any: column %d out of range", col);
  KVM_REQUIRE(!(h->flags & (KVM_CSV_NEEDS_CSV_MODULE)), "kvm_csv_match_any: file needs the csv-module path");
  std::vector<std::string> opts(options, options + noptions);
  parallel_rows((int64_t)h->row_start.size(), threads, [&](int, int64_t lo, int64_t hi) {
    std::vector<Field> fl;
    std::string tmp;
    for (int64_t i = lo; i < hi; ++i) {
      split_row(h->data + h->row_start[i], h->data + h->row_end[i], h->row_quoted[i] != 0, fl);
      unquote(fl[col], tmp);
      uint8_t m = 0;
      for (const std::string& o : opts) m |= (uint8_t)(tmp == o);
      out[i] = m;
    }
  });
  return KVM_OK;
}

// analyze.py:597-604: header + "is_cold_start" (appended, or replaced in place when the column
// already exists), every row with its coerced cells re-serialised as repr(float) - "200" ->
// "200.0", "1001.000" -> "1001.0" - everything else as csv.writer would write the parsed value.
extern "C" int kvm_csv_write_classified(const kvm_csv* h, const char* out_path, const uint8_t* is_cold,
                                        const int32_t* coerced_cols, int32_t ncoerced, int32_t threads) {
  KVM_REQUIRE(h && out_path && (is_cold || h->row_start.empty()), "kvm_csv_write_classified: null argument");
  KVM_REQUIRE(!(h->flags & (KVM_CSV_NEEDS_CSV_MODULE)), "kvm_csv_write_classified: file needs the csv-module path");
  const int32_t ncols = (int32_t)h->header.size();
  std::vector<uint8_t> coerced((size_t)ncols, 0);
  for (int32_t k = 0; k < ncoerced; ++k)
    if (coerced_cols[k] >= 0 && coerced_cols[k] < ncols) coerced[coerced_cols[k]] = 1;
  const int32_t cold_col = kvm_csv_find_col(h, "is_cold_start");
  const int64_t rows = (int64_t)h->row_start.size();
  FILE* f = fopen(out_path, "wb");
  if (!f) {
    kvm_set_error("kvm_csv_write_classified: cannot create %s", out_path);
    return KVM_ERR_INVALID;
  }
  int rc = KVM_OK;
  if (rows > 0) {  // the reference writes nothing at all for an empty log
    std::string head;
    for (int32_t c = 0; c < ncols; ++c) {
      if (c) head.push_back(',');
      write_field(h->header[c], head);
    }
    if (cold_col < 0) head += ",is_cold_start";
    head += "\r\n";
    fwrite(head.data(), 1, head.size(), f);
    const int t = worker_count(threads, rows);
    // bounded memory: blocks of rows are rendered in parallel, then written in order
    const int64_t block = 1 << 15;   // rows per worker and round: small enough that a 1M-row file keeps every core busy
    std::vector<std::string> bufs((size_t)t);
    int irregular = 0;
    for (int64_t b0 = 0; b0 < rows && rc == KVM_OK; b0 += block * t) {
      const int64_t b1 = std::min(rows, b0 + block * t);
      std::vector<std::thread> pool;
      std::vector<int> bad((size_t)t, 0);
      for (int w = 0; w < t; ++w) {
        const int64_t lo = std::min(b1, b0 + block * w), hi = std::min(b1, lo + block);
        pool.emplace_back([&, w, lo, hi] {
          std::string& o = bufs[w];
          o.clear();
          std::vector<Field> fl;
          std::string tmp;
          for (int64_t i = lo; i < hi; ++i) {
            split_row(h->data + h->row_start[i], h->data + h->row_end[i], h->row_quoted[i] != 0, fl);
            for (int32_t c = 0; c < ncols; ++c) {
              if (c) o.push_back(',');
              if (c == cold_col) {
                o += is_cold[i] ? "True" : "False";
                continue;
              }
              unquote(fl[c], tmp);
              if (coerced[c]) {
                double v;
                const uint8_t k = classify_cell(tmp.data(), tmp.size(), &v);
                if (k == KVM_CSV_CELL_NUMBER) {
                  py_float_repr(v, o);
                  continue;
                }
                if (k == KVM_CSV_CELL_IRREGULAR) bad[w] = 1;  // float() may or may not take it: not ours to say
              }
              write_field(tmp, o);
            }
            if (cold_col < 0) o += is_cold[i] ? ",True" : ",False";
            o += "\r\n";
          }
        });
      }
      for (auto& th : pool) th.join();
      for (int w = 0; w < t; ++w) {
        if (bad[w]) irregular = 1;
        if (!bufs[w].empty() && fwrite(bufs[w].data(), 1, bufs[w].size(), f) != bufs[w].size()) {
          kvm_set_error("kvm_csv_write_classified: short write to %s", out_path);
          rc = KVM_ERR_INVALID;
        }
      }
    }
    if (rc == KVM_OK && irregular) {
      kvm_set_error("kvm_csv_write_classified: a coerced cell is not a plain decimal; use the csv-module path");
      rc = KVM_ERR_UNSUPPORTED;
    }
  }
  fclose(f);
  return rc;
}
