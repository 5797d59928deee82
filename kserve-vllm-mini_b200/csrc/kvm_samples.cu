// kvm_samples.cu - columnar ingest of power samples (host only, no device code).  SURVEY.md 8f F3.
//
// Replaces the list-of-objects step in front of the energy kernels:
//   energy/collector.py:152-177  load_power_samples: json.load of power.json ({"samples": [{"ts" |
//                                "ts_s": t, "watts": w | null}, ...]} or the bare list), one
//                                PowerSample object per sample
//   analyze.py:219-247 / energy/collector.py:203-251  Prometheus query / query_range results
//                                ("values": [[t, "v"], ...] per series) turned into Python lists
// At 1 kHz x 24 h x 8 GPUs that is 6.9e8 objects; here the text is scanned once, in parallel, into
// two fp64 columns per series (ts, value; NaN = null), ready to be copied to the device.
// Numbers go through std::from_chars (correctly rounded: the same double as Python's float()).
// Anything the scanner does not model (escapes in keys, NaN / Infinity literals, nested values in
// a sample, a null timestamp ...) returns KVM_ERR_UNSUPPORTED and the caller uses the json module -
// the fast path never guesses.

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <charconv>
#include <cmath>
#include <cstring>
#include <limits>
#include <string>
#include <thread>
#include <vector>

#include "kvm_common.cuh"

struct kvm_samples {
  std::vector<std::vector<double>> ts, val;   // per series
};

namespace {

const double kNaN = std::numeric_limits<double>::quiet_NaN();

inline const char* skip_ws(const char* p, const char* e) {
  while (p < e && (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r')) ++p;
  return p;
}

// JSON number -> double; returns the end or nullptr (also for the literals Python's json accepts
// beyond JSON: NaN, Infinity)
inline const char* parse_number(const char* p, const char* e, double* out) {
  if (p >= e) return nullptr;
  const char c = *p;
  if (!(c == '-' || (c >= '0' && c <= '9'))) return nullptr;
  if (c == '-' && p + 1 < e && !(p[1] >= '0' && p[1] <= '9')) return nullptr;   // -Infinity
  const auto r = std::from_chars(p, e, *out);
  if (r.ec != std::errc() || r.ptr == p) return nullptr;
  return r.ptr;
}

// a plain string (no escapes): returns the position after the closing quote, sets [b, q)
inline const char* parse_string(const char* p, const char* e, const char** b, const char** q) {
  if (p >= e || *p != '"') return nullptr;
  ++p;
  *b = p;
  while (p < e && *p != '"') {
    if (*p == '\\') return nullptr;
    ++p;
  }
  if (p >= e) return nullptr;
  *q = p;
  return p + 1;
}

inline bool key_is(const char* b, const char* q, const char* name) {
  const size_t n = strlen(name);
  return (size_t)(q - b) == n && memcmp(b, name, n) == 0;
}

// one flat object {"k": v, ...} of a power.json sample list; p at '{'
const char* parse_sample(const char* p, const char* e, double* ts, double* watts) {
  bool have_ts = false, have_ts_s = false, have_w = false;
  double t = 0.0, t_s = 0.0, w = kNaN;
  p = skip_ws(p + 1, e);
  if (p < e && *p == '}') {   // {}: ts defaults to 0, it["watts"] raises KeyError in the reference
    return nullptr;
  }
  for (;;) {
    const char *kb, *kq;
    p = parse_string(skip_ws(p, e), e, &kb, &kq);
    if (!p) return nullptr;
    p = skip_ws(p, e);
    if (p >= e || *p != ':') return nullptr;
    p = skip_ws(p + 1, e);
    if (p >= e) return nullptr;
    const bool is_ts = key_is(kb, kq, "ts"), is_ts_s = key_is(kb, kq, "ts_s"), is_w = key_is(kb, kq, "watts");
    double v = 0.0;
    bool is_null = false, is_num = false;
    if (*p == '"') {   // a string value: numeric strings are what float() accepts too
      const char *sb, *sq;
      const char* nx = parse_string(p, e, &sb, &sq);
      if (!nx) return nullptr;
      if (is_ts || is_ts_s || is_w) {
        const char* ne = parse_number(sb, sq, &v);
        if (ne != sq) return nullptr;   // " 1.5", "1e3x", "nan": leave to float()
        is_num = true;
      }
      p = nx;
    } else if (*p == 'n') {
      if (e - p < 4 || memcmp(p, "null", 4) != 0) return nullptr;
      is_null = true;
      p += 4;
    } else if (*p == 't' || *p == 'f') {
      const size_t n = *p == 't' ? 4 : 5;
      if ((size_t)(e - p) < n || memcmp(p, *p == 't' ? "true" : "false", n) != 0) return nullptr;
      if (is_ts || is_ts_s || is_w) return nullptr;   // float(True) is 1.0 in Python: not worth modelling
      p += n;
    } else if (*p == '{' || *p == '[') {
      return nullptr;   // nested value in a sample
    } else {
      p = parse_number(p, e, &v);
      if (!p) return nullptr;
      is_num = true;
    }
    if (is_ts_s) {
      if (!is_num) return nullptr;   // float(None) raises
      t_s = v;
      have_ts_s = true;
    } else if (is_ts) {
      if (!is_num) {
        if (!is_null) return nullptr;
        // "ts": null only matters when there is no ts_s: decided below
        have_ts = true;
        t = kNaN;
      } else {
        t = v;
        have_ts = true;
      }
    } else if (is_w) {
      have_w = true;
      w = is_null ? kNaN : v;
      if (!is_null && !is_num) return nullptr;
    }
    p = skip_ws(p, e);
    if (p >= e) return nullptr;
    if (*p == ',') {
      ++p;
      continue;
    }
    if (*p == '}') {
      ++p;
      break;
    }
    return nullptr;
  }
  if (!have_w) return nullptr;                            // it["watts"] -> KeyError in the reference
  if (have_ts_s) *ts = t_s;                               // it.get("ts_s", it.get("ts", 0))
  else if (have_ts) {
    if (t != t) return nullptr;                           // float(None) -> TypeError
    *ts = t;
  } else *ts = 0.0;
  if (w != w && false) return nullptr;
  *watts = w;
  return p;
}

// [t, "v"] (Prometheus sample); p at '['
const char* parse_prom_pair(const char* p, const char* e, double* ts, double* val) {
  p = skip_ws(p + 1, e);
  p = parse_number(p, e, ts);
  if (!p) return nullptr;
  p = skip_ws(p, e);
  if (p >= e || *p != ',') return nullptr;
  p = skip_ws(p + 1, e);
  const char *sb, *sq;
  p = parse_string(p, e, &sb, &sq);
  if (!p) return nullptr;
  // Prometheus writes "NaN", "+Inf", "-Inf" for special values: float() accepts them, so do we
  const size_t n = (size_t)(sq - sb);
  if (n == 3 && memcmp(sb, "NaN", 3) == 0) *val = kNaN;
  else if (n == 4 && memcmp(sb, "+Inf", 4) == 0) *val = std::numeric_limits<double>::infinity();
  else if (n == 4 && memcmp(sb, "-Inf", 4) == 0) *val = -std::numeric_limits<double>::infinity();
  else if (parse_number(sb, sq, val) != sq) return nullptr;
  p = skip_ws(p, e);
  if (p >= e || *p != ']') return nullptr;
  return p + 1;
}

// Elements of an array (p after '['), each parsed by `one` (which starts at `open`); the array is
// cut into `threads` pieces at element boundaries (close, optional whitespace, ',', optional
// whitespace, open).  Returns the position after ']' or nullptr.
template <class One>
const char* parse_array_parallel(const char* p, const char* e, char open, char close, int threads, One one,
                                 std::vector<double>* a, std::vector<double>* b) {
  p = skip_ws(p, e);
  if (p < e && *p == ']') return p + 1;
  // find the end of the array: elements are flat, so the first ']' that follows a `close` (or the
  // first one at all for the bracket kind) ends it - scan element-wise from the back is not
  // possible, so bound the region by a forward scan for the pattern close ws ']'
  const char* end = nullptr;
  for (const char* s = p; s < e; ++s) {
    if (*s == close) {
      const char* n = skip_ws(s + 1, e);
      if (n < e && *n == ']') {
        end = n;
        break;
      }
    }
  }
  if (!end) return nullptr;
  const size_t len = (size_t)(end - p);
  if (threads < 1) threads = 1;
  if (len < (1u << 20)) threads = 1;
  std::vector<const char*> cuts(threads + 1);
  cuts[0] = p;
  cuts[threads] = end;
  for (int t = 1; t < threads; ++t) {
    const char* s = p + len * t / threads;
    const char* found = end;
    for (; s < end; ++s) {
      if (*s != close) continue;
      const char* n = skip_ws(s + 1, end);
      if (n < end && *n == ',') {
        n = skip_ws(n + 1, end);
        if (n < end && *n == open) {
          found = n;
          break;
        }
      }
    }
    cuts[t] = found;
  }
  std::vector<std::vector<double>> pa(threads), pb(threads);
  std::atomic<bool> bad{false};
  auto work = [&](int t) {
    const char* q = cuts[t];
    const char* stop = cuts[t + 1];
    auto& va = pa[t];
    auto& vb = pb[t];
    va.reserve((size_t)(stop - q) / 24 + 16);
    vb.reserve((size_t)(stop - q) / 24 + 16);
    while (q < stop && !bad.load(std::memory_order_relaxed)) {
      q = skip_ws(q, stop);
      if (q >= stop) break;
      if (*q != open) {
        bad = true;
        return;
      }
      double x, y;
      q = one(q, end, &x, &y);
      if (!q) {
        bad = true;
        return;
      }
      va.push_back(x);
      vb.push_back(y);
      q = skip_ws(q, end);
      if (q < end && *q == ',') ++q;
    }
  };
  if (threads == 1) {
    work(0);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < threads; ++t) th.emplace_back(work, t);
    for (auto& x : th) x.join();
  }
  if (bad) return nullptr;
  size_t total = 0;
  for (int t = 0; t < threads; ++t) total += pa[t].size();
  a->reserve(a->size() + total);
  b->reserve(b->size() + total);
  for (int t = 0; t < threads; ++t) {
    a->insert(a->end(), pa[t].begin(), pa[t].end());
    b->insert(b->end(), pb[t].begin(), pb[t].end());
  }
  return end + 1;
}

// position after `"key"` ws ':' ws at or after p, or nullptr
const char* find_key(const char* p, const char* e, const char* key) {
  const std::string pat = std::string("\"") + key + "\"";
  for (;;) {
    const void* f = memmem(p, (size_t)(e - p), pat.data(), pat.size());
    if (!f) return nullptr;
    const char* q = skip_ws(reinterpret_cast<const char*>(f) + pat.size(), e);
    if (q < e && *q == ':') return skip_ws(q + 1, e);
    p = reinterpret_cast<const char*>(f) + 1;
  }
}

int parse_power_text(const char* b, const char* e, int threads, kvm_samples* out) {
  const char* p = skip_ws(b, e);
  if (p >= e) return KVM_ERR_UNSUPPORTED;
  if (*p == '{') {
    // {"samples": [...], ...}: the list under "samples" (collector.py:157-160); any other dict -> []
    const char* s = find_key(p, e, "samples");
    if (!s) {
      out->ts.emplace_back();
      out->val.emplace_back();
      return KVM_OK;
    }
    p = s;
  }
  if (p >= e || *p != '[') return KVM_ERR_UNSUPPORTED;
  out->ts.emplace_back();
  out->val.emplace_back();
  const char* q = parse_array_parallel(p + 1, e, '{', '}', threads, parse_sample, &out->ts[0], &out->val[0]);
  return q ? KVM_OK : KVM_ERR_UNSUPPORTED;
}

int parse_prom_text(const char* b, const char* e, int threads, kvm_samples* out) {
  // {"status":"success","data":{"resultType":"matrix","result":[{"metric":{..},"values":[[t,"v"],..]},..]}}
  const char* p = find_key(b, e, "result");
  if (!p || p >= e || *p != '[') return KVM_ERR_UNSUPPORTED;
  const char* rt = find_key(b, e, "resultType");
  const bool vector = rt && e - rt > 8 && memcmp(rt, "\"vector\"", 8) == 0;
  for (;;) {
    const char* v = find_key(p, e, vector ? "value" : "values");
    if (!v) break;
    out->ts.emplace_back();
    out->val.emplace_back();
    if (vector) {   // "value": [t, "v"] - one sample per series
      if (*v != '[') return KVM_ERR_UNSUPPORTED;
      double x, y;
      const char* q = parse_prom_pair(v, e, &x, &y);
      if (!q) return KVM_ERR_UNSUPPORTED;
      out->ts.back().push_back(x);
      out->val.back().push_back(y);
      p = q;
    } else {
      if (*v != '[') return KVM_ERR_UNSUPPORTED;
      const char* q = parse_array_parallel(v + 1, e, '[', ']', threads, parse_prom_pair, &out->ts.back(), &out->val.back());
      if (!q) return KVM_ERR_UNSUPPORTED;
      p = q;
    }
  }
  return KVM_OK;
}

}  // namespace

extern "C" int kvm_power_json_parse(const char* path, int32_t threads, kvm_samples** out) {
  KVM_REQUIRE(path && out, "kvm_power_json_parse: null argument");
  *out = nullptr;
  const int fd = open(path, O_RDONLY);
  if (fd < 0) {
    kvm_set_error("kvm_power_json_parse: cannot open %s", path);
    return KVM_ERR_INVALID;
  }
  struct stat st;
  if (fstat(fd, &st) != 0 || st.st_size == 0) {
    close(fd);
    kvm_set_error("kvm_power_json_parse: %s is empty or unreadable", path);
    return KVM_ERR_UNSUPPORTED;
  }
  void* map = mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
  close(fd);
  if (map == MAP_FAILED) {
    kvm_set_error("kvm_power_json_parse: mmap of %s failed", path);
    return KVM_ERR_INVALID;
  }
  kvm_samples* s = new kvm_samples();
  const char* b = reinterpret_cast<const char*>(map);
  const int rc = parse_power_text(b, b + st.st_size, threads, s);
  munmap(map, (size_t)st.st_size);
  if (rc != KVM_OK) {
    delete s;
    kvm_set_error("kvm_power_json_parse: %s is outside what the columnar scanner models; use the json module", path);
    return rc;
  }
  *out = s;
  return KVM_OK;
}

extern "C" int kvm_prom_result_parse(const char* text, int64_t len, int32_t threads, kvm_samples** out) {
  KVM_REQUIRE(text && out && len >= 0, "kvm_prom_result_parse: bad argument");
  *out = nullptr;
  kvm_samples* s = new kvm_samples();
  const int rc = parse_prom_text(text, text + len, threads, s);
  if (rc != KVM_OK) {
    delete s;
    kvm_set_error("kvm_prom_result_parse: response outside what the columnar scanner models; use the json module");
    return rc;
  }
  *out = s;
  return KVM_OK;
}

extern "C" int32_t kvm_samples_series(const kvm_samples* s) { return s ? (int32_t)s->ts.size() : 0; }
extern "C" int64_t kvm_samples_count(const kvm_samples* s, int32_t series) {
  return (s && series >= 0 && series < (int32_t)s->ts.size()) ? (int64_t)s->ts[series].size() : -1;
}
extern "C" int kvm_samples_copy(const kvm_samples* s, int32_t series, double* ts_host, double* values_host) {
  KVM_REQUIRE(s && series >= 0 && series < (int32_t)s->ts.size(), "kvm_samples_copy: bad series");
  const size_t n = s->ts[series].size();
  if (n) {
    KVM_REQUIRE(ts_host && values_host, "kvm_samples_copy: null output");
    memcpy(ts_host, s->ts[series].data(), n * sizeof(double));
    memcpy(values_host, s->val[series].data(), n * sizeof(double));
  }
  return KVM_OK;
}
extern "C" void kvm_samples_free(kvm_samples* s) { delete s; }
