// textio.cu -- the reference's line-per-value text formats (host only), and a binary cache of them.
//   model: BaseTahoeTest.h:267-315   data: BaseTahoeTest.h:354-388   
// Same grammar and the same leniency as the reference: each value is one line read into a
// 1024-byte buffer and parsed with atoi/atof semantics (garbage -> 0, trailing text ignored,
// float = (float)atof: double first, then narrowed); a short file repeats the last line read, as the
// reference's unchecked fgets does.  Unlike the reference an unreadable file is a status, not
// exit(1) (:273-277).
//
// The reference parses with fgets + atof on one thread (Higgs-shaped data = 280 M lines).  Here the
// file is mapped, cut into byte ranges at line starts, and the ranges are parsed by all host threads:
// pass 1 counts lines per range, a prefix sum gives every range the index of its first value, pass 2
// parses straight into the output array.  Plain decimals take Clinger's exact fast path (<= 19 digits
// that fit 2^53, |exp10| <= 22: one correctly rounded double operation, the result atof would give);
// everything else (hex floats, inf/nan, long mantissas, garbage) goes through atof itself on a
// NUL-terminated copy of the line.  Files the line model does not cover (a line of 1024+ bytes, which
// fgets would split, or an embedded NUL) fall back to the sequential fgets reader.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "forest.h"

namespace {
constexpr int kMaxLine = 1024;  // BaseTahoeTest.h:269

// ---- sequential reader: the reference's loop, kept as the fallback ------------------------------
struct LineReader {
  FILE* fp;
  char buf[kMaxLine];
  explicit LineReader(FILE* f) : fp(f) { buf[0] = 0; }
  void next() { if (!fgets(buf, kMaxLine, fp)) { /* keep previous contents */ } }
  int as_int() { next(); return atoi(buf); }
  float as_float() { next(); return (float)atof(buf); }
};

// ---- parsing one line ------------------------------------------------------------------------
inline bool is_space(char c) { return c == ' ' || (c >= '\t' && c <= '\r'); }   // isspace in the C locale

inline int slow_int(const char* p, const char* end) {
  char buf[kMaxLine];
  const size_t n = std::min<size_t>((size_t)(end - p), kMaxLine - 1);
  std::memcpy(buf, p, n);
  buf[n] = 0;
  return atoi(buf);
}
inline float slow_float(const char* p, const char* end) {
  char buf[kMaxLine];
  const size_t n = std::min<size_t>((size_t)(end - p), kMaxLine - 1);
  std::memcpy(buf, p, n);
  buf[n] = 0;
  return (float)atof(buf);
}

// atoi on the line [p, end): optional whitespace, sign, digits; anything else ends the number
inline int parse_int(const char* p, const char* end) {
  const char* s = p;
  while (s < end && is_space(*s)) ++s;
  bool neg = false;
  if (s < end && (*s == '-' || *s == '+')) { neg = *s == '-'; ++s; }
  int64_t v = 0;
  int digits = 0;
  while (s < end && *s >= '0' && *s <= '9') {
    v = v * 10 + (*s - '0');
    ++s;
    if (++digits > 9) return slow_int(p, end);      // could leave int range: whatever atoi does
  }
  return (int)(neg ? -v : v);
}

// (float)atof on the line [p, end)
inline float parse_float(const char* p, const char* end) {
  static const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                    1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
  const char* s = p;
  while (s < end && is_space(*s)) ++s;
  bool neg = false;
  if (s < end && (*s == '-' || *s == '+')) { neg = *s == '-'; ++s; }
  uint64_t mant = 0;
  int digits = 0, frac = 0;
  bool any = false;
  const char* t = s;
  while (t < end && *t >= '0' && *t <= '9') {
    if (mant || *t != '0') { mant = mant * 10 + (uint64_t)(*t - '0'); ++digits; }
    any = true;
    ++t;
  }
  if (t < end && *t == '.') {
    ++t;
    while (t < end && *t >= '0' && *t <= '9') {
      if (mant || *t != '0') { mant = mant * 10 + (uint64_t)(*t - '0'); ++digits; }
      ++frac;
      any = true;
      ++t;
    }
  }
  if (!any) return slow_float(p, end);   // "inf", "nan", ".", garbage (0): rare, let atof decide
  // "0x..." is a hex float for atof; an 'x' right after a lone leading zero
  if (t < end && (*t == 'x' || *t == 'X')) return slow_float(p, end);
  int exp10 = -frac;
  if (t < end && (*t == 'e' || *t == 'E')) {
    const char* u = t + 1;
    bool eneg = false;
    if (u < end && (*u == '-' || *u == '+')) { eneg = *u == '-'; ++u; }
    if (u < end && *u >= '0' && *u <= '9') {
      int e = 0;
      while (u < end && *u >= '0' && *u <= '9') {
        e = e * 10 + (*u - '0');
        if (e > 100000) return slow_float(p, end);
        ++u;
      }
      exp10 += eneg ? -e : e;
    }   // "1e" / "1e+" : the exponent is not part of the number
  }
  if (digits > 19 || mant > ((uint64_t)1 << 53) || exp10 < -22 || exp10 > 22) return slow_float(p, end);
  double d = (double)mant;                         // exact
  d = exp10 < 0 ? d / kPow10[-exp10] : d * kPow10[exp10];   // one correctly rounded operation
  return (float)(neg ? -d : d);
}

// ---- mapped file -----------------------------------------------------------------------------
struct Mapped {
  const char* p = nullptr;
  size_t n = 0;
  int fd = -1;
  bool open(const char* path) {
    fd = ::open(path, O_RDONLY);
    if (fd < 0) return false;
    struct stat st;
    if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) { ::close(fd); fd = -1; return false; }
    n = (size_t)st.st_size;
    if (n == 0) { p = ""; return true; }
    void* m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE | MAP_POPULATE, fd, 0);   // prefault: 8 parser threads faulting pages one by one serialise on the address-space lock
    if (m == MAP_FAILED) { ::close(fd); fd = -1; return false; }
    madvise(m, n, MADV_SEQUENTIAL);
    p = (const char*)m;
    return true;
  }
  ~Mapped() {
    if (p && n) munmap((void*)p, n);
    if (fd >= 0) ::close(fd);
  }
};

int host_threads() {
  int t = (int)std::thread::hardware_concurrency();
  if (const char* e = std::getenv("TAHOE_LOADER_THREADS")) t = std::atoi(e);
  return std::max(1, std::min(t, 64));
}

template <typename F>
void parallel_for(int n, F body) {
  if (n <= 1) { if (n == 1) body(0); return; }
  std::vector<std::thread> th;
  th.reserve((size_t)n);
  for (int i = 0; i < n; ++i) th.emplace_back([&body, i] { body(i); });
  for (auto& t : th) t.join();
}

// The file as lines: ranges cut at line starts, lines counted per range.  ok = false when the line
// model does not hold (a line of >= kMaxLine bytes incl. its newline, or a NUL byte).
struct LineIndex {
  std::vector<size_t> start;   // [ranges + 1] byte offsets, each at a line start
  std::vector<int64_t> first;  // [ranges + 1] index of the first line of each range; back() = total lines
  bool ok = true;
};

LineIndex index_lines(const Mapped& m, int threads) {
  LineIndex ix;
  size_t min_range = 1 << 16;                      // bytes; below this a thread is not worth starting
  if (const char* e = std::getenv("TAHOE_LOADER_MIN_RANGE")) min_range = (size_t)std::max(1, std::atoi(e));
  const int ranges = (int)std::max<size_t>(1, std::min<size_t>((size_t)threads, m.n / min_range + 1));
  ix.start.assign((size_t)ranges + 1, m.n);
  ix.start[0] = 0;
  for (int r = 1; r < ranges; ++r) {
    size_t pos = m.n / (size_t)ranges * (size_t)r;
    const void* nl = pos < m.n ? std::memchr(m.p + pos, '\n', m.n - pos) : nullptr;
    ix.start[r] = nl ? (size_t)((const char*)nl - m.p) + 1 : m.n;
  }
  for (int r = 1; r <= ranges; ++r) ix.start[r] = std::max(ix.start[r], ix.start[r - 1]);
  std::vector<int64_t> count((size_t)ranges, 0);
  std::atomic<bool> ok{true};
  parallel_for(ranges, [&](int r) {
    const char* p = m.p + ix.start[r];
    const char* end = m.p + ix.start[r + 1];
    int64_t lines = 0;
    while (p < end) {
      const char* nl = (const char*)std::memchr(p, '\n', (size_t)(end - p));
      const char* stop = nl ? nl + 1 : end;
      if (stop - p >= kMaxLine || std::memchr(p, 0, (size_t)(stop - p))) { ok = false; return; }
      ++lines;
      p = stop;
    }
    count[r] = lines;
  });
  ix.ok = ok;
  ix.first.assign((size_t)ranges + 1, 0);
  for (int r = 0; r < ranges; ++r) ix.first[r + 1] = ix.first[r] + count[r];
  return ix;
}

// Calls fn(line_index, begin, end_without_newline) for every line of the file, in parallel.
template <typename F>
void for_each_line(const Mapped& m, const LineIndex& ix, F fn) {
  const int ranges = (int)ix.start.size() - 1;
  parallel_for(ranges, [&](int r) {
    const char* p = m.p + ix.start[r];
    const char* end = m.p + ix.start[r + 1];
    int64_t line = ix.first[r];
    while (p < end) {
      const char* nl = (const char*)std::memchr(p, '\n', (size_t)(end - p));
      fn(line, p, nl ? nl : end);
      ++line;
      p = nl ? nl + 1 : end;
    }
  });
}

// the text of line `i` (the last line for i past the end: the reference's buffer keeps it)
struct LineFetch {
  const Mapped& m;
  std::vector<std::pair<const char*, const char*>> head;   // the first few lines (headers)
  std::pair<const char*, const char*> last{nullptr, nullptr};
  LineFetch(const Mapped& mm, int want) : m(mm) {
    const char* p = m.p;
    const char* end = m.p + m.n;
    while (p < end && (int)head.size() < want) {
      const char* nl = (const char*)std::memchr(p, '\n', (size_t)(end - p));
      head.emplace_back(p, nl ? nl : end);
      p = nl ? nl + 1 : end;
    }
    // last line of the file
    if (m.n) {
      const char* e = end;
      if (e[-1] == '\n') --e;
      const char* b = e;
      while (b > m.p && b[-1] != '\n') --b;
      last = {b, e};
    }
  }
  std::pair<const char*, const char*> line(int i) const {
    if (i < (int)head.size()) return head[(size_t)i];
    if (last.first) return last;
    return {"", ""};
  }
};

int load_model_sequential(const char* path, tahoe_dense_node_t** nodes, int* num_trees, int* depth) {
  using namespace tahoe;
  FILE* fp = fopen(path, "r");
  if (!fp) { set_error(std::string("fail to read ") + path); return TAHOE_ERR_IO; }
  LineReader rd(fp);
  const int T = rd.as_int();
  const int d = rd.as_int() - 1;  // line 2 holds LEVELS = depth + 1 (:281-282)
  if (T < 0 || d < 0 || d > 26) { fclose(fp); set_error("bad model header"); return TAHOE_ERR_INVALID; }
  const int64_t total = (int64_t)T * (((int64_t)1 << (d + 1)) - 1);
  tahoe_dense_node_t* out = (tahoe_dense_node_t*)calloc((size_t)(total ? total : 1), sizeof(*out));
  if (!out) { fclose(fp); set_error("out of host memory"); return TAHOE_ERR_NOMEM; }
  for (int64_t i = 0; i < total; ++i) {
    const int fid = rd.as_int();
    const float value = rd.as_float();
    const bool def_left = rd.as_int() != 0;
    const float weight = rd.as_float();
    const bool is_leaf = rd.as_int() != 0;
    out[i].weight = weight;
    out[i].val = value;
    out[i].bits = (fid & 0x3FFFFFFF) | (def_left ? 0x40000000 : 0) | (is_leaf ? (int32_t)0x80000000 : 0);
  }
  fclose(fp);
  *nodes = out; *num_trees = T; *depth = d;
  return TAHOE_OK;
}

int load_data_sequential(const char* path, float** data, int* num_rows, int* num_cols, float* missing) {
  using namespace tahoe;
  FILE* fp = fopen(path, "r");
  if (!fp) { set_error(std::string("fail to read ") + path); return TAHOE_ERR_IO; }
  LineReader rd(fp);
  const int R = rd.as_int();
  const int C = rd.as_int();
  const float miss = rd.as_float();
  if (R < 0 || C < 0) { fclose(fp); set_error("bad data header"); return TAHOE_ERR_INVALID; }
  const int64_t total = (int64_t)R * C;  // 64-bit: the reference multiplies two ints
  float* out = (float*)calloc((size_t)(total ? total : 1), sizeof(float));
  if (!out) { fclose(fp); set_error("out of host memory"); return TAHOE_ERR_NOMEM; }
  for (int64_t i = 0; i < total; ++i) out[i] = rd.as_float();
  fclose(fp);
  *data = out; *num_rows = R; *num_cols = C; *missing = miss;
  return TAHOE_OK;
}

bool sequential_forced() {
  const char* e = std::getenv("TAHOE_LOADER_SEQUENTIAL");
  return e && std::atoi(e) != 0;
}

// ---- binary cache ----------------------------------------------------------------------------
// 64-byte header + the raw little-endian payload (12-byte AoS nodes / float32 rows) + nothing else.
struct CacheHeader {
  char magic[8];        // "TAHOEB2\0"
  uint32_t version;     // 1
  uint32_t kind;        // 1 = model (dense AoS nodes), 2 = data (row-major float32)
  int32_t a, b;         // model: num_trees, depth     data: num_rows, num_cols
  float missing;        // data only
  uint32_t reserved;
  uint64_t payload_bytes;
  uint64_t checksum;    // FNV-1a 64 over the payload, 8 bytes at a time
  uint64_t pad[2];
};
static_assert(sizeof(CacheHeader) == 64, "cache header is 64 bytes");
constexpr char kMagic[8] = {'T', 'A', 'H', 'O', 'E', 'B', '2', 0};

uint64_t checksum64(const void* data, size_t bytes) {
  const uint8_t* p = (const uint8_t*)data;
  uint64_t h = 0xcbf29ce484222325ull;
  size_t i = 0;
  for (; i + 8 <= bytes; i += 8) {
    uint64_t w;
    std::memcpy(&w, p + i, 8);
    h = (h ^ w) * 0x100000001b3ull;
  }
  for (; i < bytes; ++i) h = (h ^ p[i]) * 0x100000001b3ull;
  return h;
}

int save_cache(const char* path, uint32_t kind, int a, int b, float missing, const void* payload, size_t bytes) {
  using namespace tahoe;
  CacheHeader h{};
  std::memcpy(h.magic, kMagic, 8);
  h.version = 1; h.kind = kind; h.a = a; h.b = b; h.missing = missing;
  h.payload_bytes = bytes;
  h.checksum = checksum64(payload, bytes);
  const std::string tmp = std::string(path) + ".tmp";
  FILE* fp = fopen(tmp.c_str(), "wb");
  if (!fp) { set_error(std::string("fail to write ") + tmp); return TAHOE_ERR_IO; }
  const bool ok = fwrite(&h, sizeof(h), 1, fp) == 1 && (bytes == 0 || fwrite(payload, 1, bytes, fp) == bytes);
  if (fclose(fp) != 0 || !ok || std::rename(tmp.c_str(), path) != 0) {
    std::remove(tmp.c_str());
    set_error(std::string("fail to write ") + path);
    return TAHOE_ERR_IO;
  }
  return TAHOE_OK;
}

int load_cache(const char* path, uint32_t kind, CacheHeader* h, void** payload) {
  using namespace tahoe;
  FILE* fp = fopen(path, "rb");
  if (!fp) { set_error(std::string("fail to read ") + path); return TAHOE_ERR_IO; }
  if (fread(h, sizeof(*h), 1, fp) != 1 || std::memcmp(h->magic, kMagic, 8) != 0) {
    fclose(fp); set_error("not a tahoe_b200 binary cache"); return TAHOE_ERR_INVALID;
  }
  if (h->version != 1 || h->kind != kind) {
    fclose(fp); set_error("binary cache of another version or kind"); return TAHOE_ERR_INVALID;
  }
  void* buf = malloc(h->payload_bytes ? h->payload_bytes : 1);
  if (!buf) { fclose(fp); set_error("out of host memory"); return TAHOE_ERR_NOMEM; }
  const bool ok = h->payload_bytes == 0 || fread(buf, 1, h->payload_bytes, fp) == h->payload_bytes;
  char extra;
  const bool tail = fread(&extra, 1, 1, fp) == 1;
  fclose(fp);
  if (!ok || tail || checksum64(buf, h->payload_bytes) != h->checksum) {
    free(buf); set_error("binary cache is truncated or corrupt"); return TAHOE_ERR_INVALID;
  }
  *payload = buf;
  return TAHOE_OK;
}
}  // namespace

extern "C" int tahoe_load_model_text(const char* path, tahoe_dense_node_t** nodes, int* num_trees,
                                     int* depth) {
  using namespace tahoe;
  if (!path || !nodes || !num_trees || !depth) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  if (sequential_forced()) return load_model_sequential(path, nodes, num_trees, depth);
  Mapped m;
  if (!m.open(path)) { set_error(std::string("fail to read ") + path); return TAHOE_ERR_IO; }
  const LineIndex ix = index_lines(m, host_threads());
  if (!ix.ok) return load_model_sequential(path, nodes, num_trees, depth);
  const LineFetch lf(m, 2);
  const int T = parse_int(lf.line(0).first, lf.line(0).second);
  const int d = parse_int(lf.line(1).first, lf.line(1).second) - 1;  // line 2 holds LEVELS = depth + 1 (:281-282)
  if (T < 0 || d < 0 || d > 26) { set_error("bad model header"); return TAHOE_ERR_INVALID; }
  const int64_t total = (int64_t)T * (((int64_t)1 << (d + 1)) - 1);
  tahoe_dense_node_t* out = (tahoe_dense_node_t*)calloc((size_t)(total ? total : 1), sizeof(*out));
  if (!out) { set_error("out of host memory"); return TAHOE_ERR_NOMEM; }
  // value line L >= 2 is field (L - 2) % 5 of node (L - 2) / 5: fid, value, default_left, weight, is_leaf
  auto put = [&](int64_t v, const char* b, const char* e) {
    const int64_t node = v / 5;
    if (node >= total) return;
    switch ((int)(v % 5)) {
      case 0: out[node].bits |= parse_int(b, e) & 0x3FFFFFFF; break;
      case 1: out[node].val = parse_float(b, e); break;
      case 2: if (parse_int(b, e) != 0) out[node].bits |= 0x40000000; break;
      case 3: out[node].weight = parse_float(b, e); break;
      default: if (parse_int(b, e) != 0) out[node].bits |= (int32_t)0x80000000; break;
    }
  };
  // each node's `bits` is assembled by ORs from three lines that may sit in two ranges: different
  // threads then touch the same word, so those ORs go through an atomic
  auto put_mt = [&](int64_t v, const char* b, const char* e) {
    const int64_t node = v / 5;
    if (node >= total) return;
    const int field = (int)(v % 5);
    if (field == 1) { out[node].val = parse_float(b, e); return; }
    if (field == 3) { out[node].weight = parse_float(b, e); return; }
    const int x = parse_int(b, e);
    const int32_t bits = field == 0 ? (x & 0x3FFFFFFF) : (x != 0 ? (field == 2 ? 0x40000000 : (int32_t)0x80000000) : 0);
    if (bits) __atomic_fetch_or(&out[node].bits, bits, __ATOMIC_RELAXED);
  };
  for_each_line(m, ix, [&](int64_t line, const char* b, const char* e) { if (line >= 2) put_mt(line - 2, b, e); });
  // a short file: the remaining values all re-read the last line (the reference's fgets fails and
  // leaves the buffer as it was)
  const int64_t have = std::max<int64_t>(ix.first.back() - 2, 0);
  const std::pair<const char*, const char*> last = lf.line(1 << 30);
  for (int64_t v = have; v < total * 5; ++v) put(v, last.first, last.second);
  *nodes = out; *num_trees = T; *depth = d;
  return TAHOE_OK;
}

extern "C" int tahoe_load_data_text(const char* path, float** data, int* num_rows, int* num_cols,
                                    float* missing) {
  using namespace tahoe;
  if (!path || !data || !num_rows || !num_cols || !missing) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  if (sequential_forced()) return load_data_sequential(path, data, num_rows, num_cols, missing);
  Mapped m;
  if (!m.open(path)) { set_error(std::string("fail to read ") + path); return TAHOE_ERR_IO; }
  const LineIndex ix = index_lines(m, host_threads());
  if (!ix.ok) return load_data_sequential(path, data, num_rows, num_cols, missing);
  const LineFetch lf(m, 3);
  const int R = parse_int(lf.line(0).first, lf.line(0).second);
  const int C = parse_int(lf.line(1).first, lf.line(1).second);
  const float miss = parse_float(lf.line(2).first, lf.line(2).second);
  if (R < 0 || C < 0) { set_error("bad data header"); return TAHOE_ERR_INVALID; }
  const int64_t total = (int64_t)R * C;  // 64-bit: the reference multiplies two ints
  float* out = (float*)malloc((size_t)(total ? total : 1) * sizeof(float));
  if (!out) { set_error("out of host memory"); return TAHOE_ERR_NOMEM; }
  for_each_line(m, ix, [&](int64_t line, const char* b, const char* e) {
    if (line >= 3 && line - 3 < total) out[line - 3] = parse_float(b, e);
  });
  const int64_t have = std::min<int64_t>(std::max<int64_t>(ix.first.back() - 3, 0), total);
  if (have < total) {
    const std::pair<const char*, const char*> last = lf.line(1 << 30);
    const float v = parse_float(last.first, last.second);
    for (int64_t i = have; i < total; ++i) out[i] = v;
  }
  *data = out; *num_rows = R; *num_cols = C; *missing = miss;
  return TAHOE_OK;
}

extern "C" int tahoe_save_model_binary(const char* path, const tahoe_dense_node_t* nodes, int num_trees, int depth) {
  using namespace tahoe;
  if (!path || num_trees < 0 || depth < 0 || depth > 26 || (!nodes && num_trees)) { set_error("invalid argument"); return TAHOE_ERR_INVALID; }
  const size_t total = (size_t)num_trees * (((size_t)1 << (depth + 1)) - 1);
  return save_cache(path, 1, num_trees, depth, 0.0f, nodes, total * sizeof(tahoe_dense_node_t));
}

extern "C" int tahoe_load_model_binary(const char* path, tahoe_dense_node_t** nodes, int* num_trees, int* depth) {
  using namespace tahoe;
  if (!path || !nodes || !num_trees || !depth) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  CacheHeader h;
  void* buf = nullptr;
  int rc = load_cache(path, 1, &h, &buf);
  if (rc) return rc;
  const bool sane = h.a >= 0 && h.b >= 0 && h.b <= 26 &&
                    h.payload_bytes == (uint64_t)h.a * ((((uint64_t)1) << (h.b + 1)) - 1) * sizeof(tahoe_dense_node_t);
  if (!sane) { free(buf); set_error("binary model cache: header and payload disagree"); return TAHOE_ERR_INVALID; }
  *nodes = (tahoe_dense_node_t*)buf; *num_trees = h.a; *depth = h.b;
  return TAHOE_OK;
}

extern "C" int tahoe_save_data_binary(const char* path, const float* data, int num_rows, int num_cols, float missing) {
  using namespace tahoe;
  if (!path || num_rows < 0 || num_cols < 0 || (!data && num_rows && num_cols)) { set_error("invalid argument"); return TAHOE_ERR_INVALID; }
  return save_cache(path, 2, num_rows, num_cols, missing, data, (size_t)num_rows * (size_t)num_cols * sizeof(float));
}

extern "C" int tahoe_load_data_binary(const char* path, float** data, int* num_rows, int* num_cols, float* missing) {
  using namespace tahoe;
  if (!path || !data || !num_rows || !num_cols || !missing) { set_error("null argument"); return TAHOE_ERR_INVALID; }
  CacheHeader h;
  void* buf = nullptr;
  int rc = load_cache(path, 2, &h, &buf);
  if (rc) return rc;
  if (h.a < 0 || h.b < 0 || h.payload_bytes != (uint64_t)h.a * (uint64_t)h.b * sizeof(float)) {
    free(buf); set_error("binary data cache: header and payload disagree"); return TAHOE_ERR_INVALID;
  }
  *data = (float*)buf; *num_rows = h.a; *num_cols = h.b; *missing = h.missing;
  return TAHOE_OK;
}

extern "C" void tahoe_free_host(void* p) { free(p); }
