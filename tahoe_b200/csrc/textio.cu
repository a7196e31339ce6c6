// textio.cu -- the reference's line-per-value text formats (host only).
//   model: BaseTahoeTest.h:267-315   data: BaseTahoeTest.h:354-388   
// Same grammar and the same leniency: each value is one line read into a 1024-byte
// buffer and parsed with atoi/atof semantics (garbage -> 0, trailing text ignored); a
// short file repeats the last line read, as the reference's unchecked fgets does.
// Unlike the reference an unreadable file is a status, not exit(1) (:273-277).
#include <cstdio>
#include <cstdlib>

#include "forest.h"

namespace {
constexpr int kMaxLine = 1024;  // BaseTahoeTest.h:269
struct LineReader {
  FILE* fp;
  char buf[kMaxLine];
  explicit LineReader(FILE* f) : fp(f) { buf[0] = 0; }
  void next() { if (!fgets(buf, kMaxLine, fp)) { /* keep previous contents */ } }
  int as_int() { next(); return atoi(buf); }
  float as_float() { next(); return (float)atof(buf); }
};
}  // namespace

extern "C" int tahoe_load_model_text(const char* path, tahoe_dense_node_t** nodes, int* num_trees,
                                     int* depth) {
  using namespace tahoe;
  if (!path || !nodes || !num_trees || !depth) { set_error("null argument"); return TAHOE_ERR_INVALID; }
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

extern "C" int tahoe_load_data_text(const char* path, float** data, int* num_rows, int* num_cols,
                                    float* missing) {
  using namespace tahoe;
  if (!path || !data || !num_rows || !num_cols || !missing) { set_error("null argument"); return TAHOE_ERR_INVALID; }
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

extern "C" void tahoe_free_host(void* p) { free(p); }
