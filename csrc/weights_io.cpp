// weights_io.cpp — parameter-file layout, reader/writer, deterministic
// random-init generator, the DP partitioner and the error slot.
//
// Layout source: ref:375-407 slices one flat float buffer at OFFSET0..OFFSET26
// (macros from the unshipped namegen.h).  The tensor order there is
//   embedding | W_i{r,z,n} layer0, layer1 | W_h{r,z,n} layer0, layer1 |
//   b_i{r,z,n} layer0, layer1 | b_h{r,z,n} layer0, layer1 | W_fc | b_fc
// and the offsets are taken to be contiguous.  Unlike the reference, which
// has every rank read the file and never frees it (ref:369-372), the Python
// layer reads once per process and frees after upload.
#include <cerrno>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "errors.hpp"
#include "layout.hpp"

namespace b200gru {
std::string &last_error_slot() {
  static thread_local std::string slot;
  return slot;
}

static int check_cfg(const b200gru_config *c) {
  if (!c) return fail(B200GRU_ERR_ARG, "config is null");
  if (c->num_char < 2 || c->num_char > 65536)
    return fail(B200GRU_ERR_ARG, "num_char=%d out of range", c->num_char);
  if (c->emb_dim < 1 || c->hidden < 1)
    return fail(B200GRU_ERR_ARG, "emb_dim/hidden must be positive");
  if (c->num_layers < 1 || c->num_layers > B200GRU_MAX_LAYERS)
    return fail(B200GRU_ERR_ARG, "num_layers=%d not in [1,%d]", c->num_layers,
                B200GRU_MAX_LAYERS);
  if (c->max_len < 1) return fail(B200GRU_ERR_ARG, "max_len must be >= 1");
  if (c->sos < 0 || c->sos >= c->num_char || c->eos < 0 ||
      c->eos >= c->num_char)
    return fail(B200GRU_ERR_ARG, "sos/eos outside [0,num_char)");
  return 0;
}

std::vector<TensorSpec> build_layout(const b200gru_config &c) {
  std::vector<TensorSpec> v;
  int64_t off = 0;
  auto push = [&](std::string name, int rows, int cols) {
    int64_t sz = (int64_t)rows * cols;
    v.push_back({std::move(name), off, sz, rows, cols});
    off += sz;
  };
  const char *gates = "rzn";
  push("character_embedding", c.num_char, c.emb_dim);
  for (int l = 0; l < c.num_layers; ++l)
    for (int g = 0; g < 3; ++g)
      push(std::string("W_i") + gates[g] + std::to_string(l), c.hidden,
           l == 0 ? c.emb_dim : c.hidden);
  for (int l = 0; l < c.num_layers; ++l)
    for (int g = 0; g < 3; ++g)
      push(std::string("W_h") + gates[g] + std::to_string(l), c.hidden,
           c.hidden);
  for (int l = 0; l < c.num_layers; ++l)
    for (int g = 0; g < 3; ++g)
      push(std::string("b_i") + gates[g] + std::to_string(l), c.hidden, 1);
  for (int l = 0; l < c.num_layers; ++l)
    for (int g = 0; g < 3; ++g)
      push(std::string("b_h") + gates[g] + std::to_string(l), c.hidden, 1);
  push("W_fc", c.num_char, c.hidden);
  push("b_fc", c.num_char, 1);
  return v;
}

int validate_config(const b200gru_config *c) { return check_cfg(c); }

static inline uint64_t splitmix64(uint64_t &s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
// 24 random mantissa bits -> [0,1)
static inline float u01(uint64_t &s) {
  return (float)(splitmix64(s) >> 40) * (1.0f / 16777216.0f);
}
}  // namespace b200gru

using namespace b200gru;

extern "C" {

const char *b200gru_last_error(void) { return last_error_slot().c_str(); }
const char *b200gru_version(void) { return "b200gru 0.1 (sm_100a)"; }

int b200gru_num_tensors(const b200gru_config *cfg) {
  if (int rc = check_cfg(cfg)) return rc;
  return 3 + 12 * cfg->num_layers;
}

int64_t b200gru_param_count(const b200gru_config *cfg) {
  if (int rc = check_cfg(cfg)) return rc;
  auto v = build_layout(*cfg);
  return v.back().offset + v.back().size;
}

int b200gru_param_layout(const b200gru_config *cfg, int64_t *offsets,
                         int64_t *sizes, int32_t *rows, int32_t *cols) {
  if (int rc = check_cfg(cfg)) return rc;
  auto v = build_layout(*cfg);
  for (size_t i = 0; i < v.size(); ++i) {
    if (offsets) offsets[i] = v[i].offset;
    if (sizes) sizes[i] = v[i].size;
    if (rows) rows[i] = v[i].rows;
    if (cols) cols[i] = v[i].cols;
  }
  return 0;
}

const char *b200gru_tensor_name(const b200gru_config *cfg, int k) {
  static thread_local std::string name;
  if (check_cfg(cfg)) return nullptr;
  auto v = build_layout(*cfg);
  if (k < 0 || k >= (int)v.size()) {
    fail(B200GRU_ERR_ARG, "tensor index %d out of range", k);
    return nullptr;
  }
  name = v[k].name;
  return name.c_str();
}

int b200gru_read_binary(const char *path, void **out_buf, size_t *out_bytes) {
  if (!path || !out_buf || !out_bytes)
    return fail(B200GRU_ERR_ARG, "null argument");
  FILE *f = fopen(path, "rb");
  if (!f) return fail(B200GRU_ERR_IO, "open %s: %s", path, strerror(errno));
  fseek(f, 0, SEEK_END);
  long sz = ftell(f);
  rewind(f);
  if (sz < 0) {
    fclose(f);
    return fail(B200GRU_ERR_IO, "ftell %s failed", path);
  }
  void *buf = malloc(sz > 0 ? (size_t)sz : 1);
  if (!buf) {
    fclose(f);
    return fail(B200GRU_ERR_IO, "malloc of %ld bytes failed", sz);
  }
  size_t got = fread(buf, 1, (size_t)sz, f);
  fclose(f);
  if (got != (size_t)sz) {
    free(buf);
    return fail(B200GRU_ERR_IO, "short read on %s: %zu of %ld", path, got, sz);
  }
  *out_buf = buf;
  *out_bytes = (size_t)sz;
  return 0;
}

int b200gru_write_binary(const char *path, const void *buf, size_t bytes) {
  if (!path || (!buf && bytes)) return fail(B200GRU_ERR_ARG, "null argument");
  FILE *f = fopen(path, "wb");
  if (!f) return fail(B200GRU_ERR_IO, "open %s: %s", path, strerror(errno));
  size_t put = fwrite(buf, 1, bytes, f);
  int rc = fclose(f);
  if (put != bytes || rc != 0)
    return fail(B200GRU_ERR_IO, "short write on %s", path);
  return 0;
}

void b200gru_free_host(void *p) { free(p); }

int b200gru_random_init(const b200gru_config *cfg, uint64_t seed, float *out,
                        int64_t n_floats) {
  if (int rc = check_cfg(cfg)) return rc;
  int64_t need = b200gru_param_count(cfg);
  if (!out || n_floats != need)
    return fail(B200GRU_ERR_ARG, "buffer holds %lld floats, layout needs %lld",
                (long long)n_floats, (long long)need);
  auto v = build_layout(*cfg);
  const float bound = 1.0f / sqrtf((float)cfg->hidden);
  for (size_t t = 0; t < v.size(); ++t) {
    uint64_t s = seed * 0x100000001B3ull + 0x51ED270B7F4A7C15ull * (t + 1);
    // embedding rows are N(0,1)-ish in torch; a wider uniform keeps the same
    // scale without needing a normal generator.
    const float b = (t == 0) ? 1.7320508f : bound;
    float *dst = out + v[t].offset;
    for (int64_t i = 0; i < v[t].size; ++i) dst[i] = (2.0f * u01(s) - 1.0f) * b;
  }
  return 0;
}

int b200gru_random_floats(uint64_t seed, float *out, int64_t n) {
  if (!out && n) return fail(B200GRU_ERR_ARG, "null output");
  uint64_t s = seed ^ 0xD1B54A32D192ED03ull;
  for (int64_t i = 0; i < n; ++i) out[i] = u01(s);
  return 0;
}

void b200gru_partition(int n_names, int world_size, int rank, int *jpp,
                       int *start, int *end) {
  int per = world_size > 0 ? n_names / world_size : 0;
  if (jpp) *jpp = per;
  if (start) *start = per * rank;
  if (end) *end = per * rank + per;
}

}  // extern "C"
