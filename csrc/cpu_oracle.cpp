// cpu_oracle.cpp — Oracle A: the reference's operator semantics on the CPU.
//
// The reference's GPU softmax is racy (ref:294-300), so its binary is not a
// usable oracle.  What is: the CPU bodies left as comments next to each kernel
// (ref:119-125, 140-145, 160-166, 180-185, 199-205, 219-225, 243-253, 302-313)
// and the live CPU sampler (ref:322-333).  This file restates those semantics
// in the same association order as the launch sequence ref:664-873:
//   r  = sigmoid(((W_ir x + b_ir) + W_hr h) + b_hr)
//   z  = sigmoid(((W_iz x + b_iz) + W_hz h) + b_hz)
//   n  = tanh((W_in x + b_in) + r * (W_hn h + b_hn))
//   h' = (1 - z) * n + z * h            (two rounded products, one add)
//   p  = exp(f) / sum_i exp(f_i)        (index-order sum, no max subtraction)
// Dot products accumulate left to right over K with a fused multiply-add, as
// the reference's live matvec kernel does once nvcc contracts `c += a*b`
// (ref:237-239).  Transcendentals come from exact_math.h by default so the
// fp32 CUDA kernel can match bit for bit; B200GRU_ORACLE_LIBM switches to libm.
#include <cmath>
#include <cstring>
#include <vector>

#include "errors.hpp"
#include "exact_math.h"
#include "layout.hpp"

namespace {

struct Math {
  bool libm;
  float exp(float x) const { return libm ? expf(x) : b200_expf(x); }
  float tanh(float x) const { return libm ? tanhf(x) : b200_tanhf(x); }
  float sigmoid(float x) const {
    return libm ? (float)(1.0 / (1.0 + (double)expf(-x))) : b200_sigmoidf(x);
  }
};

inline float dot_lr(const float *w, const float *x, int k) {
  float c = 0.0f;
  for (int j = 0; j < k; ++j) c = fmaf(w[j], x[j], c);
  return c;
}

// One GRU layer step for one name.  w_ih/w_hh are [3H, in]/[3H, H] with gate
// blocks r,z,n stacked (the file stores the three blocks back to back, so a
// pointer to W_ir is a pointer to the stacked matrix).
void gru_cell(const Math &m, const float *w_ih, const float *w_hh,
              const float *b_ih, const float *b_hh, const float *x, float *h,
              int in_dim, int H, float *scratch /* [H] */) {
  for (int i = 0; i < H; ++i) {
    float ir = dot_lr(w_ih + (size_t)(0 * H + i) * in_dim, x, in_dim);
    float hr = dot_lr(w_hh + (size_t)(0 * H + i) * H, h, H);
    float r = m.sigmoid(((ir + b_ih[0 * H + i]) + hr) + b_hh[0 * H + i]);
    float iz = dot_lr(w_ih + (size_t)(1 * H + i) * in_dim, x, in_dim);
    float hz = dot_lr(w_hh + (size_t)(1 * H + i) * H, h, H);
    float z = m.sigmoid(((iz + b_ih[1 * H + i]) + hz) + b_hh[1 * H + i]);
    float in_ = dot_lr(w_ih + (size_t)(2 * H + i) * in_dim, x, in_dim);
    float hn = dot_lr(w_hh + (size_t)(2 * H + i) * H, h, H);
    float n = m.tanh((in_ + b_ih[2 * H + i]) + r * (hn + b_hh[2 * H + i]));
    float a = 1.0f - z;  // ref:157 computes 1.0 - x in fp64; same rounding
    float b = a * n;
    float c = z * h[i];
    scratch[i] = b + c;
  }
  memcpy(h, scratch, sizeof(float) * H);
}

void softmax(const Math &m, const float *x, float *y, int n) {
  float sum = 0.0f;
  for (int i = 0; i < n; ++i) sum += m.exp(x[i]);
  for (int i = 0; i < n; ++i) y[i] = m.exp(x[i]) / sum;
}

}  // namespace

using namespace b200gru;

extern "C" {

int b200gru_random_select(const float *probs, int n, float r) {
  float psum = 0.0f;
  for (int i = 0; i < n; ++i) {
    psum += probs[i];
    if (psum > r) return i;
  }
  return n - 1;
}

int b200gru_oracle_matvec(const float *w, const float *x, float *y, int n,
                          int k) {
  B200_REQUIRE(w && x && y && n > 0 && k > 0, "bad matvec arguments");
  for (int i = 0; i < n; ++i) y[i] = dot_lr(w + (size_t)i * k, x, k);
  return 0;
}

int b200gru_oracle_softmax(const float *x, float *y, int n, uint32_t flags) {
  B200_REQUIRE(x && y && n > 0, "bad softmax arguments");
  Math m{(flags & B200GRU_ORACLE_LIBM) != 0};
  softmax(m, x, y, n);
  return 0;
}

int b200gru_oracle_gru_cell(const float *w_ih, const float *w_hh,
                            const float *b_ih, const float *b_hh,
                            const float *x, float *h, int in_dim, int hidden,
                            uint32_t flags) {
  B200_REQUIRE(w_ih && w_hh && b_ih && b_hh && x && h, "null argument");
  B200_REQUIRE(in_dim > 0 && hidden > 0, "bad dims");
  Math m{(flags & B200GRU_ORACLE_LIBM) != 0};
  std::vector<float> scratch(hidden);
  gru_cell(m, w_ih, w_hh, b_ih, b_hh, x, h, in_dim, hidden, scratch.data());
  return 0;
}

int b200gru_oracle_generate(const b200gru_config *cfg, const float *params,
                            int n_names, const float *rfloats, char *output,
                            float *logits_out, uint32_t flags) {
  if (int rc = validate_config(cfg)) return rc;
  B200_REQUIRE(params && rfloats && output && n_names >= 0, "null argument");
  const b200gru_config &c = *cfg;
  auto lay = build_layout(c);
  LayoutIndex ix(c.num_layers);
  Math m{(flags & B200GRU_ORACLE_LIBM) != 0};
  const int H = c.hidden, E = c.emb_dim, C = c.num_char, T = c.max_len;

  memset(output, 0, (size_t)n_names * (T + 1));  // ref:643
  if (logits_out)
    memset(logits_out, 0, sizeof(float) * (size_t)n_names * T * C);

  std::vector<float> h((size_t)c.num_layers * H), scratch(H), f(C), p(C);
  const float *emb = params + lay[ix.emb()].offset;
  const float *w_fc = params + lay[ix.w_fc()].offset;
  const float *b_fc = params + lay[ix.b_fc()].offset;

  for (int n = 0; n < n_names; ++n) {  // ref:649
    int ch = c.sos;                    // ref:652
    std::fill(h.begin(), h.end(), 0.0f);
    for (int l = 0; l < T; ++l) {      // ref:661
      const float *x = emb + (size_t)ch * E;  // ref:112-118
      int in_dim = E;
      for (int layer = 0; layer < c.num_layers; ++layer) {
        float *hl = h.data() + (size_t)layer * H;
        gru_cell(m, params + lay[ix.w_i(layer, 0)].offset,
                 params + lay[ix.w_h(layer, 0)].offset,
                 params + lay[ix.b_i(layer, 0)].offset,
                 params + lay[ix.b_h(layer, 0)].offset, x, hl, in_dim, H,
                 scratch.data());
        x = hl;  // next layer consumes the *updated* hidden (ref:773)
        in_dim = H;
      }
      for (int i = 0; i < C; ++i)  // ref:861-865
        f[i] = dot_lr(w_fc + (size_t)i * H, x, H) + b_fc[i];
      if (logits_out)
        memcpy(logits_out + ((size_t)n * T + l) * C, f.data(),
               sizeof(float) * C);
      softmax(m, f.data(), p.data(), C);  // ref:302-313
      ch = b200gru_random_select(p.data(), C, rfloats[(size_t)n * T + l]);
      output[(size_t)n * (T + 1) + l] = (char)ch;  // ref:877
      if (ch == c.eos) break;                      // ref:881-882
    }
  }
  return 0;
}

}  // extern "C"
